// kernels_band.cu -- the hot kernel, banded: affine-gap global DP (Gotoh) with SeqAn's exact tie-breaking,
// computed only over the diagonals dLo <= c - r <= dHi that contain the active alignment edges, and EXACT.
//
// Replaces the same call as kernels_dp.cu (seqan::globalAlignment as LaRA instantiates it,
// src/subgradient_solver.hpp:223-225 -> SeqAn:align/dp_algorithm_impl.h:1349-1425, dp_formula_affine.h:62-125).
//
// Why this is exact although the reference fills the whole matrix: every cell outside the
// band is inactive (pssm = -inf), so it is a pure gap cell, and the values of pure gap regions are closed forms
// in running maxima over the band's boundary diagonals (derivation, executable specification and randomized
// check against the full matrix: tools/band_proto.py, tests/test_band_proto.py):
//
//   region U (c - r > dHi), per row r, boundary cell (cb, r) = (r + dHi, r):
//        H(cb+1, r) = max(H(cb,r) + ge, S(cb,r) + go)                          (first cell right of the band)
//        V(c, r)    = (c + r - 1) ge + go + M(r),   M(r) = max_{r' < r} (T(r') - r' ge),
//        T(r)       = max(H(cb,r) + ge, S(cb,r) + go) - (cb + 1) ge
//   region L (c - r < dLo): the same with rows/columns and H/V exchanged.
//
// The lane that owns the top (bottom) diagonal keeps the running maximum in a register, feeds the inflow from U
// (L) into its cells and records TH/TS/M per row (TV/TS/M per column) so that a traceback that leaves the band
// can be continued exactly (kernels_post.cu).  Requires go < ge (opening dearer than extending).
//
// Mapping: a band of W = 2 G Q diagonals is cut into 2 G blocks of Q neighbouring diagonals; lane g of the G
// lanes of a pair owns blocks A = 2g and B = 2g + 1, so 32 / G pairs share a warp (G = 1: one THREAD per pair, the
// inter-sequence parallelism of the reference's SIMD kernel; larger G when a batch has too few pairs to fill
// the machine or the band is wide).  Block b works on row r in time step 2 r + b: in loop iteration i lane g
// computes row r = i - g + 1 of block A and then of block B, each a run of Q cells of one matrix row:
//     left neighbour of the run   = last cell of the previous block in the same row
//                                   (A: block B of lane g-1, one iteration ago, by shfl_up;   B: own block A)
//     upper neighbours            = the cells one diagonal higher in row r - 1: own registers, and for the last
//                                   cell the first diagonal of the next block (A: own block B;  B: block A of lane
//                                   g+1 from this iteration, by shfl_down)
//     diagonal neighbour          = previous cell of the same diagonal (same register)
// so a lane keeps S and V of the latest cell of each of its 2 Q diagonals in registers and two shuffles serve Q
// cells.  The horizontal chain along a run is shortened to IADD + VIMNMX per cell: with go < ge,
//     H(c,r) = max(H(c-1,r) + ge, S(c-1,r) + go) = max(H(c-1,r) + ge, max(D(c-1,r), V(c-1,r)) + go)
// and the "H opened" decision bit (first operand wins ties) is unchanged by dropping H(c-1,r) from the second
// operand, because H + ge > H + go.
//
// Memory: position specific scores come from a dense band array vals[row][diagonal], stored in 16-byte chunks of 4
// diagonals and interleaved over the pairs of a warp ([row][chunk][pair slot]) so that a warp-wide 128-bit load is
// contiguous; four decision bits per cell go to trace units of Q cells, stored as four bit planes ([row][block][pair
// slot]).  One thread per pair (G = 1): the two units of a row are written as one 16-byte record and two rows share a
// 32-byte sector ([row / 2][pair slot][row % 2]), because the traceback reads a single pair's rows and pays per sector.
#include "lara_gpu.cuh"

#include <cstdlib>

#ifndef LARA_BAND_DPX
#define LARA_BAND_DPX 0
#endif
#ifndef LARA_BAND_INFLOW_SELECT
#define LARA_BAND_INFLOW_SELECT 1      /* measured: fill launch of a small set 0.308 -> 0.290 ms, all of C3 unchanged */
#endif


namespace lara
{

namespace
{

constexpr unsigned kFull = 0xffffffffu;

// 16-byte store under a predicate, without a branch
__device__ __forceinline__ void store_if(bool on, int4 * p, int4 v)
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %0, 0;\n\t@q st.global.v4.s32 [%1], {%2, %3, %4, %5};\n\t}"
                 :: "r"((int)on), "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// One trace unit = the decisions of one run of Q cells as four BIT PLANES of Q bits (plane k = decision bit k of
// lara_gpu.cuh: kTrGap, kTrFromH, kTrVOpen, kTrHOpen; bit j of a plane = cell j of the run): 2, 4 or 8 bytes.
//   Q = 16, 12: { P0 | P1 << 16, P2 | P3 << 16 }      Q = 8: P0 | P1 << 8 | P2 << 16 | P3 << 24      Q = 4: 4 bits each
// The traceback tests ONE decision of a cell at a time (kernels_post.cu), which a plane answers with one shift; and a
// plane grows by one funnel shift per cell with no chain between the four decisions of a cell.
template <int Q>
__device__ __forceinline__ uint2 pack_unit(uint32_t const (&pl)[4])
{
    if constexpr (Q == 4) return make_uint2(pl[0] | (pl[1] << 4) | (pl[2] << 8) | (pl[3] << 12), 0u);
    else if constexpr (Q == 8) return make_uint2(pl[0] | (pl[1] << 8) | (pl[2] << 16) | (pl[3] << 24), 0u);
    else return make_uint2(pl[0] | (pl[1] << 16), pl[2] | (pl[3] << 16));
}
template <int Q>
__device__ __forceinline__ void store_unit(void * base, size_t unitIdx, uint32_t const (&pl)[4])
{
    uint2 const u = pack_unit<Q>(pl);
    if constexpr (Q == 4) reinterpret_cast<uint16_t *>(base)[unitIdx] = (uint16_t)u.x;
    else if constexpr (Q == 8) reinterpret_cast<uint32_t *>(base)[unitIdx] = u.x;
    else reinterpret_cast<uint2 *>(base)[unitIdx] = u;
}

// One run of Q cells of row r on the diagonals of one block.  S[j], V[j]: latest cell of diagonal j (row r - 1 on
// entry, row r on exit).  sub[j]: position specific score of cell j.  (hL, sL): H and S of the left
// neighbour of the run; (sU, vU): S and V of the upper neighbour of the run's last cell.  Returns H of the run's last
// cell.  pl: the four decision planes of the run, cell j at bit j.
// Only H is carried along the run, IADD + VIMNMX per cell (see the header: the second operand of H is max(D, V) + go
// of the cell before); the diagonal and vertical candidates come from row r - 1, and each of the four decisions goes
// into its own plane with one funnel shift, so the cells of a run overlap freely in the instruction stream.
// kLast: additionally captures S, V, H of cell jEnd (the end cell of the matrix, for SeqAn's _correctTraceValue).
template <int Q, bool kLast>
__device__ __forceinline__ int32_t band_run(int32_t (&S)[Q], int32_t (&V)[Q], int32_t const (&sub)[Q], int32_t hL, int32_t sL,
                                            int32_t sU, int32_t vU, int32_t go, int32_t ge, uint32_t (&pl)[4],
                                            int jEnd, int32_t & sE, int32_t & vE, int32_t & hE)
{
    uint32_t p0 = 0u, p1 = 0u, p2 = 0u, p3 = 0u;
    int32_t h = hL, m = sL + go;
    #pragma unroll
    for (int j = 0; j < Q; ++j)
    {
        int32_t const us = (j + 1 < Q) ? S[j + 1] : sU;
        int32_t const uv = (j + 1 < Q) ? V[j + 1] : vU;
        int32_t const dd = S[j] + sub[j];
#if LARA_BAND_DPX
        // three-operand forms (VIADDMNMX, IADD3): max(x + ge, y) in one instruction, and the decision as the sign of
        // (x + ge) - max, which is negative exactly when the opening won (extension wins ties) -- 15 instructions per
        // cell instead of 17, and the H chain along the run is one instruction per cell.  (The DPX max-with-predicate
        // __vibmax_s32 is ISETP + SEL on sm_100a: no gain over subtraction + funnel shift.)
        int32_t const vo = us + go;
        int32_t const v = __viaddmax_s32(uv, ge, vo);
        p2 = __funnelshift_l((uint32_t)(uv + ge - v), p2, 1);           // V opened   (extension wins ties)
        int32_t const hn = __viaddmax_s32(h, ge, m);
        p3 = __funnelshift_l((uint32_t)(h + ge - hn), p3, 1);           // H opened   (extension wins ties)
#else
        int32_t const ve = uv + ge, vo = us + go;
        int32_t const v = max(ve, vo);
        p2 = __funnelshift_l((uint32_t)(ve - vo), p2, 1);               // V opened   (extension wins ties)
        int32_t const he = h + ge;
        int32_t const hn = max(he, m);
        p3 = __funnelshift_l((uint32_t)(he - m), p3, 1);                // H opened   (extension wins ties)
#endif
        int32_t const mx = max(dd, v);
        int32_t const s = max(mx, hn);
        p1 = __funnelshift_l((uint32_t)(v - hn), p1, 1);                // gap from H (vertical wins ties)
        p0 = __funnelshift_l((uint32_t)(dd - s), p0, 1);                // S from gap (diagonal wins ties): gap > dd <=> s > dd
        m = mx + go;                                                    // what the next cell's H opens from
        h = hn;
        S[j] = s;
        V[j] = v;
        if (kLast && j == jEnd) { sE = s; vE = v; hE = hn; }
    }
    // the funnel shifts pushed cell j to bit Q - 1 - j
    pl[0] = __brev(p0) >> (32 - Q); pl[1] = __brev(p1) >> (32 - Q); pl[2] = __brev(p2) >> (32 - Q); pl[3] = __brev(p3) >> (32 - Q);
    return h;
}

// The fill of the warp groups of ONE band shape: warp `warp` of the shape's part of the grid works on the slots
// [warp * 32 / G, (warp + 1) * 32 / G) of pairIdx[0 .. nSlots).
template <int G, int Q>
__device__ __forceinline__ void
band_body(int const warp, int nSlots, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
          PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    constexpr int PPW = 32 / G;                                 // pairs per warp
    constexpr int W = 2 * G * Q;                                // diagonals of the band
    constexpr int NCH = W / 4;                                  // 16-byte value chunks per row
    constexpr int NU = 2 * G;                                   // trace units per row
    constexpr int CQ = Q / 4;                                   // value chunks per block
    constexpr int kPrefetchRows = 48 / Q;                       // L2 prefetch distance: about the same time for every Q
    int const lane = threadIdx.x & 31;
    if (warp * PPW >= nSlots)
        return;
    int const q = lane / G, g = lane % G;
    int const slot = warp * PPW + q;
    int p = slot < nSlots ? pairIdx[slot] : -1;
    if (p >= 0 && state[p].done)
        p = -1;
    bool const live = p >= 0;
    PairDesc const & d = desc[live ? p : 0];
    int const la = live ? d.la : -1, lb = live ? d.lb : -1;
    int const dLo = live ? d.dLo : 0;
    int lbMax = lb;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) lbMax = max(lbMax, __shfl_xor_sync(kFull, lbMax, o));
    if (lbMax < 1)
        return;

    // the pair's own slot in its group's interleaved arrays (== q until the slot list has been compacted)
    int const gs = live ? d.bandSlot : 0;
    int4 const * __restrict__ vals = reinterpret_cast<int4 const *>(arena + d.offBVals) + gs;
    void * const trace = arena + d.offBTrace;
    // closed-form records of the gap regions, interleaved over the pairs of the group: U per row, L per column
    int4 * __restrict__ uRec = reinterpret_cast<int4 *>(arena + d.offU) + gs;      // { TH, TS, M, 0 } at [r * PPW]
    int4 * __restrict__ lRec = reinterpret_cast<int4 *>(arena + d.offL) + gs;      // { TV, TS, M, 0 } at [c * PPW]

    // end cell (la, lb): block and diagonal inside the block, if it lies in the band
    int const idxEnd = la - lb - dLo;
    int const bEnd = (live && idxEnd >= 0 && idxEnd < W) ? idxEnd / Q : -1;
    int const jEnd = idxEnd - bEnd * Q;

    // ---- row 0 (dp_formula_affine.h:245-281): S(c,0) = H(c,0) = go + (c-1) ge, V = -inf; origin S = 0 -----------
    int32_t SA[Q], VA[Q], SB[Q], VB[Q];
    int const cA0 = dLo + 2 * g * Q, cB0 = cA0 + Q;               // column of the blocks' first diagonal in row 0
    #pragma unroll
    for (int j = 0; j < Q; ++j)
    {
        int const ca = cA0 + j, cb = cB0 + j;
        SA[j] = ca > 0 ? go + (ca - 1) * ge : (ca == 0 ? 0 : kBandNeg);
        SB[j] = cb > 0 ? go + (cb - 1) * ge : (cb == 0 ? 0 : kBandNeg);
        VA[j] = kBandNeg;
        VB[j] = kBandNeg;
    }
    int32_t HA = (cA0 + Q - 1 > 0) ? SA[Q - 1] : kBandNeg;        // H of the last cell of the block's latest run
    int32_t HB = (cB0 + Q - 1 > 0) ? SB[Q - 1] : kBandNeg;
    int32_t runU = kDpNeg, runL = kDpNeg;                         // running maxima M(r) (lane G-1) / M'(c) (lane 0)
    int32_t sE = 0, vE = 0, hE = 0;

    // scores of one run: CQ 16-byte chunks (software pipelined: block B's are fetched while block A computes,
    // the next row's block A while block B computes)
    auto load_run = [&](int32_t (&sub)[Q], int row, int block)
    {
        int4 const * __restrict__ src = vals + ((size_t)row * NCH + (size_t)block * CQ) * PPW;
        #pragma unroll
        for (int k = 0; k < CQ; ++k)
        {
            int4 const x = __ldg(src + (size_t)k * PPW);
            sub[4 * k] = x.x; sub[4 * k + 1] = x.y; sub[4 * k + 2] = x.z; sub[4 * k + 3] = x.w;
        }
    };
    // Blocks of up to 8 diagonals are short: half a row of arithmetic does not cover an L2 hit, and such launches have
    // too few warps to hide it (measured: the same 2,000 cycles per row as one thread per pair).  There the register
    // pipeline is one row deeper: nxtA / nxtB hold the scores of the row after the one in subA / subB.
    constexpr bool kDeep = Q <= 8;
    int32_t subA[Q], subB[Q];
    int32_t nxtA[kDeep ? Q : 1], nxtB[kDeep ? Q : 1];
    #pragma unroll
    for (int j = 0; j < Q; ++j) { subA[j] = kDpNeg; subB[j] = kDpNeg; }
    if (1 - g >= 1 && 1 - g <= lb) load_run(subA, 1 - g, 2 * g);
    if constexpr (kDeep)
    {
        #pragma unroll
        for (int j = 0; j < Q; ++j) { nxtA[j] = kDpNeg; nxtB[j] = kDpNeg; }
        if (1 - g >= 1 && 1 - g <= lb) load_run(subB, 1 - g, 2 * g + 1);
        if (2 - g >= 1 && 2 - g <= lb) load_run(nxtA, 2 - g, 2 * g);
    }

    int const nIter = lbMax + G;
    uint2 unitA = make_uint2(0u, 0u);                               // G == 1: block A's unit waits for block B's
    for (int i = 0; i < nIter; ++i)
    {
        int const r = i - g + 1;
        bool const rowValid = r >= 1 && r <= lb;
        if constexpr (kDeep) { if (r + 1 >= 1 && r + 1 <= lb) load_run(nxtB, r + 1, 2 * g + 1); }
        else if (rowValid) load_run(subB, r, 2 * g + 1);
        // the register pipeline is half a row deep; under load HBM answers later than that, so the rows
        // kPrefetchRows ahead are pulled into L2 meanwhile (a warp's 16-byte chunks of one row are contiguous)
        if (r + kPrefetchRows <= lb && r + kPrefetchRows >= 1)
        {
            int4 const * __restrict__ src = vals + ((size_t)(r + kPrefetchRows) * NCH + (size_t)(2 * g) * CQ) * PPW;
            #pragma unroll
            for (int k = 0; k < 2 * CQ; ++k)
                asm volatile("prefetch.global.L2 [%0];" :: "l"(src + (size_t)k * PPW));
        }
        // ------------------------------ block A ----------------------------------------------------------------
        {
            int32_t hL = __shfl_up_sync(kFull, HB, 1, G);
            int32_t sL = __shfl_up_sync(kFull, SB[Q - 1], 1, G);
#if LARA_BAND_INFLOW_SELECT
            {
                // branch-free form: every lane computes the inflow, lane 0 of the pair takes it, the record is stored under a
                // predicate (the divergent block and its branches were an eighth of a small launch: profiles/README.md r2m)
                int const c = r + dLo;
                bool const on = g == 0 && rowValid && c >= 1 && c - 1 <= la;
                int32_t const tv = VA[0] + ge, ts = SA[0] + go;
                int32_t const vl = max(tv, ts);
                int32_t const hx = (c - 1 >= 1) ? (c + r - 2) * ge + go + runL : kBandNeg;
                int32_t const sx = max(vl, hx);
                if (g == 0) { hL = on ? hx : kBandNeg; sL = on ? sx : kBandNeg; }
                store_if(on, lRec + (size_t)(on ? c - 1 : 0) * PPW, make_int4(tv - r * ge, ts - r * ge, runL, 0));
                runL = on ? max(runL, vl - r * ge - (c - 1) * ge) : runL;
            }
#else
            if (g == 0)
            {
                // the left neighbour (c-1, r) of the run lies in region L, directly below the boundary cell
                // (c-1, r-1) = latest cell of diagonal dLo = (SA[0], VA[0])
                int const c = r + dLo;
                hL = kBandNeg; sL = kBandNeg;
                if (rowValid && c >= 1 && c - 1 <= la)
                {
                    int32_t const tv = VA[0] + ge, ts = SA[0] + go;
                    int32_t const vl = max(tv, ts);                              // first L cell of column c-1
                    hL = (c - 1 >= 1) ? (c + r - 2) * ge + go + runL : kBandNeg;
                    sL = max(vl, hL);
                    lRec[(size_t)(c - 1) * PPW] = make_int4(tv - r * ge, ts - r * ge, runL, 0);
                    runL = max(runL, vl - r * ge - (c - 1) * ge);
                }
            }
#endif
            if (rowValid)
            {
                uint32_t pl[4];
                if (r == lb && bEnd == 2 * g)
                    HA = band_run<Q, true>(SA, VA, subA, hL, sL, SB[0], VB[0], go, ge, pl, jEnd, sE, vE, hE);
                else
                    HA = band_run<Q, false>(SA, VA, subA, hL, sL, SB[0], VB[0], go, ge, pl, 0, sE, vE, hE);
                if constexpr (G == 1) unitA = pack_unit<Q>(pl);
                else store_unit<Q>(trace, ((size_t)r * NU + (size_t)(2 * g)) * PPW + gs, pl);
            }
        }
        if constexpr (kDeep)
        {
            #pragma unroll
            for (int j = 0; j < Q; ++j) subA[j] = nxtA[j];
            if (r + 2 >= 1 && r + 2 <= lb) load_run(nxtA, r + 2, 2 * g);
        }
        else if (r + 1 >= 1 && r + 1 <= lb) load_run(subA, r + 1, 2 * g);
        // ------------------------------ block B ----------------------------------------------------------------
        {
            int32_t sU = __shfl_down_sync(kFull, SA[0], 1, G);
            int32_t vU = __shfl_down_sync(kFull, VA[0], 1, G);
#if LARA_BAND_INFLOW_SELECT
            {
                int const c = r + dLo + W - 1;
                bool const on = g == G - 1 && live && r >= 1 && r - 1 <= lb && c - 1 <= la;
                int32_t const th = HB + ge, ts = SB[Q - 1] + go;
                int32_t const hu = max(th, ts);
                int32_t const vx = (r - 1 >= 1) ? (c + r - 2) * ge + go + runU : kBandNeg;
                int32_t const sx = max(vx, hu);
                if (g == G - 1) { vU = on ? vx : kBandNeg; sU = on ? sx : kBandNeg; }
                store_if(on, uRec + (size_t)(on ? r - 1 : 0) * PPW, make_int4(th - c * ge, ts - c * ge, runU, 0));
                runU = on ? max(runU, hu - c * ge - (r - 1) * ge) : runU;
            }
#else
            if (g == G - 1)
            {
                // the upper neighbour (c, r-1) of the run's last cell lies in region U, right of the boundary cell
                // (c-1, r-1) = latest cell of diagonal dHi = (SB[Q-1], HB); row lb is closed in iteration r = lb + 1
                int const c = r + dLo + W - 1;
                sU = kBandNeg; vU = kBandNeg;
                if (live && r >= 1 && r - 1 <= lb && c - 1 <= la)
                {
                    int32_t const th = HB + ge, ts = SB[Q - 1] + go;
                    int32_t const hu = max(th, ts);                              // first U cell of row r-1
                    vU = (r - 1 >= 1) ? (c + r - 2) * ge + go + runU : kBandNeg;
                    sU = max(vU, hu);
                    uRec[(size_t)(r - 1) * PPW] = make_int4(th - c * ge, ts - c * ge, runU, 0);
                    runU = max(runU, hu - c * ge - (r - 1) * ge);
                }
            }
#endif
            if (rowValid)
            {
                uint32_t pl[4];
                if (r == lb && bEnd == 2 * g + 1)
                    HB = band_run<Q, true>(SB, VB, subB, HA, SA[Q - 1], sU, vU, go, ge, pl, jEnd, sE, vE, hE);
                else
                    HB = band_run<Q, false>(SB, VB, subB, HA, SA[Q - 1], sU, vU, go, ge, pl, 0, sE, vE, hE);
                if constexpr (G == 1)
                {
                    static_assert(G != 1 || Q >= 12, "one thread per pair: units of 8 bytes");
                    uint2 const unitB = pack_unit<Q>(pl);
                    reinterpret_cast<uint4 *>(trace)[(((size_t)(r >> 1) * PPW + gs) << 1) + (r & 1)] =
                        make_uint4(unitA.x, unitA.y, unitB.x, unitB.y);
                }
                else store_unit<Q>(trace, ((size_t)r * NU + (size_t)(2 * g + 1)) * PPW + gs, pl);
            }
        }
        if constexpr (kDeep)
        {
            #pragma unroll
            for (int j = 0; j < Q; ++j) subB[j] = nxtB[j];
        }
    }
    if (bEnd >= 0 && (bEnd >> 1) == g)
    {
        PairState & st = state[p];
        st.currentUpper = sE; st.sEnd = sE; st.vEnd = vE; st.hEnd = hE;
    }
}

// Two runs side by side in one instruction stream: block B of row r and block A of row r + 1 (one thread per pair).
// They are independent but for one neighbour cell: the upper neighbour of A's last cell is B's first cell of row r,
// which B produces first.  B's left neighbour (hLB, sLB) = last cell of A's row r, A's left neighbour (hLA, sLA) =
// inflow from region L.  Two H chains and eight decision planes advance together, so a warp that has a scheduler
// to itself (the usual case: one thread per pair leaves the machine with about one warp per scheduler) does not wait
// for its own previous instruction.
// The scores of the NEXT iteration (block B of row r + 1 from nextB, block A of row r + 2 from nextA; nullptr: none)
// are loaded into subB / subA chunk by chunk as soon as this iteration has used a chunk: a whole iteration of lead
// time without a second set of registers.
// kLast: captures S, V, H of cell jEndB of block B or jEndA of block A (-1: none).
template <int Q, bool kLast>
__device__ __forceinline__ void band_run2(int32_t (&SB)[Q], int32_t (&VB)[Q], int32_t (&subB)[Q], int32_t hLB, int32_t sLB,
                                          int32_t sUB, int32_t vUB, uint32_t (&plB)[4], int32_t & HB,
                                          int32_t (&SA)[Q], int32_t (&VA)[Q], int32_t (&subA)[Q], int32_t hLA, int32_t sLA,
                                          uint32_t (&plA)[4], int32_t & HA, int32_t go, int32_t ge,
                                          int4 const * __restrict__ nextB, int4 const * __restrict__ nextA,
                                          int jEndB, int jEndA, int32_t & sE, int32_t & vE, int32_t & hE)
{
    uint32_t b0 = 0u, b1 = 0u, b2 = 0u, b3 = 0u, a0 = 0u, a1 = 0u, a2 = 0u, a3 = 0u;
    int32_t hB = hLB, mB = sLB + go, hA = hLA, mA = sLA + go;
    #pragma unroll
    for (int j = 0; j < Q; ++j)
    {
        {
            int32_t const us = (j + 1 < Q) ? SB[j + 1] : sUB;
            int32_t const uv = (j + 1 < Q) ? VB[j + 1] : vUB;
            int32_t const dd = SB[j] + subB[j];
            int32_t const ve = uv + ge, vo = us + go;
            int32_t const v = max(ve, vo);
            b2 = __funnelshift_l((uint32_t)(ve - vo), b2, 1);
            int32_t const he = hB + ge;
            int32_t const hn = max(he, mB);
            b3 = __funnelshift_l((uint32_t)(he - mB), b3, 1);
            int32_t const mx = max(dd, v);
            int32_t const s = max(mx, hn);
            b1 = __funnelshift_l((uint32_t)(v - hn), b1, 1);
            b0 = __funnelshift_l((uint32_t)(dd - s), b0, 1);
            mB = mx + go;
            hB = hn;
            SB[j] = s;
            VB[j] = v;
            if (kLast && j == jEndB) { sE = s; vE = v; hE = hn; }
        }
        {
            // the upper neighbour of A's last cell: B's first cell of this row, written above at j = 0
            int32_t const us = (j + 1 < Q) ? SA[j + 1] : SB[0];
            int32_t const uv = (j + 1 < Q) ? VA[j + 1] : VB[0];
            int32_t const dd = SA[j] + subA[j];
            int32_t const ve = uv + ge, vo = us + go;
            int32_t const v = max(ve, vo);
            a2 = __funnelshift_l((uint32_t)(ve - vo), a2, 1);
            int32_t const he = hA + ge;
            int32_t const hn = max(he, mA);
            a3 = __funnelshift_l((uint32_t)(he - mA), a3, 1);
            int32_t const mx = max(dd, v);
            int32_t const s = max(mx, hn);
            a1 = __funnelshift_l((uint32_t)(v - hn), a1, 1);
            a0 = __funnelshift_l((uint32_t)(dd - s), a0, 1);
            mA = mx + go;
            hA = hn;
            SA[j] = s;
            VA[j] = v;
            if (kLast && j == jEndA) { sE = s; vE = v; hE = hn; }
        }
        if ((j & 3) == 3)
        {
            // this chunk of scores has been used: fetch the next iteration's into the same registers
            int const k = j >> 2;
            if (nextB != nullptr)
            {
                int4 const x = __ldg(nextB + (size_t)k * 32);
                subB[4 * k] = x.x; subB[4 * k + 1] = x.y; subB[4 * k + 2] = x.z; subB[4 * k + 3] = x.w;
            }
            if (nextA != nullptr)
            {
                int4 const x = __ldg(nextA + (size_t)k * 32);
                subA[4 * k] = x.x; subA[4 * k + 1] = x.y; subA[4 * k + 2] = x.z; subA[4 * k + 3] = x.w;
            }
        }
    }
    plB[0] = __brev(b0) >> (32 - Q); plB[1] = __brev(b1) >> (32 - Q); plB[2] = __brev(b2) >> (32 - Q); plB[3] = __brev(b3) >> (32 - Q);
    plA[0] = __brev(a0) >> (32 - Q); plA[1] = __brev(a1) >> (32 - Q); plA[2] = __brev(a2) >> (32 - Q); plA[3] = __brev(a3) >> (32 - Q);
    HB = hB;
    HA = hA;
}

// One thread per pair (G = 1, band of 2 Q diagonals): the same fill as band_body<1, Q>, with the loop turned by half
// a row: an iteration computes block B of row r together with block A of row r + 1 (band_run2).  In the iteration of
// the pair's last row the A half works on nothing (its results are not used).
template <int Q>
__device__ __forceinline__ void
band_body_one(int const warp, int nSlots, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
              PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    constexpr int PPW = 32;
    constexpr int W = 2 * Q;
    constexpr int NCH = W / 4;
    constexpr int CQ = Q / 4;
    constexpr int kPrefetchRows = 48 / Q + 1;
    int const lane = threadIdx.x & 31;
    if (warp * PPW >= nSlots)
        return;
    int const slot = warp * PPW + lane;
    int p = slot < nSlots ? pairIdx[slot] : -1;
    if (p >= 0 && state[p].done)
        p = -1;
    bool const live = p >= 0;
    PairDesc const & d = desc[live ? p : 0];
    int const la = live ? d.la : -1, lb = live ? d.lb : -1;
    int const dLo = live ? d.dLo : 0;
    int lbMax = lb;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) lbMax = max(lbMax, __shfl_xor_sync(kFull, lbMax, o));
    if (lbMax < 1)
        return;
    int const gs = live ? d.bandSlot : 0;
    int4 const * __restrict__ vals = reinterpret_cast<int4 const *>(arena + d.offBVals) + gs;
    uint4 * const trace = reinterpret_cast<uint4 *>(arena + d.offBTrace);
    int4 * __restrict__ uRec = reinterpret_cast<int4 *>(arena + d.offU) + gs;
    int4 * __restrict__ lRec = reinterpret_cast<int4 *>(arena + d.offL) + gs;
    int const idxEnd = la - lb - dLo;
    int const bEnd = (live && idxEnd >= 0 && idxEnd < W) ? idxEnd / Q : -1;
    int const jEnd = idxEnd - bEnd * Q;

    // row 0
    int32_t SA[Q], VA[Q], SB[Q], VB[Q];
    #pragma unroll
    for (int j = 0; j < Q; ++j)
    {
        int const ca = dLo + j, cb = dLo + Q + j;
        SA[j] = ca > 0 ? go + (ca - 1) * ge : (ca == 0 ? 0 : kBandNeg);
        SB[j] = cb > 0 ? go + (cb - 1) * ge : (cb == 0 ? 0 : kBandNeg);
        VA[j] = kBandNeg;
        VB[j] = kBandNeg;
    }
    int32_t HA = (dLo + Q - 1 > 0) ? SA[Q - 1] : kBandNeg;
    int32_t HB = (dLo + 2 * Q - 1 > 0) ? SB[Q - 1] : kBandNeg;
    int32_t runU = kDpNeg, runL = kDpNeg;
    int32_t sE = 0, vE = 0, hE = 0;

    auto run_ptr = [&](int row, int block) -> int4 const *
    {
        return (row >= 1 && row <= lb) ? vals + ((size_t)row * NCH + (size_t)block * CQ) * PPW : nullptr;
    };
    auto load_run = [&](int32_t (&sub)[Q], int row, int block)
    {
        int4 const * __restrict__ src = run_ptr(row, block);
        if (src == nullptr) return;
        #pragma unroll
        for (int k = 0; k < CQ; ++k)
        {
            int4 const x = __ldg(src + (size_t)k * PPW);
            sub[4 * k] = x.x; sub[4 * k + 1] = x.y; sub[4 * k + 2] = x.z; sub[4 * k + 3] = x.w;
        }
    };
    // inflow from region L into block A of row r (left neighbour of its first cell), record of column c - 1
    auto inflow_l = [&](int r, int32_t & hL, int32_t & sL)
    {
        int const c = r + dLo;
        hL = kBandNeg; sL = kBandNeg;
        if (r >= 1 && r <= lb && c >= 1 && c - 1 <= la)
        {
            int32_t const tv = VA[0] + ge, ts = SA[0] + go;
            int32_t const vl = max(tv, ts);
            hL = (c - 1 >= 1) ? (c + r - 2) * ge + go + runL : kBandNeg;
            sL = max(vl, hL);
            lRec[(size_t)(c - 1) * PPW] = make_int4(tv - r * ge, ts - r * ge, runL, 0);
            runL = max(runL, vl - r * ge - (c - 1) * ge);
        }
    };
    // inflow from region U into block B of row r (upper neighbour of its last cell), record of row r - 1
    auto inflow_u = [&](int r, int32_t & sU, int32_t & vU)
    {
        int const c = r + dLo + W - 1;
        sU = kBandNeg; vU = kBandNeg;
        if (live && r >= 1 && r - 1 <= lb && c - 1 <= la)
        {
            int32_t const th = HB + ge, ts = SB[Q - 1] + go;
            int32_t const hu = max(th, ts);
            vU = (r - 1 >= 1) ? (c + r - 2) * ge + go + runU : kBandNeg;
            sU = max(vU, hu);
            uRec[(size_t)(r - 1) * PPW] = make_int4(th - c * ge, ts - c * ge, runU, 0);
            runU = max(runU, hu - c * ge - (r - 1) * ge);
        }
    };

    int32_t subA[Q], subB[Q];
    #pragma unroll
    for (int j = 0; j < Q; ++j) { subA[j] = kDpNeg; subB[j] = kDpNeg; }
    load_run(subA, 1, 0);
    load_run(subB, 1, 1);
    // ---- block A of row 1 -------------------------------------------------------------------------------------------
    uint2 unitA = make_uint2(0u, 0u);
    {
        int32_t hL, sL;
        inflow_l(1, hL, sL);
        if (lb >= 1)
        {
            uint32_t pl[4];
            HA = band_run<Q, true>(SA, VA, subA, hL, sL, SB[0], VB[0], go, ge, pl, (lb == 1 && bEnd == 0) ? jEnd : -1, sE, vE, hE);
            unitA = pack_unit<Q>(pl);
        }
    }
    load_run(subA, 2, 0);
    // ---- iteration r: block B of row r and block A of row r + 1; row lb + 1 only closes region U's record of row lb --
    for (int r = 1; r <= lbMax + 1; ++r)
    {
        if (r + kPrefetchRows <= lb)
        {
            int4 const * __restrict__ src = vals + (size_t)(r + kPrefetchRows) * NCH * PPW;
            #pragma unroll
            for (int k = 0; k < 2 * CQ; ++k)
                asm volatile("prefetch.global.L2 [%0];" :: "l"(src + (size_t)k * PPW));
        }
        int32_t sU, vU, hL, sL;
        inflow_u(r, sU, vU);                     // from row r - 1 of diagonal dHi
        if (r > lb) continue;
        inflow_l(r + 1, hL, sL);                 // from row r of diagonal dLo
        uint32_t plB[4], plA[4];
        int4 const * const nB = run_ptr(r + 1, 1), * const nA = run_ptr(r + 2, 0);
        if (r + 1 < lb)
            band_run2<Q, false>(SB, VB, subB, HA, SA[Q - 1], sU, vU, plB, HB, SA, VA, subA, hL, sL, plA, HA, go, ge, nB, nA, -1, -1, sE, vE, hE);
        else
            band_run2<Q, true>(SB, VB, subB, HA, SA[Q - 1], sU, vU, plB, HB, SA, VA, subA, hL, sL, plA, HA, go, ge, nB, nA,
                               (r == lb && bEnd == 1) ? jEnd : -1, (r + 1 == lb && bEnd == 0) ? jEnd : -1, sE, vE, hE);
        uint2 const unitB = pack_unit<Q>(plB);
        trace[(((size_t)(r >> 1) * PPW + gs) << 1) + (r & 1)] = make_uint4(unitA.x, unitA.y, unitB.x, unitB.y);
        unitA = pack_unit<Q>(plA);
    }
    if (bEnd >= 0)
    {
        PairState & st = state[p];
        st.currentUpper = sE; st.sEnd = sE; st.vEnd = vE; st.hEnd = hE;
    }
}

// One launch per iteration for ALL band shapes of a batch: the grid is the concatenation of the shapes' parts (the
// shape with the most slots first), a block looks up its part in the class table and runs that shape's body.  Separate
// launches per shape had to run side by side on streams (fork / join events every iteration), and a shape with a
// handful of pairs then set the latency of the whole iteration step.
#if defined(LARA_BAND_DEV) && LARA_BAND_DEV == 1          /* developer builds for SASS inspection: one shape only */
#define LARA_BAND_SHAPES(X) X(2, 8)
#elif defined(LARA_BAND_DEV) && LARA_BAND_DEV == 2
#define LARA_BAND_SHAPES(X) X(1, 16)
#else
#define LARA_BAND_SHAPES(X) \
    X(1, 16) X(2, 16) X(4, 16) X(8, 16) X(16, 16) X(32, 16) \
    X(1, 12) \
    X(2, 8) X(4, 8) X(8, 8) X(16, 8) X(32, 8) \
    X(4, 4) X(8, 4) X(16, 4) X(32, 4)
#endif

// kOne: the launch has one-thread-per-pair classes in their two-rows-at-a-time form (more registers per thread; a
// launch without such classes runs the instantiation without them and keeps its occupancy).
template <bool kOne>
__global__ void __launch_bounds__(128)
gotoh_band_kernel(BandClassTable const tab, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
                  PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    int c = 0;
    while (c + 1 < tab.n && (int)blockIdx.x >= tab.firstBlock[c + 1]) ++c;
    int const warp = ((int)blockIdx.x - tab.firstBlock[c]) * (int)(blockDim.x >> 5) + (int)(threadIdx.x >> 5);
    int32_t const * __restrict__ slots = pairIdx + tab.slotOffset[c];
    int const nSlots = tab.nSlots[c];
    switch (tab.shape[c])
    {
        // one thread per pair: the form that works on two rows at a time
        case (0x10000 | (1 << 8) | 16): if constexpr (kOne) band_body_one<16>(warp, nSlots, slots, desc, state, arena, go, ge); break;
        case (0x10000 | (1 << 8) | 12): if constexpr (kOne) band_body_one<12>(warp, nSlots, slots, desc, state, arena, go, ge); break;
#define LARA_CASE(g_, q_) case ((g_) << 8) | (q_): band_body<g_, q_>(warp, nSlots, slots, desc, state, arena, go, ge); break;
        LARA_BAND_SHAPES(LARA_CASE)
#undef LARA_CASE
        default: break;
    }
}

} // namespace

bool band_shape_supported(int G, int Q)
{
#define LARA_CASE(g_, q_) if (G == g_ && Q == q_) return true;
    LARA_BAND_SHAPES(LARA_CASE)
#undef LARA_CASE
    return false;
}

cudaError_t launch_gotoh_band_multi(int nClasses, BandLaunch const * cls, int32_t const * pairIdx, PairDesc const * desc,
                                    PairState * state, int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream)
{
    BandClassTable tab{};
    int blocks = 0;
    bool anyOne = false;
    // warps per block (LARA_B200_FILLWARPS = 1, 2, 4; default 4).  Smaller blocks would find room on an SM sooner while the
    // post kernel of another resident set holds its registers: measured no gain, alone or with the fill on a stream of the
    // highest priority (profiles/README.md r2m)
    static int const wpb = []() { char const * e = std::getenv("LARA_B200_FILLWARPS"); int v = e ? std::atoi(e) : 4;
                                  return v == 1 || v == 2 ? v : 4; }();
    for (int k = 0; k < nClasses; ++k)
    {
        if (cls[k].nSlots <= 0) continue;
        if (tab.n >= kMaxBandClasses || !band_shape_supported(cls[k].G, cls[k].Q)) return cudaErrorInvalidValue;
        int const slotsPerBlock = wpb * (32 / cls[k].G);
        // LARA_B200_FUSED=1: the one-thread-per-pair shapes run the two-rows-at-a-time body (same results; measured no
        // faster on B200: profiles/README.md r2, so the generic body with its lower register count is the default)
        static bool const fused = []() { char const * e = std::getenv("LARA_B200_FUSED"); return e && std::atoi(e) != 0; }();
        tab.shape[tab.n] = (cls[k].G << 8) | cls[k].Q | ((fused && cls[k].G == 1) ? 0x10000 : 0);
        anyOne = anyOne || (fused && cls[k].G == 1);
        tab.firstBlock[tab.n] = blocks;
        tab.slotOffset[tab.n] = cls[k].slotOffset;
        tab.nSlots[tab.n] = cls[k].nSlots;
        blocks += (cls[k].nSlots + slotsPerBlock - 1) / slotsPerBlock;
        ++tab.n;
    }
    if (blocks == 0) return cudaSuccess;
    tab.firstBlock[tab.n] = blocks;
    if (anyOne) gotoh_band_kernel<true><<<blocks, 32 * wpb, 0, stream>>>(tab, pairIdx, desc, state, arena, go, ge);
    else gotoh_band_kernel<false><<<blocks, 32 * wpb, 0, stream>>>(tab, pairIdx, desc, state, arena, go, ge);
    return cudaGetLastError();
}

cudaError_t launch_gotoh_band(int G, int Q, int nSlots, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream)
{
    BandLaunch one{G, Q, nSlots, 0};
    return launch_gotoh_band_multi(1, &one, pairIdx, desc, state, arena, go, ge, stream);
}

} // namespace lara
