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
// contiguous; four decision bits per cell go to trace units of Q nibbles ([row][block][pair slot]).  One thread per
// pair (G = 1): the two units of a row are written as one 16-byte record and two rows share a 32-byte sector
// ([row / 2][pair slot][row % 2]), because the traceback reads a single pair's rows and pays per sector.
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

constexpr unsigned kFull = 0xffffffffu;

// one trace unit = the Q decision nibbles of one run: 2, 4 or 8 bytes (Q = 12: 8 bytes, 12 nibbles used)
template <int Q>
__device__ __forceinline__ void store_unit(void * base, size_t unitIdx, uint32_t const (&tw)[(Q + 7) / 8])
{
    if constexpr (Q == 4) reinterpret_cast<uint16_t *>(base)[unitIdx] = (uint16_t)(tw[0] >> 16);
    else if constexpr (Q == 8) reinterpret_cast<uint32_t *>(base)[unitIdx] = tw[0];
    else if constexpr (Q == 12) reinterpret_cast<uint2 *>(base)[unitIdx] = make_uint2(tw[0], tw[1] >> 16);   // 12 of 16 nibbles
    else reinterpret_cast<uint2 *>(base)[unitIdx] = make_uint2(tw[0], tw[1]);
}

// One run of Q cells of row r on the diagonals of one block.  S[j], V[j]: latest cell of diagonal j (row r - 1 on
// entry, row r on exit).  (hL, sL): H and S of the left neighbour of the run; (sU, vU): S and V of the upper
// neighbour of the run's last cell.  Returns H of the run's last cell.  tw: Q decision nibbles, cell j in
// word j / 8 at bits 4 (j % 8) (Q = 4: in the upper half word).
// kLast: additionally captures S, V, H of cell jEnd (the end cell of the matrix, for SeqAn's _correctTraceValue).
template <int Q, bool kLast>
__device__ __forceinline__ int32_t band_run(int32_t (&S)[Q], int32_t (&V)[Q], int32_t const (&sub)[Q], int32_t hL, int32_t sL,
                                            int32_t sU, int32_t vU, int32_t go, int32_t ge, uint32_t (&tw)[(Q + 7) / 8],
                                            int jEnd, int32_t & sE, int32_t & vE, int32_t & hE)
{
    int32_t h = hL, m = sL + go;
    #pragma unroll
    for (int j = 0; j < Q; ++j)
    {
        int32_t const us = (j + 1 < Q) ? S[j + 1] : sU;
        int32_t const uv = (j + 1 < Q) ? V[j + 1] : vU;
        int32_t const dd = S[j] + sub[j];
        int32_t const ve = uv + ge, vo = us + go;
        int32_t const v = max(ve, vo);
        int32_t const he = h + ge;
        int32_t const hn = max(he, m);
        int32_t const gg = max(v, hn);
        int32_t const s = max(dd, gg);
        uint32_t n = (uint32_t)(he - m) >> 31;                          // bit 3: H opened   (extension wins ties)
        n = __funnelshift_l((uint32_t)(ve - vo), n, 1);                 // bit 2: V opened   (extension wins ties)
        n = __funnelshift_l((uint32_t)(v - hn), n, 1);                  // bit 1: gap from H (vertical wins ties)
        n = __funnelshift_l((uint32_t)(dd - gg), n, 1);                 // bit 0: S from gap (diagonal wins ties)
        tw[j >> 3] = __funnelshift_r(tw[j >> 3], n, 4);
        m = max(dd, v) + go;
        h = hn;
        S[j] = s;
        V[j] = v;
        if (kLast && j == jEnd) { sE = s; vE = v; hE = hn; }
    }
    return h;
}

template <int G, int Q>
__global__ void __launch_bounds__(128)
gotoh_band_kernel(int nSlots, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
                  PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    constexpr int PPW = 32 / G;                                 // pairs per warp
    constexpr int W = 2 * G * Q;                                // diagonals of the band
    constexpr int NCH = W / 4;                                  // 16-byte value chunks per row
    constexpr int NU = 2 * G;                                   // trace units per row
    constexpr int CQ = Q / 4;                                   // value chunks per block
    constexpr int TW = (Q + 7) / 8;
    constexpr int kPrefetchRows = 48 / Q;                       // L2 prefetch distance: about the same time for every Q
    int const warp = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
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
    uint32_t twA[2] = {0u, 0u};                                     // G == 1: block A's unit waits for block B's
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
            if (rowValid)
            {
                uint32_t tw[TW];
                #pragma unroll
                for (int k = 0; k < TW; ++k) tw[k] = 0u;
                if (r == lb && bEnd == 2 * g)
                    HA = band_run<Q, true>(SA, VA, subA, hL, sL, SB[0], VB[0], go, ge, tw, jEnd, sE, vE, hE);
                else
                    HA = band_run<Q, false>(SA, VA, subA, hL, sL, SB[0], VB[0], go, ge, tw, 0, sE, vE, hE);
                if constexpr (G == 1) { twA[0] = tw[0]; twA[1] = tw[TW - 1]; }
                else store_unit<Q>(trace, ((size_t)r * NU + (size_t)(2 * g)) * PPW + gs, tw);
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
            if (rowValid)
            {
                uint32_t tw[TW];
                #pragma unroll
                for (int k = 0; k < TW; ++k) tw[k] = 0u;
                if (r == lb && bEnd == 2 * g + 1)
                    HB = band_run<Q, true>(SB, VB, subB, HA, SA[Q - 1], sU, vU, go, ge, tw, jEnd, sE, vE, hE);
                else
                    HB = band_run<Q, false>(SB, VB, subB, HA, SA[Q - 1], sU, vU, go, ge, tw, 0, sE, vE, hE);
                if constexpr (G == 1)
                {
                    static_assert(G != 1 || TW == 2, "one thread per pair: units of 8 bytes");
                    uint4 const rec = (Q == 12) ? make_uint4(twA[0], twA[1] >> 16, tw[0], tw[TW - 1] >> 16)
                                                : make_uint4(twA[0], twA[1], tw[0], tw[TW - 1]);
                    reinterpret_cast<uint4 *>(trace)[(((size_t)(r >> 1) * PPW + gs) << 1) + (r & 1)] = rec;
                }
                else store_unit<Q>(trace, ((size_t)r * NU + (size_t)(2 * g + 1)) * PPW + gs, tw);
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

// Dense list of the slots whose pair is still iterating, in the old order (stable: the pairs of a warp group that
// all go on stay neighbours, so the DP's loads stay coalesced).  One block; n is a batch's slot count.
__global__ void __launch_bounds__(1024)
compact_slots_kernel(int n, int32_t const * __restrict__ in, PairState const * __restrict__ state,
                     int32_t * __restrict__ out, int32_t * __restrict__ count)
{
    __shared__ int warpCount[32];
    __shared__ int base;
    int const tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) base = 0;
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += 1024)
    {
        int const i = i0 + tid;
        int const p = i < n ? in[i] : -1;
        bool const live = p >= 0 && state[p].done == 0;
        unsigned const bal = __ballot_sync(kFull, live);
        if (lane == 0) warpCount[w] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
        #pragma unroll
        for (int k = 0; k < 32; ++k)
        {
            int const c = warpCount[k];
            before += k < w ? c : 0;
            total += c;
        }
        int const b = base;
        if (live) out[b + before + __popc(bal & ((1u << lane) - 1u))] = p;
        __syncthreads();
        if (tid == 0) base = b + total;
        __syncthreads();
    }
    if (tid == 0) *count = base;
}

template <int G, int Q>
cudaError_t launch_gq(int nSlots, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                      int32_t go, int32_t ge, cudaStream_t stream)
{
    int const slotsPerBlock = 4 * (32 / G);
    int const blocks = (nSlots + slotsPerBlock - 1) / slotsPerBlock;
    gotoh_band_kernel<G, Q><<<blocks, 128, 0, stream>>>(nSlots, pairIdx, desc, state, arena, go, ge);
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_compact_slots(int n, int32_t const * in, PairState const * state, int32_t * out, int32_t * count,
                                 cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    compact_slots_kernel<<<1, 1024, 0, stream>>>(n, in, state, out, count);
    return cudaGetLastError();
}

bool band_shape_supported(int G, int Q)
{
    if (Q == 16) return G == 1 || G == 2 || G == 4 || G == 8 || G == 16 || G == 32;
    if (Q == 12) return G == 1;
    if (Q == 8) return G == 2 || G == 4 || G == 8 || G == 16 || G == 32;
    if (Q == 4) return G == 4 || G == 8 || G == 16 || G == 32;
    return false;
}

cudaError_t launch_gotoh_band(int G, int Q, int nSlots, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream)
{
    if (nSlots <= 0)
        return cudaSuccess;
#define LARA_GQ(g_, q_) if (G == g_ && Q == q_) return launch_gq<g_, q_>(nSlots, pairIdx, desc, state, arena, go, ge, stream);
    LARA_GQ(1, 16) LARA_GQ(2, 16) LARA_GQ(4, 16) LARA_GQ(8, 16) LARA_GQ(16, 16) LARA_GQ(32, 16)
    LARA_GQ(1, 12)
    LARA_GQ(2, 8) LARA_GQ(4, 8) LARA_GQ(8, 8) LARA_GQ(16, 8) LARA_GQ(32, 8)
    LARA_GQ(4, 4) LARA_GQ(8, 4) LARA_GQ(16, 4) LARA_GQ(32, 4)
#undef LARA_GQ
    return cudaErrorInvalidValue;
}

} // namespace lara
