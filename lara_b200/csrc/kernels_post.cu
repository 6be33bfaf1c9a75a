// kernels_post.cu -- everything of one subgradient iteration that follows the DP fill, one warp per pair:
//
//   traceback            SeqAn:align/dp_traceback_impl.h:190-418,456-546 + dp_algorithm_impl.h:1168-1185
//   evaluateLines        src/lagrange.hpp:147-221
//   valid_solution       src/lagrange.hpp:325-423
//   greedy matching      src/matching.hpp:99-192 (lookahead-k with exact conflict resolution)
//   bounds / step / dual src/subgradient_solver.hpp:247-286  (scalar build)
//                        src/subgradient_solver_simd.hpp:292-350 (SIMD build, wrapping int32 step)
//   updateScores         src/lagrange.hpp:312-323
//   getStructureLines    src/lagrange.hpp:437-487
//
// The reference's per-edge std::set priority queues are not materialised: the top of queue l is
// max_m key(l,m) (kept in pv[l]) with key(l,m) = seqScore(l) + struct(l,m) + dual(l->m), and its
// iteration order among tied tops is ascending partner id (rows are stored in that order).
#include "lara_gpu.cuh"

#include <cstdlib>

namespace lara
{

namespace
{

constexpr unsigned kFull = 0xffffffffu;

constexpr int kMaxLookahead = 8;

// pair slots (x,y), x<y<8 in lexicographic order == order of the reference's conflicts vector
__constant__ uint8_t cSlotX[28] = {0,0,0,0,0,0,0, 1,1,1,1,1,1, 2,2,2,2,2, 3,3,3,3, 4,4,4, 5,5, 6};
__constant__ uint8_t cSlotY[28] = {1,2,3,4,5,6,7, 2,3,4,5,6,7, 3,4,5,6,7, 4,5,6,7, 5,6,7, 6,7, 7};

// involve mask of selection item x: the pair slots that contain x
constexpr uint32_t involve_mask(int x)
{
    constexpr uint8_t sx[28] = {0,0,0,0,0,0,0, 1,1,1,1,1,1, 2,2,2,2,2, 3,3,3,3, 4,4,4, 5,5, 6};
    constexpr uint8_t sy[28] = {1,2,3,4,5,6,7, 2,3,4,5,6,7, 3,4,5,6,7, 4,5,6,7, 5,6,7, 6,7, 7};
    uint32_t m = 0;
    for (int s = 0; s < 28; ++s)
        if (sx[s] == x || sy[s] == x)
            m |= 1u << s;
    return m;
}
__constant__ uint32_t cInvolve[8] = {involve_mask(0), involve_mask(1), involve_mask(2), involve_mask(3),
                                     involve_mask(4), involve_mask(5), involve_mask(6), involve_mask(7)};

// Matching::solveConflicts (src/matching.hpp:99-141) on bit masks: conf = set of conflicting pair slots,
// result = set of eliminated selection items.  In every conflict (first, second) the reference's edgeS is the
// later queue entry (second) and edgeL the earlier one (first), because the selection is in queue order.
__device__ uint32_t solve_conflicts(uint32_t conf, int32_t & weight, int32_t const * w, uint32_t const * inv)
{
    if (conf == 0u)
        return 0u;
    int const s = __ffs(conf) - 1;
    int const edgeL = cSlotX[s], edgeS = cSlotY[s];
    int32_t weightS = w[edgeS], weightL = w[edgeL];
    if (__popc(conf) == 1)
    {
        weight += weightS;
        return 1u << edgeS;
    }
    uint32_t const remS = conf & ~inv[edgeS];
    uint32_t elimS = solve_conflicts(remS, weightS, w, inv);
    if (weightS > weightL)
    {
        uint32_t const remL = conf & ~inv[edgeL];
        if (remS != remL)
        {
            uint32_t elimL = solve_conflicts(remL, weightL, w, inv);
            if (weightS > weightL)
            {
                weight += weightL;
                return elimL | (1u << edgeL);
            }
        }
    }
    weight += weightS;
    return elimS | (1u << edgeS);
}

struct WarpSmem
{
    int16_t * lineOfA;      // [maxLa]  line index of A position or -1
    uint32_t * linesRev;    // [maxLb]  (a << 16) | b in traceback order (last line first)
    int32_t * curE;         // [maxLb]  compact edge id of the k-th line that has interactions (alignment order)
    int16_t * workLine;     // [maxLb]  ... and its line index
    uint8_t * matched;      // [maxLb]
    uint8_t * seqB;         // [512]    Rna5 codes of the second sequence, in the upper half of win (free until step 3)
    unsigned long long * cand;  // [smemCand]
    int32_t * misc;         // [32]: 0 nList, 1 nCand, 2 matchScore, 3..10 selection weights, 11 error, 16..23 selection
    uint32_t * win;         // [256] traceback window (full mode: 32 rows x 32 nibbles; band mode: 8 word rows x L)
};

__device__ __forceinline__ uint32_t trace_nibble(int32_t const * __restrict__ tr, int R, int W, int nSteps, int c, int r)
{
    int const k = (r - 1) / R, j = (r - 1) - k * R;
    int const t = c - 1 + k;
    int const wj = j >> 3;
    int const nInWord = min(8, R - 8 * wj);
    uint32_t const word = (uint32_t)tr[(size_t)k * ((size_t)nSteps * W) + (size_t)t * W + wj];
    return (word >> (4 * (nInWord - 1 - (j & 7)))) & 15u;
}

// SeqAn trace byte of an inner cell from our nibble (SeqAn:align/dp_profile.h:125-134)
__device__ __forceinline__ uint32_t seqan_tv(uint32_t nib)
{
    uint32_t tv = (nib & kTrGap) ? ((nib & kTrFromH) ? 32u : 64u) : 1u;
    tv |= (nib & kTrVOpen) ? 16u : 4u;
    tv |= (nib & kTrHOpen) ? 8u : 2u;
    return tv;
}

// Traceback + evaluateLines, warp-cooperative and warp-uniform.
// The trace of the 32 x 32 cell window whose bottom-right corner is the walker's position is fetched by all
// lanes at once (lane i packs the nibbles of window row r - i into four words of shared memory: one round
// trip to L2/HBM per window instead of one per step).  Inside the window the walk advances by SEGMENTS, each
// found with one ballot over the 32 lanes:
//   S     lane i looks at the diagonal cell (c - i, r - i); the leading lanes whose cell took the diagonal are
//         one run of matches (_doTraceback, dp_traceback_impl.h:376-418) and write their lines in parallel;
//   VRUN  lane i looks at (c, r - i): `while ((tv & 4) && r != 1) step; step` of _doTracebackMaxFromVertical
//         (:223-264) = the leading lanes whose cell extended V, plus the opening step;
//   HRUN  lane i looks at (c - i, r), symmetric (:302-341).
// All lanes carry the same (c, r, state).  Returns the number of lines; lines in reverse order in linesRev;
// gapOut = gap score of evaluateLines (src/lagrange.hpp:147-221).
__device__ int traceback_walk(PairDesc const & d, PairState const & st, int32_t const * __restrict__ arena,
                              int32_t go, int32_t ge, uint32_t * linesRev, uint32_t * win, int32_t & gapOut, int lane)
{
    int32_t const * __restrict__ tr = arena + d.offTrace;
    int const R = d.R, W = d.W, nSteps = d.nSteps;
    int c = d.la, r = d.lb;
    int n = 0;
    int32_t gap = 0;
    int prev = 0;                       // 0 match, 1 vertical (gap in A), 2 horizontal (gap in B)
    int state = 0;                      // 0 S, 1 VRUN, 2 HRUN
    // end cell: _correctTraceValue (dp_algorithm_impl.h:1168-1185) + _retrieveInitialTraceDirection
    // (dp_traceback_impl.h:456-483, PreferGapsAtEnd_): 1 = forced vertical, 2 = forced horizontal
    int endFix = (st.vEnd == st.sEnd) ? 1 : ((st.hEnd == st.sEnd) ? 2 : 0);
    bool first = true;
    while (c > 0 && r > 0)
    {
        // ---- window rows [r-31, r], columns [c-31, c] --------------------------------------------
        int const cHi = c, rHi = r;
        {
            int const rr = rHi - lane;
            uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
            if (rr >= 1)
            {
                int const k = (rr - 1) / R, j = (rr - 1) - k * R;
                int const wj = j >> 3;
                int const nInWord = min(8, R - 8 * wj);
                int const sh = 4 * (nInWord - 1 - (j & 7));
                int32_t const * __restrict__ src = tr + (size_t)k * ((size_t)nSteps * W) + wj;
                int const tHi = cHi - 1 + k;                    // step of column cHi in strip k
                #pragma unroll
                for (int x = 0; x < 32; ++x)                    // window column x <-> matrix column cHi - x
                {
                    int const t = tHi - x;
                    uint32_t nib = 0;
                    if (cHi - x >= 1)
                        nib = ((uint32_t)__ldg(src + (size_t)t * W) >> sh) & 15u;
                    if (x < 8) w0 |= nib << (4 * x);
                    else if (x < 16) w1 |= nib << (4 * (x - 8));
                    else if (x < 24) w2 |= nib << (4 * (x - 16));
                    else w3 |= nib << (4 * (x - 24));
                }
            }
            __syncwarp();
            *reinterpret_cast<uint4 *>(win + 4 * lane) = make_uint4(w0, w1, w2, w3);
            __syncwarp();
        }
        while (c > 0 && r > 0 && cHi - c < 32 && rHi - r < 32)
        {
            int const rho = rHi - r, x = cHi - c;
            if (state == 0)
            {
                int const ri = rho + lane, xi = x + lane;
                bool const inw = ri < 32 && xi < 32 && c - lane >= 1 && r - lane >= 1;
                uint32_t nib = inw ? ((win[4 * ri + (xi >> 3)] >> (4 * (xi & 7))) & 15u) : 0u;
                if (first && lane == 0)
                {
                    if (endFix == 1) nib = (nib | kTrGap) & ~kTrFromH;
                    else if (endFix == 2) nib = nib | kTrGap | kTrFromH;
                }
                first = false;
                unsigned const diagMask = __ballot_sync(kFull, inw && !(nib & kTrGap));
                int const run = (diagMask == kFull) ? 32 : (__ffs(~diagMask) - 1);
                if (run > 0)
                {
                    if (lane < run)
                        linesRev[n + lane] = ((uint32_t)(c - lane - 1) << 16) | (uint32_t)(r - lane - 1);
                    n += run; c -= run; r -= run; prev = 0;
                    continue;
                }
                // lane 0's cell (always inside the window) took a gap: enter the run at this same cell
                uint32_t const nib0 = __shfl_sync(kFull, nib, 0);
                state = (nib0 & kTrFromH) ? 2 : 1;
            }
            if (state == 1)
            {
                int const ri = rho + lane;
                bool const inw = ri < 32 && r - lane >= 1;
                uint32_t const nib = inw ? ((win[4 * ri + (x >> 3)] >> (4 * (x & 7))) & 15u) : 0u;
                unsigned const extMask = __ballot_sync(kFull, inw && !(nib & kTrVOpen) && (r - lane) != 1);
                int const m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                // the cell that ends the run is lane m's; if it lies outside the window, stay in the run
                bool const closes = m < 32 && rho + m < 32;
                int const steps = closes ? m + 1 : m;
                if (steps > 0)
                {
                    gap += ((prev == 1) ? ge : go) + (steps - 1) * ge;
                    prev = 1;
                    r -= steps;
                }
                state = closes ? 0 : 1;
            }
            else
            {
                int const xi = x + lane;
                bool const inw = xi < 32 && c - lane >= 1;
                uint32_t const nib = inw ? ((win[4 * rho + (xi >> 3)] >> (4 * (xi & 7))) & 15u) : 0u;
                unsigned const extMask = __ballot_sync(kFull, inw && !(nib & kTrHOpen) && (c - lane) != 1);
                int const m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                bool const closes = m < 32 && x + m < 32;
                int const steps = closes ? m + 1 : m;
                if (steps > 0)
                {
                    gap += ((prev == 2) ? ge : go) + (steps - 1) * ge;
                    prev = 2;
                    c -= steps;
                }
                state = closes ? 0 : 2;
            }
        }
    }
    // leading gaps (dp_traceback_impl.h:541-544): vertical first, then horizontal
    if (r > 0) { gap += ((prev == 1) ? ge : go) + (r - 1) * ge; prev = 1; }
    if (c > 0) { gap += ((prev == 2) ? ge : go) + (c - 1) * ge; prev = 2; }
    gapOut = gap;
    return n;
}

// Traceback + evaluateLines over the dense decision matrix of the row-by-row kernel (kernels_rows.cu: one byte per cell):
// the plain walk of dp_traceback_impl.h:190-418, every lane the same steps (the pairs that come here are few and long).
__device__ int traceback_dense(PairDesc const & d, PairState const & st, int32_t const * __restrict__ arena,
                               int32_t go, int32_t ge, uint32_t * linesRev, int32_t & gapOut, int lane)
{
    uint8_t const * __restrict__ tr = reinterpret_cast<uint8_t const *>(arena + d.offTrace);
    size_t const ncol = (size_t)d.la + 1;
    int c = d.la, r = d.lb;
    int n = 0;
    int32_t gap = 0;
    int prev = 0;
    int const endFix = (st.vEnd == st.sEnd) ? 1 : ((st.hEnd == st.sEnd) ? 2 : 0);
    bool first = true;
    while (c > 0 && r > 0)
    {
        uint32_t nib = tr[(size_t)r * ncol + c];
        if (first)
        {
            if (endFix == 1) nib = (nib | kTrGap) & ~kTrFromH;
            else if (endFix == 2) nib = nib | kTrGap | kTrFromH;
            first = false;
        }
        if (!(nib & kTrGap))
        {
            if (lane == 0) linesRev[n] = ((uint32_t)(c - 1) << 16) | (uint32_t)(r - 1);
            ++n; --c; --r; prev = 0;
        }
        else if (!(nib & kTrFromH))
        {
            // vertical run: cells that extended V, closed by the cell that opened it (or by row 1)
            int steps = 0;
            for (;;)
            {
                uint32_t const x = tr[(size_t)(r - steps) * ncol + c];
                bool const ext = !(x & kTrVOpen) && (r - steps) != 1;
                ++steps;
                if (!ext) break;
            }
            gap += ((prev == 1) ? ge : go) + (steps - 1) * ge;
            prev = 1;
            r -= steps;
        }
        else
        {
            int steps = 0;
            for (;;)
            {
                uint32_t const x = tr[(size_t)r * ncol + (c - steps)];
                bool const ext = !(x & kTrHOpen) && (c - steps) != 1;
                ++steps;
                if (!ext) break;
            }
            gap += ((prev == 2) ? ge : go) + (steps - 1) * ge;
            prev = 2;
            c -= steps;
        }
    }
    if (r > 0) { gap += ((prev == 1) ? ge : go) + (r - 1) * ge; prev = 1; }
    if (c > 0) { gap += ((prev == 2) ? ge : go) + (c - 1) * ge; prev = 2; }
    __syncwarp();
    gapOut = gap;
    return n;
}

// ---- band mode (kernels_band.cu) ---------------------------------------------------------------------------
// Trace decisions of ANY cell of the matrix: inside the band from the stored nibbles, outside from the closed
// forms of the pure gap regions U (c - r > dHi) and L (c - r < dLo); see tools/band_proto.py (u_nib / l_nib).
struct BandView
{
    int la, lb, dLo, dHi;
    int W, Q, ppw, slot;                    // band geometry and position in the warp group (lara_gpu.cuh: PairDesc)
    int32_t go, ge;
    uint16_t const * trace;                 // group base of the trace, in half words (4 nibbles each)
    int4 const * uRec;                      // per row    { TH, TS, M, 0 } at [r * ppw]   (already offset by the slot)
    int4 const * lRec;                      // per column { TV, TS, M, 0 } at [c * ppw]
};

__device__ __forceinline__ int4 band_u(BandView const & v, int r) { return __ldg(v.uRec + (size_t)r * v.ppw); }
__device__ __forceinline__ int4 band_l(BandView const & v, int c) { return __ldg(v.lRec + (size_t)c * v.ppw); }
__device__ __forceinline__ int32_t band_uV(BandView const & v, int4 const & u, int c, int r)
{
    return r == 0 ? kBandNeg : (c + r - 1) * v.ge + v.go + u.z;
}
__device__ __forceinline__ int32_t band_uH(BandView const & v, int4 const & u, int c, int r)
{
    int32_t const t = max(u.x, u.y);
    bool const first = (c == r + v.dHi + 1) || r == 0;
    return c * v.ge + (first ? t : max(t, u.z + 2 * v.go + (r - 2) * v.ge));
}
__device__ __forceinline__ int32_t band_lH(BandView const & v, int4 const & l, int c, int r)
{
    return c == 0 ? kBandNeg : (c + r - 1) * v.ge + v.go + l.z;
}
__device__ __forceinline__ int32_t band_lV(BandView const & v, int4 const & l, int c, int r)
{
    int32_t const t = max(l.x, l.y);
    bool const first = (r == c - v.dLo + 1) || c == 0;
    return r * v.ge + (first ? t : max(t, l.z + 2 * v.go + (c - 2) * v.ge));
}
// nibble of a cell outside the band (c >= 1, r >= 1): the diagonal never wins there
__device__ uint32_t band_outside_nib(BandView const & v, int c, int r)
{
    uint32_t nib = kTrGap;
    if (c - r > v.dHi)
    {
        int4 const u = band_u(v, r), up = band_u(v, r - 1);
        int32_t const hCur = band_uH(v, u, c, r), vCur = band_uV(v, u, c, r);
        if (c == r + v.dHi + 1) { if (u.x < u.y) nib |= kTrHOpen; }
        else
        {
            int32_t const hl = band_uH(v, u, c - 1, r), vl = band_uV(v, u, c - 1, r);
            if (hl + v.ge < max(hl, vl) + v.go) nib |= kTrHOpen;
        }
        int32_t const hu = band_uH(v, up, c, r - 1), vu = band_uV(v, up, c, r - 1);
        if (vu + v.ge < max(hu, vu) + v.go) nib |= kTrVOpen;
        if (vCur < hCur) nib |= kTrFromH;
    }
    else
    {
        int4 const l = band_l(v, c), lp = band_l(v, c - 1);
        int32_t const hCur = band_lH(v, l, c, r), vCur = band_lV(v, l, c, r);
        if (r == c - v.dLo + 1) { if (l.x < l.y) nib |= kTrVOpen; }
        else
        {
            int32_t const hu = band_lH(v, l, c, r - 1), vu = band_lV(v, l, c, r - 1);
            if (vu + v.ge < max(hu, vu) + v.go) nib |= kTrVOpen;
        }
        int32_t const hl = band_lH(v, lp, c - 1, r), vl = band_lV(v, lp, c - 1, r);
        if (hl + v.ge < max(hl, vl) + v.go) nib |= kTrHOpen;
        if (vCur < hCur) nib |= kTrFromH;
    }
    return nib;
}

// S, V, H of the end cell when it lies outside the band
__device__ void band_end_values(BandView const & v, int32_t & sE, int32_t & vE, int32_t & hE)
{
    int const c = v.la, r = v.lb;
    if (c - r > v.dHi) { int4 const u = band_u(v, r); vE = band_uV(v, u, c, r); hE = band_uH(v, u, c, r); }
    else { int4 const l = band_l(v, c); vE = band_lV(v, l, c, r); hE = band_lH(v, l, c, r); }
    sE = max(vE, hE);
}

// Traceback + evaluateLines for band mode: the segment walker of traceback_walk over a polymorphic cell fetch.
// Window = 32 rows x 32 band diagonals around the walker, in shared memory as four decision planes per row: lane i
// fetches row rHi - i (the trace units of kernels_band.cu are bit planes already) and assembles plane k of
// that row, window diagonal x at bit x, stored plane by plane (win[32 k + i]).  The walk asks for one decision of a
// cell at a time.
__device__ int traceback_band(BandView const & v, int32_t sEnd, int32_t vEnd, int32_t hEnd, uint32_t * linesRev,
                              uint32_t * win, int32_t & gapOut, int lane)
{
    int32_t const go = v.go, ge = v.ge;
    int c = v.la, r = v.lb;
    int n = 0;
    int32_t gap = 0;
    int prev = 0, state = 0;
    int endFix = (vEnd == sEnd) ? 1 : ((hEnd == sEnd) ? 2 : 0);
    bool first = true;
    int const nUnits = v.W / v.Q;
    while (c > 0 && r > 0)
    {
        // ---- window: rows [rHi - 31, rHi], diagonal indices [xLo, xLo + 32) ----------------------------------
        int const rHi = r;
        int xLo = 0;
        if (v.W > 32)
        {
            int const x = min(max(c - r - v.dLo, 0), v.W - 1);
            xLo = min(max((x - 8) & ~15, 0), v.W - 32);
        }
        __syncwarp();
        {
            int const rr = rHi - lane;
            uint32_t w[4] = {0u, 0u, 0u, 0u};
            if (rr >= 1)
            {
                if (v.ppw == 32 || v.Q == 16)
                {
                    // units of 8 bytes { P0 | P1 << 16, P2 | P3 << 16 }; one thread per pair (Q = 16 or 12): both units of
                    // a row in one 16-byte record, two rows per sector; else two neighbouring units (xLo is a multiple of 16)
                    uint4 u;
                    if (v.ppw == 32)
                        u = __ldg(reinterpret_cast<uint4 const *>(v.trace) + ((((size_t)(rr >> 1) * 32 + v.slot) << 1) + (rr & 1)));
                    else
                    {
                        uint2 const * __restrict__ src = reinterpret_cast<uint2 const *>(v.trace) +
                                                         ((size_t)rr * nUnits + (xLo >> 4)) * v.ppw + v.slot;
                        uint2 const u0 = __ldg(src), u1 = __ldg(src + v.ppw);
                        u = make_uint4(u0.x, u0.y, u1.x, u1.y);
                    }
                    if (v.Q == 16)
                    {
                        w[0] = __byte_perm(u.x, u.z, 0x5410); w[1] = __byte_perm(u.x, u.z, 0x7632);
                        w[2] = __byte_perm(u.y, u.w, 0x5410); w[3] = __byte_perm(u.y, u.w, 0x7632);
                    }
                    else
                    {
                        // 12-bit planes: block B's continue at bit 12
                        w[0] = (u.x & 0xfffu) | ((u.z & 0xfffu) << 12); w[1] = ((u.x >> 16) & 0xfffu) | (((u.z >> 16) & 0xfffu) << 12);
                        w[2] = (u.y & 0xfffu) | ((u.w & 0xfffu) << 12); w[3] = ((u.y >> 16) & 0xfffu) | (((u.w >> 16) & 0xfffu) << 12);
                    }
                }
                else if (v.Q == 8)
                {
                    // units of one word P0 | P1 << 8 | P2 << 16 | P3 << 24: plane k of the window = byte k of four units
                    uint32_t const * __restrict__ src = reinterpret_cast<uint32_t const *>(v.trace) +
                                                        ((size_t)rr * nUnits + (xLo >> 3)) * v.ppw + v.slot;
                    uint32_t const a0 = __ldg(src), a1 = __ldg(src + v.ppw), a2 = __ldg(src + 2 * (size_t)v.ppw), a3 = __ldg(src + 3 * (size_t)v.ppw);
                    uint32_t const lo01 = __byte_perm(a0, a1, 0x5140), hi01 = __byte_perm(a0, a1, 0x7362);   // { a0.b0, a1.b0, a0.b1, a1.b1 }, { b2, b2, b3, b3 }
                    uint32_t const lo23 = __byte_perm(a2, a3, 0x5140), hi23 = __byte_perm(a2, a3, 0x7362);
                    w[0] = __byte_perm(lo01, lo23, 0x5410); w[1] = __byte_perm(lo01, lo23, 0x7632);
                    w[2] = __byte_perm(hi01, hi23, 0x5410); w[3] = __byte_perm(hi01, hi23, 0x7632);
                }
                else
                {
                    // units of one half word P0 | P1 << 4 | P2 << 8 | P3 << 12: plane k of the window = nibble k of eight units
                    uint16_t const * __restrict__ src = v.trace + ((size_t)rr * nUnits + (xLo >> 2)) * v.ppw + v.slot;
                    #pragma unroll
                    for (int h = 0; h < 8; ++h)
                    {
                        uint32_t const u = (uint32_t)__ldg(src + (size_t)h * v.ppw);
                        #pragma unroll
                        for (int k = 0; k < 4; ++k) w[k] |= ((u >> (4 * k)) & 15u) << (4 * h);
                    }
                }
            }
            // plane-major: the 32 lanes of a diagonal or vertical look-up read 32 neighbouring words (no bank conflicts)
            #pragma unroll
            for (int k = 0; k < 4; ++k) win[32 * k + lane] = w[k];
        }
        __syncwarp();
        // One decision (bit `plane` of the trace nibble) of cell (ci, ri).
        // known: the decision can be evaluated now; outside: the cell lies in a closed-form region and was not
        // evaluated (lazy == true)
        auto fetch = [&](int ci, int ri, int plane, bool & known, bool & outside, bool lazy) -> uint32_t
        {
            known = false; outside = false;
            if (ci < 1 || ri < 1) return 0u;
            int const dd = ci - ri;
            if (dd > v.dHi || dd < v.dLo)
            {
                if (lazy) { outside = true; return 0u; }
                known = true;
                return (band_outside_nib(v, ci, ri) >> plane) & 1u;
            }
            int const x = dd - v.dLo - xLo, row = rHi - ri;
            if (row > 31 || x < 0 || x > 31) return 0u;
            known = true;
            return (win[32 * plane + row] >> x) & 1u;
        };
        for (;;)
        {
            if (c <= 0 || r <= 0) break;
            {
                int const dd = c - r, x = dd - v.dLo - xLo;
                if (dd <= v.dHi && dd >= v.dLo && (rHi - r > 31 || x < 0 || x > 31)) break;   // current cell left the window
            }
            if (state == 0)
            {
                bool known, outside;
                uint32_t isGap = fetch(c - lane, r - lane, 0, known, outside, false);   // one diagonal: all in or all out
                if (first && lane == 0 && endFix != 0) isGap = 1u;
                unsigned const diagMask = __ballot_sync(kFull, known && !isGap);
                int const run = (diagMask == kFull) ? 32 : (__ffs(~diagMask) - 1);
                if (run > 0)
                {
                    if (lane < run)
                        linesRev[n + lane] = ((uint32_t)(c - lane - 1) << 16) | (uint32_t)(r - lane - 1);
                    n += run; c -= run; r -= run; prev = 0;
                    first = false;
                    continue;
                }
                // the walker's own cell took a gap: from the horizontal or the vertical matrix?
                bool k2, o2;
                uint32_t fromH = fetch(c, r, 1, k2, o2, false);
                if (first) { if (endFix == 1) fromH = 0u; else if (endFix == 2) fromH = 1u; }
                first = false;
                state = fromH ? 2 : 1;
            }
            if (state == 1)
            {
                // closed-form cells (beyond the band) are only evaluated when the run reaches them
                bool known, outside;
                uint32_t opened = fetch(c, r - lane, 2, known, outside, true);
                unsigned knownMask = __ballot_sync(kFull, known);
                unsigned extMask = __ballot_sync(kFull, known && !opened && (r - lane) != 1);
                int m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                if (m < 32 && ((__ballot_sync(kFull, outside) >> m) & 1u))
                {
                    if (outside) { opened = (band_outside_nib(v, c, r - lane) >> 2) & 1u; known = true; }
                    knownMask = __ballot_sync(kFull, known);
                    extMask = __ballot_sync(kFull, known && !opened && (r - lane) != 1);
                    m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                }
                bool const closes = m < 32 && ((knownMask >> m) & 1u);
                int const steps = closes ? m + 1 : m;
                if (steps > 0)
                {
                    gap += ((prev == 1) ? ge : go) + (steps - 1) * ge;
                    prev = 1;
                    r -= steps;
                }
                state = closes ? 0 : 1;
                if (steps == 0) break;                      // next cell not in the window: reload
            }
            else
            {
                bool known, outside;
                uint32_t opened = fetch(c - lane, r, 3, known, outside, true);
                unsigned knownMask = __ballot_sync(kFull, known);
                unsigned extMask = __ballot_sync(kFull, known && !opened && (c - lane) != 1);
                int m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                if (m < 32 && ((__ballot_sync(kFull, outside) >> m) & 1u))
                {
                    if (outside) { opened = (band_outside_nib(v, c - lane, r) >> 3) & 1u; known = true; }
                    knownMask = __ballot_sync(kFull, known);
                    extMask = __ballot_sync(kFull, known && !opened && (c - lane) != 1);
                    m = (extMask == kFull) ? 32 : (__ffs(~extMask) - 1);
                }
                bool const closes = m < 32 && ((knownMask >> m) & 1u);
                int const steps = closes ? m + 1 : m;
                if (steps > 0)
                {
                    gap += ((prev == 2) ? ge : go) + (steps - 1) * ge;
                    prev = 2;
                    c -= steps;
                }
                state = closes ? 0 : 2;
                if (steps == 0) break;
            }
        }
    }
    if (r > 0) { gap += ((prev == 1) ? ge : go) + (r - 1) * ge; prev = 1; }
    if (c > 0) { gap += ((prev == 2) ? ge : go) + (c - 1) * ge; prev = 2; }
    gapOut = gap;
    return n;
}

__device__ __forceinline__ int32_t warp_sum(int32_t v)
{
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
__device__ __forceinline__ int32_t warp_min(int32_t v)
{
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ int32_t warp_max(int32_t v)
{
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(kFull, v, o));
    return v;
}

// bitonic sort of P (power of two, >= 64 or == 32) 64-bit keys by one warp; every lane owns one compare-exchange per turn
__device__ void warp_bitonic_sort(unsigned long long * buf, int P, int lane)
{
    for (int k = 2; k <= P; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1)
        {
            for (int t = lane; t < (P >> 1); t += 32)
            {
                int const i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), ixj = i | j;
                unsigned long long const a = buf[i], b = buf[ixj];
                bool const asc = (i & k) == 0;
                if ((a > b) == asc) { buf[i] = b; buf[ixj] = a; }
            }
            __syncwarp();
        }
}

// Matching::computeGreedyMatching (src/matching.hpp:143-192) over K sorted candidates.
// key = [~(sc ^ 0x80000000) : 32][line index l : 16][line index r : 16]  (ascending == queue order
// (-2*sc, lineL, lineR) because line indices grow with the edge ids).
__device__ int32_t greedy_matching(unsigned long long const * keys, int K, int lookahead, uint8_t * matched,
                                   int32_t * misc, int lane)
{
    int L = lookahead;
    if (L > K) L = K; else if (L == 0) L = 5;                  // :153-156
    uint32_t const * const inv = cInvolve;
    int32_t score = 0;
    int pos = 0;
    while (pos < K)
    {
        // selection: the next L candidates whose endpoints are both unmatched (:165-167)
        int ns = 0;
        int mySel = -1;
        while (ns < L && pos < K)
        {
            int const i = pos + lane;
            bool ok = false;
            if (i < K)
            {
                unsigned long long const key = keys[i];
                ok = !matched[(key >> 16) & 0xffffu] && !matched[key & 0xffffu];
            }
            unsigned const bal = __ballot_sync(kFull, ok);
            int const have = __popc(bal), need = L - ns;
            int const take = min(have, need);
            // the lane of the k-th unmatched candidate hands it to selection slot ns + k
            int const rank = __popc(bal & ((1u << lane) - 1u));
            if (ok && rank < take) misc[16 + ns + rank] = i;
            if (have >= need)
                pos = pos + __ffs(__ballot_sync(kFull, ok && rank == need - 1));
            else
                pos = min(pos + 32, K);
            ns += take;
        }
        __syncwarp();
        if (lane < ns) mySel = misc[16 + lane];
        // my selected interaction
        int32_t w = 0;
        int il = -1, ir = -2;
        if (lane < ns)
        {
            unsigned long long const key = keys[mySel];
            uint32_t const hi = (uint32_t)(key >> 32);
            int32_t const sc = (int32_t)((~hi) ^ 0x80000000u);
            w = 2 * sc;                                         // -get<0>(interaction)
            il = (int)((key >> 16) & 0xffffu);
            ir = (int)(key & 0xffffu);
            misc[3 + lane] = w;
        }
        __syncwarp();
        // conflicts (:170-174): lane s tests pair slot s
        bool cf = false;
        if (lane < 28)
        {
            int const x = cSlotX[lane], y = cSlotY[lane];
            int const lx = __shfl_sync(0x0fffffffu, il, x), rx = __shfl_sync(0x0fffffffu, ir, x);
            int const ly = __shfl_sync(0x0fffffffu, il, y), ry = __shfl_sync(0x0fffffffu, ir, y);
            cf = (y < ns) && (lx == ly || lx == ry || rx == ly || rx == ry);
        }
        uint32_t const conf = __ballot_sync(kFull, cf);
        uint32_t elim = 0;
        if (conf)
        {
            if (lane == 0)
            {
                int32_t weight = 0;
                elim = solve_conflicts(conf, weight, misc + 3, inv);
            }
            elim = __shfl_sync(kFull, elim, 0);
        }
        // commit (:181-189)
        int32_t add = 0;
        if (lane < ns && !((elim >> lane) & 1u))
        {
            matched[il] = 1;
            matched[ir] = 1;
            add = w;
        }
        score += add;
        __syncwarp();
    }
    return warp_sum(score);
}

__device__ __forceinline__ int edge_of(PairDesc const & d, int32_t const * __restrict__ arena, int a, int b, bool & hasInter,
                                       int & codeA)
{
    codeA = -1;                                 // without a row record: read the sequence
    hasInter = true;                            // without a row record: unknown, the edge record tells
    if (d.offRowRec)
    {
        // narrow band: the row's active cells fit one word, stored next to the id of the row's first edge and the
        // word of the cells that have interactions
        int4 const rec = __ldg(reinterpret_cast<int4 const *>(arena + d.offRowRec) + a);
        int const k = b - (a - d.dHi);
        codeA = rec.w;
        if (k < 0 || k > 31 || !(((uint32_t)rec.y >> k) & 1u)) return -1;
        hasInter = ((uint32_t)rec.z >> k) & 1u;
        return rec.x + __popc((uint32_t)rec.y & ((1u << k) - 1u));
    }
    uint32_t const * bits = (uint32_t const *)(arena + d.offBits) + (size_t)a * d.bitWords;
    int e = arena[d.offRowStart + a];
    int const wb = b >> 5;
    // band mode: every active cell of row a lies at b >= a - dHi, the words before it are empty
    int const w0 = d.bandW > 0 ? (max(a - d.dHi, 0) >> 5) : 0;
    for (int w = w0; w < wb; ++w) e += __popc(bits[w]);
    uint32_t const word = bits[wb];
    if (!((word >> (b & 31)) & 1u)) return -1;
    return e + __popc(word & ((1u << (b & 31)) - 1u));
}
__device__ __forceinline__ int edge_of(PairDesc const & d, int32_t const * __restrict__ arena, int a, int b)
{
    bool h;
    int c;
    return edge_of(d, arena, a, b, h, c);
}

// 80 registers (6 x 128 threads): with one-warp blocks 24 pairs per SM.  More pairs per SM (72 or 64 registers, or
// more shared memory at the cost of L1) measured slower, fewer (90 registers, what ptxas takes unbounded) too.
template <bool kTracebackOnly>
__device__ __forceinline__ void
iterate_post_pair(int const p, PairDesc const * __restrict__ desc, PairState * __restrict__ state,
                  int32_t * __restrict__ arena, DevParams const & P, PairOutHdr * __restrict__ outHdr,
                  int32_t * __restrict__ outLines, int64_t const * __restrict__ outOff,
                  int32_t * __restrict__ finishedList, int maxLa, int maxLb, int smemCand, int tune, DevInputs const & in)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    int const warpInBlock = threadIdx.x >> 5;
    int const lane = threadIdx.x & 31;
    PairState & st = state[p];
    if (st.done)
        return;
    PairDesc const & d = desc[p];

    // carve the per-warp shared memory
    size_t const candBytes = (size_t)smemCand * 8;
    size_t const lb4 = ((size_t)maxLb * 4 + 15) / 16 * 16;
    size_t const perWarp = candBytes + 128 + 1024 + 2 * lb4 + ((size_t)maxLa * 2 + 15) / 16 * 16 + ((size_t)maxLb * 2 + 15) / 16 * 16 +
                           ((size_t)maxLb + 15) / 16 * 16;
    unsigned char * base = smemRaw + (size_t)warpInBlock * perWarp;
    WarpSmem sm;
    sm.cand = (unsigned long long *)base;                     base += candBytes;
    sm.misc = (int32_t *)base;                                base += 128;
    sm.win = (uint32_t *)base;                                base += 1024;
    sm.linesRev = (uint32_t *)base;                           base += lb4;
    sm.curE = (int32_t *)base;                                base += lb4;
    sm.lineOfA = (int16_t *)base;                             base += ((size_t)maxLa * 2 + 15) / 16 * 16;
    sm.workLine = (int16_t *)base;                            base += ((size_t)maxLb * 2 + 15) / 16 * 16;
    sm.matched = (uint8_t *)base;
    sm.seqB = (uint8_t *)(sm.win + 128);

    int const la = d.la, lb = d.lb;
    (void)lb;
    // bases of the second sequence for the sequence scores of step 2 (the row records carry those of the first); no
    // shared memory of their own: one more KB per warp costs the kernel its L1
    bool const stagedB = d.offRowRec != 0 && d.lb <= 512;
    if (stagedB)
    {
        uint8_t const * __restrict__ sb = in.seqCat + in.seqOff[d.seqSecond];
        for (int i = lane; i < d.lb; i += 32) sm.seqB[i] = sb[i];
    }
    if (d.offRowRec && !(tune & 4))
    {
        // the row records of the edge lookup (step 2; 16 bytes per row): into L2 while the traceback runs
        int4 const * rr = reinterpret_cast<int4 const *>(arena + d.offRowRec);
        for (int a = 2 * lane; a < la; a += 64)
            asm volatile("prefetch.global.L2 [%0];" :: "l"(rr + a));
    }

    // ---- 1. traceback + evaluateLines ------------------------------------------------------------
    int32_t gapScore = 0;
    int nLines;
    if (d.bandW > 0)
    {
        BandView bv;
        bv.la = d.la; bv.lb = d.lb; bv.dLo = d.dLo; bv.dHi = d.dHi; bv.go = P.go; bv.ge = P.ge;
        bv.W = d.bandW; bv.Q = d.bandQ; bv.ppw = 32 / d.bandG; bv.slot = d.bandSlot;
        bv.trace = reinterpret_cast<uint16_t const *>(arena + d.offBTrace);
        bv.uRec = reinterpret_cast<int4 const *>(arena + d.offU) + d.bandSlot;
        bv.lRec = reinterpret_cast<int4 const *>(arena + d.offL) + d.bandSlot;
        if (bv.ppw == 32 && !(tune & 1))
        {
            // one-thread-per-pair shapes: the whole trace of the pair is 16 bytes per row, two rows per sector; pull it
            // into L2 now (last rows first), so that only the first window of the walk waits for HBM
            uint4 const * tb = reinterpret_cast<uint4 const *>(bv.trace) + ((size_t)bv.slot << 1);
            for (int s = (d.lb >> 1) - lane; s >= 0; s -= 32)
                asm volatile("prefetch.global.L2 [%0];" :: "l"(tb + (size_t)s * 64));
        }
        int32_t sE = st.sEnd, vE = st.vEnd, hE = st.hEnd;
        int const dEnd = d.la - d.lb;
        if (dEnd > d.dHi || dEnd < d.dLo)
        {
            // the end cell lies in a pure gap region: its values are closed forms too
            band_end_values(bv, sE, vE, hE);
            __syncwarp();
            if (lane == 0) { st.currentUpper = sE; st.sEnd = sE; st.vEnd = vE; st.hEnd = hE; }
            __syncwarp();
        }
        nLines = traceback_band(bv, sE, vE, hE, sm.linesRev, sm.win, gapScore, lane);
    }
    else if (d.dense)
        nLines = traceback_dense(d, st, arena, P.go, P.ge, sm.linesRev, gapScore, lane);
    else
        nLines = traceback_walk(d, st, arena, P.go, P.ge, sm.linesRev, sm.win, gapScore, lane);
    if (lane == 0) { sm.misc[0] = 0; sm.misc[1] = 0; sm.misc[2] = 0; sm.misc[11] = 0; }
    for (int a = lane; a < la; a += 32) sm.lineOfA[a] = -1;
    __syncwarp();
    if (kTracebackOnly)
    {
        // unit entry point: lines in alignment order to offLines, count and gap score to the state
        for (int i = lane; i < nLines; i += 32)
            arena[d.offLines + i] = (int32_t)sm.linesRev[nLines - 1 - i];
        if (lane == 0) { st.nLines = nLines; st.gapScore = gapScore; }
        return;
    }

    // per edge { seqScore, pv, ioff, degree }; per interaction the hot record { reverse index, partner cell, structScore, dual }
    // (16 bytes: the interactions of a row are neighbours, four to a 64-byte fetch) and, in an array of its own, the
    // cold one { partner edge, seqScore of the partner, owner, 0 }  (lara_gpu.cuh: offEdge, offInter, offInterCold)
    int4 * __restrict__ edge = reinterpret_cast<int4 *>(arena + d.offEdge);
    int4 * __restrict__ inter = reinterpret_cast<int4 *>(arena + d.offInter);
    int4 const * __restrict__ cold = reinterpret_cast<int4 const *>(arena + d.offInterCold);
    int32_t * __restrict__ list = arena + d.offList;
    // candidates are collected in shared memory when their upper bound fits, else in the pair's global buffer
    // (and moved to shared memory for the sort when the actual number fits)
    bool const candInSmem = d.kCap <= smemCand;
    unsigned long long * cand = candInSmem ? sm.cand : (unsigned long long *)(arena + d.offCand);

    // ---- 2. lines -> edge ids --------------------------------------------------------------------
    // A line whose edge has no interaction only adds its sequence score to the lower bound (:343-349), a function
    // of the two bases (setup_count_kernel); the lines that have interactions are gathered, in alignment order, for
    // step 3, which then reads the edge records of those lines only.
    int32_t lbSeq = 0;
    int nWork = 0;
    {
        uint8_t const * __restrict__ seqA = in.seqCat + in.seqOff[d.seqFirst];
        uint8_t const * __restrict__ seqB = in.seqCat + in.seqOff[d.seqSecond];
        float const scale = st.scale;
        // the row record of a line is loaded one round ahead of its use (the records were pulled into L2 when the kernel
        // started, an L2 hit still takes longer than a round)
        bool const hasRec = d.offRowRec != 0;
        int4 const * __restrict__ rowRec = reinterpret_cast<int4 const *>(arena + d.offRowRec);
        uint32_t abNext = lane < nLines ? sm.linesRev[nLines - 1 - lane] : 0u;
        int4 recNext = make_int4(0, 0, 0, 0);
        if (hasRec && lane < nLines) recNext = __ldg(rowRec + (abNext >> 16));
        for (int i0 = 0; i0 < nLines; i0 += 32)
        {
            int const i = i0 + lane;
            bool work = false;
            int e = -1;
            uint32_t const ab = abNext;
            int4 const rec = recNext;
            if (i + 32 < nLines)
            {
                abNext = sm.linesRev[nLines - 1 - (i + 32)];
                if (hasRec) recNext = __ldg(rowRec + (abNext >> 16));
            }
            if (i < nLines)
            {
                int const a = (int)(ab >> 16), b = (int)(ab & 0xffffu);
                sm.lineOfA[a] = (int16_t)i;
                int codeA = -1;
                if (hasRec)
                {
                    // narrow band: the row's active cells fit one word (edge_of)
                    int const k = b - (a - d.dHi);
                    codeA = rec.w;
                    if (k >= 0 && k <= 31 && (((uint32_t)rec.y >> k) & 1u))
                    {
                        work = ((uint32_t)rec.z >> k) & 1u;
                        e = rec.x + __popc((uint32_t)rec.y & ((1u << k) - 1u));
                    }
                }
                else
                    e = edge_of(d, arena, a, b, work, codeA);
                if (e < 0) { sm.misc[11] = 1; work = false; }   // "Alignment match where no alignment edge is defined!"
                else if (!work)
                {
                    // the row record carries the base of A, the bases of B were staged when the kernel started
                    int const ca = codeA >= 0 ? codeA : (int)seqA[a];
                    int const cb = stagedB ? (int)sm.seqB[b] : (int)seqB[b];
                    lbSeq += __float2int_rz(__fmul_rn(scale, __int2float_rn(P.tab[ca * 5 + cb])));
                }
                sm.matched[i] = 0;
            }
            unsigned const bal = __ballot_sync(kFull, work);
            if (work)
            {
                int const k = nWork + __popc(bal & ((1u << lane) - 1u));
                sm.curE[k] = e;
                sm.workLine[k] = (int16_t)i;
            }
            nWork += __popc(bal);
        }
    }
    __syncwarp();

    // ---- 3. subgradient marks, sequence part of the lower bound, matching candidates (:343-391) ----
    // 32 lines at a time: their interactions are numbered consecutively (prefix sum of the degrees) and dealt to
    // the lanes, kItems per lane and round, so that the record loads of a round are all in flight together.
    constexpr int kItems = 4;
    constexpr int kBitWords = 111;
    // sm.win is 256 words: [0, 33) offsets, then two bitmaps with one bit per interaction of the pair
    uint32_t * const sOff = sm.win;                                 // [33] first item of each of the 32 lines
    uint32_t * const tieBits = sm.win + 34;                         // "is a tied top of its line"
    uint32_t * const metBits = tieBits + kBitWords;                 // "one of x and its reverse has been seen" (bit min(x, rev))
    // The cycle test of a tied top x of line l whose partner m is in the solution (:359-366) asks whether the reverse
    // interaction is a tied top of m.  m is a line too, so this very loop also walks m's row.  Whichever of x and its
    // reverse is seen second (an atomic on the pair's bit decides) finds the other one's answer in the bitmap and
    // settles both: no record of m is read.  Pairs with more interactions than the bitmaps hold read the two records
    // instead (tie_direct).
    bool const useBits = d.D <= 32 * kBitWords && !(tune & 8);
    if (useBits)
        for (int k = lane; k < 2 * kBitWords; k += 32) tieBits[k] = 0u;
    auto mark = [&](int x, int rx)
    {
        int const at = atomicAdd(&sm.misc[0], 2);
        list[at] = x;                                           // subgradient +1
        list[at + 1] = (int32_t)((uint32_t)rx | 0x80000000u);   // subgradient -1
    };
    auto tie_direct = [&](int x, int rx, int32_t s, bool inSol)
    {
        bool cycle = false;
        if (inSol)
        {
            int4 const r1 = cold[(size_t)x];           // { partner edge, its seqScore, owner, 0 }
            cycle = (r1.y + s + inter[(size_t)rx].w == edge[r1.x].y);
        }
        if (!cycle) mark(x, rx);
    };
    // the edge record of a line is loaded one group of 32 lines ahead of its use
    int4 erNext = make_int4(0, 0, 0, 0);
    if (lane < nWork) erNext = edge[sm.curE[lane]];
    for (int base = 0; base < nWork; base += 32)
    {
        int const e = base + lane < nWork ? sm.curE[base + lane] : -1;
        int const myLine = e >= 0 ? sm.workLine[base + lane] : 0;
        uint32_t const myCell = e >= 0 ? sm.linesRev[nLines - 1 - myLine] : 0u;
        int4 er = erNext;
        erNext = make_int4(0, 0, 0, 0);
        if (base + 32 + lane < nWork) erNext = edge[sm.curE[base + 32 + lane]];
        lbSeq += er.x;
        int off = er.w;
        #pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            int const y = __shfl_up_sync(kFull, off, o);
            if (lane >= o) off += y;
        }
        int const total = __shfl_sync(kFull, off, 31);
        off -= er.w;
        __syncwarp();
        sOff[lane] = (uint32_t)off;
        if (lane == 0) sOff[32] = (uint32_t)total;
        __syncwarp();
        for (int t0 = 0; t0 < total; t0 += 32 * kItems)
        {
            int4 r0[kItems];
            int xs[kItems], os[kItems];
            #pragma unroll
            for (int u = 0; u < kItems; ++u)
            {
                if (t0 + 32 * u >= total) break;                // uniform
                int const t = t0 + 32 * u + lane;
                int o = 0;                                      // owner: the last line that starts at or before item t
                #pragma unroll
                for (int step = 16; step > 0; step >>= 1)
                    if (sOff[o + step] <= (uint32_t)t) o += step;
                os[u] = o;
                xs[u] = __shfl_sync(kFull, er.z, o) + (t - __shfl_sync(kFull, off, o));
                r0[u] = make_int4(0, 0, 0, 0);
                if (t < total) r0[u] = inter[(size_t)xs[u]];    // { reverse interaction, partner cell, structScore, dual }
            }
            #pragma unroll
            for (int u = 0; u < kItems; ++u)
            {
                if (t0 + 32 * u >= total) break;                // uniform
                int const o = os[u];
                uint32_t const cellO = __shfl_sync(kFull, myCell, o);
                int const lineO = __shfl_sync(kFull, myLine, o);
                int32_t const sq = __shfl_sync(kFull, er.x, o), top = __shfl_sync(kFull, er.y, o);
                bool const valid = t0 + 32 * u + lane < total;
                int const x = xs[u], rx = r0[u].x;
                uint32_t const pp = (uint32_t)r0[u].y;
                int32_t const s = r0[u].z;
                int li = -1;
                bool inSol = false, tied = false;
                if (valid)
                {
                    li = sm.lineOfA[pp >> 16];
                    inSol = li >= 0 && (sm.linesRev[nLines - 1 - li] & 0xffffu) == (pp & 0xffffu);
                    tied = sq + s + r0[u].w == top;
                }
                if (useBits)
                {
                    // every tied bit of this round is set before any lane asks for one
                    if (tied && inSol) atomicOr(&tieBits[x >> 5], 1u << (x & 31));
                    __syncwarp();
                    if (inSol)
                    {
                        int const lo = min(x, rx);
                        uint32_t const old = atomicOr(&metBits[lo >> 5], 1u << (lo & 31));
                        if ((old >> (lo & 31)) & 1u)
                        {
                            bool const tiedR = (tieBits[rx >> 5] >> (rx & 31)) & 1u;
                            if (tied && !tiedR) mark(x, rx);
                            else if (tiedR && !tied) mark(rx, x);
                        }
                    }
                    else if (tied) mark(x, rx);
                }
                else if (tied) tie_direct(x, rx, s, inSol);
                // ids ascend in cell order: "line < partner" (:372) is a comparison of the two cells
                if (inSol && cellO < pp)
                {
                    int const at = atomicAdd(&sm.misc[1], 1);
                    unsigned long long const key = ((unsigned long long)(~((uint32_t)s ^ 0x80000000u)) << 32) |
                                                   ((unsigned long long)(uint32_t)lineO << 16) | (unsigned long long)li;
                    if (at < d.kCap) cand[at] = key; else sm.misc[11] = 2;
                }
            }
            __syncwarp();
        }
    }
    lbSeq = warp_sum(lbSeq);
    __syncwarp();
    int const nList = sm.misc[0];
    int const nCand = min(sm.misc[1], d.kCap);
    if (!candInSmem && nCand <= smemCand)
    {
        for (int i = lane; i < nCand; i += 32) sm.cand[i] = cand[i];
        cand = sm.cand;
        __syncwarp();
    }

    // ---- 4. structural part of the lower bound ---------------------------------------------------
    int32_t lbStruct = 0;
    if (nList > 0)
    {
        int Pn = 32;
        while (Pn < nCand) Pn <<= 1;
        for (int i = nCand + lane; i < Pn; i += 32) cand[i] = ~0ull;
        __syncwarp();
        if (nCand > 1) warp_bitonic_sort(cand, Pn, lane);
        lbStruct = greedy_matching(cand, nCand, P.lookahead, sm.matched, sm.misc, lane);
    }
    else
    {
        // every top is part of a cycle (:397-409): add the structure score of the first tied top
        int32_t add = 0;
        for (int k = lane; k < nWork; k += 32)
        {
            int const e = sm.curE[k];
            int const i = sm.workLine[k];
            uint32_t const myCell = sm.linesRev[nLines - 1 - i];
            int4 const er = edge[e];
            int32_t const sq = er.x, top = er.y;
            int found = -1;
            bool selfFirst = false;
            int4 rf = make_int4(0, 0, 0, 0);
            for (int x = er.z; x < er.z + er.w && found < 0 && !selfFirst; ++x)
            {
                int4 const r0 = inter[(size_t)x];
                if ((uint32_t)r0.y > myCell && sq == top) { selfFirst = true; break; }
                if (sq + r0.z + r0.w == top) { found = x; rf = r0; }
            }
            if (found >= 0 && !selfFirst)
            {
                add += rf.z;
                sm.matched[i] = 1;
                int const li = sm.lineOfA[(uint32_t)rf.y >> 16];
                if (li >= 0) sm.matched[li] = 1;
            }
        }
        lbStruct = warp_sum(add);
    }
    __syncwarp();

    // ---- 5. primal value, best structural alignment (:412-421) -----------------------------------
    int32_t const primal = lbSeq + lbStruct + gapScore;
    // one lane decides and broadcasts: lane 0 overwrites st.bestScore below, and a lane that runs late must
    // not compare against the new value
    bool const improved = __shfl_sync(kFull, (lane == 0 && primal > st.bestScore) ? 1 : 0, 0) != 0;
    if (improved)
    {
        for (int i = lane; i < nLines; i += 32)
            arena[d.offBest + i] = (int32_t)(sm.linesRev[nLines - 1 - i] | (sm.matched[i] ? 0x80000000u : 0u));
    }

    // ---- 6. bounds, step size, termination (solver :247-286 / simd :292-350) ----------------------
    float stepSize = 0.f;
    int finished = 0;
    if (lane == 0)
    {
        if (improved) { st.bestScore = primal; st.nBest = nLines; }
        int32_t bestUpper = st.bestUpper, bestLower = st.bestLower;
        uint32_t nd = st.nondecreasing;
        int32_t const cu = st.currentUpper;
        st.currentLower = primal;
        if (cu < bestUpper) { bestUpper = cu; nd = 0; }
        if (primal > bestLower) { bestLower = primal; nd = 0; }
        if (P.stepMode == LARA_STEP_SCALAR_FLOAT)
        {
            float f = st.stepFactorF;
            if (nd++ >= (uint32_t)P.maxNondecreasing) { f = __fdiv_rn(f, 2.0f); nd = 0; }
            st.stepFactorF = f;
            stepSize = __fdiv_rn(__fmul_rn(f, __int2float_rn(bestUpper - bestLower)), __int2float_rn(nList));
        }
        else
        {
            int32_t f = st.stepFactorI;
            nd += 1;
            if (nd > (uint32_t)P.maxNondecreasing) { f = f / 2; nd = 0; }
            st.stepFactorI = f;
            int32_t const prod = (int32_t)((uint32_t)f * (uint32_t)(bestUpper - bestLower));
            stepSize = __fdiv_rn(__fdiv_rn(__int2float_rn(prod), LARA_FACTOR2INT), __int2float_rn(nList));
        }
        st.bestUpper = bestUpper; st.bestLower = bestLower; st.nondecreasing = nd;
        st.nLines = nLines; st.gapScore = gapScore; st.nList = nList;
        st.remaining -= 1;
        finished = (bestUpper == bestLower || st.remaining == 0) ? 1 : 0;
        if (sm.misc[11]) st.error = sm.misc[11];
    }
    stepSize = __shfl_sync(kFull, stepSize, 0);
    finished = __shfl_sync(kFull, finished, 0);

    // ---- 7. multiplier update (:267-271) -----------------------------------------------------------
    // Bit 30 of a list entry records whether the multiplier went DOWN.  With a positive step that is the +1 entry of a
    // pair and never the -1 entry, but the step can be negative (the bounds may cross, and the SIMD build's int32
    // product wraps: src/subgradient_solver_simd.hpp:313) and the float round trip of a large multiplier can move it
    // either way, so the direction is taken from the values.
    for (int i = lane; i < nList; i += 32)
    {
        uint32_t const x = (uint32_t)list[i];
        size_t const di = (size_t)(x & 0x3fffffffu);
        int32_t * const dual = &reinterpret_cast<int32_t *>(inter + di)[3];
        int32_t const before = *dual;
        float const f = __int2float_rn(before);
        int32_t const after = __float2int_rz((x >> 31) ? __fadd_rn(f, stepSize) : __fsub_rn(f, stepSize));
        *dual = after;
        list[i] = (int32_t)((x & 0xbfffffffu) | (after < before ? 0x40000000u : 0u));
    }
    __syncwarp();

    if (!finished)
    {
        // ---- 8a. updateScores (:312-323): refresh the queue top of every touched edge -------------
        int32_t * __restrict__ vals = arena + (d.bandW > 0 ? d.offBVals : d.offVals);
        int32_t const * __restrict__ perm = arena + d.offPerm;
        // A multiplier that went down may have been the maximum of its row (the +1 entries were tied tops): the row's
        // maximum has to be found again.  A multiplier that went up can only raise its row's maximum: max(top, new
        // value), an atomic that commutes with the plain store of a full rescan of the same row.
        int const nPairs = nList >> 1;
        if (tune & 2)
        {
            // measurement switch: every entry rescans its row (the round-1d form)
            for (int i = lane; i < nList; i += 32)
            {
                size_t const di = (size_t)((uint32_t)list[i] & 0x3fffffffu);
                int const l = cold[di].z;
                int4 const er = edge[l];
                int32_t const sq = er.x;
                int32_t top = sq;
                for (int x = er.z; x < er.z + er.w; ++x)
                {
                    int4 const r0 = inter[(size_t)x];
                    top = max(top, sq + r0.z + r0.w);
                }
                reinterpret_cast<int32_t *>(edge + l)[1] = top;
                vals[perm[l]] = top;
            }
            return;
        }
        for (int k = lane; k < nPairs; k += 32)
        {
            #pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                uint32_t const le = (uint32_t)list[2 * k + h];
                if (!(le & 0x40000000u)) continue;
                size_t const di = (size_t)(le & 0x3fffffffu);
                int const l = cold[di].z;                   // owner of interaction di
                int4 const er = edge[l];
                int32_t const sq = er.x;
                int32_t top = sq;
                #pragma unroll 4
                for (int x = er.z; x < er.z + er.w; ++x)
                {
                    int4 const r0 = inter[(size_t)x];
                    top = max(top, sq + r0.z + r0.w);
                }
                reinterpret_cast<int32_t *>(edge + l)[1] = top;
                vals[perm[l]] = top;
            }
        }
        for (int k = lane; k < nPairs; k += 32)
        {
            #pragma unroll
            for (int h = 0; h < 2; ++h)
            {
                uint32_t const le = (uint32_t)list[2 * k + h];
                if (le & 0x40000000u) continue;
                size_t const di = (size_t)(le & 0x3fffffffu);
                int4 const r0 = inter[di];
                int const l = cold[di].z;
                int32_t const v = edge[l].x + r0.z + r0.w;
                atomicMax(&reinterpret_cast<int32_t *>(edge + l)[1], v);
                atomicMax(&vals[perm[l]], v);
            }
        }
        return;
    }

    // ---- 8b. finished: getStructureLines (:437-487) -> result slot ---------------------------------
    __syncwarp();
    int const nBest = __shfl_sync(kFull, lane == 0 ? st.nBest : 0, 0);
    int32_t minScore = 0, maxScore = 0;
    float div = 0.f;
    if (P.libLinear)
    {
        // minmax over ALL lenA*lenB queues (:445-448); an inactive edge's empty std::set reads as 0 there
        int32_t mn = INT32_MAX, mx = INT32_MIN;
        for (int e = lane; e < d.E; e += 32) { int32_t const t = edge[e].y; mn = min(mn, t); mx = max(mx, t); }
        if ((int64_t)d.E < (int64_t)d.la * d.lb) { mn = min(mn, 0); mx = max(mx, 0); }
        minScore = warp_min(mn);
        maxScore = warp_max(mx);
        div = __fdiv_rn(__fmul_rn(1.f, __uint2float_rn(P.libMax - P.libMin)), __int2float_rn(maxScore - minScore));
    }
    bool const swapIdx = d.seqFirst > d.seqSecond;
    // the pair's result slot: its position in the device's prepared pair list (the batch arrays are in admission order)
    int const oi = d.outIdx;
    int32_t * __restrict__ out = outLines + outOff[oi];
    for (int i = lane; i < nBest; i += 32)
    {
        uint32_t const v = (uint32_t)arena[d.offBest + i];
        int const a = (int)((v >> 16) & 0x7fffu), b = (int)(v & 0xffffu);
        uint32_t const isMatched = v >> 31;
        uint32_t w;
        if (P.libLinear)
        {
            int const e = edge_of(d, arena, a, b);
            unsigned long long const prod = (unsigned long long)isMatched * (unsigned long long)(long long)(edge[e].y - minScore);
            int32_t const val = __float2int_rz(__fmul_rn(__ull2float_rn(prod), div));
            w = P.libMin + (uint32_t)val;
        }
        else
            w = isMatched * 500u + 500u;
        out[i] = swapIdx ? b : a;
        out[d.lb + i] = swapIdx ? a : b;
        out[2 * d.lb + i] = (int32_t)w;
    }
    if (lane == 0)
    {
        PairOutHdr h;
        h.bestLower = st.bestLower; h.bestUpper = st.bestUpper;
        h.iterations = P.numIterations - st.remaining; h.nLines = nBest;
        outHdr[oi] = h;
        st.done = 1;
        // tell the host: finished[0] counts, finished[1 ..] lists the finished pairs (its slot can take the next pair)
        __threadfence();
        int const at = atomicAdd(finishedList, 1);
        finishedList[1 + at] = p;
    }
}

// Warp w of the grid takes slot w of the batch's slot table (pairIdx; nullptr: pair w); empty slots and pairs that
// have finished return at once.
#ifndef LARA_POST_MINBLOCKS
#define LARA_POST_MINBLOCKS 6       /* 6: 80 registers, 24 one-warp blocks per SM.  7 (72 registers, 28 blocks) and 8 (64, 32): measured */
#endif
template <bool kTracebackOnly>
__global__ void __launch_bounds__(128, LARA_POST_MINBLOCKS)
iterate_post_kernel(int n, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc, PairState * __restrict__ state,
                    int32_t * __restrict__ arena, DevParams P, PairOutHdr * __restrict__ outHdr,
                    int32_t * __restrict__ outLines, int64_t const * __restrict__ outOff,
                    int32_t * __restrict__ finished, int maxLa, int maxLb, int smemCand, int tune, DevInputs in)
{
    int const w = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    if (w >= n)
        return;
    int const p = pairIdx ? pairIdx[w] : w;
    if (p < 0)
        return;
    iterate_post_pair<kTracebackOnly>(p, desc, state, arena, P, outHdr, outLines, outOff, finished, maxLa, maxLb, smemCand, tune, in);
}

size_t post_smem_per_warp(int maxLa, int maxLb, int smemCand)
{
    return (size_t)smemCand * 8 + 128 + 1024 + 2 * (((size_t)maxLb * 4 + 15) / 16 * 16) + ((size_t)maxLa * 2 + 15) / 16 * 16 +
           ((size_t)maxLb * 2 + 15) / 16 * 16 + ((size_t)maxLb + 15) / 16 * 16;
}

} // namespace

size_t post_smem_bytes_per_warp(int maxLa, int maxLb, int smemCand) { return post_smem_per_warp(maxLa, maxLb, smemCand); }

// opt in to the full shared memory carve-out for both instantiations (per device, at context creation)
cudaError_t post_configure_device()
{
    int dev = 0, maxOptin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&maxOptin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(iterate_post_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxOptin);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(iterate_post_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, maxOptin);
}

cudaError_t launch_iterate_post(int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                                DevParams const & P, PairOutHdr * outHdr, int32_t * outLines, int64_t const * outOff,
                                int32_t * finished, int maxLa, int maxLb, int smemCand, DevInputs const & in,
                                cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    // one warp (= one pair) per block: pairs differ a lot in run time (finished pairs return at once, alignments
    // differ in length), and single-warp blocks let the SM refill per pair (measured: 4 % faster step than 4 warps)
    static int const warpsPerBlock = []() { char const * e = std::getenv("LARA_B200_POSTWARPS"); int v = e ? std::atoi(e) : 1;
                                            return v == 2 || v == 4 ? v : 1; }();
    // measurement switches: bit 0 = no L2 prefetch of the trace, bit 1 = updateScores rescans every touched row,
    // bit 2 = no prefetch of the row records, bit 3 = cycle test always from the two records (no tie bitmap)
    static int const tune = []() { char const * e = std::getenv("LARA_B200_POSTTUNE"); return e ? std::atoi(e) : 0; }();
    size_t const smem = post_smem_per_warp(maxLa, maxLb, smemCand) * warpsPerBlock;
    int const blocks = (n + warpsPerBlock - 1) / warpsPerBlock;
    iterate_post_kernel<false><<<blocks, warpsPerBlock * 32, smem, stream>>>(n, pairIdx, desc, state, arena, P, outHdr, outLines, outOff,
                                                                             finished, maxLa, maxLb, smemCand, tune, in);
    return cudaGetLastError();
}

// Result gather: the lines of pair q (slot of 3 * lb ints: posLow | posHigh | weight) -> three dense arrays at dstOff[q]
__global__ void pack_lines_kernel(int n, PairOutHdr const * __restrict__ hdr, int32_t const * __restrict__ outLines,
                                  int64_t const * __restrict__ outOff, int64_t const * __restrict__ dstOff,
                                  int32_t * __restrict__ low, int32_t * __restrict__ high, int32_t * __restrict__ weight)
{
    int const q = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    if (q >= n) return;
    int64_t const o = outOff[q];
    int const lb = (int)((outOff[q + 1] - o) / 3), nl = hdr[q].nLines;
    int32_t const * __restrict__ src = outLines + o;
    int64_t const d = dstOff[q];
    for (int i = lane; i < nl; i += 32)
    {
        low[d + i] = src[i];
        high[d + i] = src[lb + i];
        weight[d + i] = src[2 * lb + i];
    }
}

cudaError_t launch_pack_lines(int n, PairOutHdr const * hdr, int32_t const * outLines, int64_t const * outOff,
                              int64_t const * dstOff, int32_t * low, int32_t * high, int32_t * weight, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    pack_lines_kernel<<<(n + 7) / 8, 256, 0, stream>>>(n, hdr, outLines, outOff, dstOff, low, high, weight);
    return cudaGetLastError();
}

cudaError_t launch_traceback_only(int n, PairDesc const * desc, PairState * state, int32_t * arena,
                                  int32_t go, int32_t ge, int maxLa, int maxLb, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    int const warpsPerBlock = 4;
    size_t const smem = post_smem_per_warp(maxLa, maxLb, 0) * warpsPerBlock;
    DevParams P{};
    P.go = go; P.ge = ge;
    int const blocks = (n + warpsPerBlock - 1) / warpsPerBlock;
    iterate_post_kernel<true><<<blocks, warpsPerBlock * 32, smem, stream>>>(n, nullptr, desc, state, arena, P, nullptr, nullptr,
                                                                            nullptr, nullptr, maxLa, maxLb, 0, 0, DevInputs{});
    return cudaGetLastError();
}

} // namespace lara
