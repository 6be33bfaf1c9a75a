// kernels_rows.cu -- affine-gap global DP (Gotoh) with SeqAn's exact tie-breaking over the FULL matrix of a pair of any
// size, row by row, one thread block per pair: the fill for the pairs no other kernel takes (shorter sequence beyond the
// 640 nt of the single-pass full-matrix kernel AND active edges on more diagonals than the widest band).
//
// Replaces the same call as kernels_band.cu / kernels_dp.cu (seqan::globalAlignment as LaRA instantiates it,
// src/subgradient_solver.hpp:223-225 -> SeqAn:align/dp_algorithm_impl.h:1349-1425, dp_formula_affine.h:62-125, tie rule
// SeqAn:align/dp_formula.h:153-163).
//
// A row depends on the row above through the diagonal and vertical candidates only; along the row only H is carried:
//     H(c) = max(H(c-1) + ge, m(c-1)),   m(k) = max(D(k), V(k)) + go   (go < ge, see kernels_band.cu)
//          = (c - 1) ge + max_{k < c} (m(k) - k ge)
// an exclusive prefix maximum over the row, which the block computes with one scan per row (each thread owns a run of
// neighbouring columns).  "H opened" (extension wins ties) is P(c-1) < m(c-1) - (c-1) ge with P the same prefix maximum.
// S and V of the previous row live in two global row buffers (L2 resident), the position specific scores in a dense
// (lb + 1) x (la + 1) array, one decision nibble per cell goes to a byte matrix of the same shape.
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

constexpr int kRowsThreads = 256;
constexpr long long kNoneG = -(1ll << 60);          // prefix maximum over no column

__device__ __forceinline__ long long warp_incl_max(long long v, int lane)
{
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        long long const y = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v = max(v, y);
    }
    return v;
}

// kMode 0: affine gap costs with go <= ge (the recursion above).
// kMode 1: linear gap costs: the scalar reference hands gap open == gap extend to SeqAn's Needleman-Wunsch
//          (SeqAn:align/global_alignment_unbanded.h:775-778, recursion dp_formula_linear.h:62-99): one matrix,
//          S(c) = max(D(c), V(c), S(c-1) + g) with V(c) = S(c, r-1) + g; vertical wins the tie against horizontal, the
//          diagonal the tie against the better gap; no end-cell correction; the traceback moves one cell at a time.
// kMode 2: affine gap costs with go > ge (opening cheaper than extending): S >= V, H makes every gap cell an opening,
//          V(c) = S(c, r-1) + go, H(c) = S(c-1) + go: the values and tie rules of mode 1 with g = go, every "opened" bit
//          set, plus the affine end-cell correction (S, V, H of the end cell are reported).
// In modes 1 and 2 the row is S(c) = c g + max_{k <= c} (max(D(k), V(k)) - k g): the same scan.
template <int kMode>
__global__ void __launch_bounds__(kRowsThreads)
gotoh_rows_kernel(int n, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
                  PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    int32_t const gl = go;                          // modes 1, 2: cost of every gap position
    __shared__ long long sIncl[kRowsThreads];       // inclusive prefix maximum of g over the threads' runs
    __shared__ long long sExLast[kRowsThreads];     // maximum of g over a thread's run without its last column
    __shared__ long long sLastG[kRowsThreads];      // g of a thread's last column
    __shared__ long long sWarp[kRowsThreads / 32];
    if ((int)blockIdx.x >= n) return;
    int const p = pairIdx[blockIdx.x];
    if (p < 0 || state[p].done) return;             // uniform over the block
    PairDesc const & d = desc[p];
    int const la = d.la, lb = d.lb, ncol = la + 1;
    int32_t const * __restrict__ vals = arena + d.offVals;                      // [(lb + 1)][ncol]
    uint8_t * __restrict__ trace = reinterpret_cast<uint8_t *>(arena + d.offTrace);   // [(lb + 1)][ncol]
    int32_t * const buf = arena + d.offFlags;                                   // 4 rows of ncol: S, V of rows r - 1 and r
    int const tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int const K = (ncol + kRowsThreads - 1) / kRowsThreads;
    int const c0 = min(tid * K, ncol), c1 = min(c0 + K, ncol);

    // row 0 (dp_formula_affine.h:245-281): S(c,0) = go + (c-1) ge, V = -inf; origin S = 0
    for (int c = c0; c < c1; ++c)
    {
        buf[c] = c == 0 ? 0 : (kMode == 0 ? go + (c - 1) * ge : c * gl);
        buf[2 * ncol + c] = kDpNeg;
    }
    __syncthreads();
    for (int r = 1; r <= lb; ++r)
    {
        int32_t const * __restrict__ Sp = buf + ((r - 1) & 1) * ncol;
        int32_t const * __restrict__ Vp = buf + 2 * ncol + ((r - 1) & 1) * ncol;
        int32_t * __restrict__ Sc = buf + (r & 1) * ncol;
        int32_t * __restrict__ Vc = buf + 2 * ncol + (r & 1) * ncol;
        int32_t const * __restrict__ sub = vals + (size_t)r * ncol;
        int32_t const s0 = kMode == 0 ? go + (r - 1) * ge : r * gl;     // column 0 (dp_formula_affine.h:291-325): S = V, H = -inf
        // ---- pass 1: g(k) = m(k) - k ge of the thread's columns -> maximum, maximum without the last column, last
        long long runMax = kNoneG, exLast = kNoneG, lastG = kNoneG;
        for (int c = c0; c < c1; ++c)
        {
            long long g;
            if (kMode == 0)
            {
                int32_t m;
                if (c == 0) m = s0 + go;
                else
                {
                    int32_t const dd = Sp[c - 1] + sub[c];
                    int32_t const v = max(Vp[c] + ge, Sp[c] + go);
                    m = max(dd, v) + go;
                }
                g = (long long)m - (long long)c * ge;
            }
            else
            {
                int32_t x;
                if (c == 0) x = s0;
                else x = max(Sp[c - 1] + sub[c], Sp[c] + gl);
                g = (long long)x - (long long)c * gl;
            }
            exLast = runMax;
            runMax = max(runMax, g);
            lastG = g;
        }
        // ---- scan over the threads ------------------------------------------------------------------------------
        long long incl = warp_incl_max(runMax, lane);
        if (lane == 31) sWarp[wid] = incl;
        sExLast[tid] = exLast;
        sLastG[tid] = lastG;
        __syncthreads();
        {
            long long before = kNoneG;
            for (int w = 0; w < wid; ++w) before = max(before, sWarp[w]);
            incl = max(incl, before);
        }
        sIncl[tid] = incl;
        __syncthreads();
        // ---- pass 2: the cells ----------------------------------------------------------------------------------
        // P(c) = max_{k < c} g(k); at the thread's first column: the threads before it
        long long P = tid > 0 ? sIncl[tid - 1] : kNoneG;                        // P(c0)
        long long Pm1 = kNoneG, gPrev = kNoneG;                                  // P(c0 - 1), g(c0 - 1)
        if (tid > 0 && c0 < ncol)
        {
            // the column before c0 is the last one of thread tid - 1 (every thread before the end owns K >= 1 columns)
            gPrev = sLastG[tid - 1];
            Pm1 = max(tid > 1 ? sIncl[tid - 2] : kNoneG, sExLast[tid - 1]);
        }
        for (int c = c0; c < c1; ++c)
        {
            if (c == 0)
            {
                Sc[0] = s0; Vc[0] = s0;
                trace[(size_t)r * ncol] = 0;
                Pm1 = kNoneG; gPrev = kMode == 0 ? (long long)(s0 + go) : (long long)s0; P = gPrev;     // P(1) = g(0)
                continue;
            }
            int32_t const dd = Sp[c - 1] + sub[c];
            int32_t v, h;
            uint32_t nib = 0u;
            if (kMode == 0)
            {
                int32_t const ve = Vp[c] + ge, vo = Sp[c] + go;
                v = max(ve, vo);
                h = (int32_t)((long long)(c - 1) * ge + P);
                if (ve < vo) nib |= kTrVOpen;           // V opened   (extension wins ties)
                if (Pm1 < gPrev) nib |= kTrHOpen;       // H opened   (extension wins ties): H(c-1) + ge < m(c-1)
            }
            else
            {
                v = Sp[c] + gl;
                h = (int32_t)((long long)c * gl + P);   // S(c-1) + g
                nib = kTrVOpen | kTrHOpen;              // every gap cell stands alone: the walk takes one step at a time
            }
            int32_t const mx = max(dd, v);
            int32_t const s = max(mx, h);
            if (dd < s) nib |= kTrGap;                  // S from gap (diagonal wins ties)
            if (v < h) nib |= kTrFromH;                 // gap from H (vertical wins ties)
            Sc[c] = s; Vc[c] = v;
            trace[(size_t)r * ncol + c] = (uint8_t)nib;
            if (r == lb && c == la)
            {
                PairState & st = state[p];
                st.currentUpper = s; st.sEnd = s;
                // linear gap costs: SeqAn corrects nothing at the end cell (values that never equal S)
                st.vEnd = kMode == 1 ? s - 1 : v; st.hEnd = kMode == 1 ? s - 1 : h;
            }
            long long const g = kMode == 0 ? (long long)(mx + go) - (long long)c * ge : (long long)mx - (long long)c * gl;
            Pm1 = P; gPrev = g; P = max(P, g);
        }
        __syncthreads();
    }
}

} // namespace

cudaError_t launch_gotoh_rows(int mode, int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                              int32_t go, int32_t ge, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    if (mode == 0) gotoh_rows_kernel<0><<<n, kRowsThreads, 0, stream>>>(n, pairIdx, desc, state, arena, go, ge);
    else if (mode == 1) gotoh_rows_kernel<1><<<n, kRowsThreads, 0, stream>>>(n, pairIdx, desc, state, arena, go, ge);
    else gotoh_rows_kernel<2><<<n, kRowsThreads, 0, stream>>>(n, pairIdx, desc, state, arena, go, ge);
    return cudaGetLastError();
}

} // namespace lara
