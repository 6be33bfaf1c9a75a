// kernels_dp.cu -- the hot kernel: affine-gap global DP (Gotoh) over the full lenA x lenB matrix with
// SeqAn's exact tie-breaking, one warp per sequence pair.
//
// Replaces: seqan::globalAlignment(gapsH, gapsV, pssm) as instantiated by LaRA
//   (src/subgradient_solver.hpp:223-225 -> SeqAn:align/dp_algorithm_impl.h:1349-1425,
//    recursion SeqAn:align/dp_formula_affine.h:62-125, tie rule SeqAn:align/dp_formula.h:153-163).
//
// Mapping (DESIGN.md "DP kernel"): lane k owns the R = ceil(lenB/32) consecutive rows
// [kR, kR+R) and sweeps the columns; lane k works on column c = t - k + 1 in step t, so the value a
// lane needs from the row above its strip was produced by lane k-1 one step earlier and arrives with
// one __shfl_up per step.  S and H of the previous column stay in registers (2R per lane).
// The position specific scores come as a per-lane "strip-dense" stream: a step whose strip holds no
// active alignment edge (about 80 % of them) is flagged 0 and needs no load at all.
// Four decision bits per cell go to a per-lane trace stream, written with 128-bit stores.
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

// One SeqAn _maxScore step: returns max(a, b) where a wins ties, and pushes the decision into the
// trace accumulator: bit 0 of tw becomes 1 iff b won (a < b).  The sign bit of a - b is shifted in with a
// single funnel shift (SHF.L.W), so a decision costs IADD + VIMNMX + SHF and no predicate traffic.
// |a - b| < 2^31 always holds: scores are bounded by ~2^27 and the two "minus infinities" are -2^30.
__device__ __forceinline__ int32_t max_push(int32_t a, int32_t b, uint32_t & tw)
{
    tw = __funnelshift_l((uint32_t)(a - b), tw, 1);
    return max(a, b);
}

template <int R>
struct TraceGeom
{
    static constexpr int W = (R <= 8) ? 1 : (R <= 16 ? 2 : 4);    // 32-bit trace words per step
    static constexpr int G = 4 / W;                                // steps per 128-bit store
};

template <int R>
__global__ void __launch_bounds__(128)
gotoh_fill_kernel(int n, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
                  PairState * __restrict__ state, int32_t * __restrict__ arena,
                  int32_t const * __restrict__ negBuf, int32_t go, int32_t ge)
{
    constexpr int W = TraceGeom<R>::W;
    constexpr int G = TraceGeom<R>::G;
    int const warp = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    if (warp >= n)
        return;
    int const k = threadIdx.x & 31;
    int const p = pairIdx ? pairIdx[warp] : warp;
    if (state[p].done)
        return;
    PairDesc const & d = desc[p];
    int const la = d.la, lb = d.lb;
    int const nSteps = d.nSteps;
    int const nFlagWords = (nSteps + 31) >> 5;
    uint32_t const * __restrict__ flags = (uint32_t const *)(arena + d.offFlags) + (size_t)k * nFlagWords;
    int32_t const * __restrict__ vp = arena + d.offVals + arena[d.offLanePtr + k];
    uint4 * __restrict__ trace = (uint4 *)(arena + d.offTrace) + (size_t)k * (nSteps / G);

    // ---- column 0 (RecursionDirectionVertical, dp_formula_affine.h:291-325): values only ----
    int32_t S[R], H[R];
    int32_t topDiag;                    // S(c-1, kR): row above the strip, previous column
    if (go <= ge)
    {
        // V(0,r) = go + (r-1) ge: the extension always wins (or ties) from row 2 on
        #pragma unroll
        for (int j = 0; j < R; ++j)
            S[j] = go + (k * R + j) * ge;
        topDiag = (k == 0) ? 0 : go + (k * R - 1) * ge;
    }
    else
    {
        int32_t v = kDpNeg, s = 0;
        topDiag = 0;
        for (int r = 1; r <= (k + 1) * R; ++r)
        {
            int32_t ve = v + ge, vo = s + go;
            v = ve >= vo ? ve : vo;
            s = v;
            if (r == k * R) topDiag = s;
            #pragma unroll
            for (int j = 0; j < R; ++j)
                if (r == k * R + j + 1) S[j] = s;
        }
    }
    #pragma unroll
    for (int j = 0; j < R; ++j)
        H[j] = kDpNeg;

    int32_t s0 = 0, h0 = kDpNeg;        // row 0 (lane 0 only): S(c,0) = H(c,0), dp_formula_affine.h:245-281
    int32_t botS = 0, botV = 0;         // own bottom row, last computed column
    int32_t upS = 0, upV = 0;           // row above the strip, current column (kept for the end cell)
    uint32_t fw = 0;

    for (int t0 = 0; t0 < nSteps; t0 += G)
    {
        if ((t0 & 31) == 0)
            fw = flags[t0 >> 5];
        uint32_t tb[4] = {0u, 0u, 0u, 0u};
        #pragma unroll
        for (int g = 0; g < G; ++g)
        {
            int const t = t0 + g;
            int const c = t - k + 1;                           // 1-based column of this lane
            int32_t recvS = __shfl_up_sync(0xffffffffu, botS, 1);
            int32_t recvV = __shfl_up_sync(0xffffffffu, botV, 1);
            if (c >= 1 && c <= la)
            {
                if (k == 0)
                {
                    int32_t he = h0 + ge, ho = s0 + go;
                    h0 = he >= ho ? he : ho;
                    s0 = h0;
                    recvS = s0;
                    recvV = kDpNeg;
                }
                // flagged step: the next R stream entries; otherwise R "minus infinity" entries from a
                // constant buffer (same load instructions, no select per cell)
                bool const flagged = (fw >> (t & 31)) & 1u;
                int32_t const * __restrict__ src = flagged ? vp : negBuf;
                int32_t sub[R];
                #pragma unroll
                for (int j = 0; j < R; ++j)
                    sub[j] = __ldg(src + j);
                vp += flagged ? R : 0;
                int32_t diag = topDiag;
                topDiag = recvS;
                upS = recvS;
                upV = recvV;
                int32_t us = recvS, uv = recvV;
                #pragma unroll
                for (int j = 0; j < R; ++j)
                {
                    uint32_t & tw = tb[g * W + (j >> 3)];
                    int32_t dd = diag + sub[j];
                    int32_t h = max_push(H[j] + ge, S[j] + go, tw);   // bit 3: H opened   (extension wins ties)
                    int32_t v = max_push(uv + ge, us + go, tw);       // bit 2: V opened   (extension wins ties)
                    int32_t gg = max_push(v, h, tw);                  // bit 1: gap from H (vertical wins ties)
                    int32_t s = max_push(dd, gg, tw);                 // bit 0: S from gap (diagonal wins ties)
                    diag = S[j];
                    S[j] = s;
                    H[j] = h;
                    us = s;
                    uv = v;
                }
                botS = us;
                botV = uv;
            }
        }
        trace[t0 / G] = make_uint4(tb[0], tb[1], tb[2], tb[3]);
    }

    // ---- last cell (la, lb): score and the values SeqAn's _correctTraceValue compares ----
    int const kEnd = (lb - 1) / R, jEnd = (lb - 1) - kEnd * R;
    if (k == kEnd)
    {
        // re-run the vertical chain of the final column down to row jEnd (S[] holds column la)
        int32_t v = upV, us = upS, sEnd = 0, hEnd = 0, vEnd = 0;
        #pragma unroll
        for (int j = 0; j < R; ++j)
        {
            int32_t ve = v + ge, vo = us + go;
            v = ve >= vo ? ve : vo;
            if (j == jEnd) { vEnd = v; sEnd = S[j]; hEnd = H[j]; }
            us = S[j];
        }
        PairState & st = state[p];
        st.currentUpper = sEnd;
        st.sEnd = sEnd;
        st.vEnd = vEnd;
        st.hEnd = hEnd;
    }
}

template <int R>
cudaError_t launch_r(int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                     int32_t const * negBuf, int32_t go, int32_t ge, cudaStream_t stream)
{
    int const warpsPerBlock = 4;
    int const blocks = (n + warpsPerBlock - 1) / warpsPerBlock;
    gotoh_fill_kernel<R><<<blocks, warpsPerBlock * 32, 0, stream>>>(n, pairIdx, desc, state, arena, negBuf, go, ge);
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_gotoh_fill(int R, int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t const * negBuf, int32_t go, int32_t ge, cudaStream_t stream)
{
    if (n <= 0)
        return cudaSuccess;
    switch (R)
    {
#define LARA_R_CASE(r) case r: return launch_r<r>(n, pairIdx, desc, state, arena, negBuf, go, ge, stream);
        LARA_R_CASE(1) LARA_R_CASE(2) LARA_R_CASE(3) LARA_R_CASE(4) LARA_R_CASE(5)
        LARA_R_CASE(6) LARA_R_CASE(7) LARA_R_CASE(8) LARA_R_CASE(9) LARA_R_CASE(10)
        LARA_R_CASE(11) LARA_R_CASE(12) LARA_R_CASE(13) LARA_R_CASE(14) LARA_R_CASE(15)
        LARA_R_CASE(16) LARA_R_CASE(17) LARA_R_CASE(18) LARA_R_CASE(19) LARA_R_CASE(20)
#undef LARA_R_CASE
        default: return cudaErrorInvalidValue;
    }
}

} // namespace lara
