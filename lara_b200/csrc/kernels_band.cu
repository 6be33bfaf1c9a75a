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
// (L) into its cell and records TH/TS/M per row (TV/TS/M per column) so that a traceback that leaves the band can
// be continued exactly (kernels_post.cu).  Requires go < ge (opening dearer than extending).
//
// Mapping: anti-diagonal wavefront.  A pair uses L lanes (16 or 32; 32 / L pairs per warp); lane l owns the
// two diagonals dA = dLo + 2l and dB = dA + 1.  Step s handles the cells with c + r = s: even steps the A
// diagonals, odd steps the B diagonals (dLo is even), so every lane updates exactly one cell per step:
//     A cell: left neighbour = B cell of lane l-1 (shfl_up), upper neighbour = own B cell, diagonal = own A cell
//     B cell: left neighbour = own A cell, upper neighbour = A cell of lane l+1 (shfl_down), diagonal = own B
// Position specific scores come from a dense band stream vals[s * L + l] (coalesced), four decision bits per
// cell go to trace[(s >> 3) * L + l] (one word per lane and 8 steps, coalesced).
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

constexpr unsigned kFull = 0xffffffffu;

// One SeqAn _maxScore step (see kernels_dp.cu): max(a, b) with a winning ties; decision bit = "b won".
__device__ __forceinline__ int32_t max_push(int32_t a, int32_t b, uint32_t & tw)
{
    tw = __funnelshift_l((uint32_t)(a - b), tw, 1);
    return max(a, b);
}

template <int L>
__global__ void __launch_bounds__(128)
gotoh_band_kernel(int n, int32_t const * __restrict__ pairIdx, PairDesc const * __restrict__ desc,
                  PairState * __restrict__ state, int32_t * __restrict__ arena, int32_t go, int32_t ge)
{
    constexpr int PPW = 32 / L;                                 // pairs per warp
    int const warp = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    int const l = lane % L;                                     // lane inside the pair
    int const slot = warp * PPW + lane / L;
    if (warp * PPW >= n)
        return;
    bool live = slot < n;
    int p = 0;
    if (live)
    {
        p = pairIdx ? pairIdx[slot] : slot;
        live = state[p].done == 0;
    }
    PairDesc const & d = desc[p];
    // a dead sub-group runs along with la = lb = -2: no cell is ever valid, nothing is stored
    int const la = live ? d.la : -2, lb = live ? d.lb : -2;
    int const dLo = live ? d.dLo : 0;
    int const dHi = dLo + 2 * L - 1;
    int nSpad = live ? d.nSpad : 0;
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1) nSpad = max(nSpad, __shfl_xor_sync(kFull, nSpad, o));
    int32_t const * __restrict__ vals = arena + d.offVals + l;
    int32_t * __restrict__ trace = arena + d.offBTrace + l;
    int32_t * __restrict__ uTH = arena + d.offU, * uTS = uTH + (lb + 1), * uM = uTS + (lb + 1);
    int32_t * __restrict__ lTV = arena + d.offL, * lTS = lTV + (la + 1), * lM = lTS + (la + 1);

    int const dA = dLo + 2 * l;
    int32_t SA = kBandNeg, HA = kBandNeg, VA = kBandNeg;       // latest cell on diagonal dA
    int32_t SB = kBandNeg, HB = kBandNeg, VB = kBandNeg;       // latest cell on diagonal dB = dA + 1
    int32_t runU = kDpNeg, runL = kDpNeg;                       // running maxima M(r) (lane L-1) / M'(c) (lane 0)
    uint32_t tw = 0;
    int32_t subA = live ? __ldg(vals) : kDpNeg;                 // s = 0
    int32_t subB = live ? __ldg(vals + L) : kDpNeg;             // s = 1

    for (int s = 0; s < nSpad; s += 2)
    {
        // prefetch the scores of the next double step
        int32_t nextA = kDpNeg, nextB = kDpNeg;
        if (live && s + 2 < d.nSpad)
        {
            nextA = __ldg(vals + (size_t)(s + 2) * L);
            nextB = __ldg(vals + (size_t)(s + 3) * L);
        }
        // ------------------------------ A step: cells with c + r = s on the even diagonals -------------------
        {
            int const c = (s + dA) >> 1, r = (s - dA) >> 1;
            bool const valid = (c | r) >= 0 && c <= la && r <= lb;
            int32_t Hl = __shfl_up_sync(kFull, HB, 1, L);
            int32_t Sl = __shfl_up_sync(kFull, SB, 1, L);
            if (l == 0)
            {
                // left neighbour (c-1, r) lies in region L, directly below the boundary cell (c-1, r-1) = (SA, VA)
                Hl = kBandNeg; Sl = kBandNeg;
                if (c >= 1 && c - 1 <= la && r - 1 <= lb)
                {
                    int32_t const tv = VA + ge, ts = SA + go;
                    int32_t const Vl = max(tv, ts);                              // first L cell of column c-1
                    Hl = (c - 1 >= 1) ? (c + r - 2) * ge + go + runL : kBandNeg;
                    Sl = max(Vl, Hl);
                    lTV[c - 1] = tv - r * ge; lTS[c - 1] = ts - r * ge; lM[c - 1] = runL;
                    runL = max(runL, Vl - r * ge - (c - 1) * ge);
                }
            }
            int32_t const dd = SA + subA;
            int32_t const h = max_push(Hl + ge, Sl + go, tw);     // bit 3: H opened   (extension wins ties)
            int32_t const v = max_push(VB + ge, SB + go, tw);     // bit 2: V opened   (extension wins ties)
            int32_t const g = max_push(v, h, tw);                 // bit 1: gap from H (vertical wins ties)
            int32_t const sc = max_push(dd, g, tw);               // bit 0: S from gap (diagonal wins ties)
            if (valid) { SA = sc; HA = h; VA = v; }
            if (s == 0 && dA == 0 && live) { SA = 0; HA = kBandNeg; VA = kBandNeg; }     // origin
            if (valid && c == la && r == lb)
            {
                PairState & st = state[p];
                st.currentUpper = sc; st.sEnd = sc; st.vEnd = v; st.hEnd = h;
            }
        }
        // ------------------------------ B step: cells with c + r = s + 1 on the odd diagonals ---------------
        {
            int const c = (s + 1 + dA + 1) >> 1, r = (s + 1 - dA - 1) >> 1;
            bool const valid = (c | r) >= 0 && c <= la && r <= lb;
            int32_t Vu = __shfl_down_sync(kFull, VA, 1, L);
            int32_t Su = __shfl_down_sync(kFull, SA, 1, L);
            if (l == L - 1)
            {
                // upper neighbour (c, r-1) lies in region U, right of the boundary cell (c-1, r-1) = (SB, HB)
                Vu = kBandNeg; Su = kBandNeg;
                if (r >= 1 && c - 1 <= la && r - 1 <= lb)
                {
                    int32_t const th = HB + ge, ts = SB + go;
                    int32_t const Hu = max(th, ts);                              // first U cell of row r-1
                    Vu = (r - 1 >= 1) ? (c + r - 2) * ge + go + runU : kBandNeg;
                    Su = max(Vu, Hu);
                    uTH[r - 1] = th - c * ge; uTS[r - 1] = ts - c * ge; uM[r - 1] = runU;
                    runU = max(runU, Hu - c * ge - (r - 1) * ge);
                }
            }
            int32_t const dd = SB + subB;
            int32_t const h = max_push(HA + ge, SA + go, tw);
            int32_t const v = max_push(Vu + ge, Su + go, tw);
            int32_t const g = max_push(v, h, tw);
            int32_t const sc = max_push(dd, g, tw);
            if (valid) { SB = sc; HB = h; VB = v; }
            if (valid && c == la && r == lb)
            {
                PairState & st = state[p];
                st.currentUpper = sc; st.sEnd = sc; st.vEnd = v; st.hEnd = h;
            }
        }
        if ((s & 7) == 6 && live && s < d.nSpad)
            trace[(size_t)(s >> 3) * L] = (int32_t)tw;
        subA = nextA; subB = nextB;
    }
}

template <int L>
cudaError_t launch_l(int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                     int32_t go, int32_t ge, cudaStream_t stream)
{
    int const pairsPerBlock = 4 * (32 / L);
    int const blocks = (n + pairsPerBlock - 1) / pairsPerBlock;
    gotoh_band_kernel<L><<<blocks, 128, 0, stream>>>(n, pairIdx, desc, state, arena, go, ge);
    return cudaGetLastError();
}

} // namespace

cudaError_t launch_gotoh_band(int L, int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream)
{
    if (n <= 0)
        return cudaSuccess;
    if (L == 16) return launch_l<16>(n, pairIdx, desc, state, arena, go, ge, stream);
    if (L == 32) return launch_l<32>(n, pairIdx, desc, state, arena, go, ge, stream);
    return cudaErrorInvalidValue;
}

} // namespace lara
