// kernels_setup.cu -- per-pair construction of the Lagrange problem on the device (once per pair):
//
//   stage 1  edge_filter_kernel   generateEdges / PairwiseGotoh            src/edge_filter.hpp:52-158
//   stage 2  setup_count_kernel   edge ids, getSeqScore, interaction count  src/lagrange.hpp:247-300 (counting pass)
//   stage 3  setup_fill_kernel    interactions, reverse index, initial pssm src/lagrange.hpp:265-309
//
// One warp per pair.  Between the stages the host reads back E resp. (D, nVals) per pair and lays out
// the next group of arrays in the batch arena.
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

constexpr unsigned kFull = 0xffffffffu;
constexpr int32_t kEfNeg = -LARA_INFINITY;  // "-infinity" of src/edge_filter.hpp:81-96

__device__ __forceinline__ int32_t max3(int32_t a, int32_t b, int32_t c) { return max(a, max(b, c)); }

// One forward or backward PairwiseGotoh sweep.  Matrix cell (c, r), 1 <= c <= la, 1 <= r <= lb:
//   M(c,r) = P(c-1,r-1) + tab[..]                      P = max(M, X, Y)   (src/edge_filter.hpp:104-106,119-123)
//   X(c,r) = max(M+go, X+ge, Y+go)(c, r-1)             ("matrixH", :108-110)
//   Y(c,r) = max(M+go, X+go, Y+ge)(c-1, r)             ("matrixV", :112-114)
// Lane k owns rows pass*32*RF + RF*k + 1 .. + RF, sweeps the columns skewed by k; rowBuf carries the last row of
// a 32*RF-row block to the next pass (RF is chosen per launch so that the usual case is a single pass).  The 5 x 5
// score table sits in shared memory: the lanes of a step are in different columns, i.e. read different entries.  kBackward: sequences reversed, and instead of storing the prefix
// score the activity test of generateEdges (:146-154) is evaluated on the fly.
template <bool kBackward, int RF>
__device__ int32_t filter_sweep(int la, int lb, uint8_t const * __restrict__ A, uint8_t const * __restrict__ B,
                                int32_t go, int32_t ge, int32_t const * __restrict__ sTab, int32_t * __restrict__ fwdP,
                                int32_t * __restrict__ rowBuf, uint32_t * __restrict__ bits, int bitWords, int32_t threshold,
                                int lane)
{
    int const nPass = (lb + 32 * RF - 1) / (32 * RF);
    int32_t optOut = 0;
    for (int pass = 0; pass < nPass; ++pass)
    {
        int const rBase = pass * 32 * RF + lane * RF;      // rows rBase+1 .. rBase+RF
        int32_t Ml[RF], Xl[RF], Yl[RF], Pl[RF];
        int cb[RF];
        #pragma unroll
        for (int j = 0; j < RF; ++j)
        {
            int const r = rBase + j + 1;                    // column 0 (:92-97): M = X = go + ge (r-1), Y = -inf
            int32_t const v = (int32_t)((uint32_t)go + (uint32_t)ge * (uint32_t)(r - 1));
            Ml[j] = v; Xl[j] = v; Yl[j] = kEfNeg; Pl[j] = max3(v, v, kEfNeg);
            int const bi = kBackward ? lb - r : r - 1;
            cb[j] = (r <= lb) ? (int)B[bi] : 4;
        }
        // P of the cell above the strip in column 0
        int32_t diagTop;
        if (rBase == 0) diagTop = 0;                        // (0,0): M = 0
        else diagTop = (int32_t)((uint32_t)go + (uint32_t)ge * (uint32_t)(rBase - 1));
        int32_t botM = 0, botX = 0, botY = 0;
        int32_t * __restrict__ rbIn = rowBuf + (size_t)(pass & 1) * 3 * (la + 1);
        int32_t * __restrict__ rbOut = rowBuf + (size_t)((pass + 1) & 1) * 3 * (la + 1);
        int const nSteps = la + 31;
        for (int t = 0; t < nSteps; ++t)
        {
            int const c = t - lane + 1;
            int32_t uM = __shfl_up_sync(kFull, botM, 1);
            int32_t uX = __shfl_up_sync(kFull, botX, 1);
            int32_t uY = __shfl_up_sync(kFull, botY, 1);
            if (c >= 1 && c <= la)
            {
                if (lane == 0)
                {
                    if (pass == 0)
                    {   // row 0 (:85-90): M = Y = go + ge (c-1), X = -inf
                        int32_t const v = (int32_t)((uint32_t)go + (uint32_t)ge * (uint32_t)(c - 1));
                        uM = v; uX = kEfNeg; uY = v;
                    }
                    else { uM = rbIn[c]; uX = rbIn[(la + 1) + c]; uY = rbIn[2 * (la + 1) + c]; }
                }
                int const ai = kBackward ? la - c : c - 1;
                int const ca = (int)A[ai];
                // backward sweep: the forward values of the step's cells are fetched together, before the cell loop (between
                // the loop's atomics the compiler keeps the loads in program order: one exposed latency per CELL), and the
                // step's active bits (at most two words of bitmap row a) go out as at most two atomics after it
                int32_t fw[kBackward ? RF : 1];
                uint32_t actHi = 0u, actLo = 0u;
                int const bTop = lb - (rBase + 1);                  // b of the strip's first row; the words bTop >> 5 and the one below
                if (kBackward)
                {
                    #pragma unroll
                    for (int j = 0; j < RF; ++j)
                    {
                        int const b = bTop - j;
                        fw[j] = (b >= 0) ? fwdP[(size_t)(la - c) * lb + b] : 0;
                    }
                }
                int32_t diag = diagTop;
                diagTop = max3(uM, uX, uY);
                #pragma unroll
                for (int j = 0; j < RF; ++j)
                {
                    int const r = rBase + j + 1;
                    int32_t const sub = sTab[ca * 5 + cb[j]];
                    int32_t const m = diag + sub;
                    int32_t const x = max3(uM + go, uX + ge, uY + go);
                    int32_t const y = max3(Ml[j] + go, Xl[j] + go, Yl[j] + ge);
                    if (r <= lb)
                    {
                        if (!kBackward)
                            fwdP[(size_t)(c - 1) * lb + (r - 1)] = diag;
                        else
                        {
                            int const b = lb - r;
                            if (fw[j] + sub + diag >= threshold)
                            {
                                if ((b >> 5) == (bTop >> 5)) actHi |= 1u << (b & 31); else actLo |= 1u << (b & 31);
                            }
                        }
                    }
                    diag = Pl[j];
                    Ml[j] = m; Xl[j] = x; Yl[j] = y; Pl[j] = max3(m, x, y);
                    uM = m; uX = x; uY = y;
                }
                if (kBackward)
                {
                    uint32_t * const row = bits + (size_t)(la - c) * bitWords;
                    if (actHi) atomicOr(row + (bTop >> 5), actHi);
                    if (actLo) atomicOr(row + (bTop >> 5) - 1, actLo);
                }
                botM = uM; botX = uX; botY = uY;
                if (lane == 31 && pass + 1 < nPass)
                { rbOut[c] = botM; rbOut[(la + 1) + c] = botX; rbOut[2 * (la + 1) + c] = botY; }
            }
        }
        // optimum = P(la, lb): held by the lane that owns row lb after its last column
        int const rEnd = lb - 1 - pass * 32 * RF;          // 0-based row inside this pass
        if (rEnd >= 0 && rEnd < 32 * RF)
        {
            int const owner = rEnd / RF, jj = rEnd % RF;
            int32_t mine = 0;
            #pragma unroll
            for (int j = 0; j < RF; ++j) if (j == jj) mine = Pl[j];
            optOut = __shfl_sync(kFull, mine, owner);
        }
        __syncwarp();
    }
    return optOut;
}

template <int RF>
__global__ void __launch_bounds__(128)
edge_filter_kernel(int n, PairDesc * __restrict__ desc, PairState * __restrict__ state, int32_t * __restrict__ arena,
                   DevInputs in, DevParams P)
{
    __shared__ int32_t sTab[25];
    if (threadIdx.x < 25) sTab[threadIdx.x] = P.tab[threadIdx.x];
    __syncthreads();
    int const p = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    if (p >= n) return;
    PairDesc & d = desc[p];
    int const la = d.la, lb = d.lb;
    uint8_t const * A = in.seqCat + in.seqOff[d.seqFirst];
    uint8_t const * B = in.seqCat + in.seqOff[d.seqSecond];
    int32_t * fwdP = arena + d.offScratch;
    int32_t * rowBuf = fwdP + (size_t)la * lb;
    uint32_t * bits = (uint32_t *)(arena + d.offBits);
    int const nWords = la * d.bitWords;
    for (int i = lane; i < nWords; i += 32) bits[i] = 0u;
    __syncwarp();
    int32_t const opt = filter_sweep<false, RF>(la, lb, A, B, P.go, P.ge, sTab, fwdP, rowBuf, bits, d.bitWords, 0, lane);
    __syncwarp();
    int32_t const threshold = opt - P.subopt;                  // src/edge_filter.hpp:142
    filter_sweep<true, RF>(la, lb, A, B, P.go, P.ge, sTab, fwdP, rowBuf, bits, d.bitWords, threshold, lane);
    __syncwarp();
    int cnt = 0;
    for (int i = lane; i < nWords; i += 32) cnt += __popc(bits[i]);
    // diagonal range a - b of the active edges (band geometry of kernels_band.cu)
    int dmin = INT32_MAX, dmax = INT32_MIN;
    for (int a = lane; a < la; a += 32)
    {
        int bFirst = -1, bLast = -1;
        for (int w = 0; w < d.bitWords; ++w)
        {
            uint32_t const m = bits[(size_t)a * d.bitWords + w];
            if (m)
            {
                if (bFirst < 0) bFirst = (w << 5) + __ffs(m) - 1;
                bLast = (w << 5) + 31 - __clz(m);
            }
        }
        if (bFirst >= 0) { dmin = min(dmin, a - bLast); dmax = max(dmax, a - bFirst); }
    }
    #pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        cnt += __shfl_xor_sync(kFull, cnt, o);
        dmin = min(dmin, __shfl_xor_sync(kFull, dmin, o));
        dmax = max(dmax, __shfl_xor_sync(kFull, dmax, o));
    }
    if (lane == 0)
    {
        d.E = cnt;
        d.dMin = dmin; d.dMax = dmax;
        PairState & st = state[p];
        st.optScore = opt;
        // avgId (src/edge_filter.hpp:157) and sequenceScaleFactor (src/lagrange.hpp:252), no FMA contraction
        float const avg = __fdiv_rn(__fdiv_rn(__int2float_rn(opt), LARA_FACTOR2INT), __int2float_rn(max(la, lb)));
        st.scale = __fadd_rn(__fmul_rn(P.sequenceScale, avg), 1.0f);
    }
}

__device__ __forceinline__ bool bit_at(uint32_t const * __restrict__ bits, int bitWords, int a, int b)
{
    return (bits[(size_t)a * bitWords + (b >> 5)] >> (b & 31)) & 1u;
}

__device__ __forceinline__ int edge_id(uint32_t const * __restrict__ bits, int bitWords, int32_t const * __restrict__ rowStart,
                                       int a, int b)
{
    uint32_t const * row = bits + (size_t)a * bitWords;
    int e = rowStart[a];
    int const wb = b >> 5;
    for (int w = 0; w < wb; ++w) e += __popc(row[w]);
    return e + __popc(row[wb] & ((1u << (b & 31)) - 1u));
}

// bits [lo, lo+len) of bitmap row (len <= 32)
__device__ __forceinline__ uint32_t strip_bits(uint32_t const * __restrict__ row, int bitWords, int lo, int len)
{
    int const w = lo >> 5, s = lo & 31;
    if (w >= bitWords) return 0u;
    uint32_t const w0 = row[w];
    uint32_t const w1 = (w + 1 < bitWords) ? row[w + 1] : 0u;
    uint32_t const v = __funnelshift_r(w0, w1, s);
    return len >= 32 ? v : (v & ((1u << len) - 1u));
}

__device__ __forceinline__ int warp_excl_scan(int v, int lane, int & total)
{
    int x = v;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        int const y = __shfl_up_sync(kFull, x, o);
        if (lane >= o) x += y;
    }
    total = __shfl_sync(kFull, x, 31);
    return x - v;
}

template <typename F>
__device__ __forceinline__ void for_each_interaction(PairDesc const & d, DevInputs const & in, uint32_t const * __restrict__ bits,
                                                     int a, int b, F && f)
{
    int64_t const gA = in.seqOff[d.seqFirst] + a, gB = in.seqOff[d.seqSecond] + b;
    int64_t const a0 = in.bpOff[gA], a1 = in.bpOff[gA + 1], b0 = in.bpOff[gB], b1 = in.bpOff[gB + 1];
    for (int64_t x = a0; x < a1; ++x)
    {
        int const ap = in.bpPartner[x];
        int const wa = in.bpWeight[x];
        for (int64_t y = b0; y < b1; ++y)
        {
            int const bp = in.bpPartner[y];
            bool const nonCrossing = (a < ap && b < bp) || (ap < a && bp < b);      // src/lagrange.hpp:101-106
            if (nonCrossing && bit_at(bits, d.bitWords, ap, bp))
                f(ap, bp, wa + in.bpWeight[y]);
        }
    }
}

__global__ void __launch_bounds__(128)
setup_count_kernel(int n, PairDesc * __restrict__ desc, PairState * __restrict__ state, int32_t * __restrict__ arena,
                   DevInputs in, DevParams P)
{
    int const p = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    if (p >= n) return;
    PairDesc & d = desc[p];
    int const la = d.la, lb = d.lb, R = d.R, bitWords = d.bitWords;
    uint32_t const * __restrict__ bits = (uint32_t const *)(arena + d.offBits);
    int32_t * __restrict__ rowStart = arena + d.offRowStart;
    uint8_t const * A = in.seqCat + in.seqOff[d.seqFirst];
    uint8_t const * B = in.seqCat + in.seqOff[d.seqSecond];

    // 1. rowStart = exclusive prefix of the row popcounts
    int carry = 0;
    for (int a0 = 0; a0 < la; a0 += 32)
    {
        int const a = a0 + lane;
        int c = 0;
        if (a < la)
            for (int w = 0; w < bitWords; ++w) c += __popc(bits[(size_t)a * bitWords + w]);
        int total;
        int const ex = warp_excl_scan(c, lane, total);
        if (a < la)
        {
            rowStart[a] = carry + ex;
            if (d.offRowRec)
            {
                // the row's window of the band as one word (kernels_post.cu: edge_of)
                int const lo = a - d.dHi;
                uint32_t word = 0u;
                if (lo >= 0) word = strip_bits(bits + (size_t)a * bitWords, bitWords, lo, 32);
                else if (lo > -32) word = strip_bits(bits + (size_t)a * bitWords, bitWords, 0, 32) << (-lo);
                // third word: the cells that have interactions, set by setup_fill_kernel; fourth: the base of A
                reinterpret_cast<int4 *>(arena + d.offRowRec)[a] = make_int4(carry + ex, (int)word, 0, (int)A[a]);
            }
        }
        carry += total;
    }
    if (lane == 0) rowStart[la] = carry;
    __syncwarp();

    // 2. cell ids and sequence scores (src/lagrange.hpp:223-226, :270)
    int32_t * __restrict__ cell = arena + d.offCell;
    int32_t * __restrict__ seqScore = arena + d.offSeqScore;
    float const scale = state[p].scale;
    for (int a = lane; a < la; a += 32)
    {
        int e = rowStart[a];
        for (int w = 0; w < bitWords; ++w)
        {
            uint32_t m = bits[(size_t)a * bitWords + w];
            while (m)
            {
                int const b = (w << 5) + __ffs(m) - 1;
                m &= m - 1;
                cell[e] = (a << 16) | b;
                seqScore[e] = __float2int_rz(__fmul_rn(scale, __int2float_rn(P.tab[(int)A[a] * 5 + (int)B[b]])));
                ++e;
            }
        }
    }
    __syncwarp();

    // 3. interaction counts -> CSR offsets
    int32_t * __restrict__ ioff = arena + d.offIoff;
    int const E = d.E;
    carry = 0;
    for (int e0 = 0; e0 < E; e0 += 32)
    {
        int const e = e0 + lane;
        int c = 0;
        if (e < E)
        {
            int const ab = cell[e];
            for_each_interaction(d, in, bits, ab >> 16, ab & 0xffff, [&](int, int, int) { ++c; });
        }
        int total;
        int const ex = warp_excl_scan(c, lane, total);
        if (e < E) ioff[e] = carry + ex;
        carry += total;
    }
    if (lane == 0) { ioff[E] = carry; d.D = carry; }

    // 4. geometry of the value stream
    if (d.bandW > 0)
    {
        // band mode: the dense band array of the pair's warp group is laid out by the host (offBVals)
        if (lane == 0) d.nVals = 0;
        return;
    }
    if (d.dense)
    {
        // row-by-row kernel: dense (lb + 1) x (la + 1) score array
        if (lane == 0) d.nVals = (lb + 1) * (la + 1);
        return;
    }
    // full-matrix mode: strip-dense stream: step flags per lane, stream start per lane
    int const nFlagWords = (d.nSteps + 31) >> 5;
    uint32_t * __restrict__ flags = (uint32_t *)(arena + d.offFlags) + (size_t)lane * nFlagWords;
    for (int w = 0; w < nFlagWords; ++w) flags[w] = 0u;
    int mine = 0;
    for (int a = 0; a < la; ++a)
    {
        if (strip_bits(bits + (size_t)a * bitWords, bitWords, lane * R, min(R, max(0, lb - lane * R))))
        {
            int const t = a + lane;
            flags[t >> 5] |= 1u << (t & 31);
            ++mine;
        }
    }
    int total;
    int const ex = warp_excl_scan(mine * R, lane, total);
    arena[d.offLanePtr + lane] = ex;
    if (lane == 0) d.nVals = total;
}

__global__ void __launch_bounds__(128)
setup_fill_kernel(int n, PairDesc * __restrict__ desc, PairState * __restrict__ state, int32_t * __restrict__ arena,
                  DevInputs in, DevParams P)
{
    int const p = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    if (p >= n) return;
    PairDesc const & d = desc[p];
    int const la = d.la, lb = d.lb, R = d.R, bitWords = d.bitWords, E = d.E;
    uint32_t const * __restrict__ bits = (uint32_t const *)(arena + d.offBits);
    int32_t const * __restrict__ rowStart = arena + d.offRowStart;
    int32_t const * __restrict__ cell = arena + d.offCell;
    int32_t const * __restrict__ seqScore = arena + d.offSeqScore;
    int32_t const * __restrict__ ioff = arena + d.offIoff;
    int32_t * __restrict__ part = arena + d.offPart;
    int32_t * __restrict__ ppos = arena + d.offPpos;
    int32_t * __restrict__ sc = arena + d.offSc;
    int32_t * __restrict__ dual = arena + d.offDual;
    int32_t * __restrict__ rev = arena + d.offRev;
    int32_t * __restrict__ pv = arena + d.offPv;

    // 1. interactions in ascending partner order (src/lagrange.hpp:278-299)
    for (int e = lane; e < E; e += 32)
    {
        int const ab = cell[e];
        int x = ioff[e];
        int32_t top = seqScore[e];
        int32_t const sq = top;
        for_each_interaction(d, in, bits, ab >> 16, ab & 0xffff, [&](int ap, int bp, int wsum)
        {
            // structScore = balance * (head.first + tail.first) / 2 * factor2int  (:289), float, truncated
            int32_t const s = __float2int_rz(__fmul_rn(__fdiv_rn(__fmul_rn(P.balance, __int2float_rn(wsum)), 2.0f), LARA_FACTOR2INT));
            part[x] = edge_id(bits, bitWords, rowStart, ap, bp);
            ppos[x] = (ap << 16) | bp;
            sc[x] = s;
            dual[x] = 0;
            top = max(top, sq + s);
            ++x;
        });
        pv[e] = top;                                             // :305-309
    }
    __syncwarp();
    // 2. reverse interaction index
    for (int e = lane; e < E; e += 32)
        for (int x = ioff[e]; x < ioff[e + 1]; ++x)
        {
            int const m = part[x];
            int r = -1;
            for (int y = ioff[m]; y < ioff[m + 1]; ++y)
                if (part[y] == e) { r = y; break; }
            rev[x] = r;
            if (r < 0) state[p].error = 3;
        }
    // 3. value stream + permutation
    int32_t * __restrict__ vals = arena + d.offVals;
    int32_t * __restrict__ perm = arena + d.offPerm;
    if (d.bandW > 0)
    {
        // band mode: matrix cell (c, r) = (a + 1, b + 1) sits in row r at diagonal index x = a - b - dLo of the
        // group's interleaved band array (lara_gpu.cuh: band_val_index); rows 1..lb of this pair's slot
        int const ppw = 32 / d.bandG, nch = d.bandW >> 2;
        int4 * __restrict__ bv = reinterpret_cast<int4 *>(arena + d.offBVals) + d.bandSlot;
        __syncwarp();
        for (int i = lane; i < lb * nch; i += 32)
            bv[((size_t)nch + (size_t)i) * ppw] = make_int4(kDpNeg, kDpNeg, kDpNeg, kDpNeg);
        __syncwarp();
        vals = arena + d.offBVals;
        for (int e = lane; e < E; e += 32)
        {
            int const ab = cell[e];
            int const a = ab >> 16, b = ab & 0xffff;
            int const idx = (int)band_val_index(d.bandW, ppw, d.bandSlot, b + 1, a - b - d.dLo);
            vals[idx] = pv[e];
            perm[e] = idx;
        }
    }
    else if (d.dense)
    {
        // dense array of the row-by-row kernel: cell (c, r) = (a + 1, b + 1) at r * (la + 1) + c
        int const ncol = la + 1;
        int const total = (lb + 1) * ncol;
        for (int i = lane; i < total; i += 32) vals[i] = kDpNeg;
        __syncwarp();
        for (int e = lane; e < E; e += 32)
        {
            int const ab = cell[e];
            int const idx = ((ab & 0xffff) + 1) * ncol + (ab >> 16) + 1;
            vals[idx] = pv[e];
            perm[e] = idx;
        }
    }
    else
    {
    int ptr = arena[d.offLanePtr + lane];
    int const lo = lane * R, len = min(R, max(0, lb - lo));
    for (int a = 0; a < la; ++a)
    {
        uint32_t const sb = strip_bits(bits + (size_t)a * bitWords, bitWords, lo, len);
        if (!sb) continue;
        int e = edge_id(bits, bitWords, rowStart, a, lo);        // id of the first active cell at or after (a, lo)
        for (int j = 0; j < R; ++j)
        {
            if ((sb >> j) & 1u) { vals[ptr + j] = pv[e]; perm[e] = ptr + j; ++e; }
            else vals[ptr + j] = kDpNeg;
        }
        ptr += R;
    }
    }
    // 4. records of the iteration kernel (lara_gpu.cuh: offEdge, offInter, offInterCold)
    __syncwarp();
    {
        int4 * __restrict__ edge = reinterpret_cast<int4 *>(arena + d.offEdge);
        int4 * __restrict__ inter = reinterpret_cast<int4 *>(arena + d.offInter);
        int4 * __restrict__ cold = reinterpret_cast<int4 *>(arena + d.offInterCold);
        for (int e = lane; e < E; e += 32)
        {
            int const x0 = ioff[e], x1 = ioff[e + 1];
            edge[e] = make_int4(seqScore[e], pv[e], x0, x1 - x0);
            if (d.offRowRec && x1 > x0)
            {
                int const ab = cell[e], a = ab >> 16, b = ab & 0xffff;
                atomicOr(&reinterpret_cast<int32_t *>(arena + d.offRowRec)[4 * a + 2], (int)(1u << (b - (a - d.dHi))));
            }
            for (int x = x0; x < x1; ++x)
            {
                int const m = part[x];
                inter[(size_t)x] = make_int4(rev[x], ppos[x], sc[x], 0);
                cold[(size_t)x] = make_int4(m, seqScore[m], e, 0);
            }
        }
    }
    // 5. solver state (src/subgradient_solver.hpp:86-100, simd :153,387-393)
    if (lane == 0)
    {
        PairState & st = state[p];
        st.bestLower = -kInf; st.bestUpper = kInf; st.currentLower = -kInf; st.currentUpper = kInf;
        st.nondecreasing = 0; st.stepFactorF = P.stepSizeFactor;
        st.stepFactorI = __float2int_rz(__fmul_rn(P.stepSizeFactor, LARA_FACTOR2INT));
        st.remaining = P.numIterations; st.done = 0; st.bestScore = -kInf; st.nBest = 0;
        st.nLines = 0; st.gapScore = 0; st.nList = 0;
    }
}

// unit entry point: every cell active, values straight from a dense pssm
__global__ void dense_to_stream_kernel(int n, PairDesc const * __restrict__ desc, int32_t * __restrict__ arena,
                                       int32_t const * __restrict__ pssm)
{
    int const p = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    int const lane = threadIdx.x & 31;
    if (p >= n) return;
    PairDesc const & d = desc[p];
    int const la = d.la, lb = d.lb, R = d.R;
    int const nFlagWords = (d.nSteps + 31) >> 5;
    uint32_t * flags = (uint32_t *)(arena + d.offFlags) + (size_t)lane * nFlagWords;
    for (int w = 0; w < nFlagWords; ++w) flags[w] = 0xffffffffu;
    arena[d.offLanePtr + lane] = lane * la * R;
    int32_t * vals = arena + d.offVals + (size_t)lane * la * R;
    int32_t const * src = pssm + (size_t)p * la * lb;
    for (int a = 0; a < la; ++a)
        for (int j = 0; j < R; ++j)
        {
            int const b = lane * R + j;
            vals[a * R + j] = b < lb ? src[(size_t)a * lb + b] : kDpNeg;
        }
}

} // namespace

cudaError_t launch_edge_filter(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                               DevParams const & P, int maxLb, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    // rows per lane: the smallest instantiation that covers the longest second sequence of the launch in one pass
    if (maxLb <= 128) edge_filter_kernel<4><<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    else if (maxLb <= 256) edge_filter_kernel<8><<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    else if (maxLb <= 384) edge_filter_kernel<12><<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    else edge_filter_kernel<16><<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    return cudaGetLastError();
}

cudaError_t launch_setup_count(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                               DevParams const & P, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    setup_count_kernel<<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    return cudaGetLastError();
}

cudaError_t launch_setup_fill(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                              DevParams const & P, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    setup_fill_kernel<<<(n + 3) / 4, 128, 0, stream>>>(n, desc, state, arena, in, P);
    return cudaGetLastError();
}

cudaError_t launch_dense_to_stream(int n, PairDesc const * desc, int32_t * arena, int32_t const * pssm,
                                   cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    dense_to_stream_kernel<<<(n + 3) / 4, 128, 0, stream>>>(n, desc, arena, pssm);
    return cudaGetLastError();
}

} // namespace lara
