// lara_gpu.cuh -- shared device-side definitions of the B200 LaRA hot path.
//
// Execution model: a batch of sequence pairs advances through the subgradient
// iterations in lock step; every iteration is
//   1. gotoh_band_kernel<G,Q> G lanes per pair (one thread per pair at G = 1): the affine-gap DP over the band of
//                             diagonals that holds the pair's active edges, exact (kernels_band.cu); pairs without a
//                             usable band: gotoh_fill_kernel<R>, one warp per pair over the full matrix (kernels_dp.cu)
//   2. iterate_post_kernel    one warp per pair: traceback, valid_solution + greedy matching,
//                             bounds / step size / multiplier update, pssm refresh
// and the per-pair Lagrange problem lives in HBM in the flat layout described below.
#ifndef LARA_GPU_CUH_
#define LARA_GPU_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lara_b200.h"

namespace lara
{

// SeqAn:align/dp_cell.h:143-146  (lowest()/2) -- "minus infinity" of the DP borders.
constexpr int32_t kDpNeg = INT32_MIN / 2;
// src/score.hpp:92 -- pssm value of inactive cells.  Only "never wins against a gap path" matters.
constexpr int32_t kPssmInit = -1431655764;
// src/data_types.hpp:77
constexpr int32_t kInf = LARA_INFINITY;
// band DP: value of cells that do not exist (yet); never wins, never overflows when a gap cost or kDpNeg is added
constexpr int32_t kBandNeg = -(1 << 29);

constexpr int kWarp = 32;
// Largest rows-per-lane the single-pass DP kernel is instantiated for: lenB <= 32 * kMaxR.
constexpr int kMaxR = 20;

// Trace nibble of one DP cell: the four decisions of SeqAn's _computeScore, each
// bit set when the RIGHT operand of _maxScore won (the left one wins ties).  A lane pushes the nibbles of
// its R cells MSB-first into ceil(R/8) words per step: cell j sits in word j/8 at bit 4*(n-1-(j%8)) where n
// is the number of cells in that word.
enum : uint32_t
{
    kTrGap   = 1u,  // S = best gap score (else: diagonal)                 -> !DIAGONAL
    kTrFromH = 2u,  // best gap matrix is H (else: V)                      -> MAX_FROM_HORIZONTAL_MATRIX
    kTrVOpen = 4u,  // V opened from S (else: extended)                    -> VERTICAL_OPEN
    kTrHOpen = 8u   // H opened from S (else: extended)                    -> HORIZONTAL_OPEN
};

// Static description of one pair in the batch (written by the host after the set-up counts are known).
// All off* are offsets in 4-byte units into the batch arena.
struct PairDesc
{
    int32_t la, lb;             // lengths; A = first = longer sequence = DP columns (H), B = rows (V)
    int32_t seqFirst, seqSecond;// sequence indices (first is the longer one, src/subgradient_solver.hpp:117-120)
    int32_t R;                  // rows per lane of the DP: ceil(lb / 32)
    int32_t W;                  // trace words per lane and step: ceil(R / 8)
    int32_t nSteps;             // la + 31 rounded up to a multiple of 4
    int32_t E;                  // active alignment edges
    int32_t D;                  // interactions == dual variables (Lagrange::dimension)
    int32_t nVals;              // entries of the strip-dense pssm stream
    int32_t bitWords;           // words per row of the active bitmap: ceil(lb / 32)
    int32_t kCap;               // upper bound of matching candidates per iteration
    int32_t outIdx;             // position of the pair in the device's prepared pair list (== its result slot)
    int32_t dense;              // 1: row-by-row full-matrix kernel (kernels_rows.cu): offVals = dense (lb+1) x (la+1) scores,
                                //    offTrace = one decision byte per cell, offFlags = four rows of la + 1 ints (S, V buffers)
    int64_t offBits;            // active bitmap   [la][bitWords]
    int64_t offRowStart;        // first compact edge id of bitmap row a  [la + 1]
    int64_t offFlags;           // per lane step flags: bit (t & 31) of word [lane][t >> 5], nFlagWords words per lane
    int64_t offLanePtr;         // [32] start of each lane's value stream (relative to offVals)
    int64_t offVals;            // strip-dense pssm stream [nVals]
    int64_t offTrace;           // trace words [32 lanes][nSteps][W]
    int64_t offCell;            // per edge: (a << 16) | b          [E]
    int64_t offSeqScore;        // per edge: getSeqScore            [E]
    int64_t offPv;              // per edge: top of its priority queue == pssm value [E]
    int64_t offPerm;            // per edge: position in the value stream [E]
    int64_t offIoff;            // per edge: CSR row start in the interaction arrays [E + 1]
    int64_t offPart;            // per interaction: partner edge id  [D]
    int64_t offPpos;            // per interaction: partner (a << 16) | b [D]
    int64_t offSc;              // per interaction: structure score  [D]
    int64_t offDual;            // per interaction: Lagrange multiplier [D]
    int64_t offRev;             // per interaction: index of the reverse interaction [D]
    int64_t offList;            // subgradient index list (sign in bit 31) [D]
    // records read by the iteration kernel (written at the end of the set-up from the arrays above, which stay
    // as the build-time form; pv and dual live only here afterwards):
    int64_t offEdge;            // per edge, 16 bytes: { seqScore, pv (queue top == pssm value), ioff, degree }   [E]
    int64_t offInter;           // per interaction, 16 bytes (hot): { reverse index, partner cell, structScore, dual }  [D]
    int64_t offInterCold;       // per interaction, 16 bytes (cold): { partner edge, seqScore of the partner, owner edge, 0 }  [D]
    int64_t offRowRec;          // bands of <= 32 diagonals, per row a of the bitmap, 16 bytes: { id of the row's first edge,
                                //   activity bits of b = a - dHi .. a - dHi + 31, of those the cells with interactions,
                                //   Rna5 code of A[a] }   [la]   (0: not present)
    int64_t offLines;           // current lines  (a << 16) | b  [lb]
    int64_t offBest;            // best lines     (a << 16) | b | matched << 31  [lb]
    int64_t offCand;            // global candidate buffer for the matching sort, 2 words each [2 * kCap] (8-byte aligned)
    int64_t offScratch;         // set-up scratch: forward prefix scores [la * lb]
    // ---- band mode (kernels_band.cu): the DP runs only over the diagonals dLo <= c - r <= dHi ------------
    int32_t dMin, dMax;         // diagonal range (a - b) of the active edges (edge filter); dMin > dMax: none
    int32_t bandW;              // diagonals of the band = 2 * bandG * bandQ (24 or a multiple of 32); 0 = full-matrix kernel
    int32_t bandG, bandQ;       // lanes per pair (32 / bandG pairs share a warp) and diagonals per block
    int32_t bandSlot;           // position of the pair inside its warp group: 0 .. 32 / bandG - 1
    int32_t dLo, dHi;           // band (dLo <= 0 <= dHi, dHi = dLo + bandW - 1)
    // band arrays of the pair's warp group (PPW = 32 / bandG pairs, rows 0 .. max lb of the group), interleaved:
    //   score of (row r, diagonal index x = c - r - dLo):  int  [((r * bandW/4 + x/4) * PPW + slot) * 4 + x%4]
    //   trace nibble of the same cell: unit (r * 2 bandG + x / bandQ) * PPW + slot of bandQ nibbles (bandQ / 2 bytes),
    //   nibble x % bandQ at bits 4 * (x % bandQ) (little endian)
    int64_t offBVals;           // group base of the scores
    int64_t offBTrace;          // group base of the trace
    int64_t offU;               // region U (c - r > dHi): TH[lb + 1], TS[lb + 1], M[lb + 1]
    int64_t offL;               // region L (c - r < dLo): TV[la + 1], TS[la + 1], M[la + 1]
};

// Mutable per-pair solver state (BoundInfo + SubgradientSolver fields, src/subgradient_solver.hpp:64-108).
struct PairState
{
    int32_t bestLower, bestUpper, currentLower, currentUpper;
    uint32_t nondecreasing;
    float stepFactorF;          // scalar build: float stepSizeFactor
    int32_t stepFactorI;        // SIMD build: int32 step factor (src/subgradient_solver_simd.hpp:153)
    int32_t remaining;          // remainingIterations
    int32_t done;               // 1 once the pair has finished (result written)
    int32_t sEnd, vEnd, hEnd;   // S, V, H of the last DP cell (end-cell trace correction)
    int32_t bestScore;          // Lagrange::bestStructuralAlignmentScore
    int32_t nBest;              // number of best lines
    int32_t nLines, gapScore;   // last alignment
    int32_t nList;              // subgradientIndices.size()
    int32_t optScore;           // edge filter optimum (forward == backward)
    float   scale;              // sequenceScaleFactor (src/lagrange.hpp:252)
    int32_t error;              // non-zero: internal invariant violated
    int32_t pad_[3];
};

// Solver parameters in device constant form.
struct DevParams
{
    int32_t numIterations, maxNondecreasing, lookahead, stepMode;
    float stepSizeFactor, balance, sequenceScale;
    int32_t subopt;             // int32(suboptimalDiff * 8192)
    int32_t go, ge;
    int32_t tab[25];
    uint32_t libMin, libMax;
    int32_t libLinear;
};

// Read-only inputs shared by all pairs of a context.
struct DevInputs
{
    uint8_t const * seqCat;
    int64_t const * seqOff;
    int32_t const * bpPartner;
    int32_t const * bpWeight;
    int64_t const * bpOff;
};

// Output slot of one pair (copied back to the host as is).
struct PairOutHdr
{
    int32_t bestLower, bestUpper, iterations, nLines;
};

// ---- launchers (one translation unit each) -------------------------------------------------
// DP fill over the pairs listed in pairIdx[0..n) that all have the same R.
cudaError_t launch_gotoh_fill(int R, int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t const * negBuf, int32_t go, int32_t ge, cudaStream_t stream);
// full-matrix DP fill, row by row, one block per pair of any size (kernels_rows.cu): pairs listed in pairIdx[0..n)
// mode 0: affine gap costs with go <= ge, 1: linear gap costs (Needleman-Wunsch), 2: affine with go > ge
cudaError_t launch_gotoh_rows(int mode, int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                              int32_t go, int32_t ge, cudaStream_t stream);
// band DP fill (kernels_band.cu) over warp groups of 32 / G pairs with the same band shape (G lanes per pair, blocks
// of Q diagonals): pairIdx[0..nSlots) lists the groups slot by slot, -1 = empty slot
cudaError_t launch_gotoh_band(int G, int Q, int nSlots, int32_t const * pairIdx, PairDesc const * desc, PairState * state,
                              int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream);
bool band_shape_supported(int G, int Q);
// all band shapes of a batch in one launch: class k works on pairIdx[slotOffset .. slotOffset + nSlots)
constexpr int kMaxBandClasses = 16;
struct BandLaunch { int G, Q, nSlots, slotOffset; };
struct BandClassTable
{
    int32_t n;
    int32_t shape[kMaxBandClasses];             // (G << 8) | Q
    int32_t firstBlock[kMaxBandClasses + 1];    // first block of the class's part of the grid
    int32_t slotOffset[kMaxBandClasses];
    int32_t nSlots[kMaxBandClasses];
};
cudaError_t launch_gotoh_band_multi(int nClasses, BandLaunch const * cls, int32_t const * pairIdx, PairDesc const * desc,
                                    PairState * state, int32_t * arena, int32_t go, int32_t ge, cudaStream_t stream);
// position (in ints, relative to offBVals) of the score of band cell (row r, diagonal index x)
__host__ __device__ inline int64_t band_val_index(int bandW, int ppw, int slot, int r, int x)
{
    return (((int64_t)r * (bandW >> 2) + (x >> 2)) * ppw + slot) * 4 + (x & 3);
}
// traceback + valid_solution + matching + bounds + update for the pairs in the n slots of the batch's slot table
// (pairIdx; -1 = empty slot).  finished[0] counts the pairs that have finished, finished[1 ..] lists them.
cudaError_t launch_iterate_post(int n, int32_t const * pairIdx, PairDesc const * desc, PairState * state, int32_t * arena,
                                DevParams const & P, PairOutHdr * outHdr, int32_t * outLines, int64_t const * outOff,
                                int32_t * finished, int maxLa, int maxLb, int smemCand, DevInputs const & in,
                                cudaStream_t stream);
// traceback only (unit entry point lara_gpu_gotoh_batch)
cudaError_t launch_traceback_only(int n, PairDesc const * desc, PairState * state, int32_t * arena,
                                  int32_t go, int32_t ge, int maxLa, int maxLb, cudaStream_t stream);
// result gather: the used part of every pair's output slot -> dense posLow / posHigh / weight arrays at dstOff[q]
cudaError_t launch_pack_lines(int n, PairOutHdr const * hdr, int32_t const * outLines, int64_t const * outOff,
                              int64_t const * dstOff, int32_t * low, int32_t * high, int32_t * weight, cudaStream_t stream);
// dynamic shared memory one warp of the post kernels needs for pairs up to maxLa x maxLb
size_t post_smem_bytes_per_warp(int maxLa, int maxLb, int smemCand);
// per-device one-time configuration of the post kernels (shared memory opt-in)
cudaError_t post_configure_device();
// set-up stage 1: edge filter -> bitmap, optimum score, E
cudaError_t launch_edge_filter(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                               DevParams const & P, int maxLb, cudaStream_t stream);
// set-up stage 2: rowStart, cell ids, sequence scores, interaction counts -> D, stream geometry -> nVals
cudaError_t launch_setup_count(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                               DevParams const & P, cudaStream_t stream);
// set-up stage 3: interactions, reverse indices, initial pssm, value stream
cudaError_t launch_setup_fill(int n, PairDesc * desc, PairState * state, int32_t * arena, DevInputs in,
                              DevParams const & P, cudaStream_t stream);
// dense pssm -> strip-dense stream with all cells active (unit entry point)
cudaError_t launch_dense_to_stream(int n, PairDesc const * desc, int32_t * arena, int32_t const * pssm,
                                   cudaStream_t stream);
// integer issue-rate microbenchmark
cudaError_t launch_int_issue(int kind, int blocks, int threads, int iters, int32_t * out, cudaStream_t stream);

} // namespace lara
#endif
