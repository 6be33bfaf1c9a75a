/*
 * lara_b200.h -- C ABI of the B200-native LaRA 2 hot path.
 *
 * One seam of the reference is replaced:
 *
 *     void lara::solve(OutputLibrary & results, InputStorage const & store, Parameters & params)
 *         reference: src/subgradient_solver.hpp:110 (scalar, the default build)
 *                    src/subgradient_solver_simd.hpp:108 (SEQAN_SIMD_ENABLED build)
 *         called once from src/lara.cpp:57
 *
 * Everything that function does per sequence pair (edge filter, Lagrange problem
 * set-up, Gotoh DP + traceback, valid_solution + greedy matching, bound tracking,
 * subgradient step, multiplier update, library-line extraction) runs on the GPU
 * behind the entry points below.  Plain C, flat arrays, no C++/torch types, no
 * exceptions across the boundary.  All functions return 0 on success or a negative
 * LARA_E_* code; lara_gpu_last_error() gives the text.
 *
 * There is NO CPU fallback: lara_gpu_create() fails with LARA_E_NO_DEVICE when no
 * sm_100 device is usable.
 */
#ifndef LARA_B200_H_
#define LARA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LARA_ABI_VERSION 1

/* error codes */
#define LARA_OK            0
#define LARA_E_INVALID    -1   /* bad argument */
#define LARA_E_NO_DEVICE  -2   /* no usable CUDA device / driver */
#define LARA_E_CUDA       -3   /* CUDA runtime error (text in last_error) */
#define LARA_E_UNSUPPORTED -4  /* option outside the accelerated path (e.g. -m 0 LEMON) */
#define LARA_E_NOMEM      -5
#define LARA_E_IO         -6   /* host helper: file could not be read/written/parsed */

/* Fixed-point scale of every score: src/data_types.hpp:78 (factor2int). */
#define LARA_FACTOR2INT 8192.0f
/* src/data_types.hpp:77 (infinity = INT32_MAX / 3 * 2). */
#define LARA_INFINITY 1431655764

/* Which reference build's step-size arithmetic to reproduce. */
#define LARA_STEP_SCALAR_FLOAT 0   /* src/subgradient_solver.hpp:259-271  (default CMake build) */
#define LARA_STEP_SIMD_INT32   1   /* src/subgradient_solver_simd.hpp:303-335 (wrapping int32 step) */

/*
 * POD mirror of the lara::Parameters fields solve() consumes
 * (src/parameters.hpp:66-92; conversion to x8192 ints at :303-314).
 */
typedef struct lara_params
{
    int32_t  num_iterations;        /* -n  numIterations        (default 500) */
    int32_t  max_nondecreasing;     /* -a  maxNondecrIterations (default 50)  */
    int32_t  matching;              /* -m  greedy lookahead (default 5; 0 = LEMON: unsupported) */
    int32_t  step_mode;             /* LARA_STEP_*                              */
    float    step_size_factor;      /* -f  stepSizeFactor (1.0)                 */
    float    suboptimal_diff;       /* -u  suboptimalDiff (40.0)                */
    float    balance;               /* -b  balance (1.0)                        */
    float    sequence_scale;        /* -c  sequenceScale (0.0)                  */
    int32_t  gap_open;              /* rnaScore.data_gap_open   = -y * 8192 (default -49152) */
    int32_t  gap_extend;            /* rnaScore.data_gap_extend = -x * 8192 (default -16384) */
    int32_t  score_matrix[25];      /* rnaScore.data_tab, Rna5 x Rna5, x8192 truncated */
    uint32_t lib_score_min;         /* -l MIN (500)                             */
    uint32_t lib_score_max;         /* -l MAX (1000)                            */
    int32_t  lib_score_linear;      /* libraryScoreIsLinear = isSet(-l)         */
    int32_t  reserved[9];
} lara_params;

/* Fill *p with the reference defaults (Ribosum65N, -6/-2 gaps, 500 iterations ...). */
void lara_params_default(lara_params * p);

/*
 * Result of one pair == one results.addAlignment(getStructureLines(), bestLower)
 * call of the reference (src/subgradient_solver.hpp:292, src/lagrange.hpp:437-487,
 * src/io.hpp:380-383).
 */
typedef struct lara_pair_result
{
    int32_t idx_low;        /* min(sequence index)  -> WeightedAlignedColumns.first.first  */
    int32_t idx_high;       /* max(sequence index)                                          */
    int32_t best_lower;     /* bounds.bestLower (x8192), the score printed by -o pairs      */
    int32_t best_upper;     /* bounds.bestUpper                                              */
    int32_t iterations;     /* numIterations - remainingIterations at finish                */
    int32_t n_lines;        /* number of (posLow,posHigh,weight) triples                    */
    int64_t lines_off;      /* offset of the first triple in lara_results.line_*            */
} lara_pair_result;

typedef struct lara_results
{
    int64_t            n_pairs;
    lara_pair_result * pairs;       /* n_pairs entries, in the order the pairs were requested */
    int64_t            n_lines_total;
    int32_t *          line_pos_low;   /* 0-based position in sequence idx_low  */
    int32_t *          line_pos_high;  /* 0-based position in sequence idx_high */
    uint32_t *         line_weight;    /* T-Coffee library weight                */
    void *             owner_;         /* private */
} lara_results;

/* Device-side timing and work counters of the last solve call. */
typedef struct lara_stats
{
    double  ms_total;          /* wall time of the solve call (host clock)            */
    double  ms_h2d;            /* host->device copies (CUDA events)                    */
    double  ms_setup;          /* edge filter + Lagrange set-up kernels                */
    double  ms_dp;             /* Gotoh fill kernel, summed over launches              */
    double  ms_traceback;      /* traceback + evaluateLines kernel                     */
    double  ms_bounds;         /* valid_solution + matching + bounds + dual update     */
    double  ms_d2h;            /* result copies                                        */
    int64_t dp_cells;          /* sum over pairs and iterations of lenA*lenB           */
    int64_t dp_launches;       /* number of DP kernel launches                         */
    int64_t kernel_launches;   /* all kernels launched by the library                  */
    int64_t filter_cells;      /* 2*lenA*lenB per pair                                  */
    int64_t h2d_bytes;
    int64_t d2h_bytes;
    int64_t n_pairs;
    int64_t sum_iterations;
    int64_t n_active_edges;    /* sum of E                                              */
    int64_t n_interactions;    /* sum of D (dual dimension)                             */
    double  ms_device;         /* CUDA-event span of the solve on the library's stream (max over devices) */
    int64_t reserved[16];      /* [0] concurrent workers per device, [1] pairs solved by the band kernels, [2] of those in
                                  blocks of 16 diagonals, [3] sum of their band widths, [4] algorithmic HBM bytes of the
                                  DP launches, [5] DP cells actually computed (band cells), [6] batches run,
                                  [7] iteration steps run (one DP launch + one post launch each), [8] pair slots of the
                                  DP launches, summed over the steps, [9] of those the slots whose pair was still
                                  iterating (live-lane fraction = [9] / [8]), [10] algorithmic HBM bytes of ONE DP launch
                                  with every pair of the first batch live, [11] pairs of that launch, [12] pairs taken
                                  into freed slots of a running batch (slot refill), [13..15] unused */
} lara_stats;

typedef struct lara_ctx lara_ctx;

int          lara_gpu_version(void);
/* number of visible CUDA devices (0 when there is no driver / device) */
int          lara_gpu_visible_devices(void);
/* n_devices==0 / device_ids==NULL -> current device only. */
int          lara_gpu_create(lara_ctx ** ctx, lara_params const * params, int n_devices, int const * device_ids);
void         lara_gpu_destroy(lara_ctx * ctx);
char const * lara_gpu_last_error(lara_ctx * ctx);   /* ctx may be NULL for create() errors */

/*
 * The part of InputStorage solve() reads (src/lagrange.hpp:233-262):
 *   seq_cat/seq_off : Rna5 codes 0..4 (A C G U N) of all n sequences, concatenated;
 *   bp_*            : per base (global index in seq_cat) a CSR list of partner positions
 *                     (0-based, local to the sequence) and weights ALREADY truncated to int
 *                     the way std::pair<ScoreType,size_t> truncates them (src/lagrange.hpp:135).
 * Inputs are copied; the caller may free them on return.
 */
int lara_gpu_set_sequences(lara_ctx * ctx, int32_t n_seq,
                           uint8_t const * seq_cat, int64_t const * seq_off,
                           int32_t const * bp_partner, int32_t const * bp_weight, int64_t const * bp_off);

/*
 * Solve the given pairs.  (idx_a[i], idx_b[i]) are sequence indices in any order; the library
 * orients each pair longer-sequence-first exactly like src/subgradient_solver.hpp:115-120.
 * Results come back in request order.  out is owned by the library until lara_gpu_free_results().
 */
int lara_gpu_solve_pairs(lara_ctx * ctx, int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b,
                         lara_results * out);
/* All i<j pairs, in (i,j) lexicographic order == the order OutputLibrary prints them (src/io.hpp:372). */
int lara_gpu_solve_all_pairs(lara_ctx * ctx, lara_results * out);
void lara_gpu_free_results(lara_results * out);
int lara_gpu_get_stats(lara_ctx * ctx, lara_stats * stats);

/*
 * Device-resident variant used for kernel-only timing: uploads and prepares once
 * (prepare), then every run() re-solves the same pairs from HBM-resident inputs and
 * leaves results in HBM; fetch() copies the last results out.
 */
int lara_gpu_prepare_pairs(lara_ctx * ctx, int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b);
int lara_gpu_run_prepared(lara_ctx * ctx);
int lara_gpu_fetch_prepared(lara_ctx * ctx, lara_results * out);

/* ---- unit-level entry points (parity tests of single kernels) ------------------------------ */

/*
 * Affine-gap global alignment with SeqAn's exact tie-breaking of a batch of
 * problems that share lenA/lenB, each over its own dense position-specific score matrix
 * pssm[p][a*lenB+b] (src/score.hpp:94-99).  Outputs per problem: score, and the aligned
 * columns as match lines: n_lines[p], lines_a/lines_b[p*lenB + k] (alignment order).
 * Also gap_score[p] as Lagrange::evaluateLines computes it (src/lagrange.hpp:147-221).
 */
int lara_gpu_gotoh_batch(lara_ctx * ctx, int32_t n_problems, int32_t lenA, int32_t lenB,
                         int32_t const * pssm, int32_t gap_open, int32_t gap_extend,
                         int32_t * score, int32_t * n_lines, int32_t * lines_a, int32_t * lines_b,
                         int32_t * gap_score);

/*
 * The same alignment computed by the BAND kernels (lara_b200/csrc/kernels_band.cu): only the diagonals
 * d_lo <= a - b <= d_lo + band_width - 1 are filled (d_lo <= 0 <= d_lo + band_width - 1), the rest of the matrix is
 * covered by closed forms.  `lanes` threads work on one problem, each on two blocks of band_width / (2 * lanes)
 * diagonals; supported: blocks of 16 diagonals with 1, 2, 4, 8, 16 or 32 lanes (band_width 32 .. 1024), blocks of 8
 * with 2 .. 32 lanes (band_width 32 .. 512) and blocks of 4 with 4 .. 32 lanes (band_width 32 .. 256).  Every active cell (pssm > -2^30) must lie inside the
 * band; the result is then identical to lara_gpu_gotoh_batch.  Needs gap_open < gap_extend.
 */
int lara_gpu_gotoh_band_batch(lara_ctx * ctx, int32_t n_problems, int32_t lenA, int32_t lenB,
                              int32_t const * pssm, int32_t d_lo, int32_t band_width, int32_t lanes,
                              int32_t gap_open, int32_t gap_extend,
                              int32_t * score, int32_t * n_lines, int32_t * lines_a, int32_t * lines_b,
                              int32_t * gap_score);

/*
 * generateEdges (src/edge_filter.hpp:131-158) for one pair of Rna5-coded sequences:
 * active[a*lenB+b] in {0,1}; *opt_score = forward optimal score.
 */
int lara_gpu_edge_filter(lara_ctx * ctx, uint8_t const * seqA, int32_t lenA, uint8_t const * seqB, int32_t lenB,
                         uint8_t * active, int32_t * opt_score);

/*
 * lara_gpu_edge_filter for callers that cannot let the library allocate (CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same computation and same results as lara_gpu_edge_filter, but
 *   - every pointer is a DEVICE pointer of the context's first device, allocated by the caller:
 *     d_seqA[lenA], d_seqB[lenB] (Rna5 codes), d_active[lenA*lenB] (one byte per cell), d_opt_score[1] (may be NULL);
 *   - the pair record, the bitmap and the filter's scratch (lenA*lenB + 6*(lenA+1) ints) live in ONE file-scope
 *     `static __device__` array of LARA_NOALLOC_SCRATCH_BYTES that every GPU which loads the library holds until the
 *     process exits; the call performs no cudaMalloc / cudaFree / stream-ordered allocation (the first call on a device
 *     looks the array's device address up with cudaGetSymbolAddress, later calls only launch kernels);
 *   - sizes whose scratch does not fit (about lenA*lenB > 1.0 M cells, e.g. beyond 1,000 x 1,000 nt) are refused with
 *     LARA_E_UNSUPPORTED and a message that states the need;
 *   - all calls share the array: a call overwrites what the previous one left (nothing needs to be cleared by the
 *     caller), and two calls that overlap on one GPU (two threads, two contexts) RACE - serialise them.
 * Returns after the context's stream has been synchronised.
 */
#define LARA_NOALLOC_SCRATCH_BYTES (4 * 1024 * 1024)
int lara_gpu_edge_filter_noalloc(lara_ctx * ctx, uint8_t const * d_seqA, int32_t lenA, uint8_t const * d_seqB, int32_t lenB,
                                 uint8_t * d_active, int32_t * d_opt_score);

/* Integer-issue microbenchmark for the roofline denominator: returns Gop/s (lane-ops) of
 * dependent-free IADD/VIMNMX streams on the current device. kind: 0 = IADD3, 1 = VIMNMX, 2 = mixed DP mix */
int lara_gpu_int_issue_rate(lara_ctx * ctx, int kind, double * gops);

#ifdef __cplusplus
}
#endif
#endif /* LARA_B200_H_ */
