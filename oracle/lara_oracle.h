/*
 * lara_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's all-pairs Lagrangian subgradient loop
 * (seqan/lara, src/subgradient_solver.hpp, src/lagrange.hpp, src/edge_filter.hpp,
 * src/matching.hpp and the SeqAn affine-gap DP it instantiates).  It is the checker
 * for the CUDA path: only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product library
 * (lara_b200/csrc) never links or calls anything in oracle/.
 *
 * Parity is PINNED: tests/test_oracle_vs_reference.py compares this restatement with
 * the outputs of the unmodified reference binary (oracle/_ref/lara_scalar, lara_simd,
 * built from /root/reference by oracle/build_ref.sh) on the committed fixtures under
 * tests/golden/, and with SeqAn's own Gotoh known-answer tests.
 */
#ifndef LARA_ORACLE_H_
#define LARA_ORACLE_H_

#include "../include/lara_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Same input layout as lara_gpu_set_sequences + lara_gpu_solve_pairs.  n_threads OpenMP
 * threads over pairs (dynamic), like the reference's -j.  Results in request order.
 * The caller frees with lara_oracle_free_results. */
int lara_oracle_solve_pairs(lara_params const * params, int32_t n_seq,
                            uint8_t const * seq_cat, int64_t const * seq_off,
                            int32_t const * bp_partner, int32_t const * bp_weight, int64_t const * bp_off,
                            int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b,
                            int n_threads, lara_results * out, lara_stats * stats);
void lara_oracle_free_results(lara_results * out);

/* SeqAn-exact affine global alignment on a dense position-specific matrix (one problem). */
int lara_oracle_gotoh(int32_t lenA, int32_t lenB, int32_t const * pssm, int32_t gap_open, int32_t gap_extend,
                      int32_t * score, int32_t * n_lines, int32_t * lines_a, int32_t * lines_b,
                      int32_t * gap_score, uint8_t * trace_out /* (lenA+1)*(lenB+1) or NULL */);

/* generateEdges of src/edge_filter.hpp:131-158. */
int lara_oracle_edge_filter(lara_params const * params, uint8_t const * seqA, int32_t lenA,
                            uint8_t const * seqB, int32_t lenB, uint8_t * active, int32_t * opt_score);

/* Matching::computeGreedyMatching (src/matching.hpp:143-192) on explicit interactions
 * (weight w[k], endpoints l[k] < r[k]); returns the score, matched[] flags per vertex id < n_vertices. */
int32_t lara_oracle_greedy_matching(int32_t n, int32_t const * w, int32_t const * l, int32_t const * r,
                                    int32_t lookahead, int32_t n_vertices, uint8_t * matched);

#ifdef __cplusplus
}
#endif
#endif
