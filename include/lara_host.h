/*
 * lara_host.h -- C ABI of the thin host side that stays around the GPU hot path:
 * the reference's input readers, option table and output writers.
 *
 *   InputStorage            src/io.hpp:71-339   (readRnaFile, extractBppFromDotplot, computeStructure)
 *   Parameters              src/parameters.hpp:54-333
 *   OutputLibrary::print    src/io.hpp:387-489  (T-Coffee lib, MAFFT "pairs", fasta)
 *
 * The host code is plain C++ (no CUDA, no torch); the CLI binary `lara_b200` and the
 * Python harness (ctypes) both go through these entry points, then hand the flat arrays to
 * lara_gpu_set_sequences / lara_gpu_solve_* (lara_b200.h).
 */
#ifndef LARA_HOST_H_
#define LARA_HOST_H_

#include "lara_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lara_input lara_input;

/* Mirror of the non-solver fields of lara::Parameters (src/parameters.hpp:66-79). */
typedef struct lara_cli_options
{
    int32_t  verbose;            /* -v */
    int32_t  threads;            /* -j (number of GPUs to use here; 0 = all visible) */
    char     in_file[1024];      /* -i */
    char     ref_file[1024];     /* -r */
    char     out_file[1024];     /* -w */
    char     out_format[16];     /* -o lib|pairs|fasta */
    int32_t  n_dotplots;         /* -d (repeatable) */
    char **  dotplot_files;      /* owned by the options struct; free with lara_cli_options_free */
    float    epsilon;            /* -e (parsed, unused: src/parameters.hpp:290) */
    int32_t  structure_scoring;  /* -p */
    char     score_matrix_file[1024]; /* -s */
} lara_cli_options;

/* Parse argv like src/parameters.hpp:101-332.  Returns 2 (CONTINUE), 0 (EXIT_OK: help/version printed)
 * or 1 (EXIT_ERROR), i.e. lara::Parameters::Status. */
int  lara_parse_args(int argc, char const * const * argv, lara_params * params, lara_cli_options * opts);
void lara_cli_options_free(lara_cli_options * opts);

/* InputStorage */
int  lara_input_create(lara_input ** in);
void lara_input_destroy(lara_input * in);
/* readRnaFile (src/io.hpp:135-200): format from the extension: .ebpseq .dbn .bpseq .ct; anything else = FASTA */
int  lara_input_read_file(lara_input * in, char const * path);
/* extractBppFromDotplot (src/io.hpp:202-284) */
int  lara_input_read_dotplot(lara_input * in, char const * path);
/* computeStructure for every record (src/io.hpp:286-338) + the "at least two sequences" check (:115-119).
 * A record without any structure is an error (the reference needs ViennaRNA there, which is out of scope). */
int  lara_input_finish(lara_input * in);
/* add one record from memory (used by tests/bench): seq as characters, pairs as (i,j,weight) triples with
 * float weights that get truncated exactly like src/lagrange.hpp:135; is_bpp=0 means fixed structure (x10). */
int  lara_input_add_record(lara_input * in, char const * name, char const * seq, int32_t n_pairs,
                           int32_t const * pi, int32_t const * pj, double const * w, int is_bpp);

int32_t      lara_input_size(lara_input const * in);
char const * lara_input_name(lara_input const * in, int32_t i);
int32_t      lara_input_length(lara_input const * in, int32_t i);
/* Flat arrays in the layout lara_gpu_set_sequences takes; pointers stay valid until the input changes. */
int  lara_input_flatten(lara_input * in, uint8_t const ** seq_cat, int64_t const ** seq_off,
                        int32_t const ** bp_partner, int32_t const ** bp_weight, int64_t const ** bp_off);

/* OutputLibrary::print: results are sorted by (idx_low, idx_high) like the std::set of src/io.hpp:372.
 * path NULL or "" = stdout.  format: "lib", "pairs" or "fasta". */
int  lara_output_write(lara_input const * in, lara_results const * res, char const * format, char const * path);

char const * lara_host_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* LARA_HOST_H_ */
