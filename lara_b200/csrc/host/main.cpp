// lara_b200 -- command line front end: same flow as the reference's main (src/lara.cpp:44-61):
// parse parameters -> read inputs -> solve (GPU) -> print.
#include <chrono>
#include <cstdio>
#include <string>
#include <vector>

#include "lara_host.h"

extern "C" int lara_gpu_visible_devices(void);

int main(int argc, char const ** argv)
{
    auto const t0 = std::chrono::steady_clock::now();
    lara_params params;
    lara_cli_options opts;
    int const status = lara_parse_args(argc, argv, &params, &opts);
    if (status != 2)
        return status;
    auto log = [&](int level, char const * fmt, auto... a) { if (opts.verbose >= level) std::fprintf(stderr, fmt, a...); };
    log(1, "1) Parse parameters...\n   * number of GPUs requested: %d\n", opts.threads);

    // Read input files and prepare structured sequences (src/io.hpp:74-125).
    log(1, "2) Read input files...\n");
    lara_input * store = nullptr;
    lara_input_create(&store);
    int rc = lara_input_read_file(store, opts.in_file);
    if (rc == 0) rc = lara_input_read_file(store, opts.ref_file);
    for (int k = 0; rc == 0 && k < opts.n_dotplots; ++k)
        rc = lara_input_read_dotplot(store, opts.dotplot_files[k]);
    if (rc == 0) rc = lara_input_finish(store);
    if (rc != 0)
    {
        std::fprintf(stderr, "%s\n", lara_host_last_error());
        return 1;
    }
    if (lara_input_size(store) > 2 && std::string(opts.out_format) == "fasta")
        std::fprintf(stderr, "WARNING: We are computing more than one pairwise alignment and you have selected fasta output.\n");

    uint8_t const * seqCat; int64_t const * seqOff; int32_t const * bpPartner; int32_t const * bpWeight; int64_t const * bpOff;
    lara_input_flatten(store, &seqCat, &seqOff, &bpPartner, &bpWeight, &bpOff);

    // solve(outlib, store, params) on the GPU(s)
    int nDev = opts.threads;
    std::vector<int> ids;
    {
        // -j counts GPUs here; clamp to what is visible
        int const vis = lara_gpu_visible_devices();
        if (vis <= 0) { std::fprintf(stderr, "ERROR: no CUDA device available (lara_b200 has no CPU fallback)\n"); return 1; }
        if (nDev > vis) nDev = vis;
        for (int k = 0; k < nDev; ++k) ids.push_back(k);
    }
    lara_ctx * ctx = nullptr;
    rc = lara_gpu_create(&ctx, &params, nDev, ids.data());
    if (rc != 0) { std::fprintf(stderr, "ERROR: %s\n", lara_gpu_last_error(nullptr)); return 1; }
    rc = lara_gpu_set_sequences(ctx, lara_input_size(store), seqCat, seqOff, bpPartner, bpWeight, bpOff);
    lara_results res{};
    long const nSeq = lara_input_size(store);
    log(1, "3) Solve %ld structural alignments on %d GPU(s)...\n", nSeq * (nSeq - 1) / 2, nDev);
    if (rc == 0) rc = lara_gpu_solve_all_pairs(ctx, &res);
    if (rc != 0) { std::fprintf(stderr, "ERROR: %s\n", lara_gpu_last_error(ctx)); return 1; }
    lara_stats st{};
    lara_gpu_get_stats(ctx, &st);
    log(1, "   * solve -> %.0fms (setup %.0fms, DP %.0fms, post %.0fms; %.2f GCUPS)\n", st.ms_total, st.ms_setup, st.ms_dp,
        st.ms_bounds, st.ms_total > 0 ? (double)st.dp_cells / st.ms_total / 1e6 : 0.0);
    if (opts.verbose >= 2)
        for (int64_t k = 0; k < res.n_pairs; ++k)
            std::fprintf(stderr, "     finished alignment %d/%d after %d iterations.\n", res.pairs[k].idx_low,
                         res.pairs[k].idx_high, res.pairs[k].iterations);

    log(1, "4) Write results...\n");
    rc = lara_output_write(store, &res, opts.out_format, opts.out_file);
    if (rc != 0) std::fprintf(stderr, "%s\n", lara_host_last_error());
    lara_gpu_free_results(&res);
    lara_gpu_destroy(ctx);
    lara_input_destroy(store);
    lara_cli_options_free(&opts);
    double const sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    log(1, "LaRA has run for %.0f seconds.\n", sec);
    return 0;
}
