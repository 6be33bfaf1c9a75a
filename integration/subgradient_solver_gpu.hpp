// subgradient_solver_gpu.hpp -- the reference-side binding of liblara_b200.so.
//
// A third definition of lara::solve(results, store, params) next to the reference's two
// (src/subgradient_solver.hpp:110, src/subgradient_solver_simd.hpp:108; chosen in src/lara.cpp:38-42): it flattens
// what Lagrange::Lagrange reads from the InputStorage (src/lagrange.hpp:130-138, 259-262), hands it to the GPU
// library through the C ABI of include/lara_b200.h and feeds the answers to OutputLibrary::addAlignment
// (src/io.hpp:380-383).  Everything else -- SeqAn's argument parser, the RNA file readers, the writers -- is the
// reference's own code, compiled from the reference tree.
//
// Build: integration/build.sh (needs the reference sources; the result runs wherever liblara_b200.so does).
// -j is the number of GPUs here.  LARA_B200_STEP_MODE=1 selects the step-size arithmetic of the reference's SIMD
// build (src/subgradient_solver_simd.hpp:303-330) instead of the scalar build's.
#pragma once

#include <cstdint>
#include <cstdlib>
#include <numeric>
#include <stdexcept>
#include <string>
#include <vector>

#include <lara_b200.h>

#include "data_types.hpp"
#include "io.hpp"
#include "parameters.hpp"

namespace lara
{

inline void solve(OutputLibrary & results, InputStorage const & store, Parameters & params)
{
    lara_params p;
    lara_params_default(&p);
    p.num_iterations    = static_cast<int32_t>(params.numIterations);
    p.max_nondecreasing = static_cast<int32_t>(params.maxNondecrIterations);
    p.matching          = static_cast<int32_t>(params.matching);
    p.step_mode         = LARA_STEP_SCALAR_FLOAT;
    if (char const * e = std::getenv("LARA_B200_STEP_MODE"))
        p.step_mode = std::atoi(e) == 1 ? LARA_STEP_SIMD_INT32 : LARA_STEP_SCALAR_FLOAT;
    p.step_size_factor  = params.stepSizeFactor;
    p.suboptimal_diff   = params.suboptimalDiff;
    p.balance           = params.balance;
    p.sequence_scale    = params.sequenceScale;
    p.gap_open          = params.rnaScore.data_gap_open;            // already x8192, src/parameters.hpp:303-304
    p.gap_extend        = params.rnaScore.data_gap_extend;
    for (int i = 0; i < 25; ++i)
        p.score_matrix[i] = params.rnaScore.data_tab[i];
    p.lib_score_min     = params.libraryScoreMin;
    p.lib_score_max     = params.libraryScoreMax;
    p.lib_score_linear  = params.libraryScoreIsLinear ? 1 : 0;

    // Rna5 codes and, per base, the partners of the structure graph a Lagrange object would read
    std::vector<uint8_t> seq;
    std::vector<int64_t> seqOff{0}, bpOff{0};
    std::vector<int32_t> partner, weight;
    for (seqan::RnaRecord const & rec : store)
    {
        seqan::RnaStructureGraph const & graph = seqan::front(seqan::empty(rec.bppMatrGraphs) ? rec.fixedGraphs
                                                                                              : rec.bppMatrGraphs);
        for (size_t i = 0; i < seqan::length(rec.sequence); ++i)
        {
            seq.push_back(static_cast<uint8_t>(seqan::ordValue(rec.sequence[i])));
            for (seqan::RnaAdjacencyIterator it(graph.inter, i); !seqan::atEnd(it); seqan::goNext(it))
            {
                size_t const other = seqan::value(it);
                float const probability = seqan::cargo(seqan::findEdge(graph.inter, i, other));
                partner.push_back(static_cast<int32_t>(other));
                weight.push_back(static_cast<ScoreType>(probability));       // Contact truncates, src/lagrange.hpp:135
            }
            bpOff.push_back(static_cast<int64_t>(partner.size()));
        }
        seqOff.push_back(static_cast<int64_t>(seq.size()));
    }

    int nDev = static_cast<int>(params.threads);
    int const visible = lara_gpu_visible_devices();
    if (nDev > visible) nDev = visible;
    if (nDev < 1) nDev = 1;
    std::vector<int> devices(static_cast<size_t>(nDev));
    std::iota(devices.begin(), devices.end(), 0);

    _LOG(1, "3) Solve " << store.size() * (store.size() - 1) / 2 << " structural alignments on " << nDev
            << " GPU(s)..." << std::endl);
    lara_ctx * ctx = nullptr;
    if (lara_gpu_create(&ctx, &p, nDev, devices.data()) != LARA_OK)
        throw std::runtime_error(std::string("lara_b200: ") + lara_gpu_last_error(nullptr));
    lara_results res{};
    if (lara_gpu_set_sequences(ctx, static_cast<int32_t>(store.size()), seq.data(), seqOff.data(), partner.data(),
                               weight.data(), bpOff.data()) != LARA_OK
        || lara_gpu_solve_all_pairs(ctx, &res) != LARA_OK)
    {
        std::string const msg = std::string("lara_b200: ") + lara_gpu_last_error(ctx);
        lara_gpu_destroy(ctx);
        throw std::runtime_error(msg);
    }
    for (int64_t k = 0; k < res.n_pairs; ++k)
    {
        lara_pair_result const & r = res.pairs[k];
        WeightedAlignedColumns columns;
        columns.first = PosPair{static_cast<size_t>(r.idx_low), static_cast<size_t>(r.idx_high)};
        columns.second.reserve(static_cast<size_t>(r.n_lines));
        for (int32_t x = 0; x < r.n_lines; ++x)
            columns.second.emplace_back(static_cast<size_t>(res.line_pos_low[r.lines_off + x]),
                                        static_cast<size_t>(res.line_pos_high[r.lines_off + x]),
                                        static_cast<unsigned>(res.line_weight[r.lines_off + x]));
        results.addAlignment(columns, r.best_lower);
    }
    lara_stats st{};
    lara_gpu_get_stats(ctx, &st);
    _LOG(1, "   * solve -> " << st.ms_total << "ms on the device(s)" << std::endl);
    lara_gpu_free_results(&res);
    lara_gpu_destroy(ctx);
}

} // namespace lara
