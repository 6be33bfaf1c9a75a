// lara_gpu_main.cpp -- the reference's program with the GPU solver at its one seam.
//
// The reference tree is read-only here, so instead of a third branch in src/lara.cpp:38-42 this translation unit
// includes the reference's own headers (argument parser, readers, writers: all unmodified, from the reference
// tree) together with integration/subgradient_solver_gpu.hpp and repeats the flow of the reference's main
// (src/lara.cpp:44-61): parameters -> InputStorage -> OutputLibrary -> solve -> print.
#include "data_types.hpp"
#include "io.hpp"
#include "parameters.hpp"

#include "subgradient_solver_gpu.hpp"

int main(int argc, char const ** argv)
{
    lara::Clock::time_point const started = lara::Clock::now();
    lara::Parameters params(argc, argv);
    if (params.status != lara::Parameters::Status::CONTINUE)
        return static_cast<int>(params.status);
    lara::InputStorage store(params);
    if (store.had_err())
        return 1;
    lara::OutputLibrary library(store, params.outFormat);
    try
    {
        lara::solve(library, store, params);
    }
    catch (std::exception const & e)
    {
        std::cerr << "ERROR: " << e.what() << std::endl;
        return 1;
    }
    library.print(params.outFile);
    _LOG(1, "LaRA has run for " << lara::timeDiff<std::chrono::seconds>(started) << " seconds.\n");
    return 0;
}
