// lara_cli.cpp -- option table of the reference (src/parameters.hpp:101-332) without SeqAn's
// ArgumentParser: same short/long names, defaults, ranges and float -> x8192 int conversions.
#include "lara_host.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

extern "C" {

void lara_params_default(lara_params * p)
{
    // Ribosum65N (src/rna_score_matrices.hpp:3960-3968; order A C G U N); converted like
    // src/parameters.hpp:311-314 (float * 8192.f, truncated by the int assignment).
    static float const ribosum65n[25] = {
        1.392427f, -0.790206f, -0.503302f, -0.467629f, -1.506506f,
        -0.790206f, 0.838942f, -1.506506f, -0.294126f, -1.506506f,
        -0.503302f, -1.506506f, 0.76724f, -0.718789f, -1.506506f,
        -0.467629f, -0.294126f, -0.718789f, 0.874874f, -1.506506f,
        -1.506506f, -1.506506f, -1.506506f, -1.506506f, -1.506506f};
    std::memset(p, 0, sizeof(*p));
    p->num_iterations = 500;
    p->max_nondecreasing = 50;
    p->matching = 5;
    p->step_mode = LARA_STEP_SCALAR_FLOAT;
    p->step_size_factor = 1.0f;
    p->suboptimal_diff = 40.0f;
    p->balance = 1.0f;
    p->sequence_scale = 0.0f;
    p->gap_open = (int32_t)(-6.0f * LARA_FACTOR2INT);
    p->gap_extend = (int32_t)(-2.0f * LARA_FACTOR2INT);
    for (int i = 0; i < 25; ++i)
        p->score_matrix[i] = (int32_t)(ribosum65n[i] * LARA_FACTOR2INT);
    p->lib_score_min = 500;
    p->lib_score_max = 1000;
    p->lib_score_linear = 0;
}

void lara_cli_options_free(lara_cli_options * o)
{
    if (!o) return;
    for (int32_t k = 0; k < o->n_dotplots; ++k) std::free(o->dotplot_files[k]);
    std::free(o->dotplot_files);
    o->dotplot_files = nullptr;
    o->n_dotplots = 0;
}

} // extern "C"

namespace
{

struct OptDef { char const * s; char const * l; int nargs; };
OptDef const OPTS[] = {
    {"v", "verbose", 1}, {"j", "threads", 1}, {"i", "infile", 1}, {"r", "reffile", 1}, {"d", "dotplot", 1},
    {"w", "write", 1}, {"l", "libscore", 2}, {"o", "outformat", 1}, {"n", "numiter", 1},
    {"a", "maxnondecreasing", 1}, {"f", "factor", 1}, {"e", "epsilon", 1}, {"m", "matching", 1},
    {"u", "subopt", 1}, {"b", "balance", 1}, {"c", "seqscale", 1}, {"p", "probscoremode", 1},
    {"x", "gapextend", 1}, {"y", "gapopen", 1}, {"s", "scorematrix", 1}, {"g", "stepmode", 1}};

void usage()
{
    std::fprintf(stderr,
        "lara_b200 - Lagrangian Relaxed Alignment for RNA structures (B200 GPU implementation of LaRA 2)\n"
        "    lara_b200 -i inFile [parameters]\n"
        "    lara_b200 -d dpFile -d dpFile [-d ...] [parameters]\n"
        "  -v INT  verbosity 0..3            -j INT  number of GPUs (0 = all)\n"
        "  -i IN   input file (.ebpseq .dbn) -r IN   reference input file   -d IN  RNAfold dot plot (repeatable)\n"
        "  -w OUT  output file (stdout)      -l MIN MAX  T-Coffee score range    -o lib|pairs|fasta\n"
        "  -n INT  iterations (500)          -a INT  non-decreasing iterations (50)\n"
        "  -f F    step size factor (1.0)    -e F    epsilon (0.01, unused)  -m INT greedy lookahead (5)\n"
        "  -u F    subopt (40.0)             -b F    balance (1.0)           -c F   seqscale (0.0)\n"
        "  -p INT  prob. score mode (0)      -x F    gap extend (-2.0)       -y F   gap open (-6.0)\n"
        "  -s FILE score matrix (Ribosum65N) -g INT  step arithmetic: 0 = default scalar build, 1 = SIMD build\n");
}

bool parse_int(char const * s, long & v)
{
    char * end = nullptr;
    v = std::strtol(s, &end, 10);
    return end && *end == 0 && end != s;
}
bool parse_float(char const * s, float & v)
{
    char * end = nullptr;
    v = std::strtof(s, &end);
    return end && *end == 0 && end != s;
}

// SeqAn score matrix file (SeqAn:score/score_matrix_io.h): '#' comments, a header row of letters, then
// one row per letter: letter followed by values.  Cells not given keep 0.
bool load_matrix(char const * path, float * tab)
{
    std::ifstream f(path);
    if (!f.is_open()) return false;
    auto code = [](char c) -> int {
        switch (c) { case 'A': case 'a': return 0; case 'C': case 'c': return 1; case 'G': case 'g': return 2;
                     case 'U': case 'u': case 'T': case 't': return 3; case 'N': case 'n': return 4; default: return -1; }
    };
    for (int i = 0; i < 25; ++i) tab[i] = 0.f;
    std::string line;
    std::vector<int> cols;
    while (std::getline(f, line))
    {
        if (line.empty() || line[0] == '#') continue;
        std::istringstream iss(line);
        std::string tok;
        if (cols.empty())
        {
            while (iss >> tok) cols.push_back(code(tok[0]));
            continue;
        }
        if (!(iss >> tok)) continue;
        int r = code(tok[0]);
        for (size_t k = 0; k < cols.size(); ++k)
        {
            float v;
            if (!(iss >> v)) break;
            if (r >= 0 && cols[k] >= 0) tab[r * 5 + cols[k]] = v;
        }
    }
    return !cols.empty();
}

} // namespace

extern "C" int lara_parse_args(int argc, char const * const * argv, lara_params * P, lara_cli_options * O)
{
    if (!P || !O) return 1;
    lara_params_default(P);
    std::memset(O, 0, sizeof(*O));
    O->threads = 1;
    O->epsilon = 0.01f;
    std::strcpy(O->out_format, "lib");
    float gapOpen = -6.0f, gapExtend = -2.0f;
    std::vector<std::string> dotplots;
    for (int k = 1; k < argc; ++k)
    {
        std::string a = argv[k];
        if (a == "-h" || a == "--help") { usage(); return 0; }
        if (a == "--version") { std::printf("lara_b200 version: 2.0.1 (ABI %d)\n", LARA_ABI_VERSION); return 0; }
        OptDef const * def = nullptr;
        std::string inlineVal;
        bool hasInline = false;
        if (a.size() >= 3 && a[0] == '-' && a[1] == '-')
        {
            std::string name = a.substr(2);
            size_t eq = name.find('=');
            if (eq != std::string::npos) { inlineVal = name.substr(eq + 1); name = name.substr(0, eq); hasInline = true; }
            for (OptDef const & d : OPTS) if (name == d.l) def = &d;
        }
        else if (a.size() == 2 && a[0] == '-')
        {
            for (OptDef const & d : OPTS) if (a[1] == d.s[0]) def = &d;
        }
        if (!def)
        {
            std::fprintf(stderr, "lara_b200: unknown option or stray argument '%s'\n", a.c_str());
            return 1;
        }
        std::vector<std::string> vals;
        if (hasInline) vals.push_back(inlineVal);
        while ((int)vals.size() < def->nargs)
        {
            if (k + 1 >= argc) { std::fprintf(stderr, "lara_b200: option -%s needs %d value(s)\n", def->s, def->nargs); return 1; }
            vals.push_back(argv[++k]);
        }
        long iv = 0; float fv = 0.f;
        auto needInt = [&](long lo, long hi) {
            if (!parse_int(vals[0].c_str(), iv) || iv < lo || iv > hi)
            { std::fprintf(stderr, "lara_b200: the value '%s' of -%s is not valid\n", vals[0].c_str(), def->s); return false; }
            return true; };
        auto needFloat = [&](float lo) {
            if (!parse_float(vals[0].c_str(), fv) || fv < lo)
            { std::fprintf(stderr, "lara_b200: the value '%s' of -%s is not valid\n", vals[0].c_str(), def->s); return false; }
            return true; };
        float const NEG = -3.4e38f;
        switch (def->s[0])
        {
            case 'v': if (!needInt(0, 3)) return 1; O->verbose = (int32_t)iv; break;
            case 'j': if (!needInt(0, 1 << 20)) return 1; O->threads = (int32_t)iv; break;
            case 'i': std::snprintf(O->in_file, sizeof(O->in_file), "%s", vals[0].c_str()); break;
            case 'r': std::snprintf(O->ref_file, sizeof(O->ref_file), "%s", vals[0].c_str()); break;
            case 'd': dotplots.push_back(vals[0]); break;
            case 'w': std::snprintf(O->out_file, sizeof(O->out_file), "%s", vals[0].c_str()); break;
            case 'l':
            {
                long lo, hi;
                if (!parse_int(vals[0].c_str(), lo) || !parse_int(vals[1].c_str(), hi))
                { std::fprintf(stderr, "lara_b200: -l needs two integers\n"); return 1; }
                P->lib_score_min = (uint32_t)lo; P->lib_score_max = (uint32_t)hi; P->lib_score_linear = 1;
                break;
            }
            case 'o':
                if (vals[0] != "lib" && vals[0] != "pairs" && vals[0] != "fasta")
                { std::fprintf(stderr, "lara_b200: the value '%s' of -o is not in {lib, pairs, fasta}\n", vals[0].c_str()); return 1; }
                std::snprintf(O->out_format, sizeof(O->out_format), "%s", vals[0].c_str());
                break;
            case 'n': if (!needInt(1, 1 << 30)) return 1; P->num_iterations = (int32_t)iv; break;
            case 'a': if (!needInt(0, 1 << 30)) return 1; P->max_nondecreasing = (int32_t)iv; break;
            case 'f': if (!needFloat(NEG)) return 1; P->step_size_factor = fv; break;
            case 'e': if (!needFloat(NEG)) return 1; O->epsilon = fv; break;
            case 'm': if (!needInt(0, 1 << 20)) return 1; P->matching = (int32_t)iv; break;
            case 'u': if (!needFloat(NEG)) return 1; P->suboptimal_diff = fv; break;
            case 'b': if (!needFloat(0.f)) return 1; P->balance = fv; break;
            case 'c': if (!needFloat(0.f)) return 1; P->sequence_scale = fv; break;
            case 'p': if (!needInt(0, 1)) return 1; O->structure_scoring = (int32_t)iv; break;
            case 'x': if (!needFloat(NEG)) return 1; gapExtend = fv; break;
            case 'y': if (!needFloat(NEG)) return 1; gapOpen = fv; break;
            case 's': std::snprintf(O->score_matrix_file, sizeof(O->score_matrix_file), "%s", vals[0].c_str()); break;
            case 'g': if (!needInt(0, 1)) return 1; P->step_mode = (int32_t)iv; break;
        }
    }
    if (O->threads == 0)
    {
        unsigned n = std::thread::hardware_concurrency();
        O->threads = n ? (int32_t)n : 1;
    }
    if (O->in_file[0] == 0 && dotplots.empty()) { usage(); return 1; }           // src/parameters.hpp:273-277
    P->gap_open = (int32_t)(gapOpen * LARA_FACTOR2INT);                            // :303-304
    P->gap_extend = (int32_t)(gapExtend * LARA_FACTOR2INT);
    if (O->score_matrix_file[0])
    {
        float tab[25];
        if (!load_matrix(O->score_matrix_file, tab))
        {
            std::fprintf(stderr, "Error: Matrix file could not be opened: %s\n", O->score_matrix_file);
            return 1;
        }
        for (int i = 0; i < 25; ++i) P->score_matrix[i] = (int32_t)(tab[i] * LARA_FACTOR2INT);
    }
    O->n_dotplots = (int32_t)dotplots.size();
    if (!dotplots.empty())
    {
        O->dotplot_files = (char **)std::malloc(sizeof(char *) * dotplots.size());
        for (size_t k = 0; k < dotplots.size(); ++k) O->dotplot_files[k] = strdup(dotplots[k].c_str());
    }
    return 2;
}
