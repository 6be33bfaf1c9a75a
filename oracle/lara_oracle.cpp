/*
 * lara_oracle.cpp -- TEST INFRASTRUCTURE ONLY (see lara_oracle.h).
 *
 * CPU restatement of the reference hot path with flat arrays.  Every function cites
 * the reference lines it restates; "SeqAn:" = /root/reference/include/seqan/include/seqan/.
 * Compile with -ffp-contract=off: the reference's default build has no FMA
 * (CMakeLists.txt has no -march flag), and the float expressions below must round
 * the same way (step size, dual update, struct score, scale).
 */
#include "lara_oracle.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <tuple>
#include <vector>

namespace
{

typedef int32_t ScoreType;                                  // src/data_types.hpp:66
ScoreType const INF = LARA_INFINITY;                        // src/data_types.hpp:77
ScoreType const PSSM_INIT = -1431655764;                    // src/score.hpp:92  lowest()/3*2
ScoreType const DP_NEG = INT32_MIN / 2;                     // SeqAn:align/dp_cell.h:143-146

// SeqAn:align/dp_profile.h:125-134
enum : uint8_t
{
    T_NONE = 0, T_DIAG = 1, T_H = 2, T_V = 4, T_HOPEN = 8, T_VOPEN = 16, T_MAXH = 32, T_MAXV = 64
};

// ---------------------------------------------------------------------------------------------
// Edge filter: src/edge_filter.hpp:52-129 (PairwiseGotoh) and :131-158 (generateEdges)
// ---------------------------------------------------------------------------------------------
struct PrefixGotoh
{
    int la, lb;
    std::vector<ScoreType> M, H, V;
    inline size_t at(int a, int b) const { return (size_t)(lb + 1) * a + b; }       // :63-66

    // reversed == the ModReverse views of :137-140
    PrefixGotoh(uint8_t const * A, int la_, uint8_t const * B, int lb_, int32_t const * tab,
                ScoreType go, ScoreType ge, bool reversed) : la(la_), lb(lb_)
    {
        size_t n = (size_t)(la + 1) * (lb + 1);
        M.resize(n); H.resize(n); V.resize(n);
        M[at(0, 0)] = 0; H[at(0, 0)] = -INF; V[at(0, 0)] = -INF;                   // :80-82
        for (int a = 0; a < la; ++a)                                               // :85-90
        {
            ScoreType v = (ScoreType)((uint32_t)go + (uint32_t)ge * (uint32_t)a);
            M[at(a + 1, 0)] = v; H[at(a + 1, 0)] = -INF; V[at(a + 1, 0)] = v;
        }
        for (int b = 0; b < lb; ++b)                                               // :92-97
        {
            ScoreType v = (ScoreType)((uint32_t)go + (uint32_t)ge * (uint32_t)b);
            M[at(0, b + 1)] = v; H[at(0, b + 1)] = v; V[at(0, b + 1)] = -INF;
        }
        for (int a = 0; a < la; ++a)                                               // :100-116
        {
            int ca = reversed ? A[la - 1 - a] : A[a];
            for (int b = 0; b < lb; ++b)
            {
                int cb = reversed ? B[lb - 1 - b] : B[b];
                M[at(a + 1, b + 1)] = std::max({M[at(a, b)], H[at(a, b)], V[at(a, b)]}) + tab[ca * 5 + cb];
                H[at(a + 1, b + 1)] = std::max({M[at(a + 1, b)] + go, H[at(a + 1, b)] + ge, V[at(a + 1, b)] + go});
                V[at(a + 1, b + 1)] = std::max({M[at(a, b + 1)] + go, H[at(a, b + 1)] + go, V[at(a, b + 1)] + ge});
            }
        }
    }
    inline ScoreType prefix(int a, int b) const                                    // :119-123
    {
        return std::max({M[at(a, b)], H[at(a, b)], V[at(a, b)]});
    }
};

// returns avgId (:157); active is la*lb bytes
float generate_edges(uint8_t const * A, int la, uint8_t const * B, int lb, int32_t const * tab,
                     ScoreType go, ScoreType ge, ScoreType subopt, uint8_t * active, ScoreType * optOut)
{
    PrefixGotoh fwd(A, la, B, lb, tab, go, ge, false);
    PrefixGotoh bwd(A, la, B, lb, tab, go, ge, true);
    ScoreType opt = fwd.prefix(la, lb);
    ScoreType threshold = opt - subopt;                                            // :142
    for (int a = 0; a < la; ++a)
        for (int b = 0; b < lb; ++b)
            active[(size_t)lb * a + b] =
                (fwd.prefix(a, b) + tab[A[a] * 5 + B[b]] + bwd.prefix(la - a - 1, lb - b - 1) >= threshold) ? 1 : 0;
    if (optOut)
        *optOut = opt;
    return (float)opt / LARA_FACTOR2INT / (float)std::max(la, lb);                 // :157
}

// ---------------------------------------------------------------------------------------------
// SeqAn affine-gap global alignment, SingleTrace + GapsLeft, unbanded
//   fill:      SeqAn:align/dp_formula_affine.h:62-125 (all), :245-281 (row 0), :291-325 (column 0)
//   tie rule:  SeqAn:align/dp_formula.h:153-163  (left operand wins ties)
//   end cell:  SeqAn:align/dp_algorithm_impl.h:1168-1185 ; start dir: dp_traceback_impl.h:456-483
//   walk:      SeqAn:align/dp_traceback_impl.h:190-418, 533-545
// pssm index = lb*a + b (src/score.hpp:94-99 with dim = lb).  H = sequence A (columns), V = B (rows).
// Output: per alignment column (left to right) an op: 0 = match (A,B), 1 = gap in A (consumes B,
// "vertical"), 2 = gap in B (consumes A, "horizontal").
// ---------------------------------------------------------------------------------------------
// developer statistics (LARA_ORACLE_STATS=1): largest |S| of an active cell and largest |score| seen by the DP
static bool g_dbgRange = std::getenv("LARA_ORACLE_STATS") != nullptr;
static int64_t g_dbgMaxS = 0, g_dbgMaxP = 0;

ScoreType gotoh_fill(int la, int lb, ScoreType const * pssm, ScoreType go, ScoreType ge,
                     std::vector<uint8_t> & trace)
{
    size_t const stride = (size_t)lb + 1;
    trace.assign((size_t)(la + 1) * stride, 0);
    std::vector<ScoreType> S(lb + 1), H(lb + 1);
    // column 0
    S[0] = 0; H[0] = DP_NEG;
    ScoreType V = DP_NEG;
    trace[0] = T_NONE;
    for (int r = 1; r <= lb; ++r)
    {
        ScoreType ve = V + ge, vo = S[r - 1] + go;
        uint8_t t;
        if (ve < vo) { V = vo; t = T_VOPEN; } else { V = ve; t = T_V; }
        S[r] = V; H[r] = DP_NEG;
        trace[r] = t | T_MAXV;
    }
    ScoreType lastV = V, lastH = H[lb];
    for (int c = 1; c <= la; ++c)
    {
        uint8_t * tcol = &trace[(size_t)c * stride];
        // row 0: horizontal only
        ScoreType diag = S[0];
        {
            ScoreType he = H[0] + ge, ho = S[0] + go;
            uint8_t t;
            if (he < ho) { H[0] = ho; t = T_HOPEN; } else { H[0] = he; t = T_H; }
            S[0] = H[0];
            tcol[0] = t | T_MAXH;
        }
        V = DP_NEG;
        ScoreType const * prow = pssm + (size_t)(c - 1) * lb;
        for (int r = 1; r <= lb; ++r)
        {
            ScoreType d = diag + prow[r - 1];
            diag = S[r];
            uint8_t t;
            ScoreType he = H[r] + ge, ho = S[r] + go;
            ScoreType h;
            if (he < ho) { h = ho; t = T_HOPEN; } else { h = he; t = T_H; }
            ScoreType ve = V + ge, vo = S[r - 1] + go;
            if (ve < vo) { V = vo; t |= T_VOPEN; } else { V = ve; t |= T_V; }
            ScoreType g; uint8_t t2;
            if (V < h) { g = h; t2 = T_MAXH; } else { g = V; t2 = T_MAXV; }
            ScoreType s;
            if (d < g) { s = g; t = t2 | t; } else { s = d; t = T_DIAG | t; }
            S[r] = s; H[r] = h;
            tcol[r] = t;
            if (g_dbgRange) { int64_t const a = std::llabs((long long)s); if (a > g_dbgMaxS && prow[r - 1] > DP_NEG) g_dbgMaxS = a;
                              int64_t const b = std::llabs((long long)prow[r - 1]); if (prow[r - 1] > PSSM_INIT && b > g_dbgMaxP) g_dbgMaxP = b; }
        }
        lastV = V; lastH = H[lb];
    }
    // end-cell correction (SingleTrace)
    uint8_t & tv = trace[(size_t)la * stride + lb];
    if (lastV == S[lb]) { tv &= (uint8_t)~T_DIAG; tv |= T_MAXV; }
    else if (lastH == S[lb]) { tv &= (uint8_t)~T_DIAG; tv |= T_MAXH; }
    return S[lb];
}

// ops are produced right-to-left and reversed at the end
void gotoh_traceback(int la, int lb, std::vector<uint8_t> const & trace, std::vector<uint8_t> & ops)
{
    size_t const stride = (size_t)lb + 1;
    ops.clear();
    int c = la, r = lb;
    uint8_t tv = trace[(size_t)c * stride + r];
    // _retrieveInitialTraceDirection with PreferGapsAtEnd_ (affine)
    if (tv & T_MAXV) tv &= (T_V | T_VOPEN | T_MAXV);
    else if (tv & T_MAXH) tv &= (T_H | T_HOPEN | T_MAXH);
    auto at = [&](int cc, int rr) { return trace[(size_t)cc * stride + rr]; };
    while (c > 0 && r > 0 && tv != T_NONE)
    {
        if (tv & T_DIAG)
        {
            --c; --r; tv = at(c, r); ops.push_back(0);
        }
        else if ((tv & T_MAXV) && (tv & T_V))
        {
            while ((!(tv & T_VOPEN) || (tv & T_V)) && r != 1)
            {
                --r; tv = at(c, r); ops.push_back(1);
            }
            --r; tv = at(c, r); ops.push_back(1);
        }
        else if ((tv & T_MAXV) && (tv & T_VOPEN))
        {
            --r; tv = at(c, r); ops.push_back(1);
        }
        else if ((tv & T_MAXH) && (tv & T_H))
        {
            while ((!(tv & T_HOPEN) || (tv & T_H)) && c != 1)
            {
                --c; tv = at(c, r); ops.push_back(2);
            }
            --c; tv = at(c, r); ops.push_back(2);
        }
        else if ((tv & T_MAXH) && (tv & T_HOPEN))
        {
            --c; tv = at(c, r); ops.push_back(2);
        }
        else
            break; // undefined trace value (SEQAN_ASSERT_FAIL in debug builds)
    }
    for (; r > 0; --r) ops.push_back(1);   // leading vertical gaps   (:541-542)
    for (; c > 0; --c) ops.push_back(2);   // leading horizontal gaps (:543-544)
    std::reverse(ops.begin(), ops.end());
}

// ---------------------------------------------------------------------------------------------
// Linear gap costs.  SeqAn's globalAlignment switches to Needleman-Wunsch when gap open == gap extend
// (SeqAn:align/global_alignment_unbanded.h:775-778): one score matrix, recursion
// SeqAn:align/dp_formula_linear.h:62-99 (vertical wins the tie against horizontal, the diagonal wins the tie against the
// better gap), no end-cell correction (dp_algorithm_impl.h: _correctTraceValue does nothing for LinearGaps), traceback one
// cell at a time (dp_traceback_impl.h:190-262 with TGapCosts = LinearGaps, :400-418).
// ---------------------------------------------------------------------------------------------
ScoreType nw_fill(int la, int lb, ScoreType const * pssm, ScoreType g, std::vector<uint8_t> & trace)
{
    size_t const stride = (size_t)lb + 1;
    trace.assign((size_t)(la + 1) * stride, 0);
    std::vector<ScoreType> S(lb + 1);
    S[0] = 0;
    trace[0] = T_NONE;
    for (int r = 1; r <= lb; ++r) { S[r] = S[r - 1] + g; trace[r] = T_V | T_MAXV; }
    for (int c = 1; c <= la; ++c)
    {
        uint8_t * tcol = &trace[(size_t)c * stride];
        ScoreType diag = S[0];
        S[0] = S[0] + g;
        tcol[0] = T_H | T_MAXH;
        ScoreType const * prow = pssm + (size_t)(c - 1) * lb;
        for (int r = 1; r <= lb; ++r)
        {
            ScoreType const d = diag + prow[r - 1];
            diag = S[r];
            ScoreType const v = S[r - 1] + g, h = S[r] + g;
            ScoreType gm; uint8_t t;
            if (v < h) { gm = h; t = T_H | T_MAXH; } else { gm = v; t = T_V | T_MAXV; }
            if (d < gm) { S[r] = gm; tcol[r] = t; } else { S[r] = d; tcol[r] = T_DIAG; }
        }
    }
    return S[lb];
}

void nw_traceback(int la, int lb, std::vector<uint8_t> const & trace, std::vector<uint8_t> & ops)
{
    size_t const stride = (size_t)lb + 1;
    ops.clear();
    int c = la, r = lb;
    while (c > 0 && r > 0)
    {
        uint8_t const tv = trace[(size_t)c * stride + r];
        if (tv & T_DIAG) { --c; --r; ops.push_back(0); }
        else if ((tv & T_MAXV) && (tv & T_V)) { --r; ops.push_back(1); }
        else if ((tv & T_MAXH) && (tv & T_H)) { --c; ops.push_back(2); }
        else break;
    }
    for (; r > 0; --r) ops.push_back(1);
    for (; c > 0; --c) ops.push_back(2);
    std::reverse(ops.begin(), ops.end());
}

// Lagrange::evaluateLines, src/lagrange.hpp:147-221
ScoreType evaluate_lines(std::vector<uint8_t> const & ops, ScoreType go, ScoreType ge,
                         std::vector<int32_t> & la_, std::vector<int32_t> & lb_)
{
    la_.clear(); lb_.clear();
    bool open0 = false, open1 = false;
    int pa = 0, pb = 0;
    ScoreType gap = 0;
    for (uint8_t op : ops)
    {
        if (op == 1) { gap += open0 ? ge : go; open0 = true; ++pb; } else open0 = false;   // gap in A
        if (op == 2) { gap += open1 ? ge : go; open1 = true; ++pa; } else open1 = false;   // gap in B
        if (op == 0) { la_.push_back(pa); lb_.push_back(pb); ++pa; ++pb; }
    }
    return gap;
}


// ---------------------------------------------------------------------------------------------
// Matching::computeGreedyMatching + solveConflicts, src/matching.hpp:99-192.
// Interaction = (-2*weight, lineL, lineR) ordered as a std::set<std::tuple<...>> orders it.
// Iterators of the reference become indices into the sorted vector.
// ---------------------------------------------------------------------------------------------
typedef std::tuple<ScoreType, int64_t, int64_t> Interaction;
typedef std::pair<int, int> EdgeConflict;

std::set<int> solve_conflicts(std::vector<Interaction> const & q, ScoreType & weight,
                              std::vector<EdgeConflict> const & conflicts)
{
    if (conflicts.empty())
        return std::set<int>();
    auto edge_cmp = [&q](int const & a, int const & b) { return q[a] >= q[b]; };      // :104
    int edgeS = std::min(conflicts.front().first, conflicts.front().second, edge_cmp); // :105
    int edgeL = std::max(conflicts.front().first, conflicts.front().second, edge_cmp); // :106
    ScoreType weightS = -std::get<0>(q[edgeS]);
    ScoreType weightL = -std::get<0>(q[edgeL]);
    if (conflicts.size() == 1)
    {
        weight += weightS;
        return std::set<int>{edgeS};
    }
    std::vector<EdgeConflict> remainingS;
    for (EdgeConflict const & c : conflicts)
        if (edgeS != c.first && edgeS != c.second)
            remainingS.push_back(c);
    std::set<int> eliminateS = solve_conflicts(q, weightS, remainingS);
    if (weightS > weightL)                                                             // :121
    {
        std::vector<EdgeConflict> remainingL;
        for (EdgeConflict const & c : conflicts)
            if (edgeL != c.first && edgeL != c.second)
                remainingL.push_back(c);
        if (remainingS != remainingL)                                                  // :127
        {
            std::set<int> eliminateL = solve_conflicts(q, weightL, remainingL);
            if (weightS > weightL)
            {
                eliminateL.insert(edgeL);
                weight += weightL;
                return eliminateL;
            }
        }
    }
    eliminateS.insert(edgeS);
    weight += weightS;
    return eliminateS;
}

// q must be sorted ascending and duplicate-free (== iteration order of the std::set).
// matchedOf: callback marking both endpoints.
template <typename TMark, typename TIsMatched>
ScoreType greedy_matching(std::vector<Interaction> const & q, size_t lookahead, TMark mark, TIsMatched isMatched)
{
    ScoreType score = 0;
    if (lookahead > q.size())                                                          // :153-156
        lookahead = q.size();
    else if (lookahead == 0)
        lookahead = 5;
    std::vector<int> selection;
    size_t it = 0;
    while (it != q.size())
    {
        for (selection.clear(); selection.size() < lookahead && it != q.size(); ++it)  // :165-167
            if (!isMatched(std::get<1>(q[it])) && !isMatched(std::get<2>(q[it])))
                selection.push_back((int)it);
        std::vector<EdgeConflict> conflicts;                                           // :170-174
        for (size_t x = 0; x < selection.size(); ++x)
            for (size_t y = x + 1; y < selection.size(); ++y)
            {
                Interaction const & a = q[selection[x]];
                Interaction const & b = q[selection[y]];
                if (std::get<1>(a) == std::get<1>(b) || std::get<1>(a) == std::get<2>(b) ||
                    std::get<2>(a) == std::get<1>(b) || std::get<2>(a) == std::get<2>(b))
                    conflicts.emplace_back(selection[x], selection[y]);
            }
        ScoreType weight = 0;
        std::set<int> eliminate = solve_conflicts(q, weight, conflicts);
        for (int s : selection)                                                        // :181-189
            if (eliminate.count(s) == 0)
            {
                mark(std::get<1>(q[s]), std::get<2>(q[s]));
                score += -std::get<0>(q[s]);
            }
    }
    return score;
}

// ---------------------------------------------------------------------------------------------
// Lagrange (src/lagrange.hpp:64-488) with flat arrays.
//   * active alignment edges get compact ids in ascending cell order, so comparing compact ids
//     == comparing the reference's edge ids (cell = lb*a + b);
//   * priorityQ[l] is not materialised: its top is max_m key(l,m) (pv[l]) and its iteration
//     order among tied tops is ascending partner id, where key(l,m) = seqScore(l)+struct(l,m)+dual(l->m)
//     and the self entry has key seqScore(l), id l  (:271, :294, :317-321).
// ---------------------------------------------------------------------------------------------
struct SeqView
{
    uint8_t const * s; int len;
    int64_t base;                       // global index of position 0 (into bp_off)
    int32_t const * bp_partner; int32_t const * bp_weight; int64_t const * bp_off;
};

struct Lagrange
{
    int la, lb;
    SeqView A, B;
    lara_params const * P;
    std::vector<uint8_t> active;
    std::vector<int32_t> eid;           // cell -> compact edge id or -1
    std::vector<int32_t> cellOf;        // compact id -> cell
    std::vector<ScoreType> seqScore;    // getSeqScore, :223-226
    std::vector<int64_t> ioff;          // CSR rows of interactions, sorted by partner id
    std::vector<int32_t> part, rev;
    std::vector<ScoreType> sc, dual;
    std::vector<float> sub;             // subgradient
    std::vector<ScoreType> pv;          // -priorityQ[l].begin()->first
    std::vector<ScoreType> pssm;        // dense, index lb*a+b, PSSM_INIT elsewhere (src/score.hpp:68-79)
    float scale;
    ScoreType bestScore;
    std::vector<int32_t> bestLines;     // edge ids (compact)
    std::vector<uint8_t> bestMatched;   // edgeMatching.count(idx) per best line
    std::vector<uint8_t> inSol;
    std::vector<int64_t> list;          // subgradientIndices
    // developer statistics (LARA_ORACLE_STATS=1): sums over the iterations of one pair
    int64_t stLines = 0, stWork = 0, stItems = 0, stList = 0, stCand = 0, stRescan = 0;

    inline ScoreType key(int32_t e, int64_t i) const { return seqScore[e] + sc[i] + dual[i]; }

    ScoreType rowTop(int32_t e) const
    {
        ScoreType t = seqScore[e];
        for (int64_t i = ioff[e]; i < ioff[e + 1]; ++i)
            t = std::max(t, key(e, i));
        return t;
    }

    Lagrange(SeqView const & a, SeqView const & b, lara_params const * p) : la(a.len), lb(b.len), A(a), B(b), P(p)
    {
        size_t C = (size_t)la * lb;
        active.resize(C);
        float avSeqId = generate_edges(A.s, la, B.s, lb, P->score_matrix, P->gap_open, P->gap_extend,
                                       (ScoreType)(P->suboptimal_diff * LARA_FACTOR2INT), active.data(), nullptr); // :250-251
        scale = P->sequence_scale * avSeqId + 1;                                        // :252
        eid.assign(C, -1);
        for (size_t c = 0; c < C; ++c)
            if (active[c]) { eid[c] = (int32_t)cellOf.size(); cellOf.push_back((int32_t)c); }
        int32_t E = (int32_t)cellOf.size();
        seqScore.resize(E);
        ioff.assign(E + 1, 0);
        // adjacency sorted by partner so that rows come out sorted by partner id
        std::vector<std::pair<int32_t, int32_t>> ha, tb;
        for (int32_t e = 0; e < E; ++e)
        {
            int a0 = cellOf[e] / lb, b0 = cellOf[e] % lb;
            seqScore[e] = (ScoreType)(scale * (float)P->score_matrix[A.s[a0] * 5 + B.s[b0]]);   // :225
            ha.clear(); tb.clear();
            for (int64_t k = A.bp_off[A.base + a0]; k < A.bp_off[A.base + a0 + 1]; ++k)
                ha.emplace_back(A.bp_partner[k], A.bp_weight[k]);
            for (int64_t k = B.bp_off[B.base + b0]; k < B.bp_off[B.base + b0 + 1]; ++k)
                tb.emplace_back(B.bp_partner[k], B.bp_weight[k]);
            std::sort(ha.begin(), ha.end());
            std::sort(tb.begin(), tb.end());
            for (auto const & h : ha)
                for (auto const & t : tb)
                {
                    size_t m = (size_t)lb * h.first + t.first;
                    bool nonCrossing = (a0 < h.first && b0 < t.first) || (h.first < a0 && t.first < b0);     // :101-106
                    if (active[m] && nonCrossing)
                    {
                        ScoreType structScore = (ScoreType)(P->balance * (float)(h.second + t.second) / 2 * LARA_FACTOR2INT); // :289
                        part.push_back(eid[m]);
                        sc.push_back(structScore);
                    }
                }
            ioff[e + 1] = (int64_t)part.size();
        }
        int64_t D = (int64_t)part.size();
        rev.assign(D, -1);
        for (int32_t e = 0; e < E; ++e)
            for (int64_t i = ioff[e]; i < ioff[e + 1]; ++i)
            {
                int32_t m = part[i];
                for (int64_t j = ioff[m]; j < ioff[m + 1]; ++j)
                    if (part[j] == e) { rev[i] = j; break; }
            }
        dual.assign(D, 0);
        sub.assign(D, 0.0f);
        pv.resize(E);
        pssm.assign(C, PSSM_INIT);
        for (int32_t e = 0; e < E; ++e)                                                 // :305-309
        {
            pv[e] = rowTop(e);
            pssm[cellOf[e]] = pv[e];
        }
        bestScore = -INF;                                                              // :238
        inSol.assign(E, 0);
    }

    int64_t dimension() const { return (int64_t)part.size(); }

    // valid_solution, :325-423.  ops = the DP's alignment.
    ScoreType valid_solution(std::vector<uint8_t> const & ops)
    {
        std::vector<int32_t> pa, pb;
        ScoreType gapScore = evaluate_lines(ops, P->gap_open, P->gap_extend, pa, pb);
        std::vector<int32_t> cur(pa.size());
        for (size_t k = 0; k < pa.size(); ++k)
        {
            cur[k] = eid[(size_t)lb * pa[k] + pb[k]];
            if (cur[k] < 0) abort();       // "Alignment match where no alignment edge is defined!" (:338)
            inSol[cur[k]] = 1;
        }
        list.clear();
        stLines += (int64_t)cur.size();
        for (int32_t idx : cur) { stWork += ioff[idx + 1] > ioff[idx]; stItems += ioff[idx + 1] - ioff[idx]; }
        for (int32_t idx : cur)                                                        // :344-374
        {
            ScoreType top = pv[idx];
            bool selfDone = false;
            auto visit = [&](int32_t maxPE, int64_t i)
            {
                bool foundCycle = false;
                if (inSol[maxPE])
                {
                    if (maxPE == idx) foundCycle = true;                       // self entry is a tied top of Q[idx]
                    else foundCycle = (key(maxPE, rev[i]) == pv[maxPE]);        // idx among the tied tops of Q[maxPE]
                }
                if (!foundCycle)
                {
                    sub[i] = 1.0f;  list.push_back(i);
                    sub[rev[i]] = -1.0f; list.push_back(rev[i]);
                }
            };
            for (int64_t i = ioff[idx]; i < ioff[idx + 1]; ++i)
            {
                if (!selfDone && part[i] > idx) { selfDone = true; if (seqScore[idx] == top) visit(idx, -1); }
                if (key(idx, i) == top) visit(part[i], i);
            }
            if (!selfDone && seqScore[idx] == top) visit(idx, -1);
        }
        ScoreType lowerBound = 0;
        for (int32_t idx : cur)                                                        // :376-378
            lowerBound += seqScore[idx];
        std::vector<uint8_t> matched(cur.size(), 0);
        // position of an in-solution edge in cur: via its A position
        std::vector<int32_t> lineOfA(la, -1);
        for (size_t k = 0; k < pa.size(); ++k) lineOfA[pa[k]] = (int32_t)k;
        if (!list.empty())                                                             // :381-396
        {
            std::vector<Interaction> q;
            for (int32_t line : cur)
                for (int64_t i = ioff[line]; i < ioff[line + 1]; ++i)
                    if (inSol[part[i]] && line < part[i])
                        q.emplace_back(-2 * sc[i], (int64_t)cellOf[line], (int64_t)cellOf[part[i]]);
            std::sort(q.begin(), q.end());
            q.erase(std::unique(q.begin(), q.end()), q.end());
            stCand += (int64_t)q.size();
            auto lineOfCell = [&](int64_t cell) { return lineOfA[cell / lb]; };
            lowerBound += greedy_matching(q, (size_t)P->matching,
                [&](int64_t l, int64_t r) { matched[lineOfCell(l)] = 1; matched[lineOfCell(r)] = 1; },
                [&](int64_t v) { return matched[lineOfCell(v)] != 0; });
        }
        else                                                                           // :397-409
        {
            for (size_t k = 0; k < cur.size(); ++k)
            {
                int32_t idx = cur[k];
                // begin() of the queue: top key, smallest partner id among ties
                ScoreType top = pv[idx];
                int32_t maxPE = -1; int64_t at = -1;
                for (int64_t i = ioff[idx]; i < ioff[idx + 1] && maxPE < 0; ++i)
                {
                    if (part[i] > idx && seqScore[idx] == top) { maxPE = idx; break; }
                    if (key(idx, i) == top) { maxPE = part[i]; at = i; }
                }
                if (maxPE < 0) maxPE = idx;
                if (idx != maxPE)
                {
                    lowerBound += sc[at];
                    matched[k] = 1;
                    int32_t k2 = lineOfA[cellOf[maxPE] / lb];
                    if (k2 >= 0 && cur[k2] == maxPE) matched[k2] = 1;
                }
            }
        }
        for (int32_t idx : cur) inSol[idx] = 0;
        ScoreType primal = lowerBound + gapScore;                                       // :412
        if (primal > bestScore)                                                        // :416-421
        {
            bestScore = primal;
            bestLines = cur;
            bestMatched = matched;
        }
        return primal;
    }

    // updateScores, :312-323
    void update_scores()
    {
        stList += (int64_t)list.size();
        for (int64_t d : list)
        {
            // row owner of entry d: the reverse entry's partner
            int32_t l = part[rev[d]];
            stRescan += ioff[l + 1] - ioff[l];
            pv[l] = rowTop(l);
            pssm[cellOf[l]] = pv[l];
        }
    }
};

struct PairOut
{
    int32_t idxLow, idxHigh, bestLower, bestUpper, iterations;
    std::vector<int32_t> posLow, posHigh;
    std::vector<uint32_t> weight;
    int64_t E, D;
};

// One pair, first = the longer sequence (src/subgradient_solver.hpp:115-120), run to completion:
// the loop body of src/subgradient_solver.hpp:219-333 (step_mode 0) or
// src/subgradient_solver_simd.hpp:231-403 (step_mode 1) for a single slot.
void solve_pair(lara_params const * P, SeqView const & sa, SeqView const & sb, int32_t idxFirst, int32_t idxSecond,
                PairOut & out)
{
    Lagrange lag(sa, sb, P);
    ScoreType bestLower = -INF, bestUpper = INF, currentLower = -INF, currentUpper = INF;
    float stepSizeFactor = P->step_size_factor;
    ScoreType stepFactorInt = (ScoreType)(P->step_size_factor * LARA_FACTOR2INT);      // simd :153
    uint32_t nondecreasing = 0;
    uint32_t remaining = (uint32_t)P->num_iterations;
    uint32_t const maxND = (uint32_t)P->max_nondecreasing;
    std::vector<uint8_t> trace, ops;
    for (;;)
    {
        // the scalar solver lets SeqAn choose (linear gap costs when open == extend); the SIMD solver asks for AffineGaps
        // explicitly (src/subgradient_solver_simd.hpp:257-263)
        if (P->gap_open == P->gap_extend && P->step_mode == LARA_STEP_SCALAR_FLOAT)
        {
            currentUpper = nw_fill(lag.la, lag.lb, lag.pssm.data(), P->gap_extend, trace);
            nw_traceback(lag.la, lag.lb, trace, ops);
        }
        else
        {
            currentUpper = gotoh_fill(lag.la, lag.lb, lag.pssm.data(), P->gap_open, P->gap_extend, trace);
            gotoh_traceback(lag.la, lag.lb, trace, ops);
        }
        currentLower = lag.valid_solution(ops);
        if (currentUpper < bestUpper) { bestUpper = currentUpper; nondecreasing = 0; }
        if (currentLower > bestLower) { bestLower = currentLower; nondecreasing = 0; }
        float stepSize;
        if (P->step_mode == LARA_STEP_SCALAR_FLOAT)
        {
            if (nondecreasing++ >= maxND) { stepSizeFactor /= 2.0f; nondecreasing = 0; }        // :259-263
            stepSize = stepSizeFactor * (float)(bestUpper - bestLower) / (float)lag.list.size(); // :265-266
        }
        else
        {
            nondecreasing += 1;                                                        // simd :303-306
            if (nondecreasing > maxND) { stepFactorInt = stepFactorInt / 2; nondecreasing = 0; }
            ScoreType prod = (ScoreType)((uint32_t)stepFactorInt * (uint32_t)(bestUpper - bestLower)); // :313 wraps
            stepSize = (float)prod / LARA_FACTOR2INT / (float)lag.list.size();          // :330
        }
        for (int64_t si : lag.list)                                                    // :267-271
        {
            lag.dual[si] = (ScoreType)((float)lag.dual[si] - stepSize * lag.sub[si]);
            lag.sub[si] = 0.0f;
        }
        --remaining;
        if (bestUpper == bestLower || remaining == 0u)
            break;
        lag.update_scores();
    }
    if (std::getenv("LARA_ORACLE_STATS"))
    {
        double const it = (double)(P->num_iterations - (int32_t)remaining);
        std::fprintf(stderr, "[oracle] pair %d %d la %d lb %d E %zu D %lld iterations %.0f per iteration: lines %.1f with-interactions %.1f "
                             "interactions-of-lines %.1f list %.1f candidates %.1f rescanned %.1f max|S| %lld max|score| %lld\n", idxFirst, idxSecond, lag.la, lag.lb,
                     lag.cellOf.size(), (long long)lag.dimension(), it, lag.stLines / it, lag.stWork / it, lag.stItems / it,
                     lag.stList / it, lag.stCand / it, lag.stRescan / it, (long long)g_dbgMaxS, (long long)g_dbgMaxP);
    }
    // getStructureLines, src/lagrange.hpp:437-487
    bool swapIdx = idxFirst > idxSecond;
    out.idxLow = swapIdx ? idxSecond : idxFirst;
    out.idxHigh = swapIdx ? idxFirst : idxSecond;
    out.bestLower = bestLower;
    out.bestUpper = bestUpper;
    out.iterations = P->num_iterations - (int32_t)remaining;
    out.E = (int64_t)lag.cellOf.size();
    out.D = lag.dimension();
    ScoreType minScore = 0, maxScore = 0;
    float div = 0.f;
    if (P->lib_score_linear)
    {
        // minmax_element over ALL lenA*lenB queues (:445-448).  Inactive edges have empty sets; reading
        // begin()->first of an empty libstdc++ std::set yields the node count, i.e. 0 (undefined
        // behaviour in the reference, reproduced as observed).
        bool first = true;
        size_t C = (size_t)lag.la * lag.lb;
        for (size_t c = 0; c < C; ++c)
        {
            ScoreType v = lag.eid[c] >= 0 ? lag.pv[lag.eid[c]] : 0;
            if (first) { minScore = maxScore = v; first = false; }
            else { minScore = std::min(minScore, v); maxScore = std::max(maxScore, v); }
        }
        div = 1.f * (float)(P->lib_score_max - P->lib_score_min) / (float)(maxScore - minScore);
    }
    for (size_t k = 0; k < lag.bestLines.size(); ++k)
    {
        int32_t e = lag.bestLines[k];
        int a = lag.cellOf[e] / lag.lb, b = lag.cellOf[e] % lag.lb;
        out.posLow.push_back(swapIdx ? b : a);
        out.posHigh.push_back(swapIdx ? a : b);
        uint32_t w;
        if (P->lib_score_linear)
        {
            uint64_t cnt = lag.bestMatched[k] ? 1u : 0u;
            ScoreType val = (ScoreType)((float)(cnt * (uint64_t)(int64_t)(lag.pv[e] - minScore)) * div);   // :455
            w = P->lib_score_min + (uint32_t)val;
        }
        else
            w = (lag.bestMatched[k] ? 1u : 0u) * 500u + 500u;                          // :477
        out.weight.push_back(w);
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------
// C interface
// ---------------------------------------------------------------------------------------------
namespace
{
struct ResultOwner
{
    std::vector<lara_pair_result> pairs;
    std::vector<int32_t> posLow, posHigh;
    std::vector<uint32_t> weight;
};
} // namespace

extern "C" {

void lara_oracle_params_default(lara_params * p)
{
    // Ribosum65N (src/rna_score_matrices.hpp:3960-3968), order A C G U N; x8192 truncated like
    // src/parameters.hpp:311-314.
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

int lara_oracle_solve_pairs(lara_params const * params, int32_t n_seq,
                            uint8_t const * seq_cat, int64_t const * seq_off,
                            int32_t const * bp_partner, int32_t const * bp_weight, int64_t const * bp_off,
                            int64_t n_pairs, int32_t const * idx_a, int32_t const * idx_b,
                            int n_threads, lara_results * out, lara_stats * stats)
{
    if (!params || !out || n_seq < 0 || n_pairs < 0)
        return LARA_E_INVALID;
    if (params->matching == 0)
        return LARA_E_UNSUPPORTED;
    std::vector<PairOut> res((size_t)n_pairs);
    auto t0 = std::chrono::steady_clock::now();
    int bad = 0;
    if (n_threads < 1) n_threads = 1;
    #pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
    for (int64_t p = 0; p < n_pairs; ++p)
    {
        int32_t i = std::min(idx_a[p], idx_b[p]), j = std::max(idx_a[p], idx_b[p]);
        if (i < 0 || j >= n_seq || i == j) { bad = 1; continue; }
        int li = (int)(seq_off[i + 1] - seq_off[i]), lj = (int)(seq_off[j + 1] - seq_off[j]);
        int32_t first = (li >= lj) ? i : j, second = (li >= lj) ? j : i;               // solver :117-120
        SeqView a{seq_cat + seq_off[first], (int)(seq_off[first + 1] - seq_off[first]), seq_off[first],
                  bp_partner, bp_weight, bp_off};
        SeqView b{seq_cat + seq_off[second], (int)(seq_off[second + 1] - seq_off[second]), seq_off[second],
                  bp_partner, bp_weight, bp_off};
        solve_pair(params, a, b, first, second, res[(size_t)p]);
    }
    if (bad)
        return LARA_E_INVALID;
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    ResultOwner * own = new ResultOwner();
    own->pairs.resize((size_t)n_pairs);
    int64_t cells = 0, iters = 0, sumE = 0, sumD = 0, fcells = 0;
    for (int64_t p = 0; p < n_pairs; ++p)
    {
        PairOut const & r = res[(size_t)p];
        lara_pair_result & o = own->pairs[(size_t)p];
        o.idx_low = r.idxLow; o.idx_high = r.idxHigh; o.best_lower = r.bestLower; o.best_upper = r.bestUpper;
        o.iterations = r.iterations; o.n_lines = (int32_t)r.posLow.size(); o.lines_off = (int64_t)own->posLow.size();
        own->posLow.insert(own->posLow.end(), r.posLow.begin(), r.posLow.end());
        own->posHigh.insert(own->posHigh.end(), r.posHigh.begin(), r.posHigh.end());
        own->weight.insert(own->weight.end(), r.weight.begin(), r.weight.end());
        int64_t la = seq_off[r.idxLow + 1] - seq_off[r.idxLow], lb = seq_off[r.idxHigh + 1] - seq_off[r.idxHigh];
        cells += la * lb * r.iterations; fcells += 2 * la * lb; iters += r.iterations; sumE += r.E; sumD += r.D;
    }
    out->n_pairs = n_pairs;
    out->pairs = own->pairs.data();
    out->n_lines_total = (int64_t)own->posLow.size();
    out->line_pos_low = own->posLow.data();
    out->line_pos_high = own->posHigh.data();
    out->line_weight = own->weight.data();
    out->owner_ = own;
    if (stats)
    {
        std::memset(stats, 0, sizeof(*stats));
        stats->ms_total = ms; stats->dp_cells = cells; stats->filter_cells = fcells; stats->n_pairs = n_pairs;
        stats->sum_iterations = iters; stats->n_active_edges = sumE; stats->n_interactions = sumD;
    }
    return LARA_OK;
}

void lara_oracle_free_results(lara_results * out)
{
    if (out && out->owner_)
    {
        delete static_cast<ResultOwner *>(out->owner_);
        std::memset(out, 0, sizeof(*out));
    }
}

int lara_oracle_gotoh(int32_t lenA, int32_t lenB, int32_t const * pssm, int32_t gap_open, int32_t gap_extend,
                      int32_t * score, int32_t * n_lines, int32_t * lines_a, int32_t * lines_b,
                      int32_t * gap_score, uint8_t * trace_out)
{
    if (lenA < 1 || lenB < 1 || !pssm)
        return LARA_E_INVALID;
    std::vector<uint8_t> trace, ops;
    ScoreType s = gotoh_fill(lenA, lenB, pssm, gap_open, gap_extend, trace);
    gotoh_traceback(lenA, lenB, trace, ops);
    std::vector<int32_t> pa, pb;
    ScoreType g = evaluate_lines(ops, gap_open, gap_extend, pa, pb);
    if (score) *score = s;
    if (gap_score) *gap_score = g;
    if (n_lines) *n_lines = (int32_t)pa.size();
    for (size_t k = 0; k < pa.size(); ++k)
    {
        if (lines_a) lines_a[k] = pa[k];
        if (lines_b) lines_b[k] = pb[k];
    }
    if (trace_out)
        std::memcpy(trace_out, trace.data(), trace.size());
    return LARA_OK;
}

int lara_oracle_edge_filter(lara_params const * params, uint8_t const * seqA, int32_t lenA,
                            uint8_t const * seqB, int32_t lenB, uint8_t * active, int32_t * opt_score)
{
    if (!params || lenA < 1 || lenB < 1)
        return LARA_E_INVALID;
    generate_edges(seqA, lenA, seqB, lenB, params->score_matrix, params->gap_open, params->gap_extend,
                   (ScoreType)(params->suboptimal_diff * LARA_FACTOR2INT), active, opt_score);
    return LARA_OK;
}

int32_t lara_oracle_greedy_matching(int32_t n, int32_t const * w, int32_t const * l, int32_t const * r,
                                    int32_t lookahead, int32_t n_vertices, uint8_t * matched)
{
    std::vector<Interaction> q;
    for (int32_t k = 0; k < n; ++k)
        q.emplace_back(-2 * w[k], (int64_t)l[k], (int64_t)r[k]);
    std::sort(q.begin(), q.end());
    q.erase(std::unique(q.begin(), q.end()), q.end());
    std::vector<uint8_t> m((size_t)n_vertices, 0);
    int32_t s = greedy_matching(q, (size_t)lookahead,
        [&](int64_t a, int64_t b) { m[a] = 1; m[b] = 1; }, [&](int64_t v) { return m[v] != 0; });
    if (matched)
        std::memcpy(matched, m.data(), m.size());
    return s;
}

} // extern "C"
