// lara_host.cpp -- host side around the GPU hot path: input readers, option table, output writers.
// Written from scratch against the behaviour of the reference (file:line cited per function);
// no SeqAn.  See include/lara_host.h.
#include "lara_host.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

namespace
{

thread_local std::string g_err;

int fail(int code, std::string const & msg)
{
    g_err = msg;
    return code;
}

struct Edge
{
    int32_t i, j;
    double w;
};

struct Graph
{
    std::vector<Edge> edges;
};

// seqan::RnaRecord as far as LaRA reads it (SeqAn:rna_io/rna_record.h)
struct Record
{
    std::string name;
    std::string seq;                   // characters as printed by the writers (A C G U N)
    std::vector<Graph> fixedGraphs;
    std::vector<Graph> bppGraphs;
    bool skipCompute = false;          // records from -d dot plots never pass computeStructure
};

// IUPAC/char -> Rna5 (SeqAn:basic/alphabet_residue_tabs.h:270-294): A0 C1 G2 T/U3, everything else N4.
inline uint8_t rna5_code(char c)
{
    switch (c)
    {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 3;
        default: return 4;
    }
}
inline char rna5_char(uint8_t code) { return "ACGUN"[code > 4 ? 4 : code]; }

std::string to_rna5_string(std::string const & raw)
{
    std::string s(raw.size(), 'N');
    for (size_t k = 0; k < raw.size(); ++k)
        s[k] = rna5_char(rna5_code(raw[k]));
    return s;
}

bool ends_with(std::string const & s, char const * suf)
{
    size_t n = std::strlen(suf);
    if (s.size() < n) return false;
    for (size_t k = 0; k < n; ++k)
        if (std::tolower((unsigned char)s[s.size() - n + k]) != suf[k]) return false;
    return true;
}

void rstrip(std::string & s)
{
    while (!s.empty() && (s.back() == '\r' || s.back() == '\n' || s.back() == ' ' || s.back() == '\t'))
        s.pop_back();
}

std::vector<std::string> split_ws(std::string const & s)
{
    std::vector<std::string> out;
    std::istringstream iss(s);
    std::string t;
    while (iss >> t) out.push_back(t);
    return out;
}

// ---- dot-bracket: SeqAn:rna_io/dot_bracket_read_write.h:143-200 (bracket2graph), :292-345 (readRecord) ----
bool bracket2graph(std::string const & br, Graph & g, std::string & err)
{
    static std::string const open = "({<[ABCDEFGHIJKLMNOPQRSTUVWXYZ";
    static std::string const close = ")}>]abcdefghijklmnopqrstuvwxyz";
    static std::string const unpaired = ".,:_-=";
    std::vector<std::vector<int32_t>> stack(open.size());
    for (size_t k = 0; k < br.size(); ++k)
    {
        char c = br[k];
        if (unpaired.find(c) != std::string::npos) continue;
        size_t o = open.find(c);
        if (o != std::string::npos) { stack[o].push_back((int32_t)k); continue; }
        size_t cl = close.find(c);
        if (cl == std::string::npos) { err = "Invalid bracket notation: unknown symbol"; return false; }
        if (stack[cl].empty()) { err = "Invalid bracket notation: unpaired closing bracket"; return false; }
        g.edges.push_back(Edge{stack[cl].back(), (int32_t)k, 1.0});
        stack[cl].pop_back();
    }
    for (auto const & st : stack)
        if (!st.empty()) { err = "Invalid bracket notation: unpaired opening bracket"; return false; }
    return true;
}

int read_dbn(std::istream & in, std::vector<Record> & out)
{
    std::string line;
    while (std::getline(in, line))
    {
        rstrip(line);
        if (line.empty()) continue;
        if (line[0] != '>') return fail(LARA_E_IO, "dbn: expected '>' header line");
        Record rec;
        std::string header = line.substr(1);
        // name/offset convention "name/1-73" (:300-316)
        size_t pos = header.find_last_of('/');
        rec.name = header;
        if (pos != std::string::npos)
        {
            std::string posStr = header.substr(pos + 1);
            size_t dash = posStr.find('-');
            bool ok = dash != std::string::npos && dash > 0;
            for (size_t k = 0; ok && k < dash; ++k) ok = std::isdigit((unsigned char)posStr[k]) != 0;
            if (ok) rec.name = header.substr(0, pos);
        }
        if (!std::getline(in, line)) return fail(LARA_E_IO, "dbn: missing sequence line");
        rstrip(line);
        rec.seq = to_rna5_string(line);
        if (!std::getline(in, line)) return fail(LARA_E_IO, "dbn: missing bracket line");
        rstrip(line);
        std::string br = line.substr(0, line.find_first_of(" \t"));
        if (br.size() != rec.seq.size())
            return fail(LARA_E_IO, "ERROR: Bracket string (" + std::to_string(br.size()) +
                                   ") must be as long as sequence (" + std::to_string(rec.seq.size()) + ").");
        Graph g;
        std::string err;
        if (!bracket2graph(br, g, err)) return fail(LARA_E_IO, err);
        rec.fixedGraphs.push_back(g);
        out.push_back(rec);
    }
    return LARA_OK;
}

// ---- bpseq: SeqAn:rna_io/bpseq_read_write.h:81-159 ------------------------------------------------------------------
// Records follow each other; each starts with '#' lines (the first one is the name, the others are comments) and
// continues with "index nucleotide partner" lines up to the next '#' line.  An edge of weight 1 is added on the
// line of the higher index (:139-140); the first index of a record is its offset.
int read_bpseq(std::istream & in, std::vector<Record> & out)
{
    std::string line;
    bool have = false;
    auto next = [&]() -> bool
    {
        while (std::getline(in, line)) { rstrip(line); size_t p = line.find_first_not_of(" \t"); if (p != std::string::npos) { line = line.substr(p); return true; } }
        return false;
    };
    have = next();
    while (have)
    {
        Record rec;
        bool named = false;
        while (have && line[0] == '#')
        {
            size_t p = line.find_first_not_of(" \t", 1);
            if (!named) { rec.name = p == std::string::npos ? std::string() : line.substr(p); named = !rec.name.empty(); }
            have = next();
        }
        Graph g;
        long offset = -1, cur = 0;
        std::string raw;
        while (have && line[0] != '#')
        {
            std::vector<std::string> t = split_ws(line);
            if (t.size() < 3) return fail(LARA_E_IO, "bpseq: expected 'index nucleotide partner'");
            if (offset < 0)
            {
                char * end = nullptr;
                offset = std::strtol(t[0].c_str(), &end, 10);
                if (!end || *end || offset < 0) return fail(LARA_E_IO, "bpseq: bad index '" + t[0] + "'");
                cur = offset;
            }
            else
                ++cur;
            raw.push_back(t[1][0]);
            char * end = nullptr;
            long const pr = std::strtol(t[2].c_str(), &end, 10);
            if (!end || *end || pr < 0) return fail(LARA_E_IO, "bpseq: bad partner '" + t[2] + "'");
            if (pr != 0 && cur > pr)
            {
                if (pr < offset) return fail(LARA_E_IO, "bpseq: partner before the first index");
                g.edges.push_back(Edge{(int32_t)(pr - offset), (int32_t)(cur - offset), 1.0});
            }
            have = next();
        }
        if (raw.empty()) return fail(LARA_E_IO, "bpseq: record without nucleotides");
        rec.seq = to_rna5_string(raw);
        rec.fixedGraphs.push_back(g);
        out.push_back(rec);
    }
    return LARA_OK;
}

// ---- connect (.ct): SeqAn:rna_io/connect_read_write.h:112-213 --------------------------------------------------------
// "N [ENERGY = e] name" then N lines "index nucleotide index-1 index+1 partner index"; edge of weight 1 on the line of
// the higher index.
int read_ct(std::istream & in, std::vector<Record> & out)
{
    std::string line;
    auto next = [&]() -> bool
    {
        while (std::getline(in, line)) { rstrip(line); if (line.find_first_not_of(" \t") != std::string::npos) return true; }
        return false;
    };
    while (next())
    {
        Record rec;
        size_t p = line.find_first_not_of(" \t");
        size_t q = line.find_first_of(" \t", p);
        std::string const nStr = line.substr(p, q == std::string::npos ? std::string::npos : q - p);
        char * end = nullptr;
        long const n = std::strtol(nStr.c_str(), &end, 10);
        if (!end || *end || n < 1) return fail(LARA_E_IO, "ct: bad sequence length '" + nStr + "'");
        // second token up to white space or '=' (:131-148)
        p = q == std::string::npos ? std::string::npos : line.find_first_not_of(" \t", q);
        if (p != std::string::npos)
        {
            q = line.find_first_of(" \t=", p);
            std::string tok = line.substr(p, q == std::string::npos ? std::string::npos : q - p);
            if (tok.compare(0, 6, "ENERGY") == 0)
            {
                size_t eq = line.find('=', p);
                size_t e0 = eq == std::string::npos ? std::string::npos : line.find_first_not_of(" \t", eq + 1);
                size_t e1 = e0 == std::string::npos ? std::string::npos : line.find_first_of(" \t", e0);
                size_t nm = e1 == std::string::npos ? std::string::npos : line.find_first_not_of(" \t", e1);
                rec.name = nm == std::string::npos ? std::string() : line.substr(nm);
            }
            else
                rec.name = tok + (q == std::string::npos ? std::string() : line.substr(q));
        }
        Graph g;
        std::string raw;
        long offset = 1;
        for (long k = 0; k < n; ++k)
        {
            if (!next()) return fail(LARA_E_IO, "ct: record ends after " + std::to_string(k) + " of " + std::to_string(n) + " lines");
            std::vector<std::string> t = split_ws(line);
            if (t.size() < 5) return fail(LARA_E_IO, "ct: expected 'index nucleotide prev next partner index'");
            if (k == 0)
            {
                offset = std::strtol(t[0].c_str(), &end, 10);
                if (!end || *end) return fail(LARA_E_IO, "ct: bad index '" + t[0] + "'");
            }
            long const cur = offset + k;
            raw += t[1];
            long const pr = std::strtol(t[4].c_str(), &end, 10);
            if (!end || *end || pr < 0) return fail(LARA_E_IO, "ct: bad partner '" + t[4] + "'");
            if (pr != 0 && cur > pr)
            {
                if (pr < offset) return fail(LARA_E_IO, "ERROR: Incompatible pairing position in input file.");
                g.edges.push_back(Edge{(int32_t)(pr - offset), (int32_t)(cur - offset), 1.0});
            }
        }
        rec.seq = to_rna5_string(raw);
        rec.fixedGraphs.push_back(g);
        out.push_back(rec);
    }
    return LARA_OK;
}

// ---- extended bpseq: SeqAn:rna_io/ebpseq_read_write.h:112-165 (header), :173-395 (records) ----
int read_ebpseq(std::istream & in, std::vector<Record> & out)
{
    std::map<std::string, std::string> seqLabels;   // identifier -> label
    std::string line;
    bool haveLine = false;
    // header: lines starting with "##"
    while (std::getline(in, line))
    {
        rstrip(line);
        if (line.empty()) continue;
        if (line.size() >= 2 && line[0] == '#' && line[1] == '#')
        {
            size_t p = 2;
            while (p < line.size() && std::isspace((unsigned char)line[p])) ++p;
            if (p >= line.size()) continue;
            char key = line[p++];
            size_t q = p;
            while (q < line.size() && line[q] != ':' && line[q] != ' ' && line[q] != '\t') ++q;
            std::string ident = line.substr(p, q - p);
            if (q < line.size()) ++q;
            while (q < line.size() && (line[q] == ' ' || line[q] == '\t')) ++q;
            std::string label = q < line.size() ? line.substr(q) : std::string();
            if (key == 'S') seqLabels[ident] = label;
            else if (key != 'G' && key != 'F' && key != 'M' && key != 'T')
                return fail(LARA_E_IO, "ERROR: Unknown key in header.");
            continue;
        }
        haveLine = true;
        break;
    }
    while (haveLine)
    {
        // "# S1 T2" record line
        if (line.empty() || line[0] != '#') return fail(LARA_E_IO, "ebpseq: expected record line '# S<id>'");
        Record rec;
        bool haveName = false;
        for (std::string const & tok : split_ws(line.substr(1)))
        {
            if (tok[0] == 'S')
            {
                auto it = seqLabels.find(tok.substr(1));
                if (it == seqLabels.end()) return fail(LARA_E_IO, "ERROR: Unknown tag symbol " + tok);
                rec.name = it->second;
                haveName = true;
            }
            else if (tok[0] != 'T')
                return fail(LARA_E_IO, "ERROR: Unknown tag symbol " + tok);
        }
        if (!haveName) return fail(LARA_E_IO, "ERROR: Record must contain a sequence tag.");
        // "# I NT QU ... F1 M1" column line
        if (!std::getline(in, line)) return fail(LARA_E_IO, "ebpseq: missing column line");
        rstrip(line);
        if (line.empty() || line[0] != '#') return fail(LARA_E_IO, "ebpseq: expected column line");
        std::vector<std::string> cols = split_ws(line.substr(1));
        size_t nFix = 0, nBpp = 0;
        for (std::string const & c : cols)
        {
            if (c[0] == 'F') ++nFix;
            else if (c[0] == 'M') ++nBpp;
        }
        rec.fixedGraphs.resize(nFix);
        rec.bppGraphs.resize(nBpp);
        // matrix
        unsigned lineNo = 1;
        long offset = 1;
        haveLine = false;
        while (std::getline(in, line))
        {
            rstrip(line);
            if (line.empty()) continue;
            if (line[0] == '#') { haveLine = true; break; }
            size_t p = 0;
            size_t fixID = 0, bppID = 0;
            auto nextToken = [&]() {
                while (p < line.size() && (line[p] == ' ' || line[p] == '\t')) ++p;
                size_t q = p;
                while (q < line.size() && line[q] != ' ' && line[q] != '\t') ++q;
                std::string t = line.substr(p, q - p);
                p = q;
                return t;
            };
            for (std::string const & label : cols)
            {
                std::string buf = nextToken();
                if (label == "I") { if (lineNo == 1) offset = std::atol(buf.c_str()); }
                else if (label == "NT") rec.seq.push_back(rna5_char(rna5_code(buf.empty() ? 'N' : buf[0])));
                else if (label == "QU") {}
                else if (label[0] == 'F')
                {
                    Graph & g = rec.fixedGraphs[fixID++];
                    long pairPos = std::atol(buf.c_str());
                    if (pairPos != 0 && (long)lineNo + offset - 1 > pairPos)                 // :311-314
                        g.edges.push_back(Edge{(int32_t)(pairPos - offset), (int32_t)(lineNo - 1), 1.0});
                }
                else if (label[0] == 'M')
                {
                    Graph & g = rec.bppGraphs[bppID++];
                    if (buf.find('<') == std::string::npos)
                        return fail(LARA_E_IO, "ERROR: Expected a bracket pair: < >");
                    if (buf.find('>') == std::string::npos)
                    {
                        size_t q = line.find('>', p);
                        if (q == std::string::npos) q = line.size();
                        buf += line.substr(p, q - p);
                        p = std::min(q + 1, line.size());
                    }
                    // regex \d+/\d\.?\d*  (:327): digits '/' ONE digit, optional '.', digits
                    size_t k = 0;
                    while (k < buf.size())
                    {
                        if (!std::isdigit((unsigned char)buf[k])) { ++k; continue; }
                        size_t s0 = k;
                        while (k < buf.size() && std::isdigit((unsigned char)buf[k])) ++k;
                        if (k + 1 >= buf.size() || buf[k] != '/' || !std::isdigit((unsigned char)buf[k + 1]))
                            continue;
                        long pairPos = std::atol(buf.substr(s0, k - s0).c_str());
                        size_t v0 = k + 1;
                        size_t v = v0 + 1;
                        if (v < buf.size() && buf[v] == '.') ++v;
                        while (v < buf.size() && std::isdigit((unsigned char)buf[v])) ++v;
                        double prob = std::strtod(buf.substr(v0, v - v0).c_str(), nullptr);
                        if (pairPos != 0 && (long)lineNo + offset - 1 > pairPos)             // :341-344
                            g.edges.push_back(Edge{(int32_t)(pairPos - offset), (int32_t)(lineNo - 1), prob});
                        k = v;
                    }
                }
                else if (label[0] == 'R') {}
                else
                    return fail(LARA_E_IO, "ERROR: Unknown label " + label);
            }
            ++lineNo;
        }
        out.push_back(rec);
    }
    return LARA_OK;
}

// ---- FASTA (src/io.hpp:178-199): records without structure ----
int read_fasta(std::istream & in, std::vector<Record> & out)
{
    std::string line;
    Record rec;
    bool open = false;
    std::string raw;
    auto flush = [&]() {
        if (open) { rec.seq = to_rna5_string(raw); out.push_back(rec); }
        rec = Record(); raw.clear();
    };
    while (std::getline(in, line))
    {
        rstrip(line);
        if (!line.empty() && line[0] == '>') { flush(); open = true; rec.name = line.substr(1); }
        else if (open)
            for (char c : line) if (!std::isspace((unsigned char)c)) raw.push_back(c);
    }
    flush();
    return LARA_OK;
}

} // namespace

struct lara_input
{
    std::vector<Record> recs;
    // flattened view
    std::vector<uint8_t> seqCat;
    std::vector<int64_t> seqOff, bpOff;
    std::vector<int32_t> bpPartner, bpWeight;
    bool flat = false;
};

extern "C" {

char const * lara_host_last_error(void) { return g_err.c_str(); }

int lara_input_create(lara_input ** in)
{
    if (!in) return fail(LARA_E_INVALID, "null argument");
    *in = new lara_input();
    return LARA_OK;
}

void lara_input_destroy(lara_input * in) { delete in; }

int lara_input_read_file(lara_input * in, char const * path)
{
    if (!in || !path) return fail(LARA_E_INVALID, "null argument");
    if (!*path) return LARA_OK;                                         // src/io.hpp:137-138
    std::ifstream f(path);
    if (!f.is_open()) return fail(LARA_E_IO, std::string("ERROR: The file ") + path + " cannot be opened.");
    in->flat = false;
    std::string p(path);
    if (ends_with(p, ".ebpseq")) return read_ebpseq(f, in->recs);
    if (ends_with(p, ".dbn")) return read_dbn(f, in->recs);
    if (ends_with(p, ".bpseq")) return read_bpseq(f, in->recs);
    if (ends_with(p, ".ct")) return read_ct(f, in->recs);
    // .sth (Stockholm) is not read: SeqAn's reader fills only record.align and leaves record.sequence empty
    // (SeqAn:rna_io/stockholm_read_write.h:253-264), so the reference itself cannot align such input; .dbv (Vienna
    // without names) is not read either.  Everything else is treated as a sequence file, like src/io.hpp:154-199.
    return read_fasta(f, in->recs);
}

// src/io.hpp:202-284
int lara_input_read_dotplot(lara_input * in, char const * path)
{
    if (!in || !path) return fail(LARA_E_INVALID, "null argument");
    float const minProb = 0.003f;
    std::ifstream file(path);
    if (!file.is_open()) return fail(LARA_E_IO, std::string("ERROR: Cannot open dotplot file ") + path);
    std::string line, raw;
    while (std::getline(file, line))
    {
        if (line.find("/sequence") != std::string::npos)
        {
            while (std::getline(file, line) && line.find(')') == std::string::npos)
                for (char c : line) if (std::isalpha((unsigned char)c)) raw.push_back(c);
            break;
        }
    }
    if (raw.empty()) return fail(LARA_E_IO, std::string("ERROR: no sequence in dotplot file ") + path);
    Record rec;
    rec.seq = to_rna5_string(raw);
    Graph bpp, fixed;
    while (std::getline(file, line))
    {
        bool ubox = line.find("ubox") != std::string::npos;
        bool lbox = !ubox && line.find("lbox") != std::string::npos;
        if (!ubox && !lbox) continue;
        std::istringstream iss(line);
        unsigned iPos = 0, jPos = 0;
        float prob = 0;
        if (!(iss >> iPos >> jPos >> prob)) continue;
        if (ubox)
        {
            if (prob * prob > minProb)                                               // :257-258
                bpp.edges.push_back(Edge{(int32_t)iPos - 1, (int32_t)jPos - 1, (double)std::log(prob * prob / minProb)});
        }
        else
            fixed.edges.push_back(Edge{(int32_t)iPos - 1, (int32_t)jPos - 1, 1.0});
    }
    std::string p(path);
    std::string name = p.substr(p.find_last_of("/\\") + 1);
    name = name.substr(0, name.rfind(".ps"));
    name = name.substr(0, name.rfind("_dp"));                                         // :277-278 (npos = keep all)
    // note: the reference evaluates name.rfind("_dp") on the un-cut name; identical for *_dp.ps
    rec.name = name;
    if (!bpp.edges.empty()) rec.bppGraphs.push_back(bpp);
    else rec.fixedGraphs.push_back(fixed);
    rec.skipCompute = true;
    if (rec.bppGraphs.empty())                                                        // src/io.hpp:101-107
        return fail(LARA_E_IO, std::string("ERROR: The dotplot file ") + path +
                               " does not contain any base pair probabilities. Please make sure that you execute "
                               "RNAfold with -p option and specify the resulting _dp.ps file!");
    in->recs.push_back(rec);
    in->flat = false;
    return LARA_OK;
}


// computeStructure, src/io.hpp:286-338, and the size check :115-119
int lara_input_finish(lara_input * in)
{
    if (!in) return fail(LARA_E_INVALID, "null argument");
    for (Record & rec : in->recs)
    {
        if (rec.skipCompute || !rec.bppGraphs.empty()) { rec.skipCompute = true; continue; }
        if (!rec.fixedGraphs.empty())
        {
            for (Edge & e : rec.fixedGraphs.front().edges) e.w *= 10;                  // :290-296
            rec.skipCompute = true;
            continue;
        }
        return fail(LARA_E_UNSUPPORTED,
                    "Cannot compute a structure without the ViennaRNA library. Please install ViennaRNA and try again.");
    }
    if (in->recs.size() <= 1)
        return fail(LARA_E_IO, "ERROR: The given file(s) must contain at least two sequences.");
    in->flat = false;
    return LARA_OK;
}

int lara_input_add_record(lara_input * in, char const * name, char const * seq, int32_t n_pairs,
                          int32_t const * pi, int32_t const * pj, double const * w, int is_bpp)
{
    if (!in || !name || !seq) return fail(LARA_E_INVALID, "null argument");
    Record rec;
    rec.name = name;
    rec.seq = to_rna5_string(seq);
    Graph g;
    for (int32_t k = 0; k < n_pairs; ++k)
    {
        if (pi[k] < 0 || pj[k] < 0 || pi[k] >= (int32_t)rec.seq.size() || pj[k] >= (int32_t)rec.seq.size())
            return fail(LARA_E_INVALID, "base pair position out of range");
        g.edges.push_back(Edge{pi[k], pj[k], w ? w[k] : 1.0});
    }
    if (is_bpp) rec.bppGraphs.push_back(g); else rec.fixedGraphs.push_back(g);
    in->recs.push_back(rec);
    in->flat = false;
    return LARA_OK;
}

int32_t lara_input_size(lara_input const * in) { return in ? (int32_t)in->recs.size() : 0; }
char const * lara_input_name(lara_input const * in, int32_t i)
{
    return (in && i >= 0 && i < (int32_t)in->recs.size()) ? in->recs[i].name.c_str() : "";
}
int32_t lara_input_length(lara_input const * in, int32_t i)
{
    return (in && i >= 0 && i < (int32_t)in->recs.size()) ? (int32_t)in->recs[i].seq.size() : -1;
}

// The graph Lagrange reads: front(bppMatrGraphs) if any, else front(fixedGraphs) (src/lagrange.hpp:259-262);
// weights go double -> float -> int32 exactly like extractContacts (src/lagrange.hpp:135-136).
int lara_input_flatten(lara_input * in, uint8_t const ** seq_cat, int64_t const ** seq_off,
                       int32_t const ** bp_partner, int32_t const ** bp_weight, int64_t const ** bp_off)
{
    if (!in) return fail(LARA_E_INVALID, "null argument");
    if (!in->flat)
    {
        in->seqCat.clear(); in->seqOff.assign(1, 0); in->bpOff.assign(1, 0);
        in->bpPartner.clear(); in->bpWeight.clear();
        for (Record const & rec : in->recs)
        {
            int32_t L = (int32_t)rec.seq.size();
            for (char c : rec.seq) in->seqCat.push_back(rna5_code(c));
            in->seqOff.push_back((int64_t)in->seqCat.size());
            Graph const * g = !rec.bppGraphs.empty() ? &rec.bppGraphs.front()
                            : (!rec.fixedGraphs.empty() ? &rec.fixedGraphs.front() : nullptr);
            std::vector<std::vector<std::pair<int32_t, int32_t>>> adj((size_t)L);
            if (g)
                for (Edge const & e : g->edges)
                {
                    if (e.i < 0 || e.j < 0 || e.i >= L || e.j >= L || e.i == e.j) continue;
                    float probability = (float)e.w;
                    int32_t wi = (int32_t)probability;
                    bool dup = false;                       // parallel edges: first one wins (findEdge)
                    for (auto const & x : adj[e.i]) if (x.first == e.j) dup = true;
                    if (dup) continue;
                    adj[e.i].emplace_back(e.j, wi);
                    adj[e.j].emplace_back(e.i, wi);
                }
            for (int32_t p = 0; p < L; ++p)
            {
                std::sort(adj[p].begin(), adj[p].end());
                for (auto const & x : adj[p]) { in->bpPartner.push_back(x.first); in->bpWeight.push_back(x.second); }
                in->bpOff.push_back((int64_t)in->bpPartner.size());
            }
        }
        in->flat = true;
    }
    if (seq_cat) *seq_cat = in->seqCat.data();
    if (seq_off) *seq_off = in->seqOff.data();
    if (bp_partner) *bp_partner = in->bpPartner.data();
    if (bp_weight) *bp_weight = in->bpWeight.data();
    if (bp_off) *bp_off = in->bpOff.data();
    return LARA_OK;
}

// OutputLibrary::printLib / printAlignments, src/io.hpp:387-455
int lara_output_write(lara_input const * in, lara_results const * res, char const * format, char const * path)
{
    if (!in || !res || !format) return fail(LARA_E_INVALID, "null argument");
    std::string fmt(format);
    if (fmt != "lib" && fmt != "pairs" && fmt != "fasta") return fail(LARA_E_INVALID, "unknown output format " + fmt);
    std::vector<int64_t> order((size_t)res->n_pairs);
    for (int64_t k = 0; k < res->n_pairs; ++k) order[(size_t)k] = k;
    std::stable_sort(order.begin(), order.end(), [&](int64_t x, int64_t y) {
        lara_pair_result const & a = res->pairs[x];
        lara_pair_result const & b = res->pairs[y];
        return a.idx_low != b.idx_low ? a.idx_low < b.idx_low : a.idx_high < b.idx_high;
    });
    std::ofstream file;
    std::ostream * os = &std::cout;
    if (path && *path)
    {
        file.open(path, std::ios::out);
        if (!file.is_open()) return fail(LARA_E_IO, std::string("Error: Unable to open the file for writing: ") + path);
        os = &file;
    }
    std::ostream & stream = *os;
    if (fmt == "lib")
    {
        stream << "! T-COFFEE_LIB_FORMAT_01\n" << in->recs.size() << '\n';
        for (Record const & rec : in->recs)
            stream << rec.name << " " << rec.seq.size() << " " << rec.seq << '\n';
        for (int64_t k : order)
        {
            lara_pair_result const & r = res->pairs[k];
            stream << "# " << (r.idx_low + 1) << " " << (r.idx_high + 1) << '\n';
            for (int32_t x = 0; x < r.n_lines; ++x)
                stream << res->line_pos_low[r.lines_off + x] + 1 << " " << res->line_pos_high[r.lines_off + x] + 1
                       << " " << res->line_weight[r.lines_off + x] << '\n';
        }
        stream << "! SEQ_1_TO_N\n";
    }
    else
    {
        for (int64_t k : order)
        {
            lara_pair_result const & r = res->pairs[k];
            if (r.idx_low < 0 || r.idx_high >= (int32_t)in->recs.size()) return fail(LARA_E_INVALID, "pair index out of range");
            Record const & rec1 = in->recs[r.idx_low];
            Record const & rec2 = in->recs[r.idx_high];
            std::string g1, g2;
            size_t c1 = 0, c2 = 0;
            for (int32_t x = 0; x < r.n_lines; ++x)
            {
                size_t p1 = (size_t)res->line_pos_low[r.lines_off + x], p2 = (size_t)res->line_pos_high[r.lines_off + x];
                while (c1 < p1) { g1.push_back(rec1.seq[c1++]); g2.push_back('-'); }
                while (c2 < p2) { g1.push_back('-'); g2.push_back(rec2.seq[c2++]); }
                g1.push_back(rec1.seq[c1++]);
                g2.push_back(rec2.seq[c2++]);
            }
            while (c1 < rec1.seq.size()) { g1.push_back(rec1.seq[c1++]); g2.push_back('-'); }
            while (c2 < rec2.seq.size()) { g1.push_back('-'); g2.push_back(rec2.seq[c2++]); }
            if (fmt == "pairs")
            {
                stream << '>' << rec1.name << " && " << rec2.name << " (score " << (float)r.best_lower / LARA_FACTOR2INT << ")\n";
                stream << g1 << '\n' << g2 << '\n';
            }
            else
            {
                stream << '>' << rec1.name << '\n' << g1 << '\n';
                stream << '>' << rec2.name << '\n' << g2 << '\n';
            }
        }
    }
    stream.flush();
    return LARA_OK;
}

} // extern "C"
