#!/usr/bin/env python3
"""Seeded synthetic RNA inputs for the five BASELINE.json configs.

The same bytes are fed to the reference binary (oracle/_ref/lara_scalar), to the
CPU restatement (oracle/) and to the GPU library, so every format quirk of the
reference's readers is respected here:

* ``.ebpseq``: header ``## S<i>: name`` / ``## M1: ...``; per record ``# S<i>`` and
  ``# I NT M1``; the M column is ``<p/w | p/w>`` with ONE digit before the decimal
  point (reader regex ``\\d+/\\d\\.?\\d*``, SeqAn rna_io/ebpseq_read_write.h:327) and
  the reader keeps only the entry on the higher-index line (``:341``).  Weights are
  >= 1.0 because the reference truncates them to integers (src/lagrange.hpp:135).
* ``.dbn``: ``>name`` / sequence / dot-bracket line.
* RNAfold dot plots ``<name>_dp.ps``: ``/sequence { (\\`` block, then
  ``i j sqrt(p) ubox`` lines (src/io.hpp:221-260).

Usage: gen_inputs.py <config> <outdir>   with config in c1..c5 or
       gen_inputs.py custom --n N --len L [--lenmax M] --fmt ebpseq|dbn|dp --seed S <outdir>
"""
import argparse
import math
import os
import sys

import numpy as np

BASES = "ACGU"
COMP = {"A": "U", "U": "A", "G": "C", "C": "G"}


def _ancestor(rng, length, n_stems):
    """Random sequence with planted hairpin stems; returns (seq list, pairs list)."""
    seq = [BASES[i] for i in rng.integers(0, 4, size=length)]
    pairs = []
    used = np.zeros(length, dtype=bool)
    tries = 0
    while len([1 for _ in pairs]) < 10**9 and n_stems > 0 and tries < 200:
        tries += 1
        stem = int(rng.integers(4, 9))
        loop = int(rng.integers(4, 12))
        span = 2 * stem + loop
        if span + 2 > length:
            continue
        start = int(rng.integers(0, length - span))
        if used[start:start + span].any():
            continue
        used[start:start + span] = True
        for k in range(stem):
            i, j = start + k, start + span - 1 - k
            seq[j] = COMP[seq[i]]
            pairs.append((i, j))
        n_stems -= 1
    return seq, pairs


def _mutate(rng, seq, pairs, length, rate=0.18, indel=0.03):
    """Derive a family member: point mutations (compensatory inside stems), a few
    indels, then cut/pad to the requested length.  Returns (string, pairs)."""
    seq = list(seq)
    partner = {}
    for i, j in pairs:
        partner[i] = j
        partner[j] = i
    for i in range(len(seq)):
        if rng.random() < rate:
            b = BASES[int(rng.integers(0, 4))]
            seq[i] = b
            if i in partner and rng.random() < 0.8:
                seq[partner[i]] = COMP[b]
    # indels: build a map old index -> new index (or -1 when deleted)
    out = []
    newidx = [-1] * len(seq)
    for i, b in enumerate(seq):
        r = rng.random()
        if r < indel / 2 and i not in partner:
            continue  # deletion
        newidx[i] = len(out)
        out.append(b)
        if r > 1 - indel / 2:
            out.append(BASES[int(rng.integers(0, 4))])  # insertion
    # fit to length
    while len(out) < length:
        out.append(BASES[int(rng.integers(0, 4))])
    out = out[:length]
    npairs = []
    for i, j in pairs:
        a, b = newidx[i], newidx[j]
        if 0 <= a < length and 0 <= b < length and a < b:
            npairs.append((a, b))
    return "".join(out), npairs


def make_family(seed, n, lengths, related=True):
    """n sequences with the given lengths (list) -> list of (name, seq, stem_pairs)."""
    rng = np.random.default_rng(seed)
    lmax = max(lengths)
    anc, apairs = _ancestor(rng, lmax + 8, n_stems=max(3, int(rng.integers(3, 7)) * max(1, lmax // 120)))
    recs = []
    for k in range(n):
        L = int(lengths[k])
        if related:
            s, p = _mutate(rng, anc, apairs, L)
        else:
            a2, p2 = _ancestor(rng, L, n_stems=int(rng.integers(3, 7)))
            s, p = "".join(a2), p2
        recs.append(("seq%d" % (k + 1), s, p))
    return recs, rng


def bpp_from_stems(rng, L, pairs, decoys_per_base=2, dense=False):
    """Weighted partner lists: stems U(3.0,5.8), decoys U(1.0,3.0); dict (i<j) -> weight."""
    w = {}
    for i, j in pairs:
        w[(i, j)] = float(rng.uniform(3.0, 5.8))
    ndec = int(L * decoys_per_base / 2)
    if dense:
        ndec = int(L * 4.5)
    for _ in range(ndec):
        i = int(rng.integers(0, L))
        j = int(rng.integers(0, L))
        if abs(i - j) < 4:
            continue
        a, b = min(i, j), max(i, j)
        if (a, b) not in w:
            w[(a, b)] = float(rng.uniform(1.0, 3.0))
    return w


def write_ebpseq(path, recs, rng, decoys_per_base=2):
    with open(path, "w") as f:
        for k, (name, _, _) in enumerate(recs):
            f.write("## S%d: %s\n" % (k + 1, name))
        f.write("## M1: synthetic base pair weights\n")
        for k, (name, s, pairs) in enumerate(recs):
            L = len(s)
            w = bpp_from_stems(rng, L, pairs, decoys_per_base)
            adj = [[] for _ in range(L)]
            for (i, j), x in sorted(w.items()):
                adj[i].append((j, x))
                adj[j].append((i, x))
            f.write("# S%d\n# I\tNT\tM1\n" % (k + 1))
            for i in range(L):
                items = " | ".join("%d/%.2f" % (j + 1, x) for j, x in sorted(adj[i]))
                f.write("%d\t%s\t<%s>\n" % (i + 1, s[i], items))


def dotbracket(L, pairs):
    d = ["."] * L
    # keep only a nested subset (stems are planted as non-overlapping hairpins)
    for i, j in pairs:
        d[i] = "("
        d[j] = ")"
    # validate nesting; drop offending pairs if any
    st = []
    ok = True
    for c in d:
        if c == "(":
            st.append(c)
        elif c == ")":
            if not st:
                ok = False
                break
            st.pop()
    if st:
        ok = False
    if not ok:
        d = ["."] * L
        depth_ok = []
        for i, j in sorted(pairs):
            if all(not (a < i < b < j or i < a < j < b) for a, b in depth_ok):
                depth_ok.append((i, j))
        for i, j in depth_ok:
            d[i] = "("
            d[j] = ")"
    return "".join(d)


def write_dbn(path, recs):
    with open(path, "w") as f:
        for name, s, pairs in recs:
            f.write(">%s\n%s\n%s\n" % (name, s, dotbracket(len(s), pairs)))


def _one_partner_pairs(L, pairs):
    """pairs in which every position occurs at most once (bpseq / ct hold one partner per base)"""
    used = set()
    out = []
    for i, j in sorted(pairs):
        if i in used or j in used or not (0 <= i < j < L):
            continue
        used.add(i)
        used.add(j)
        out.append((i, j))
    return out


def write_bpseq(path, recs):
    """SeqAn rna_io/bpseq_read_write.h: '# name' then 'index nucleotide partner' lines (0 = unpaired)."""
    with open(path, "w") as f:
        for name, s, pairs in recs:
            partner = {}
            for i, j in _one_partner_pairs(len(s), pairs):
                partner[i] = j
                partner[j] = i
            f.write("# %s\n" % name)
            for i, c in enumerate(s):
                f.write("%d %s %d\n" % (i + 1, c, partner.get(i, -1) + 1))


def write_ct(path, recs):
    """SeqAn rna_io/connect_read_write.h: 'N [ENERGY = e] name' then 'index nucleotide index-1 index+1 partner index'."""
    with open(path, "w") as f:
        for k, (name, s, pairs) in enumerate(recs):
            partner = {}
            for i, j in _one_partner_pairs(len(s), pairs):
                partner[i] = j
                partner[j] = i
            if k % 2 == 0:
                f.write("%d\tENERGY = -%d.5\t%s\n" % (len(s), 10 + k, name))
            else:
                f.write("%5d %s\n" % (len(s), name))
            for i, c in enumerate(s):
                f.write("%5d %s %5d %5d %5d %5d\n" % (i + 1, c, i, (i + 2) if i + 1 < len(s) else 0, partner.get(i, -1) + 1, i + 1))


def write_fasta(path, recs):
    with open(path, "w") as f:
        for name, s, _ in recs:
            f.write(">%s\n%s\n" % (name, s))


def write_dotplots(outdir, recs, rng):
    """One RNAfold-style <name>_dp.ps per record; returns the file list."""
    files = []
    for name, s, pairs in recs:
        L = len(s)
        w = bpp_from_stems(rng, L, pairs, dense=True)
        path = os.path.join(outdir, "%s_dp.ps" % name)
        with open(path, "w") as f:
            f.write("%!PS-Adobe-3.0 EPSF-3.0\n%%Title: RNA Dot Plot\n/sequence { (\\\n")
            for o in range(0, L, 255):
                f.write(s[o:o + 255] + "\\\n")
            f.write(") } def\n/len { sequence length } bind def\n%start of base pair probability data\n")
            for (i, j), x in sorted(w.items()):
                # x in [1,5.8) is the target ln(p/0.003); the dot plot stores sqrt(p)
                p = min(0.999, 0.003 * math.exp(x))
                f.write("%d %d %.9f ubox\n" % (i + 1, j + 1, math.sqrt(p)))
            f.write("showpage\nend\n%%EOF\n")
        files.append(path)
    return files


def skewed_lengths(rng, n, lo=50, hi=600, mu=150.0, sigma=0.6):
    x = np.exp(rng.normal(math.log(mu), sigma, size=n))
    return np.clip(np.rint(x), lo, hi).astype(int)


CONFIGS = {
    "c1": dict(seed=1, n=10, length=120, fmt="dbn"),
    "c2": dict(seed=2, n=100, length=150, fmt="ebpseq"),
    "c3": dict(seed=3, n=500, length=300, fmt="ebpseq"),
    "c4": dict(seed=4, n=50, length=1500, fmt="dp"),
    "c5": dict(seed=5, n=2000, length=0, fmt="ebpseq"),
}


def generate(outdir, seed, n, length, fmt, lenmax=0, related=True, skew=False):
    os.makedirs(outdir, exist_ok=True)
    rng0 = np.random.default_rng(seed + 7919)
    if skew or length == 0:
        lengths = skewed_lengths(rng0, n)
    elif lenmax and lenmax > length:
        lengths = rng0.integers(length, lenmax + 1, size=n)
    else:
        lengths = [length] * n
    recs, rng = make_family(seed, n, list(lengths), related=related)
    if fmt == "dbn":
        p = os.path.join(outdir, "input.dbn")
        write_dbn(p, recs)
        return [p]
    if fmt == "ebpseq":
        p = os.path.join(outdir, "input.ebpseq")
        write_ebpseq(p, recs, rng)
        return [p]
    if fmt == "bpseq":
        p = os.path.join(outdir, "input.bpseq")
        write_bpseq(p, recs)
        return [p]
    if fmt == "ct":
        p = os.path.join(outdir, "input.ct")
        write_ct(p, recs)
        return [p]
    if fmt == "fasta":
        p = os.path.join(outdir, "input.fasta")
        write_fasta(p, recs)
        return [p]
    if fmt == "dp":
        return write_dotplots(outdir, recs, rng)
    raise ValueError(fmt)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("config")
    ap.add_argument("outdir")
    ap.add_argument("--n", type=int, default=10)
    ap.add_argument("--len", type=int, default=100)
    ap.add_argument("--lenmax", type=int, default=0)
    ap.add_argument("--fmt", default="ebpseq")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--unrelated", action="store_true")
    a = ap.parse_args(argv)
    if a.config in CONFIGS:
        c = CONFIGS[a.config]
        files = generate(a.outdir, c["seed"], c["n"], c["length"], c["fmt"], skew=(a.config == "c5"))
    else:
        files = generate(a.outdir, a.seed, a.n, a.len, a.fmt, lenmax=a.lenmax, related=not a.unrelated)
    for p in files:
        print(p)


if __name__ == "__main__":
    main()
