#!/usr/bin/env python3
"""Developer tool (CPU): diagonals a pair needs in a straight band (whole diagonals, must reach from 0 to lenA - lenB)
against a sheared band (lower bound non-decreasing with the row, tools/band_proto.py), from the oracle's edge filter on
a sample of pairs of a BASELINE config.  usage: band_need.py <c3|c5> [nseq] [npairs]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_inputs  # noqa: E402
from oracle import binding as ob  # noqa: E402
from tests import util  # noqa: E402


def needs(active, margin=1):
    """active[a, b] (lenA >= lenB).  Returns (straight, sheared) band widths incl. the margin on both sides."""
    la, lb = active.shape
    aa, bb = np.nonzero(active)
    if len(aa) == 0:
        return 1, 1
    d = aa - bb
    straight = max(d.max() + margin, 0) - min(d.min() - margin, 0) + 1
    lo = np.full(lb, 1 << 30); hi = np.full(lb, -(1 << 30))
    np.minimum.at(lo, bb, d); np.maximum.at(hi, bb, d)
    lo[0] = min(lo[0], 0); hi[0] = max(hi[0], 0)                    # the origin lies in the band
    floor = np.minimum.accumulate(lo[::-1])[::-1]                   # largest non-decreasing bound below the rows' minima
    rows = hi > -(1 << 29)
    sheared = int((hi[rows] - floor[rows]).max()) + 1 + 2 * margin
    return int(straight), sheared


def main():
    cfg = sys.argv[1] if len(sys.argv) > 1 else "c5"
    nseq = int(sys.argv[2]) if len(sys.argv) > 2 else 120
    npairs = int(sys.argv[3]) if len(sys.argv) > 3 else 400
    c = gen_inputs.CONFIGS[cfg]
    files = util.make_input(seed=c["seed"], n=nseq, length=c["length"], fmt=c["fmt"], skew=(cfg == "c5"))
    cat, off = util.load_store(files).flat()[:2]                   # host reader only: no GPU involved
    seqs = [cat[off[i]:off[i + 1]] for i in range(len(off) - 1)]
    rng = np.random.default_rng(7)
    P = ob.default_params()
    st, sh, cells_st, cells_sh, cells_p2, cells_full = [], [], 0, 0, 0, 0
    for _ in range(npairs):
        i, j = rng.choice(len(seqs), 2, replace=False)
        a, b = (seqs[i], seqs[j]) if len(seqs[i]) >= len(seqs[j]) else (seqs[j], seqs[i])
        act, _ = ob.edge_filter(P, a, b)
        s, h = needs(act)
        st.append(s); sh.append(h)
        r32 = lambda w: max(32, (w + 31) // 32 * 32) if w > 24 else 24
        cells_st += r32(s) * len(b); cells_sh += r32(h) * len(b)
        p2 = 24 if s <= 24 else 1 << int(np.ceil(np.log2(max(s, 32))))          # what choose_band takes today
        cells_p2 += min(p2 * len(b), len(a) * len(b)); cells_full += len(a) * len(b)
    st, sh = np.array(st), np.array(sh)
    q = lambda x: "p10 %d  p50 %d  p90 %d  max %d" % tuple(np.percentile(x, [10, 50, 90, 100]).astype(int))
    print("%s, first %d sequences, %d random pairs" % (cfg, nseq, npairs))
    print("straight band need: ", q(st))
    print("sheared band need:  ", q(sh))
    print("band cells per pair (widths rounded up to the kernels' 24 / 32 k): straight %.0f, sheared %.0f (%.2fx)"
          % (cells_st / npairs, cells_sh / npairs, cells_st / max(cells_sh, 1)))
    print("today (24 or a power of two >= 32, full matrix when the band is not smaller): %.0f; full matrix: %.0f"
          % (cells_p2 / npairs, cells_full / npairs))


if __name__ == "__main__":
    main()
