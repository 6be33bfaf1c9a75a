#!/usr/bin/env python3
"""Developer tool (GPU box): for one pair of a bench workload, the first iteration count n at which the GPU's result differs
from the oracle's (bounds, lines), in the given step mode.  Narrows a whole-run mismatch down to one iteration.

usage: gpu_upper_bisect.py <workload> <nseq> <i> <j> [step_mode]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import numpy as np  # noqa: E402
import bench  # noqa: E402
import lara_b200 as lb  # noqa: E402
from oracle import binding as ob  # noqa: E402


def main():
    wl, nseq, i, j = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    mode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    w, files = bench.make_workload(wl, nseq)
    store = bench.load_store(files, w["fmt"])
    flat = store.flat()
    ia, ib = np.array([i], np.int32), np.array([j], np.int32)
    lens = store.lengths()
    print("pair (%d,%d) lengths %d %d" % (i, j, lens[i], lens[j]))
    first = None
    for n in list(range(1, 40)) + list(range(40, 501, 10)):
        p = lb.Parameters()
        p.set(num_iterations=n, step_mode=mode)
        ctx = lb.Context(p)
        ctx.set_sequences(flat)
        rg = ctx.solve_pairs(ia, ib)
        g = lb.results_to_list(rg)[0]
        ctx.free(rg)
        ctx.close()
        ro, _ = ob.solve_pairs(p, flat, ia, ib, n_threads=1)
        o = lb.results_to_list(ro)[0]
        ob.free_results(ro)
        same = g.key() == o.key()
        if not same or n in (1, 2, 5, 10, 100, 500):
            print("n=%3d %s gpu lower %d upper %d it %d | oracle lower %d upper %d it %d" % (
                n, "same" if same else "DIFF", g.best_lower, g.best_upper, g.iterations, o.best_lower, o.best_upper, o.iterations))
        if not same and first is None:
            first = n
            if n > 40:
                break
    print("first differing iteration count:", first)


if __name__ == "__main__":
    main()
