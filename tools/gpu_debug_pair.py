#!/usr/bin/env python3
"""Developer tool (GPU box): for one pair of a workload sample, find the first iteration budget N at which the
GPU result differs from the oracle's."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import bench, lara_b200 as lb
from oracle import binding as ob

wl, nseq, i, j, nmax = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
w, files = bench.make_workload(wl, nseq)
store = lb.InputStorage.from_files(in_file=files[0])
flat = store.flat()
ia = np.array([i], np.int32); ib = np.array([j], np.int32)
for N in range(1, nmax + 1):
    p = lb.Parameters(); p.set(num_iterations=N)
    ctx = lb.Context(p); ctx.set_sequences(flat)
    rg = ctx.solve_pairs(ia, ib); got = lb.results_to_list(rg)[0]
    ro, _ = ob.solve_pairs(p, flat, ia, ib); want = lb.results_to_list(ro)[0]
    same = got.key() == want.key()
    print("N=%d same=%s gpu(l=%d u=%d it=%d) oracle(l=%d u=%d it=%d)" % (N, same, got.best_lower, got.best_upper, got.iterations,
          want.best_lower, want.best_upper, want.iterations))
    if not same:
        for k, (a, b) in enumerate(zip(got.lines, want.lines)):
            if a != b:
                lo = max(0, k - 2)
                print("  line %d: gpu %r oracle %r ; context gpu %r oracle %r" % (k, a, b, got.lines[lo:k + 3], want.lines[lo:k + 3]))
        break
    ctx.free(rg); ob.free_results(ro); ctx.close()
