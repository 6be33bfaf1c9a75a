#!/usr/bin/env python3
"""Developer tool (GPU box): histogram of the iterations the pairs of a bench workload actually run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np  # noqa: E402
import bench  # noqa: E402
import lara_b200 as lb  # noqa: E402
from lara_b200 import partition  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
nseq = int(sys.argv[2]) if len(sys.argv) > 2 else 100
w, files = bench.make_workload(wl, nseq)
store = lb.InputStorage.from_files(in_file=files[0])
ia, ib = partition.all_pairs(len(store))
ctx = lb.Context(lb.Parameters())
ctx.set_sequences(store.flat())
res = ctx.solve_pairs(ia.astype(np.int32), ib.astype(np.int32))
its = np.array([r.iterations for r in lb.results_to_list(res)])
h, edges = np.histogram(its, bins=[0, 25, 50, 100, 150, 200, 250, 300, 350, 400, 450, 499, 501])
print("pairs", len(its), "mean", its.mean())
for c, e0, e1 in zip(h, edges[:-1], edges[1:]):
    print("  [%3d,%3d): %6d  %.1f%%" % (e0, e1, c, 100.0 * c / len(its)))
