#!/usr/bin/env python3
"""Developer tool (GPU box): three-way comparison on a sample of a bench workload --
GPU library vs CPU restatement (oracle) vs unmodified reference binary (oracle/_ref/lara_scalar).

usage: gpu_check.py [workload] [nseq] [extra lara options as key=value, e.g. num_iterations=50]
"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import bench  # noqa: E402
import lara_b200 as lb  # noqa: E402
from lara_b200 import _lib  # noqa: E402
from oracle import binding as ob  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
    nseq = int(sys.argv[2]) if len(sys.argv) > 2 else 24
    kw = {}
    for a in sys.argv[3:]:
        k, v = a.split("=")
        kw[k] = float(v) if "." in v else int(v)
    w, files = bench.make_workload(wl, nseq)
    store = lb.InputStorage.from_files(in_file=files[0])
    flat = store.flat()
    p = lb.Parameters()
    p.set(**kw)
    ia, ib = ob.all_pairs(len(store))
    ctx = lb.Context(p)
    ctx.set_sequences(flat)
    t0 = time.perf_counter()
    rg = ctx.solve_pairs(ia, ib)
    tg = time.perf_counter() - t0
    got = lb.results_to_list(rg)
    t0 = time.perf_counter()
    ro, _ = ob.solve_pairs(p, flat, ia, ib, n_threads=os.cpu_count())
    to = time.perf_counter() - t0
    want = lb.results_to_list(ro)
    nbad = 0
    for x, y in zip(got, want):
        if x.key() != y.key():
            nbad += 1
            if nbad <= 8:
                nl = sum(1 for a, b in zip(x.lines, y.lines) if a != b)
                print("DIFF pair (%d,%d): gpu lower=%d upper=%d it=%d lines=%d | oracle lower=%d upper=%d it=%d lines=%d | differing lines %d"
                      % (x.idx_low, x.idx_high, x.best_lower, x.best_upper, x.iterations, len(x.lines),
                         y.best_lower, y.best_upper, y.iterations, len(y.lines), nl))
                if nbad <= 3:
                    for k, (a, b) in enumerate(zip(x.lines, y.lines)):
                        if a != b:
                            print("    line %d: gpu %r oracle %r" % (k, a, b))
    its = [y.iterations for y in want]
    print("pairs=%d gpu_vs_oracle_mismatches=%d gpu %.2fs oracle %.2fs iterations min/mean/max %d/%.0f/%d"
          % (len(want), nbad, tg, to, min(its), sum(its) / len(its), max(its)))
    if not kw and os.path.exists(ob.REF_SCALAR):
        mine = files[0] + ".oracle.out"
        rc = _lib.load().lara_output_write(store._h, C.byref(ro), b"lib", mine.encode())
        assert rc == 0
        t0 = time.perf_counter()
        cp = ob.run_reference(ob.REF_SCALAR, ["-i", files[0], "-j", str(os.cpu_count()), "-w", files[0] + ".ref.out"])
        tr = time.perf_counter() - t0
        A = open(mine, "rb").read()
        B = open(files[0] + ".ref.out", "rb").read()
        print("reference rc=%d %.2fs; oracle output %s reference output (%d vs %d bytes)"
              % (cp.returncode, tr, "IDENTICAL to" if A == B else "DIFFERS from", len(A), len(B)))
        if A != B:
            la, lb_ = A.decode().splitlines(), B.decode().splitlines()
            k = 0
            for i, (x, y) in enumerate(zip(la, lb_)):
                if x != y:
                    print("  line %d: oracle %r ref %r" % (i + 1, x[:80], y[:80]))
                    k += 1
                    if k > 5:
                        break
            print(cp.stderr.decode()[-500:])
    ctx.free(rg)
    ob.free_results(ro)
    ctx.close()
    return 1 if nbad else 0


if __name__ == "__main__":
    sys.exit(main())
