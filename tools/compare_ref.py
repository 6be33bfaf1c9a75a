#!/usr/bin/env python3
"""Developer tool: run the unmodified reference binary and the CPU restatement (oracle) on the same
input file(s) and compare the written output byte for byte.

usage: compare_ref.py [--simd] [--hostlib PATH] -- <lara args, e.g. -i x.ebpseq -n 50 -o pairs>
"""
import ctypes as C
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lara_b200 import _lib  # noqa: E402


def main():
    argv = sys.argv[1:]
    simd = "--simd" in argv
    hostlib = None
    if "--hostlib" in argv:
        hostlib = argv[argv.index("--hostlib") + 1]
    largs = argv[argv.index("--") + 1:]
    if hostlib:
        _lib.load(hostlib, symbols=dict(_lib.HOST_SYMBOLS, lara_params_default=_lib.GPU_SYMBOLS["lara_params_default"]))
    from lara_b200 import api
    from oracle import binding as ob
    P = api.Parameters(largs)
    assert P.status == 2, P.status
    P.c.step_mode = 1 if simd else 0
    store = api.InputStorage(P)
    flat = store.flat()
    ia, ib = ob.all_pairs(len(store))
    res, st = ob.solve_pairs(P, flat, ia, ib, n_threads=os.cpu_count())
    tmp = tempfile.mkdtemp(prefix="cmp")
    mine = os.path.join(tmp, "oracle.out")
    rc = _lib.load().lara_output_write(store._h, C.byref(res), P.out_format.encode(), mine.encode())
    assert rc == 0
    refout = os.path.join(tmp, "ref.out")
    a2 = [x for x in largs]
    if "-w" in a2:
        k = a2.index("-w"); del a2[k:k + 2]
    if "-g" in a2:
        k = a2.index("-g"); del a2[k:k + 2]
    cp = ob.run_reference(ob.REF_SIMD if simd else ob.REF_SCALAR, a2 + ["-w", refout, "-j", str(os.cpu_count())])
    if cp.returncode != 0:
        print("reference failed:", cp.stderr.decode()[-2000:])
        return 2
    A = open(mine, "rb").read()
    B = open(refout, "rb").read()
    its = [res.pairs[k].iterations for k in range(res.n_pairs)]
    print("pairs=%d oracle_ms=%.0f iters[min/mean/max]=%d/%.0f/%d E=%d D=%d" % (
        res.n_pairs, st.ms_total, min(its), sum(its) / len(its), max(its), st.n_active_edges, st.n_interactions))
    if A == B:
        print("IDENTICAL (%d bytes)" % len(A))
        return 0
    la, lb = A.decode().splitlines(), B.decode().splitlines()
    nd = 0
    for i, (x, y) in enumerate(zip(la, lb)):
        if x != y:
            nd += 1
            if nd <= 10:
                print("line %d:\n  oracle: %s\n  ref:    %s" % (i + 1, x[:200], y[:200]))
    print("DIFFERENT: %d differing lines, lengths %d vs %d (%s)" % (nd, len(la), len(lb), tmp))
    return 1


if __name__ == "__main__":
    sys.exit(main())
