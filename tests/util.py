"""Helpers shared by the tests: synthetic inputs (same generator as the benchmark), result comparison."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_inputs  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def make_input(seed, n, length, fmt="ebpseq", lenmax=0, related=True, skew=False):
    d = tempfile.mkdtemp(prefix="lara_t_")
    return gen_inputs.generate(d, seed, n, length, fmt, lenmax=lenmax, related=related, skew=skew)


def load_store(files):
    import lara_b200 as lb
    if files[0].endswith("_dp.ps"):
        return lb.InputStorage.from_files(dotplots=files)
    return lb.InputStorage.from_files(in_file=files[0])


def results_equal(a, b):
    """a, b: lists of PairResult.  Bit-exact comparison of everything solve() reports."""
    if len(a) != len(b):
        return "different number of pairs: %d vs %d" % (len(a), len(b))
    for x, y in zip(a, b):
        if x.key() != y.key():
            return "pair (%d,%d): %r vs %r" % (x.idx_low, x.idx_high, x, y)
    return None


def random_pssm(rng, n, la, lb, density=0.3, lo=-20000, hi=60000):
    """Dense position specific matrices with inactive cells at the reference's INIT value."""
    p = rng.integers(lo, hi, size=(n, la, lb), dtype=np.int64).astype(np.int32)
    mask = rng.random((n, la, lb)) < density
    p[~mask] = -1431655764
    return p


# ---- golden fixtures (tests/golden/manifest.json: outputs of the unmodified reference binaries) ----------
def golden_cases():
    import json
    with open(os.path.join(GOLDEN, "manifest.json")) as f:
        return json.load(f)["cases"]


def golden_setup(case):
    """-> (Parameters, InputStorage) for one manifest entry, parsed by the host CLI mirror exactly like the
    reference parsed its command line."""
    import lara_b200 as lb
    files = [os.path.join(GOLDEN, f) for f in case["files"]]
    argv = []
    if case["format"] == "dp":
        for f in files:
            argv += ["-d", f]
    else:
        argv += ["-i", files[0]]
    if case.get("ref_file"):
        argv += ["-r", os.path.join(GOLDEN, case["ref_file"])]
    argv += [os.path.join(GOLDEN, o[len("@golden/"):]) if o.startswith("@golden/") else o for o in case["options"]]
    if case["build"] == "simd":
        argv += ["-g", "1"]
    p = lb.Parameters(argv)
    assert p.status == lb.Parameters.CONTINUE
    return p, lb.InputStorage(p)


def write_results(store, res, fmt):
    """lara_output_write into a temp file -> bytes."""
    import ctypes as C
    from lara_b200 import _lib
    out = os.path.join(tempfile.mkdtemp(prefix="lara_o_"), "out.txt")
    rc = _lib.load().lara_output_write(store._h, C.byref(res), fmt.encode(), out.encode())
    assert rc == 0, _lib.load().lara_host_last_error()
    with open(out, "rb") as f:
        return f.read()
