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
