"""CPU test of the band-DP specification (tools/band_proto.py): closed forms for the pure gap regions outside the
band reproduce the full matrix (values, trace decisions, traceback lines), and agree with the oracle's Gotoh."""
import os
import sys

import numpy as np
import pytest

from tests import util  # noqa: F401

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import band_proto as bp  # noqa: E402


@pytest.mark.parametrize("seed", range(6))
def test_band_closed_forms_equal_full_matrix(seed):
    rng = np.random.default_rng(100 + seed)
    for it in range(40):
        la = int(rng.integers(1, 36)); lb = int(rng.integers(1, la + 1))
        dlo = -int(rng.integers(0, 6)); dhi = int(rng.integers(0, 8))
        go, ge = [(-49152, -16384), (-3 * 8192, -8192), (-20000, -19999)][it % 3]
        alpha = None if it % 2 else [-16384, 0, 16384, 32768]
        msg = bp.check_one(rng, la, lb, dlo, dhi, go, ge, float(rng.uniform(0.1, 0.95)), alpha)
        assert msg is None, "la=%d lb=%d band=[%d,%d] go=%d: %s" % (la, lb, dlo, dhi, go, msg)


@pytest.mark.parametrize("seed", range(4))
def test_sheared_band_closed_forms_equal_full_matrix(seed):
    """Specification for the next band kernel: a band whose diagonals move right with the rows
    (dlo[r], dhi[r] non-decreasing, shift < width) keeps the same closed forms for the gap regions: values, trace
    decisions, end values and traceback lines equal the full matrix for pairs with lenA well above lenB."""
    rng = np.random.default_rng(300 + seed)
    for it in range(40):
        lb = int(rng.integers(1, 30)); la = lb + int(rng.integers(0, 40))
        dlo, dhi = bp.sheared_bounds(rng, la, lb, int(rng.integers(2, 10)))
        go, ge = [(-49152, -16384), (-3 * 8192, -8192), (-20000, -19999)][it % 3]
        alpha = None if it % 2 else [-16384, 0, 16384, 32768]
        msg = bp.check_one(rng, la, lb, dlo, dhi, go, ge, float(rng.uniform(0.1, 0.95)), alpha)
        assert msg is None, "la=%d lb=%d go=%d dlo=%s: %s" % (la, lb, go, list(dlo), msg)


def test_band_walk_equals_oracle_gotoh(built):
    from oracle import binding as ob
    rng = np.random.default_rng(9)
    escapes = 0
    for it in range(60):
        la = int(rng.integers(2, 50)); lb = int(rng.integers(1, la + 1))
        dlo = -int(rng.integers(0, 5)); dhi = int(rng.integers(0, 6))
        go, ge = (-49152, -16384)
        pssm = np.full((la, lb), bp.PINF, np.int64)
        mask = rng.random((la, lb)) < 0.5
        for a in range(la):
            for b in range(lb):
                if dlo <= a - b <= dhi and mask[a, b]:
                    pssm[a, b] = int(rng.integers(-20000, 60000))
        band = bp.Band(la, lb, pssm, go, ge, dlo, dhi)
        sE, vE, hE = band.end_values()
        seen = []

        def nib_at(c, r):
            seen.append(c - r)
            return band.nib_at(c, r)
        lines = bp.walk(la, lb, nib_at, vE, hE, sE)
        escapes += any(d > dhi or d < dlo for d in seen)
        p32 = np.where(pssm == bp.PINF, -1431655764, pssm).astype(np.int32)
        sc, oa, obb, _ = ob.gotoh(p32, go, ge)
        assert sc == sE
        assert [(int(x), int(y)) for x, y in zip(oa, obb)] == lines
    assert escapes > 10          # the walk really left the band in many cases
