"""GPU parity tests of the BAND kernels through the C ABI (-m gpu): banded fill + closed-form regions + traceback
against the oracle's full-matrix Gotoh on identical inputs.  Bit-exact scores, lines and gap scores."""
import numpy as np
import pytest

from tests import util  # noqa: F401

pytestmark = pytest.mark.gpu

GO, GE = -49152, -16384
INIT = -1431655764


def band_pssm(rng, n, la, lb, d_lo, d_hi, density, alphabet=None):
    p = np.full((n, la, lb), INIT, np.int32)
    a = np.arange(la)[:, None]; b = np.arange(lb)[None, :]
    inband = (a - b >= d_lo) & (a - b <= d_hi)
    for i in range(n):
        mask = inband & (rng.random((la, lb)) < density)
        if alphabet is None:
            v = rng.integers(-20000, 60000, size=(la, lb))
        else:
            v = rng.choice(alphabet, size=(la, lb))
        p[i][mask] = v[mask]
    return p


def check(ctx, pssm, d_lo, lanes, go=GO, ge=GE):
    from oracle import binding as ob
    score, nl, la_, lb_, gap = ctx.gotoh_band_batch(pssm, d_lo, lanes, go, ge)
    for i in range(pssm.shape[0]):
        s, a, b, g = ob.gotoh(pssm[i], go, ge)
        assert score[i] == s, "problem %d: score %d vs %d" % (i, score[i], s)
        assert nl[i] == len(a), "problem %d: %d lines vs %d" % (i, nl[i], len(a))
        assert np.array_equal(la_[i, :nl[i]], a) and np.array_equal(lb_[i, :nl[i]], b), "alignment %d differs" % i
        assert gap[i] == g


@pytest.mark.parametrize("la,lb,d_lo,lanes", [
    (1, 1, 0, 16), (2, 1, 0, 16), (7, 5, -2, 16), (40, 33, -8, 16), (64, 64, -16, 16), (150, 150, -12, 16),
    (300, 300, -14, 16), (300, 290, -10, 16), (97, 65, -20, 32), (300, 300, -30, 32), (330, 257, -6, 32),
    (600, 600, -32, 32)])
def test_band_random(gpu_ctx_factory, la, lb, d_lo, lanes):
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(la * 977 + lb + lanes)
    d_hi = d_lo + 2 * lanes - 1
    for density in (0.15, 0.6):
        check(ctx, band_pssm(rng, 5, la, lb, d_lo, d_hi, density), d_lo, lanes)


@pytest.mark.parametrize("la,lb,d_lo,lanes", [(12, 12, -4, 16), (70, 50, -10, 16), (150, 140, -8, 16), (90, 90, -20, 32)])
def test_band_ties(gpu_ctx_factory, la, lb, d_lo, lanes):
    """Small score alphabet: masses of ties, inside the band and against the closed-form regions."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(5 + la)
    d_hi = d_lo + 2 * lanes - 1
    alpha = (np.array([-2, -1, 0, 1, 2]) * 16384).astype(np.int64)
    p = band_pssm(rng, 8, la, lb, d_lo, d_hi, 0.5, alpha)
    check(ctx, p, d_lo, lanes)
    check(ctx, p, d_lo, lanes, go=-3 * 8192, ge=-1 * 8192)


@pytest.mark.parametrize("la,lb,d_lo,lanes", [(200, 60, 0, 16), (200, 60, -6, 16), (400, 100, -2, 32), (120, 119, 0, 16)])
def test_band_end_cell_outside_and_escapes(gpu_ctx_factory, la, lb, d_lo, lanes):
    """la - lb beyond the band: the end cell and the start of the traceback lie in region U; d_lo = 0 pushes
    vertical detours into region L."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(la + 3 * lb + d_lo)
    d_hi = d_lo + 2 * lanes - 1
    for density in (0.05, 0.3, 0.9):
        check(ctx, band_pssm(rng, 6, la, lb, d_lo, d_hi, density), d_lo, lanes)


def test_band_all_inactive(gpu_ctx_factory):
    ctx, _ = gpu_ctx_factory()
    p = np.full((2, 50, 37), INIT, np.int32)
    check(ctx, p, -4, 16)


def test_band_rejects_active_cell_outside(gpu_ctx_factory):
    import lara_b200 as lb
    ctx, _ = gpu_ctx_factory()
    p = np.full((1, 80, 80), INIT, np.int32)
    p[0, 70, 2] = 100
    with pytest.raises(lb.LaraError):
        ctx.gotoh_band_batch(p, -4, 16, GO, GE)
