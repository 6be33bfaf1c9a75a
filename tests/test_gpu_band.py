"""GPU parity tests of the BAND kernels through the C ABI (-m gpu): banded fill + closed-form regions + traceback
against the oracle's full-matrix Gotoh on identical inputs.  Bit-exact scores, lines and gap scores.
Band shapes: width W diagonals, G lanes per problem, blocks of Q = W / (2 G) diagonals (16 or 4)."""
import numpy as np
import pytest

from tests import util  # noqa: F401

pytestmark = pytest.mark.gpu

GO, GE = -49152, -16384
INIT = -1431655764

# (band width, lanes per problem)
SHAPES = [(32, 1), (64, 2), (128, 4), (256, 8), (512, 16), (1024, 32), (32, 4), (64, 8), (128, 16), (256, 32),
          (32, 2), (64, 4), (128, 8), (256, 16), (512, 32), (24, 1)]


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


def check(ctx, pssm, d_lo, width, lanes, go=GO, ge=GE):
    from oracle import binding as ob
    score, nl, la_, lb_, gap = ctx.gotoh_band_batch(pssm, d_lo, width, lanes, go, ge)
    for i in range(pssm.shape[0]):
        s, a, b, g = ob.gotoh(pssm[i], go, ge)
        assert score[i] == s, "problem %d: score %d vs %d" % (i, score[i], s)
        assert nl[i] == len(a), "problem %d: %d lines vs %d" % (i, nl[i], len(a))
        assert np.array_equal(la_[i, :nl[i]], a) and np.array_equal(lb_[i, :nl[i]], b), "alignment %d differs" % i
        assert gap[i] == g


@pytest.mark.parametrize("width,lanes", SHAPES)
@pytest.mark.parametrize("la,lb,frac", [(1, 1, 0.0), (2, 1, 0.0), (7, 5, 0.3), (40, 33, 0.5), (64, 64, 0.5), (150, 150, 0.4),
                                        (300, 300, 0.45), (300, 290, 0.3), (330, 257, 0.2)])
def test_band_random(gpu_ctx_factory, la, lb, frac, width, lanes):
    """frac: share of the band below the main diagonal (d_lo = -frac * width)."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(la * 977 + lb + width + lanes)
    d_lo = -int(frac * width)
    d_hi = d_lo + width - 1
    # more problems than one warp group holds (32 / lanes), so that slots and partial groups are covered
    n = 5 if lanes > 8 else 37
    if la * lb * n > 4e6:
        n = max(3, int(4e6 // (la * lb)))
    for density in (0.15, 0.6):
        check(ctx, band_pssm(rng, n, la, lb, d_lo, d_hi, density), d_lo, width, lanes)


@pytest.mark.parametrize("width,lanes", [(32, 1), (24, 1), (32, 2), (32, 4), (64, 2), (64, 8)])
@pytest.mark.parametrize("la,lb,d_lo", [(12, 12, -4), (70, 50, -10), (150, 140, -8), (90, 90, -20)])
def test_band_ties(gpu_ctx_factory, la, lb, d_lo, width, lanes):
    """Small score alphabet: masses of ties, inside the band and against the closed-form regions."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(5 + la)
    d_hi = d_lo + width - 1
    alpha = (np.array([-2, -1, 0, 1, 2]) * 16384).astype(np.int64)
    p = band_pssm(rng, 8, la, lb, d_lo, d_hi, 0.5, alpha)
    check(ctx, p, d_lo, width, lanes)
    check(ctx, p, d_lo, width, lanes, go=-3 * 8192, ge=-1 * 8192)


@pytest.mark.parametrize("width,lanes", [(32, 1), (24, 1), (32, 2), (32, 4), (64, 2), (64, 8)])
@pytest.mark.parametrize("la,lb,d_lo", [(200, 60, 0), (200, 60, -6), (400, 100, -2), (120, 119, 0), (90, 80, -31)])
def test_band_end_cell_outside_and_escapes(gpu_ctx_factory, la, lb, d_lo, width, lanes):
    """la - lb beyond the band: the end cell and the start of the traceback lie in region U; d_lo = 0 pushes
    vertical detours into region L; d_lo = -31 stands for "main diagonal at the band's upper edge"."""
    if d_lo == -31:
        d_lo = -(width - 1)
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(la + 3 * lb + d_lo)
    d_hi = d_lo + width - 1
    for density in (0.05, 0.3, 0.9):
        check(ctx, band_pssm(rng, 6, la, lb, d_lo, d_hi, density), d_lo, width, lanes)


def test_band_long(gpu_ctx_factory):
    """Longer than the single-pass full-matrix kernel allows (lenB > 640)."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(77)
    la, lb, d_lo, width = 1000, 900, -40, 256
    p = band_pssm(rng, 3, la, lb, d_lo, d_lo + width - 1, 0.3)
    check(ctx, p, d_lo, width, 8)
    check(ctx, p, d_lo, width, 32)


def test_band_all_inactive(gpu_ctx_factory):
    ctx, _ = gpu_ctx_factory()
    p = np.full((2, 50, 37), INIT, np.int32)
    check(ctx, p, -4, 32, 1)
    check(ctx, p, -4, 32, 4)


def test_band_rejects_active_cell_outside(gpu_ctx_factory):
    import lara_b200 as lb
    ctx, _ = gpu_ctx_factory()
    p = np.full((1, 80, 80), INIT, np.int32)
    p[0, 70, 2] = 100
    with pytest.raises(lb.LaraError):
        ctx.gotoh_band_batch(p, -4, 32, 1, GO, GE)
