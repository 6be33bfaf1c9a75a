"""GPU parity tests of the single kernels through the C ABI (-m gpu): DP fill + traceback and edge filter
against the CPU restatement (oracle) on identical inputs.  Bit-exact."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu

GO, GE = -49152, -16384


def _check_gotoh(ctx, pssm, go=GO, ge=GE):
    from oracle import binding as ob
    score, nl, la_, lb_, gap = ctx.gotoh_batch(pssm, go, ge)
    for i in range(pssm.shape[0]):
        s, a, b, g = ob.gotoh(pssm[i], go, ge)
        assert score[i] == s, "score of problem %d: %d vs %d" % (i, score[i], s)
        assert nl[i] == len(a)
        assert np.array_equal(la_[i, :nl[i]], a) and np.array_equal(lb_[i, :nl[i]], b), "alignment %d differs" % i
        assert gap[i] == g


@pytest.mark.parametrize("la,lb", [(1, 1), (2, 1), (5, 3), (33, 32), (40, 33), (64, 64), (97, 65), (150, 150),
                                   (200, 129), (300, 300), (330, 257), (600, 600), (640, 640)])
def test_gotoh_random_sparse(gpu_ctx_factory, la, lb):
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(la * 1000 + lb)
    _check_gotoh(ctx, util.random_pssm(rng, 6, la, lb, density=0.25))


@pytest.mark.parametrize("la,lb", [(12, 12), (70, 50), (150, 140)])
def test_gotoh_ties(gpu_ctx_factory, la, lb):
    """Small score alphabet -> masses of ties: exercises every left-operand-wins rule."""
    ctx, _ = gpu_ctx_factory()
    rng = np.random.default_rng(7 + la)
    p = (rng.integers(-2, 3, size=(8, la, lb)) * 16384).astype(np.int32)
    _check_gotoh(ctx, p)
    _check_gotoh(ctx, p, go=-3 * 8192, ge=-1 * 8192)
    # extension dearer than opening: the generic column-0 path
    _check_gotoh(ctx, p, go=-1 * 8192, ge=-3 * 8192)


def test_gotoh_all_inactive(gpu_ctx_factory):
    ctx, _ = gpu_ctx_factory()
    p = np.full((2, 50, 37), -1431655764, np.int32)
    _check_gotoh(ctx, p)


@pytest.mark.parametrize("la,lb,seed", [(1, 1, 0), (30, 20, 1), (150, 150, 2), (300, 290, 3), (300, 31, 4), (700, 520, 5)])
def test_edge_filter(gpu_ctx_factory, la, lb, seed):
    from oracle import binding as ob
    ctx, p = gpu_ctx_factory()
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 4, size=la).astype(np.uint8)
    b = a[:lb].copy() if lb <= la else rng.integers(0, 4, size=lb).astype(np.uint8)
    mut = rng.random(lb) < 0.3
    b[mut] = rng.integers(0, 5, size=int(mut.sum())).astype(np.uint8)
    act_g, opt_g = ctx.edge_filter(a, b)
    act_o, opt_o = ob.edge_filter(p, a, b)
    assert opt_g == opt_o
    assert np.array_equal(act_g, act_o)
