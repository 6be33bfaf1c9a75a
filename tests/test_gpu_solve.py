"""GPU parity tests of the whole hot path through the C ABI (-m gpu): lara_gpu_solve_* against the CPU
restatement on the same seeded inputs.  Bit-exact: lines, library weights, bounds and iteration counts."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _solve_both(files, n_threads=8, pairs=None, **kw):
    import lara_b200 as lb
    from oracle import binding as ob
    store = util.load_store(files)
    p = lb.Parameters()
    p.set(**kw)
    flat = store.flat()
    ia, ib = ob.all_pairs(len(store)) if pairs is None else pairs
    ctx = lb.Context(p)
    ctx.set_sequences(flat)
    rg = ctx.solve_pairs(ia, ib)
    got = lb.results_to_list(rg)
    st = ctx.stats()
    ctx.free(rg)
    ro, _ = ob.solve_pairs(p, flat, ia, ib, n_threads=n_threads)
    want = lb.results_to_list(ro)
    ob.free_results(ro)
    ctx.close()
    return got, want, st


@pytest.mark.parametrize("kw", [
    dict(num_iterations=60, max_nondecreasing=8),
    dict(num_iterations=40, max_nondecreasing=5, step_mode=1),
    dict(num_iterations=40, matching=3),
    dict(num_iterations=40, matching=1, balance=0.7),
    dict(num_iterations=40, sequence_scale=1.5, step_size_factor=1.7),
    dict(num_iterations=40, lib_score_linear=1, lib_score_min=100, lib_score_max=900),
])
def test_solve_ebpseq_small(built, kw):
    files = util.make_input(seed=31, n=7, length=60, lenmax=90)
    got, want, st = _solve_both(files, **kw)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)
    assert st.dp_launches > 0 and st.kernel_launches > st.dp_launches


def test_solve_dbn_early_termination(built):
    files = util.make_input(seed=32, n=8, length=70, lenmax=110, fmt="dbn", related=False)
    got, want, _ = _solve_both(files)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)
    assert min(r.iterations for r in got) < 500


def test_solve_unrelated_lengths(built):
    files = util.make_input(seed=33, n=8, length=35, lenmax=160, related=False)
    got, want, _ = _solve_both(files, num_iterations=50, max_nondecreasing=6)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_dotplots(built):
    files = util.make_input(seed=34, n=4, length=100, lenmax=120, fmt="dp")
    got, want, _ = _solve_both(files, num_iterations=60)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_default_500_iterations(built):
    files = util.make_input(seed=35, n=6, length=120)
    got, want, _ = _solve_both(files)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_300nt(built):
    files = util.make_input(seed=36, n=5, length=300)
    got, want, _ = _solve_both(files, num_iterations=80, max_nondecreasing=10)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_long_sequences(built):
    """Shorter sequence beyond the single-pass full-matrix kernel (640 nt): band kernels only (BASELINE config C4)."""
    files = util.make_input(seed=37, n=3, length=800, lenmax=900, fmt="dp")
    got, want, st = _solve_both(files, num_iterations=30, max_nondecreasing=5)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)
    assert st.reserved[1] == 3                      # all three pairs ran banded


def test_solve_long_unequal_pair(built):
    """A pair whose shorter sequence exceeds the single-pass full-matrix kernel (640 nt) and whose active edges span more
    diagonals than the widest band (1,024): 2,100 x 680 nt runs on the row-by-row kernel; its neighbours in the same
    call take bands."""
    import lara_b200 as lb
    rng = np.random.default_rng(44)
    s = lb.InputStorage()
    for name, L in (("long", 2100), ("mid", 680), ("short", 660)):
        seq = "".join("ACGU"[i] for i in rng.integers(0, 4, size=L))
        pairs = []
        used = set()
        for _ in range(L // 6):
            i, j = sorted(int(x) for x in rng.integers(0, L, size=2))
            if j - i < 4 or i in used or j in used:
                continue
            used.add(i); used.add(j)
            pairs.append((i, j, float(rng.uniform(1.0, 5.5))))
        s.add_record(name, seq, pairs)
    s.finish()
    got, want = _solve_store(s, num_iterations=10, max_nondecreasing=3)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


@pytest.mark.parametrize("kw,env", [
    (dict(gap_open=-3 * 8192, gap_extend=-3 * 8192), {}),                                # scalar build: SeqAn's linear gap costs
    (dict(gap_open=-3 * 8192, gap_extend=-3 * 8192, step_mode=1), {}),                   # SIMD build: affine, everything ties
    (dict(gap_open=-3 * 8192, gap_extend=-3 * 8192, step_mode=1), dict(LARA_B200_ROWS="1")),
    (dict(gap_open=-2 * 8192, gap_extend=-4 * 8192), {}),                                # opening cheaper than extending
    (dict(gap_open=-2 * 8192, gap_extend=-4 * 8192), dict(LARA_B200_ROWS="1")),
    (dict(gap_open=-2 * 8192, gap_extend=-4 * 8192, step_mode=1), dict(LARA_B200_ROWS="1")),
])
def test_solve_unusual_gap_costs(built, kw, env, monkeypatch):
    """Gap open >= gap extend: no band kernels (their closed forms need opening to be dearer than extending); linear
    gap costs when the scalar reference would switch to Needleman-Wunsch, the three forms of the row-by-row kernel."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    files = util.make_input(seed=46, n=9, length=40, lenmax=120)
    got, want, _ = _solve_both(files, num_iterations=70, max_nondecreasing=8, **kw)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_negative_step(built):
    """SIMD-build arithmetic on skewed lengths for many iterations: the wrapping int32 product makes negative steps, so
    multipliers of +1 entries go up and those of -1 entries go down (every touched row is rescanned or raised by what
    its multiplier really did)."""
    files = util.make_input(seed=45, n=16, length=0, skew=True)
    got, want, _ = _solve_both(files, num_iterations=220, max_nondecreasing=10, step_mode=1)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_skewed_lengths(built):
    """Lognormal lengths 50..600 (BASELINE config C5): mixed band widths and launch classes in one batch."""
    files = util.make_input(seed=38, n=14, length=0, skew=True)
    got, want, _ = _solve_both(files, num_iterations=40, max_nondecreasing=6)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


@pytest.mark.parametrize("env", [
    dict(LARA_B200_BAND="0"),
    dict(LARA_B200_BANDQ="4"),
    dict(LARA_B200_BANDQ="8"),
    dict(LARA_B200_BANDQ="16"),
    dict(LARA_B200_BANDQ="16", LARA_B200_BATCHCAP="20", LARA_B200_WORKERPAIRS="20", LARA_B200_WORKERS="3"),
    dict(LARA_B200_BANDMAXW="32", LARA_B200_BATCH="33"),
    dict(LARA_B200_BANDMINWARPS="1"),                   # lets the 24-diagonal one-thread shape in on a small batch
    dict(LARA_B200_BAND24="0", LARA_B200_BANDMINWARPS="1"),
    dict(LARA_B200_BAND="0", LARA_B200_ROWS="1"),       # row-by-row full-matrix kernel (the fallback for pairs of any size)
    dict(LARA_B200_BANDMAXW="32", LARA_B200_ROWS="1", LARA_B200_FUSED="1"),
    dict(LARA_B200_SPLIT="1", LARA_B200_SPLITWARPS="1", LARA_B200_BANDMINWARPS="1", LARA_B200_BANDCLASSMINWARPS="1"),   # every class its own launch
    dict(LARA_B200_FUSED="1", LARA_B200_BANDMINWARPS="1", LARA_B200_BANDCLASSMINWARPS="1"),        # two-rows-at-a-time body
    dict(LARA_B200_BANDMINWARPS="100000"),              # low-latency form: blocks of 8 diagonals, twice the lanes per pair
    dict(LARA_B200_BANDMINWARPS="1", LARA_B200_BATCHCAP="16", LARA_B200_REFILL="2"),   # 16 slots refilled as pairs finish
])
def test_solve_kernel_variants(built, env, monkeypatch):
    """Every DP kernel family and the scheduling switches give the same bits: full matrix, band blocks of 4 / 8 / 16
    diagonals, small resident sets refilled from the queue on concurrent workers, band / full-matrix mixes."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    files = util.make_input(seed=39, n=12, length=90, lenmax=200)
    got, want, _ = _solve_both(files, num_iterations=120, max_nondecreasing=10)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


@pytest.mark.parametrize("env,workers", [
    (dict(LARA_B200_BATCHCAP="30", LARA_B200_WORKERPAIRS="60", LARA_B200_REFILL="2"), 2),
    (dict(LARA_B200_BATCHCAP="24", LARA_B200_WORKERPAIRS="40", LARA_B200_WORKERS="3", LARA_B200_REFILL="2"), 3),
    (dict(LARA_B200_BATCHCAP="50", LARA_B200_WORKERPAIRS="1000", LARA_B200_REFILL="2"), 1),
    (dict(LARA_B200_BATCHCAP="20", LARA_B200_WORKERPAIRS="1000", LARA_B200_REFILL="1"), 1),      # eager first, set by set at the end
    (dict(LARA_B200_BATCHCAP="40", LARA_B200_WORKERPAIRS="80", LARA_B200_REFILL="0"), 2),        # set by set only
    (dict(LARA_B200_MEMBUDGET="6", LARA_B200_WORKERPAIRS="80", LARA_B200_REFILL="2"), 2),       # cohorts limited by the memory budget
    (dict(LARA_B200_MEMBUDGET="4", LARA_B200_WORKERPAIRS="1000", LARA_B200_REFILL="1"), 1),
    (dict(LARA_B200_MEMBUDGET="3", LARA_B200_WORKERPAIRS="1000", LARA_B200_REFILL="0"), 1),
    # every third pool allocation is told "out of memory" once: drain the device, trim the pool, allocate again
    (dict(LARA_B200_BATCHCAP="30", LARA_B200_WORKERPAIRS="60", LARA_B200_REFILL="2", LARA_B200_FAILALLOC="3"), 2),
])
def test_solve_many_batches_per_worker(built, env, workers, monkeypatch):
    """Concurrent workers whose resident sets (24..50 slots, or what fits 3..6 MB) are refilled many times from the
    shared queue while they iterate: 231 pairs on 1..3 workers.  Pairs of different cohorts iterate side by side in one slot table, new
    pairs take the band-group slots of finished ones."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    files = util.make_input(seed=43, n=22, length=70, lenmax=150)
    got, want, st = _solve_both(files, num_iterations=60, max_nondecreasing=8)
    assert util.results_equal(got, want) is None, util.results_equal(got, want)
    assert st.reserved[0] == workers
    if "LARA_B200_BATCHCAP" in env:
        assert st.reserved[6] >= 231 // int(env["LARA_B200_BATCHCAP"])      # cohorts: first fill + refills
    assert st.reserved[6] >= 3 * workers
    assert st.reserved[12] > 0                                               # pairs taken into a running or drained set
    assert st.reserved[9] == st.sum_iterations and 0 < st.reserved[9] <= st.reserved[8]


def _solve_store(store, **kw):
    import lara_b200 as lb
    from oracle import binding as ob
    p = lb.Parameters()
    p.set(**kw)
    flat = store.flat()
    ia, ib = ob.all_pairs(len(store))
    ctx = lb.Context(p)
    ctx.set_sequences(flat)
    rg = ctx.solve_pairs(ia, ib)
    got = lb.results_to_list(rg)
    ctx.free(rg)
    ro, _ = ob.solve_pairs(p, flat, ia, ib, n_threads=4)
    want = lb.results_to_list(ro)
    ob.free_results(ro)
    ctx.close()
    return got, want


def test_solve_edge_cases(built):
    """Degenerate records in one batch: lengths 1..4, no base pairs at all, a single base pair, N residues,
    identical sequences, a long sequence against a very short one (end cell far outside every band)."""
    import lara_b200 as lb
    rng = np.random.default_rng(41)
    s = lb.InputStorage()
    long_seq = "".join("ACGU"[i] for i in rng.integers(0, 4, size=260))
    recs = [
        ("one", "A", []),
        ("two", "GC", []),
        ("four", "ACGU", [(0, 3, 2.0)]),
        ("nopairs", "ACGUACGUAAAAGGGCCC", []),
        ("onepair", "GGGAAACCC", [(0, 8, 5.0)]),
        ("withN", "ACGNNNUACGNACGU", [(0, 14, 3.0), (1, 13, 2.5)]),
        ("hairpinA", "GGGGCAAAAGCCCCAUAU", [(0, 13, 4.0), (1, 12, 4.0), (2, 11, 4.0), (3, 10, 3.0)]),
        ("hairpinB", "GGGGCAAAAGCCCCAUAU", [(0, 13, 4.0), (1, 12, 4.0), (2, 11, 4.0), (3, 10, 3.0)]),
        ("long", long_seq, [(10, 60, 4.0), (11, 59, 4.0), (12, 58, 4.0), (100, 180, 3.0), (101, 179, 3.0), (200, 250, 2.0)]),
    ]
    for name, seq, pairs in recs:
        s.add_record(name, seq, pairs)
    s.finish()
    got, want = _solve_store(s, num_iterations=60, max_nondecreasing=7)
    assert len(got) == len(recs) * (len(recs) - 1) // 2
    assert util.results_equal(got, want) is None, util.results_equal(got, want)


def test_solve_round_trip_property(built):
    """Size-independent properties at BASELINE size (300 nt): every reported line is an in-range, strictly
    increasing position pair, weights are 500 or 1000, iterations within the limit; a pair that stopped early has
    bestLower == bestUpper.  (bestLower <= bestUpper is NOT a property of the reference: its multiplier update
    truncates to int32, and the oracle reproduces the occasional crossing bit for bit -- checked below.)"""
    import lara_b200 as lb
    files = util.make_input(seed=42, n=40, length=300)
    store = util.load_store(files)
    p = lb.Parameters()
    ctx = lb.Context(p)
    ctx.set_sequences(store.flat())
    from lara_b200 import partition
    ia, ib = partition.all_pairs(len(store))
    rg = ctx.solve_pairs(ia.astype(np.int32), ib.astype(np.int32))
    res = lb.results_to_list(rg)
    ctx.free(rg)
    ctx.close()
    assert len(res) == 780
    crossed = [r for r in res if r.best_lower > r.best_upper]
    if crossed:
        from oracle import binding as ob
        ca = np.array([r.idx_low for r in crossed[:8]], np.int32); cb = np.array([r.idx_high for r in crossed[:8]], np.int32)
        ro, _ = ob.solve_pairs(p, store.flat(), ca, cb, n_threads=8)
        want = lb.results_to_list(ro)
        ob.free_results(ro)
        assert util.results_equal(crossed[:8], want) is None, util.results_equal(crossed[:8], want)
    for r in res:
        assert 1 <= r.iterations <= 500
        assert r.iterations == 500 or r.best_lower == r.best_upper
        lo = [x[0] for x in r.lines]; hi = [x[1] for x in r.lines]
        assert all(a < b for a, b in zip(lo, lo[1:])) and all(a < b for a, b in zip(hi, hi[1:]))
        assert not lo or (0 <= lo[0] and lo[-1] < 300 and 0 <= hi[0] and hi[-1] < 300)
        assert all(x[2] in (500, 1000) for x in r.lines)


@pytest.mark.parametrize("kw", [
    dict(),
    dict(balance=0.35, sequence_scale=1.7),
    dict(balance=1.9, suboptimal_diff=12.5),
    dict(sequence_scale=0.4, suboptimal_diff=75.0, step_mode=1),
])
@pytest.mark.parametrize("fmt", ["ebpseq", "dbn"])
def test_lagrange_constructor(built, kw, fmt):
    """Row a4 on its own (Lagrange::Lagrange, src/lagrange.hpp:233-309): after ONE iteration nothing but the edge filter
    and the constructor has shaped the result.  The number of active edges and of interactions (dual dimension) of every
    pair set, the first upper bound (Gotoh over the initial queue tops seqScore + max structScore, i.e. every
    structScore in the reference's float order) and the first lower bound and lines must equal the CPU restatement's, for
    balance / scale / filter parameters away from their defaults."""
    import lara_b200 as lb
    from oracle import binding as ob
    files = util.make_input(seed=71, n=9, length=60, fmt=fmt, lenmax=170)
    store = util.load_store(files)
    p = lb.Parameters()
    p.set(num_iterations=1, max_nondecreasing=1, **kw)
    flat = store.flat()
    ia, ib = ob.all_pairs(len(store))
    ctx = lb.Context(p)
    ctx.set_sequences(flat)
    rg = ctx.solve_pairs(ia, ib)
    got = lb.results_to_list(rg)
    st = ctx.stats()
    ctx.free(rg)
    ctx.close()
    ro, so = ob.solve_pairs(p, flat, ia, ib, n_threads=4)
    want = lb.results_to_list(ro)
    ob.free_results(ro)
    assert st.n_active_edges == so.n_active_edges and st.n_active_edges > 0
    assert st.n_interactions == so.n_interactions and st.n_interactions > 0
    assert util.results_equal(got, want) is None, util.results_equal(got, want)
    assert all(r.iterations == 1 for r in got)


def test_context_reused_with_other_inputs(built):
    """One context, three input sets in a row (large, small, large again): the device arrays of the inputs and of the
    result slots are kept between calls when they are large enough, and every call must still give the oracle's bits."""
    import lara_b200 as lb
    from oracle import binding as ob
    p = lb.Parameters()
    p.set(num_iterations=30, max_nondecreasing=6)
    ctx = lb.Context(p)
    for seed, n, length, lenmax in [(81, 10, 80, 160), (82, 4, 30, 50), (83, 11, 90, 170), (82, 4, 30, 50)]:
        store = util.load_store(util.make_input(seed=seed, n=n, length=length, lenmax=lenmax))
        flat = store.flat()
        ia, ib = ob.all_pairs(len(store))
        ctx.set_sequences(flat)
        rg = ctx.solve_pairs(ia, ib)
        got = lb.results_to_list(rg)
        ctx.free(rg)
        ro, _ = ob.solve_pairs(p, flat, ia, ib, n_threads=4)
        want = lb.results_to_list(ro)
        ob.free_results(ro)
        assert util.results_equal(got, want) is None, (seed, util.results_equal(got, want))
    ctx.close()
