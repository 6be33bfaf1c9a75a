"""Host-side mirror of the reference's interface around the hot path.

    reference (C++)                                   here
    ------------------------------------------------  ---------------------------------
    lara::Parameters(argc, argv)   parameters.hpp:95   Parameters(argv)
    lara::InputStorage(params)     io.hpp:74           InputStorage(params) / .from_files(...)
    lara::OutputLibrary(store,fmt) io.hpp:376          OutputLibrary(store, fmt)
    lara::solve(results, store, params)                solve(results, store, params)
                                   subgradient_solver.hpp:110
    OutputLibrary::print(file)     io.hpp:465          OutputLibrary.print(file)

Everything heavy happens in liblara_b200.so (C ABI, include/lara_b200.h): the readers/writers are host
C++, solve() runs the CUDA kernels.  No CPU fallback exists.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import LaraCliOptions, LaraParams, LaraResults, LaraStats


class LaraError(RuntimeError):
    pass


def _host_check(rc):
    if rc != 0:
        raise LaraError("lara host error %d: %s" % (rc, _lib.load().lara_host_last_error().decode()))


class Parameters:
    """lara::Parameters.  `argv` excludes the program name, e.g. ["-i", "x.ebpseq", "-n", "50"]."""

    CONTINUE, EXIT_OK, EXIT_ERROR = 2, 0, 1

    def __init__(self, argv=None):
        lib = _lib.load()
        self.c = LaraParams()
        self.opts = LaraCliOptions()
        if argv is None:
            lib.lara_params_default(C.byref(self.c))
            self.status = self.CONTINUE
            self.opts.out_format = b"lib"
            self.opts.threads = 1
        else:
            args = [b"lara_b200"] + [a.encode() if isinstance(a, str) else a for a in argv]
            arr = (C.c_char_p * len(args))(*args)
            self.status = lib.lara_parse_args(len(args), arr, C.byref(self.c), C.byref(self.opts))

    # convenient attribute access to the POD fields
    def __getattr__(self, k):
        c = self.__dict__.get("c")
        if c is not None and k in dict(LaraParams._fields_):
            return getattr(c, k)
        raise AttributeError(k)

    def set(self, **kw):
        for k, v in kw.items():
            setattr(self.c, k, v)
        return self

    @property
    def in_file(self):
        return self.opts.in_file.decode()

    @property
    def ref_file(self):
        return self.opts.ref_file.decode()

    @property
    def out_file(self):
        return self.opts.out_file.decode()

    @property
    def out_format(self):
        return self.opts.out_format.decode()

    @property
    def dotplot_files(self):
        return [self.opts.dotplot_files[k].decode() for k in range(self.opts.n_dotplots)]

    @property
    def threads(self):
        return self.opts.threads


class InputStorage:
    """lara::InputStorage: the list of records plus the flat arrays solve() consumes."""

    def __init__(self, params=None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        _host_check(self._lib.lara_input_create(C.byref(self._h)))
        self._finished = False
        if params is not None:
            self.read_file(params.in_file)
            self.read_file(params.ref_file)
            for f in params.dotplot_files:
                self.read_dotplot(f)
            self.finish()

    @classmethod
    def from_files(cls, in_file="", ref_file="", dotplots=()):
        s = cls()
        s.read_file(in_file)
        s.read_file(ref_file)
        for f in dotplots:
            s.read_dotplot(f)
        s.finish()
        return s

    def __del__(self):
        try:
            if self._h:
                self._lib.lara_input_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def read_file(self, path):
        if path:
            _host_check(self._lib.lara_input_read_file(self._h, path.encode()))

    def read_dotplot(self, path):
        _host_check(self._lib.lara_input_read_dotplot(self._h, path.encode()))

    def add_record(self, name, seq, pairs=(), is_bpp=True):
        n = len(pairs)
        pi = np.ascontiguousarray([p[0] for p in pairs], dtype=np.int32)
        pj = np.ascontiguousarray([p[1] for p in pairs], dtype=np.int32)
        w = np.ascontiguousarray([p[2] for p in pairs], dtype=np.float64)
        _host_check(self._lib.lara_input_add_record(self._h, name.encode(), seq.encode(), n,
                                                    pi.ctypes.data, pj.ctypes.data, w.ctypes.data, 1 if is_bpp else 0))

    def finish(self):
        _host_check(self._lib.lara_input_finish(self._h))
        self._finished = True

    def __len__(self):
        return self._lib.lara_input_size(self._h)

    def name(self, i):
        return self._lib.lara_input_name(self._h, i).decode()

    def length(self, i):
        return self._lib.lara_input_length(self._h, i)

    def lengths(self):
        return np.array([self.length(i) for i in range(len(self))], dtype=np.int64)

    def flat(self):
        """(seq_cat u8, seq_off i64, bp_partner i32, bp_weight i32, bp_off i64) as numpy copies."""
        ptrs = [C.c_void_p() for _ in range(5)]
        _host_check(self._lib.lara_input_flatten(self._h, *[C.byref(p) for p in ptrs]))
        n = len(self)
        seq_off = np.ctypeslib.as_array(C.cast(ptrs[1], C.POINTER(C.c_int64)), shape=(n + 1,)).copy()
        total = int(seq_off[-1])
        seq_cat = np.ctypeslib.as_array(C.cast(ptrs[0], C.POINTER(C.c_uint8)), shape=(max(total, 1),))[:total].copy()
        bp_off = np.ctypeslib.as_array(C.cast(ptrs[4], C.POINTER(C.c_int64)), shape=(total + 1,)).copy()
        nb = int(bp_off[-1])
        if nb:
            bp_partner = np.ctypeslib.as_array(C.cast(ptrs[2], C.POINTER(C.c_int32)), shape=(nb,)).copy()
            bp_weight = np.ctypeslib.as_array(C.cast(ptrs[3], C.POINTER(C.c_int32)), shape=(nb,)).copy()
        else:
            bp_partner = np.zeros(0, np.int32)
            bp_weight = np.zeros(0, np.int32)
        return seq_cat, seq_off, bp_partner, bp_weight, bp_off


class PairResult:
    __slots__ = ("idx_low", "idx_high", "best_lower", "best_upper", "iterations", "lines")

    def __repr__(self):
        return "PairResult(%d,%d lower=%d upper=%d it=%d lines=%d)" % (
            self.idx_low, self.idx_high, self.best_lower, self.best_upper, self.iterations, len(self.lines))

    def key(self):
        return (self.idx_low, self.idx_high, self.best_lower, self.best_upper, self.iterations,
                tuple(map(tuple, self.lines)))


def results_to_list(res):
    """Copy a LaraResults struct into Python objects (list of PairResult)."""
    out = []
    for k in range(res.n_pairs):
        r = res.pairs[k]
        p = PairResult()
        p.idx_low, p.idx_high = r.idx_low, r.idx_high
        p.best_lower, p.best_upper, p.iterations = r.best_lower, r.best_upper, r.iterations
        o = r.lines_off
        p.lines = [(res.line_pos_low[o + x], res.line_pos_high[o + x], res.line_weight[o + x])
                   for x in range(r.n_lines)]
        out.append(p)
    return out


class Context:
    """Owner of a lara_ctx (device state).  One per process per GPU set."""

    def __init__(self, params, device_ids=None):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        pc = params.c if isinstance(params, Parameters) else params
        if device_ids:
            arr = (C.c_int * len(device_ids))(*device_ids)
            rc = self._lib.lara_gpu_create(C.byref(self._h), C.byref(pc), len(device_ids), arr)
        else:
            rc = self._lib.lara_gpu_create(C.byref(self._h), C.byref(pc), 0, None)
        if rc != 0:
            msg = self._lib.lara_gpu_last_error(None)
            raise LaraError("lara_gpu_create failed (%d): %s" % (rc, (msg or b"").decode()))
        self._keep = None

    def _check(self, rc):
        if rc != 0:
            raise LaraError("lara gpu error %d: %s" % (rc, (self._lib.lara_gpu_last_error(self._h) or b"").decode()))

    def close(self):
        if self._h:
            self._lib.lara_gpu_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_sequences(self, flat):
        seq_cat, seq_off, bp_partner, bp_weight, bp_off = [np.ascontiguousarray(a) for a in flat]
        self._check(self._lib.lara_gpu_set_sequences(self._h, len(seq_off) - 1, seq_cat.ctypes.data, seq_off.ctypes.data,
                                                     bp_partner.ctypes.data, bp_weight.ctypes.data, bp_off.ctypes.data))

    def solve_pairs(self, idx_a, idx_b):
        a = np.ascontiguousarray(idx_a, dtype=np.int32)
        b = np.ascontiguousarray(idx_b, dtype=np.int32)
        res = LaraResults()
        self._check(self._lib.lara_gpu_solve_pairs(self._h, len(a), a.ctypes.data, b.ctypes.data, C.byref(res)))
        return res

    def solve_all_pairs(self):
        res = LaraResults()
        self._check(self._lib.lara_gpu_solve_all_pairs(self._h, C.byref(res)))
        return res

    def free(self, res):
        self._lib.lara_gpu_free_results(C.byref(res))

    def prepare_pairs(self, idx_a, idx_b):
        a = np.ascontiguousarray(idx_a, dtype=np.int32)
        b = np.ascontiguousarray(idx_b, dtype=np.int32)
        self._check(self._lib.lara_gpu_prepare_pairs(self._h, len(a), a.ctypes.data, b.ctypes.data))

    def run_prepared(self):
        self._check(self._lib.lara_gpu_run_prepared(self._h))

    def fetch_prepared(self):
        res = LaraResults()
        self._check(self._lib.lara_gpu_fetch_prepared(self._h, C.byref(res)))
        return res

    def stats(self):
        s = LaraStats()
        self._check(self._lib.lara_gpu_get_stats(self._h, C.byref(s)))
        return s

    def gotoh_batch(self, pssm, gap_open, gap_extend):
        """pssm: int32 array [n, lenA, lenB] -> (score[n], n_lines[n], lines_a[n,lenB], lines_b[n,lenB], gap[n])."""
        pssm = np.ascontiguousarray(pssm, dtype=np.int32)
        n, la, lb = pssm.shape
        score = np.zeros(n, np.int32)
        nl = np.zeros(n, np.int32)
        la_ = np.zeros((n, lb), np.int32)
        lb_ = np.zeros((n, lb), np.int32)
        gap = np.zeros(n, np.int32)
        self._check(self._lib.lara_gpu_gotoh_batch(self._h, n, la, lb, pssm.ctypes.data, gap_open, gap_extend,
                                                   score.ctypes.data, nl.ctypes.data, la_.ctypes.data, lb_.ctypes.data,
                                                   gap.ctypes.data))
        return score, nl, la_, lb_, gap

    def gotoh_band_batch(self, pssm, d_lo, band_width, lanes, gap_open, gap_extend):
        """Band kernels on the same problem as gotoh_batch; all active cells must lie inside the band
        d_lo <= a - b < d_lo + band_width; `lanes` threads per problem."""
        pssm = np.ascontiguousarray(pssm, dtype=np.int32)
        n, la, lb = pssm.shape
        score = np.zeros(n, np.int32)
        nl = np.zeros(n, np.int32)
        la_ = np.zeros((n, lb), np.int32)
        lb_ = np.zeros((n, lb), np.int32)
        gap = np.zeros(n, np.int32)
        self._check(self._lib.lara_gpu_gotoh_band_batch(self._h, n, la, lb, pssm.ctypes.data, d_lo, band_width, lanes, gap_open,
                                                        gap_extend, score.ctypes.data, nl.ctypes.data, la_.ctypes.data,
                                                        lb_.ctypes.data, gap.ctypes.data))
        return score, nl, la_, lb_, gap

    def edge_filter(self, seq_a, seq_b):
        a = np.ascontiguousarray(seq_a, dtype=np.uint8)
        b = np.ascontiguousarray(seq_b, dtype=np.uint8)
        active = np.zeros(len(a) * len(b), np.uint8)
        opt = C.c_int32()
        self._check(self._lib.lara_gpu_edge_filter(self._h, a.ctypes.data, len(a), b.ctypes.data, len(b),
                                                   active.ctypes.data, C.byref(opt)))
        return active.reshape(len(a), len(b)), opt.value

    def edge_filter_noalloc(self, d_seq_a, len_a, d_seq_b, len_b, d_active, d_opt_score=0):
        """lara_gpu_edge_filter_noalloc: DEVICE addresses (ints) of caller-allocated buffers; the library allocates nothing."""
        self._check(self._lib.lara_gpu_edge_filter_noalloc(self._h, d_seq_a, len_a, d_seq_b, len_b, d_active,
                                                           d_opt_score or None))

    def int_issue_rate(self, kind=0):
        g = C.c_double()
        self._check(self._lib.lara_gpu_int_issue_rate(self._h, kind, C.byref(g)))
        return g.value


class OutputLibrary:
    """lara::OutputLibrary: collects one alignment per pair, prints lib / pairs / fasta."""

    def __init__(self, store, out_format="lib"):
        self.store = store
        self.format = out_format
        self._res = None
        self._ctx = None

    def _set(self, ctx, res):
        self._release()
        self._ctx, self._res = ctx, res

    def _release(self):
        if self._res is not None and self._ctx is not None:
            self._ctx.free(self._res)
        self._res = None

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def alignments(self):
        return results_to_list(self._res) if self._res is not None else []

    def print(self, filename=""):
        if self._res is None:
            raise LaraError("OutputLibrary is empty: call solve() first")
        lib = _lib.load()
        _host_check(lib.lara_output_write(self.store._h, C.byref(self._res), self.format.encode(),
                                          filename.encode() if filename else None))


def solve(results, store, params, device_ids=None, ctx=None):
    """lara::solve(results, store, params): all-pairs structural alignment on the GPU(s)."""
    own = ctx is None
    if own:
        ctx = Context(params, device_ids)
    ctx.set_sequences(store.flat())
    res = ctx.solve_all_pairs()
    results._set(ctx, res)
    return ctx
