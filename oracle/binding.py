"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/liblara_oracle.so (the CPU restatement) and
helpers to run the unmodified reference binaries in oracle/_ref/.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module."""
import ctypes as C
import os
import subprocess

import numpy as np

from lara_b200._lib import LaraParams, LaraResults, LaraStats

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblara_oracle.so")
REF_SCALAR = os.path.join(_HERE, "_ref", "lara_scalar")
REF_SIMD = os.path.join(_HERE, "_ref", "lara_simd")

_lib = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
        lib.lara_oracle_params_default.argtypes = [C.POINTER(LaraParams)]
        lib.lara_oracle_params_default.restype = None
        lib.lara_oracle_solve_pairs.restype = C.c_int
        lib.lara_oracle_solve_pairs.argtypes = [C.POINTER(LaraParams), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int,
                                                C.POINTER(LaraResults), C.POINTER(LaraStats)]
        lib.lara_oracle_free_results.argtypes = [C.POINTER(LaraResults)]
        lib.lara_oracle_free_results.restype = None
        lib.lara_oracle_gotoh.restype = C.c_int
        lib.lara_oracle_gotoh.argtypes = [C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.lara_oracle_edge_filter.restype = C.c_int
        lib.lara_oracle_edge_filter.argtypes = [C.POINTER(LaraParams), C.c_void_p, C.c_int32, C.c_void_p, C.c_int32,
                                                C.c_void_p, C.POINTER(C.c_int32)]
        lib.lara_oracle_greedy_matching.restype = C.c_int32
        lib.lara_oracle_greedy_matching.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                                    C.c_int32, C.c_void_p]
        _lib = lib
    return _lib


def default_params():
    p = LaraParams()
    load().lara_oracle_params_default(C.byref(p))
    return p


def all_pairs(n):
    ia, ib = np.triu_indices(n, k=1)
    return ia.astype(np.int32), ib.astype(np.int32)


def solve_pairs(params, flat, idx_a, idx_b, n_threads=1):
    """Returns (LaraResults, LaraStats); free with free_results()."""
    lib = load()
    seq_cat, seq_off, bp_partner, bp_weight, bp_off = [np.ascontiguousarray(a) for a in flat]
    a = np.ascontiguousarray(idx_a, dtype=np.int32)
    b = np.ascontiguousarray(idx_b, dtype=np.int32)
    res = LaraResults()
    st = LaraStats()
    pc = params.c if hasattr(params, "c") else params
    rc = lib.lara_oracle_solve_pairs(C.byref(pc), len(seq_off) - 1, seq_cat.ctypes.data, seq_off.ctypes.data,
                                     bp_partner.ctypes.data, bp_weight.ctypes.data, bp_off.ctypes.data,
                                     len(a), a.ctypes.data, b.ctypes.data, n_threads, C.byref(res), C.byref(st))
    if rc != 0:
        raise RuntimeError("oracle error %d" % rc)
    return res, st


def free_results(res):
    load().lara_oracle_free_results(C.byref(res))


def gotoh(pssm, gap_open, gap_extend, want_trace=False):
    lib = load()
    pssm = np.ascontiguousarray(pssm, dtype=np.int32)
    la, lb = pssm.shape
    score = C.c_int32()
    nl = C.c_int32()
    gap = C.c_int32()
    a = np.zeros(lb, np.int32)
    b = np.zeros(lb, np.int32)
    tr = np.zeros((la + 1, lb + 1), np.uint8) if want_trace else None
    rc = lib.lara_oracle_gotoh(la, lb, pssm.ctypes.data, gap_open, gap_extend, C.byref(score), C.byref(nl),
                               a.ctypes.data, b.ctypes.data, C.byref(gap), tr.ctypes.data if want_trace else None)
    if rc != 0:
        raise RuntimeError("oracle gotoh error %d" % rc)
    n = nl.value
    out = (score.value, a[:n].copy(), b[:n].copy(), gap.value)
    return out + (tr,) if want_trace else out


def edge_filter(params, seq_a, seq_b):
    lib = load()
    a = np.ascontiguousarray(seq_a, dtype=np.uint8)
    b = np.ascontiguousarray(seq_b, dtype=np.uint8)
    active = np.zeros(len(a) * len(b), np.uint8)
    opt = C.c_int32()
    pc = params.c if hasattr(params, "c") else params
    rc = lib.lara_oracle_edge_filter(C.byref(pc), a.ctypes.data, len(a), b.ctypes.data, len(b), active.ctypes.data,
                                     C.byref(opt))
    if rc != 0:
        raise RuntimeError("oracle edge filter error %d" % rc)
    return active.reshape(len(a), len(b)), opt.value


def greedy_matching(w, l, r, lookahead, n_vertices):
    lib = load()
    w = np.ascontiguousarray(w, dtype=np.int32)
    l = np.ascontiguousarray(l, dtype=np.int32)
    r = np.ascontiguousarray(r, dtype=np.int32)
    m = np.zeros(n_vertices, np.uint8)
    s = lib.lara_oracle_greedy_matching(len(w), w.ctypes.data, l.ctypes.data, r.ctypes.data, lookahead, n_vertices,
                                        m.ctypes.data)
    return s, m


def run_reference(binary, args, timeout=3600):
    """Run an unmodified reference binary (oracle/_ref/lara_scalar|lara_simd); returns CompletedProcess."""
    if not os.path.exists(binary):
        raise FileNotFoundError(binary)
    return subprocess.run([binary] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=timeout)
