"""ctypes bindings of the C ABI declared in include/lara_b200.h and include/lara_host.h.

The product library is lara_b200/liblara_b200.so (host C++ + sm_100a CUDA kernels, built in-tree
by __graft_entry__.build()).  Nothing here falls back to another implementation: if the library is
missing, `load()` raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# LARA_B200_LIB: another build of the same library (developer A/B runs: tools/sweep_env.sh); never a different implementation
LIB_PATH = os.environ.get("LARA_B200_LIB") or os.path.join(_HERE, "liblara_b200.so")


class LaraParams(C.Structure):
    _fields_ = [
        ("num_iterations", C.c_int32),
        ("max_nondecreasing", C.c_int32),
        ("matching", C.c_int32),
        ("step_mode", C.c_int32),
        ("step_size_factor", C.c_float),
        ("suboptimal_diff", C.c_float),
        ("balance", C.c_float),
        ("sequence_scale", C.c_float),
        ("gap_open", C.c_int32),
        ("gap_extend", C.c_int32),
        ("score_matrix", C.c_int32 * 25),
        ("lib_score_min", C.c_uint32),
        ("lib_score_max", C.c_uint32),
        ("lib_score_linear", C.c_int32),
        ("reserved", C.c_int32 * 9),
    ]


class LaraPairResult(C.Structure):
    _fields_ = [
        ("idx_low", C.c_int32),
        ("idx_high", C.c_int32),
        ("best_lower", C.c_int32),
        ("best_upper", C.c_int32),
        ("iterations", C.c_int32),
        ("n_lines", C.c_int32),
        ("lines_off", C.c_int64),
    ]


class LaraResults(C.Structure):
    _fields_ = [
        ("n_pairs", C.c_int64),
        ("pairs", C.POINTER(LaraPairResult)),
        ("n_lines_total", C.c_int64),
        ("line_pos_low", C.POINTER(C.c_int32)),
        ("line_pos_high", C.POINTER(C.c_int32)),
        ("line_weight", C.POINTER(C.c_uint32)),
        ("owner_", C.c_void_p),
    ]


class LaraStats(C.Structure):
    _fields_ = [
        ("ms_total", C.c_double),
        ("ms_h2d", C.c_double),
        ("ms_setup", C.c_double),
        ("ms_dp", C.c_double),
        ("ms_traceback", C.c_double),
        ("ms_bounds", C.c_double),
        ("ms_d2h", C.c_double),
        ("dp_cells", C.c_int64),
        ("dp_launches", C.c_int64),
        ("kernel_launches", C.c_int64),
        ("filter_cells", C.c_int64),
        ("h2d_bytes", C.c_int64),
        ("d2h_bytes", C.c_int64),
        ("n_pairs", C.c_int64),
        ("sum_iterations", C.c_int64),
        ("n_active_edges", C.c_int64),
        ("n_interactions", C.c_int64),
        ("ms_device", C.c_double),
        ("reserved", C.c_int64 * 16),
    ]


class LaraCliOptions(C.Structure):
    _fields_ = [
        ("verbose", C.c_int32),
        ("threads", C.c_int32),
        ("in_file", C.c_char * 1024),
        ("ref_file", C.c_char * 1024),
        ("out_file", C.c_char * 1024),
        ("out_format", C.c_char * 16),
        ("n_dotplots", C.c_int32),
        ("dotplot_files", C.POINTER(C.c_char_p)),
        ("epsilon", C.c_float),
        ("structure_scoring", C.c_int32),
        ("score_matrix_file", C.c_char * 1024),
    ]


# every symbol include/*.h declares: name -> (restype, argtypes)
_P = C.POINTER
GPU_SYMBOLS = {
    "lara_params_default": (None, [_P(LaraParams)]),
    "lara_gpu_version": (C.c_int, []),
    "lara_gpu_visible_devices": (C.c_int, []),
    "lara_gpu_create": (C.c_int, [_P(C.c_void_p), _P(LaraParams), C.c_int, _P(C.c_int)]),
    "lara_gpu_destroy": (None, [C.c_void_p]),
    "lara_gpu_last_error": (C.c_char_p, [C.c_void_p]),
    "lara_gpu_set_sequences": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lara_gpu_solve_pairs": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, _P(LaraResults)]),
    "lara_gpu_solve_all_pairs": (C.c_int, [C.c_void_p, _P(LaraResults)]),
    "lara_gpu_free_results": (None, [_P(LaraResults)]),
    "lara_gpu_get_stats": (C.c_int, [C.c_void_p, _P(LaraStats)]),
    "lara_gpu_prepare_pairs": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "lara_gpu_run_prepared": (C.c_int, [C.c_void_p]),
    "lara_gpu_fetch_prepared": (C.c_int, [C.c_void_p, _P(LaraResults)]),
    "lara_gpu_gotoh_batch": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lara_gpu_gotoh_band_batch": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32,
                                            C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "lara_gpu_edge_filter": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, _P(C.c_int32)]),
    "lara_gpu_edge_filter_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "lara_gpu_int_issue_rate": (C.c_int, [C.c_void_p, C.c_int, _P(C.c_double)]),
}
HOST_SYMBOLS = {
    "lara_parse_args": (C.c_int, [C.c_int, _P(C.c_char_p), _P(LaraParams), _P(LaraCliOptions)]),
    "lara_cli_options_free": (None, [_P(LaraCliOptions)]),
    "lara_input_create": (C.c_int, [_P(C.c_void_p)]),
    "lara_input_destroy": (None, [C.c_void_p]),
    "lara_input_read_file": (C.c_int, [C.c_void_p, C.c_char_p]),
    "lara_input_read_dotplot": (C.c_int, [C.c_void_p, C.c_char_p]),
    "lara_input_finish": (C.c_int, [C.c_void_p]),
    "lara_input_add_record": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "lara_input_size": (C.c_int32, [C.c_void_p]),
    "lara_input_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "lara_input_length": (C.c_int32, [C.c_void_p, C.c_int32]),
    "lara_input_flatten": (C.c_int, [C.c_void_p, _P(C.c_void_p), _P(C.c_void_p), _P(C.c_void_p), _P(C.c_void_p), _P(C.c_void_p)]),
    "lara_output_write": (C.c_int, [C.c_void_p, _P(LaraResults), C.c_char_p, C.c_char_p]),
    "lara_host_last_error": (C.c_char_p, []),
}
ALL_SYMBOLS = dict(GPU_SYMBOLS)
ALL_SYMBOLS.update(HOST_SYMBOLS)

_lib = None


def load(path=None, symbols=None):
    """Load liblara_b200.so and type every exported entry point.  Raises if the library is missing.
    `symbols` (default: everything include/*.h declares) is only narrowed by developer tools."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            "lara_b200: %s not found -- build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
            "there is no CPU fallback" % p)
    lib = C.CDLL(p, mode=C.RTLD_LOCAL)
    for name, (res, args) in (symbols or ALL_SYMBOLS).items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if path is None or symbols is not None:
        _lib = lib
    return lib
