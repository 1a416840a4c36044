"""ctypes binding of libmc_b200.so (include/mc_b200.h).  No CPU fallback: if the
library is missing or fails to load, importing the compute path raises."""
from __future__ import annotations

import ctypes as C
import os
import threading

from .build import LIB_PATH

MC_ABI_VERSION = 3
MC_MAX_PCTS = 16
MC_MAX_PATH_PCTS = 8
MC_MAX_ASSETS = 32
MC_STATIC_HIST_MAX_BINS = 4096

SHOCK_BOOTSTRAP, SHOCK_NORMAL, SHOCK_INJECTED = 0, 1, 2
VAR_GARCH, VAR_CONST = 0, 1
F32, F64 = 0, 1
FLAG_STORE_PATHS, FLAG_FAN_CHART, FLAG_NO_TERMINAL_SELECT, FLAG_PATHS_VOLATILE, FLAG_TERMINAL_GIVEN = 1, 2, 4, 8, 16
FLAG_FULL_RES_NORMALS = 32

EXPORTS = (
    "mc_abi_version", "mc_create", "mc_destroy", "mc_last_error", "mc_run",
    "mc_stage_begin", "mc_stage_words", "mc_stage_max_words", "mc_stage_next", "mc_fetch_result",
    "mc_stage_reads_paths", "mc_stage_resimulate", "mc_stage_adopt", "mc_stage_shared",
    "mc_debug_shocks", "mc_debug_philox", "mc_percentiles", "mc_histogram", "mc_prepare_inputs",
    "mc_run_batch", "mc_calibrate_garch",
    "mc_peer_export", "mc_peer_attach", "mc_run_sharded", "mc_debug_set_tuning", "mc_debug_counter", "mc_debug_set_pilot", "mc_histogram_static",
)


class McParams(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("version", C.c_uint32),
        ("n_paths", C.c_uint64), ("path_offset", C.c_uint64), ("n_paths_global", C.c_uint64),
        ("paths_ld", C.c_uint64),
        ("n_steps", C.c_uint32), ("n_assets", C.c_uint32),
        ("shock", C.c_int32), ("var_model", C.c_int32), ("dtype", C.c_int32), ("flags", C.c_uint32),
        ("seed", C.c_uint64),
        ("s0", C.c_double), ("var0", C.c_double),
        ("omega", C.c_double), ("alpha", C.c_double), ("beta", C.c_double),
        ("drift", C.c_double),
        ("pool_len", C.c_uint32), ("hist_bins", C.c_uint32),
        ("hist_lo", C.c_double), ("hist_hi", C.c_double),
        ("n_pcts", C.c_uint32), ("n_path_pcts", C.c_uint32),
        ("pcts", C.c_double * MC_MAX_PCTS),
        ("path_pcts", C.c_double * MC_MAX_PATH_PCTS),
        ("tail_pct", C.c_double),
        ("n_ranks", C.c_uint32), ("pass_group", C.c_uint32),
    ]


class McPortfolio(C.Structure):
    _fields_ = [
        ("h_chol", C.POINTER(C.c_double)), ("h_sigma_step", C.POINTER(C.c_double)),
        ("h_drift_step", C.POINTER(C.c_double)), ("h_s0", C.POINTER(C.c_double)),
        ("h_weights", C.POINTER(C.c_double)),
    ]


class McResult(C.Structure):
    _fields_ = [
        ("n", C.c_uint64), ("n_above", C.c_uint64),
        ("mean", C.c_double), ("std", C.c_double), ("min", C.c_double), ("max", C.c_double),
        ("pct", C.c_double * MC_MAX_PCTS), ("pct_lo", C.c_double * MC_MAX_PCTS),
        ("pct_hi", C.c_double * MC_MAX_PCTS),
        ("tail_count", C.c_uint64), ("tail_sum", C.c_double),
        ("hist_lo", C.c_double), ("hist_hi", C.c_double),
        ("sim_ms", C.c_float), ("total_ms", C.c_float),
        ("launches", C.c_uint32), ("status", C.c_uint32),
    ]


RESULT_DTYPE = [("n", "<u8"), ("n_above", "<u8"), ("mean", "<f8"), ("std", "<f8"), ("min", "<f8"), ("max", "<f8"),
                ("pct", "<f8", (MC_MAX_PCTS,)), ("pct_lo", "<f8", (MC_MAX_PCTS,)), ("pct_hi", "<f8", (MC_MAX_PCTS,)),
                ("tail_count", "<u8"), ("tail_sum", "<f8"), ("hist_lo", "<f8"), ("hist_hi", "<f8"),
                ("sim_ms", "<f4"), ("total_ms", "<f4"), ("launches", "<u4"), ("status", "<u4")]


class McSeriesInputs(C.Structure):
    """mc_series_inputs (include/mc_b200.h): what app.py:592/606/609/1213 derive from one price history."""
    _fields_ = [("s0", C.c_double), ("var0", C.c_double), ("pool_std", C.c_double),
                ("pool_len", C.c_uint64), ("flags", C.c_uint64)]


class McBatchParams(C.Structure):
    """mc_batch_params (include/mc_b200.h)."""
    _fields_ = [
        ("struct_size", C.c_uint32), ("version", C.c_uint32),
        ("n_paths", C.c_uint64), ("seed", C.c_uint64),
        ("n_steps", C.c_uint32), ("n_series", C.c_uint32), ("min_history", C.c_uint32), ("flags", C.c_uint32),
        ("omega", C.c_double), ("alpha", C.c_double), ("beta", C.c_double),
        ("n_pcts", C.c_uint32), ("n_path_pcts", C.c_uint32),
        ("pcts", C.c_double * MC_MAX_PCTS), ("path_pcts", C.c_double * MC_MAX_PATH_PCTS),
        ("tail_pct", C.c_double),
    ]


class McGarchFit(C.Structure):
    """mc_garch_fit (include/mc_b200.h)."""
    _fields_ = [("omega", C.c_double), ("alpha", C.c_double), ("beta", C.c_double), ("objective", C.c_double),
                ("valid", C.c_uint64)]


GARCH_FIT_DTYPE = [("omega", "<f8"), ("alpha", "<f8"), ("beta", "<f8"), ("objective", "<f8"), ("valid", "<u8")]
INPUT_NAN = 1
SERIES_INPUTS_DTYPE = [("s0", "<f8"), ("var0", "<f8"), ("pool_std", "<f8"), ("pool_len", "<u8"), ("flags", "<u8")]

_lock = threading.Lock()
_lib = None


def _declare(lib):
    vp, u32, u64, i32, dbl = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32, C.c_double
    PP, PF, PR = C.POINTER(McParams), C.POINTER(McPortfolio), C.POINTER(McResult)
    lib.mc_abi_version.restype = C.c_int
    lib.mc_abi_version.argtypes = []
    lib.mc_create.restype = C.c_int
    lib.mc_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.mc_destroy.restype = C.c_int
    lib.mc_destroy.argtypes = [vp]
    lib.mc_last_error.restype = C.c_char_p
    lib.mc_last_error.argtypes = [vp]
    lib.mc_run.restype = C.c_int
    lib.mc_run.argtypes = [vp, PP, PF, vp, vp, vp, vp, PR, vp, vp, vp]
    lib.mc_stage_begin.restype = C.c_int
    lib.mc_stage_begin.argtypes = [vp, PP, PF, vp, vp, vp, vp, vp, u64, vp]
    lib.mc_stage_reads_paths.restype = C.c_int
    lib.mc_stage_reads_paths.argtypes = [vp]
    lib.mc_stage_resimulate.restype = C.c_int
    lib.mc_stage_resimulate.argtypes = [vp, vp]
    lib.mc_stage_adopt.restype = C.c_int
    lib.mc_stage_adopt.argtypes = [vp, vp, vp, u32, vp, C.POINTER(C.c_int), vp]
    lib.mc_stage_shared.restype = C.c_int
    lib.mc_stage_shared.argtypes = [vp]
    lib.mc_stage_words.restype = C.c_int64
    lib.mc_stage_words.argtypes = [vp]
    lib.mc_stage_max_words.restype = C.c_int64
    lib.mc_stage_max_words.argtypes = [PP]
    lib.mc_stage_next.restype = C.c_int
    lib.mc_stage_next.argtypes = [vp, vp, u32, vp, C.POINTER(C.c_int), vp]
    lib.mc_fetch_result.restype = C.c_int
    lib.mc_fetch_result.argtypes = [vp, PR, vp, vp, vp]
    lib.mc_debug_shocks.restype = C.c_int
    lib.mc_debug_shocks.argtypes = [vp, PP, vp, vp, vp]
    lib.mc_debug_philox.restype = C.c_int
    lib.mc_debug_philox.argtypes = [vp, vp, u32, u32, u32, vp, vp]
    lib.mc_percentiles.restype = C.c_int
    lib.mc_percentiles.argtypes = [vp, vp, i32, u64, u64, u64, C.POINTER(dbl), u32, vp, vp]
    lib.mc_histogram.restype = C.c_int
    lib.mc_histogram.argtypes = [vp, vp, i32, u64, u32, dbl, dbl, vp, vp]
    lib.mc_prepare_inputs.restype = C.c_int
    lib.mc_prepare_inputs.argtypes = [vp, vp, vp, u32, u32, u64, u32, i32, vp, u64, vp, vp]
    lib.mc_run_batch.restype = C.c_int
    lib.mc_run_batch.argtypes = [vp, C.POINTER(McBatchParams), vp, u64, vp, vp, vp, vp, vp, vp, vp]
    lib.mc_calibrate_garch.restype = C.c_int
    lib.mc_calibrate_garch.argtypes = [vp, vp, u64, vp, u32, u32, u32, vp, vp]
    lib.mc_peer_export.restype = C.c_int
    lib.mc_peer_export.argtypes = [vp, u32, vp, C.POINTER(vp), C.POINTER(u64)]
    lib.mc_peer_attach.restype = C.c_int
    lib.mc_peer_attach.argtypes = [vp, u32, u32, vp, vp]
    lib.mc_run_sharded.restype = C.c_int
    lib.mc_run_sharded.argtypes = [vp, PP, PF, vp, vp, vp, vp, vp]
    lib.mc_debug_set_tuning.restype = C.c_int
    lib.mc_debug_set_tuning.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.mc_debug_set_pilot.restype = C.c_int
    lib.mc_debug_set_pilot.argtypes = [u32]
    lib.mc_histogram_static.restype = C.c_int
    lib.mc_histogram_static.argtypes = [C.c_int, vp, i32, u64, u32, dbl, dbl, vp, vp, C.c_char_p, u32]
    lib.mc_debug_counter.restype = C.c_int64
    lib.mc_debug_counter.argtypes = [vp, C.c_int]


def load_library(path: str | None = None):
    """Load (once) and return the ctypes handle of libmc_b200.so.  Raises RuntimeError
    when the library has not been built -- there is deliberately no fallback."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = path or os.environ.get("MC_B200_LIB", LIB_PATH)
        if not os.path.exists(path):
            raise RuntimeError(
                f"montecarlo_b200: CUDA library not found at {path}. Build it with "
                "`python -m montecarlo_b200.build` (needs nvcc, targets sm_100a). "
                "There is no CPU fallback.")
        try:
            lib = C.CDLL(path)
        except OSError as exc:
            raise RuntimeError(f"montecarlo_b200: cannot load {path}: {exc}") from exc
        for name in EXPORTS:
            if not hasattr(lib, name):
                raise RuntimeError(f"montecarlo_b200: {path} does not export {name}")
        _declare(lib)
        if lib.mc_abi_version() != MC_ABI_VERSION:
            raise RuntimeError("montecarlo_b200: library ABI version mismatch; rebuild")
        _lib = lib
        return lib
