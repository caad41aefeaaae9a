"""ctypes binding of ``libassl.so`` (the C ABI declared in ``include/assl.h``) and its in-tree build.

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.environ.get("ASSL_LIB") or os.path.join(CSRC, "libassl.so")   # ASSL_LIB: an experimental build
SOURCES = ["plan.cu", "logmel.cu", "window.cu", "stats.cu", "views.cu", "kmix.cu", "pipeline.cu"]
HEADERS = ["common.cuh", "fft32p_gen.cuh", "f32x2.cuh", os.path.join("..", "..", "include", "assl.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]

_lib = None


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into ``csrc/libassl.so`` (nvcc cross-compiles without a GPU)."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("libassl.so is out of date or missing and nvcc was not found")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + SOURCES
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


class Cfg(C.Structure):
    _fields_ = [("sample_rate", C.c_int32), ("n_fft", C.c_int32), ("hop", C.c_int32),
                ("n_mels", C.c_int32), ("fmin", C.c_float), ("fmax", C.c_float)]


class ViewCfg(C.Structure):
    _fields_ = [("n_freq", C.c_int32), ("n_time", C.c_int32), ("canvas_h", C.c_int32),
                ("canvas_w", C.c_int32), ("interp", C.c_int32), ("log_mixup", C.c_int32), ("ring_pitch", C.c_int32)]


_P, _I64, _F = C.c_void_p, C.c_int64, C.c_float
SIGNATURES = {
    "assl_version": (C.c_int, []),
    "assl_status_string": (C.c_char_p, [C.c_int]),
    "assl_plan_create": (C.c_int, [C.POINTER(_P), C.POINTER(Cfg)]),
    "assl_plan_destroy": (C.c_int, [_P]),
    "assl_plan_num_frames": (_I64, [_P, _I64]),
    "assl_plan_num_tiles": (_I64, [_P, _I64]),
    "assl_plan_filterbank": (C.c_int, [_P, _P]),
    "assl_logmel_f32": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _P, _I64, _I64, _P, C.c_int32, _P]),
    "assl_logmel_pitched_f32": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _P, _I64, _I64, _I64, _P, C.c_int32, _P]),
    "assl_logmel_deferred_f32": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _P, _I64, _I64, _I64, _P, C.c_int32, _P]),
    "assl_clip_moments_f32": (C.c_int, [_P, _I64, _I64, _P, _P]),
    "assl_clip_normalize_f32": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, _P]),
    "assl_wave_l2_gain_f32": (C.c_int, [_P, _I64, _I64, _I64, _P, _P]),
    "assl_extract_windows_f32": (C.c_int, [_P, _P, _P, _P, _I64, _I64, _P, _I64, _P, _P]),
    "assl_running_stats": (C.c_int, [_P, _P, _I64, _I64, _I64, _I64, _I64, _I64, _P, _I64, _I64, _P]),
    "assl_running_norm_stats_static_f32": (C.c_int, [_P, _P, _I64, _I64, _I64, _P, _I64, _I64, _P]),
    "assl_clip_sums": (C.c_int, [_P, _I64, _I64, _P, _P]),
    "assl_exchange_create": (C.c_int, [C.POINTER(_P), C.c_int32, C.c_int32, _I64, _P]),
    "assl_exchange_connect": (C.c_int, [_P, _P]),
    "assl_exchange_destroy": (C.c_int, [_P]),
    "assl_exchange_use_staging": (C.c_int, [_P]),
    "assl_exchange_staging": (C.c_int, [_P, _I64, C.POINTER(_P), C.POINTER(_I64)]),
    "assl_exchange_set_timeout": (C.c_int, [_P, C.c_double]),
    "assl_exchange_error": (C.c_int, [_P]),
    "assl_exchange_send": (C.c_int, [_P, _P, _I64, _I64, _P]),
    "assl_running_stats_exchange_scan": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, C.c_int32, C.c_int32, _P]),
    "assl_exchange_scan_send": (C.c_int, [_P, _P, _I64, _I64, _P, _I64, _I64, C.c_int32, _P, _I64, _I64, _P]),
    "assl_running_stats_exchange": (C.c_int, [_P, _P, _P, _I64, _I64, _I64, _I64, _P, _I64, _I64, C.c_int32, _P]),
    "assl_peer_alloc": (C.c_int, [C.POINTER(_P), _I64, _P]),
    "assl_peer_free": (C.c_int, [_P]),
    "assl_peer_open": (C.c_int, [_P, C.POINTER(_P)]),
    "assl_peer_close": (C.c_int, [_P]),
    "assl_two_views_global_f32": (C.c_int, [C.POINTER(ViewCfg), _P, _P, C.c_int32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I64,
                                             _P, _P, _P]),
    "assl_two_views_f32": (C.c_int, [C.POINTER(ViewCfg), _P, _P, _I64, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "assl_two_views_ordered_f32": (C.c_int, [C.POINTER(ViewCfg), _P, _P, _I64, _P, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "assl_two_views_deferred_f32": (C.c_int, [C.POINTER(ViewCfg), _P, _P, _I64, _P, _P, _P, _P, _P, _I64, _P, _P, _P]),
    "assl_clip_clusters_f32": (C.c_int, [_P, _I64, C.c_int32, C.c_int32, _P, C.c_int32, _P, _P, _P]),
    "assl_batch_sums": (C.c_int, [_P, _I64, _P, _P]),
    "assl_normalize_apply_f32": (C.c_int, [_P, _P, _I64, _P, _P]),
    "assl_affine_norm_f32": (C.c_int, [_P, _P, _I64, _F, _F, _P]),
    "assl_pipeline_create": (C.c_int, [C.POINTER(_P), _P, _P]),
    "assl_pipeline_destroy": (C.c_int, [_P]),
    "assl_pipeline_workspace_bytes": (_I64, [_P, _I64, _I64, _I64]),
    "assl_pipeline_host_sync": (C.c_int, [_P]),
    "assl_pipeline_host_f32": (C.c_int, [_P, C.POINTER(ViewCfg), _P, _I64, _P, _I64, _I64, _I64, _I64, _I64,
                                          _P, _P, _P, _P, _P, C.POINTER(_P), C.POINTER(_P), _P]),
    "assl_pipeline_host_pcm16": (C.c_int, [_P, C.POINTER(ViewCfg), _P, _I64, _P, _I64, _I64, _I64, _I64, _I64,
                                          _P, _P, _P, _P, _P, C.POINTER(_P), C.POINTER(_P), _P]),
    "assl_pcm16_to_f32": (C.c_int, [_P, _I64, _P, _P]),
}


def lib() -> C.CDLL:
    """Load ``libassl.so`` (once).  Raises if it has not been built -- there is no other path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(the CUDA extension is the only implementation; there is no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, what: str = "assl call") -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed: {lib().assl_status_string(rc).decode()} (status {rc})")
