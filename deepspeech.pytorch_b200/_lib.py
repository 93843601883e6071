"""ctypes binding of libctc_b200.so (the C ABI declared in include/ctc_b200.h).

The library is built in-tree with nvcc for sm_100a only; there is no CPU or other-arch fallback: if the shared
object is missing or a CUDA device is absent, the op raises.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.path.join(_PKG, "libctc_b200.so")
if os.environ.get("CTC_B200_LIB"):  # A/B experiments: another in-tree build of the same sources (e.g. other -D constants)
    LIB_PATH = os.path.abspath(os.environ["CTC_B200_LIB"])
SOURCES = [os.path.join(_PKG, "csrc", f) for f in ("ctc_api.cu", "ctc_kernels.cuh", "ctc_device.cuh", "ctc_metrics.cuh",
                                                   "fc_kernels.cuh")] + [
    os.path.join(_ROOT, "include", "ctc_b200.h")
]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared",
              "-Xcompiler", "-fPIC"]

EXPORTS = ("ctc_b200_get_version", "ctc_b200_status_string", "ctc_b200_default_options", "ctc_b200_workspace_size",
           "ctc_b200_compute", "ctc_b200_create", "ctc_b200_destroy", "ctc_b200_compute_host",
           "ctc_b200_compute_decode", "ctc_b200_greedy_decode", "ctc_b200_error_counts", "ctc_b200_last_launch_count",
           "ctc_b200_comm_create", "ctc_b200_comm_handle", "ctc_b200_comm_local", "ctc_b200_comm_open",
           "ctc_b200_comm_set_peers", "ctc_b200_comm_destroy", "ctc_b200_comm_set_timeout",
           "ctc_b200_comm_timed_out", "ctc_b200_scale_grads", "ctc_b200_last_copy_bytes",
           "ctc_b200_compute_static", "ctc_b200_fc_workspace_size", "ctc_b200_fc_forward")


class CtcOptions(ctypes.Structure):
    """Mirror of ctcB200Options (include/ctc_b200.h)."""
    _fields_ = [("stream", ctypes.c_void_p), ("blank", ctypes.c_int), ("zero_infinity", ctypes.c_int),
                ("grad_scale", ctypes.c_float), ("lengths_on_device", ctypes.c_int),
                ("max_label_len", ctypes.c_int), ("phase_mask", ctypes.c_int), ("nan_to_zero", ctypes.c_int),
                ("costs_ready", ctypes.c_void_p), ("nan_count", ctypes.c_void_p),
                ("loss_sum", ctypes.c_void_p), ("comm", ctypes.c_void_p),
                ("loss_sum_all", ctypes.c_void_p), ("emissions_ready", ctypes.c_int)]


class FcLayer(ctypes.Structure):
    """Mirror of ctcB200FcLayer (include/ctc_b200.h)."""
    _fields_ = [("weight", ctypes.c_void_p), ("bn_weight", ctypes.c_void_p), ("bn_bias", ctypes.c_void_p),
                ("bn_mean", ctypes.c_void_p), ("bn_var", ctypes.c_void_p), ("bn_eps", ctypes.c_float),
                ("bn_batch_stats", ctypes.c_int), ("batch_mean_out", ctypes.c_void_p),
                ("batch_var_out", ctypes.c_void_p), ("reuse_folded", ctypes.c_int)]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libctc_b200.so")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    mt = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > mt for s in SOURCES)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/ctc_api.cu -> libctc_b200.so for sm_100a (cross-compiles without a GPU)."""
    if force or is_stale():
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH, SOURCES[0]]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
        if verbose:
            print(res.stderr)
    return LIB_PATH


_lib = None


def load():
    """dlopen the library and declare every symbol of include/ctc_b200.h.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    P, I, V = ctypes.POINTER, ctypes.c_int, ctypes.c_void_p
    lib.ctc_b200_get_version.restype = I
    lib.ctc_b200_status_string.restype = ctypes.c_char_p
    lib.ctc_b200_status_string.argtypes = [I]
    lib.ctc_b200_default_options.argtypes = [P(CtcOptions)]
    lib.ctc_b200_default_options.restype = None
    lib.ctc_b200_workspace_size.argtypes = [I, I, I, I, P(CtcOptions), P(ctypes.c_size_t)]
    lib.ctc_b200_workspace_size.restype = I
    lib.ctc_b200_compute.argtypes = [V, V, V, V, V, I, I, I, V, V, ctypes.c_size_t, P(CtcOptions)]
    lib.ctc_b200_compute.restype = I
    lib.ctc_b200_compute_static.argtypes = [V, V, V, V, V, I, I, I, V, P(CtcOptions)]
    lib.ctc_b200_compute_static.restype = I
    lib.ctc_b200_fc_workspace_size.argtypes = [I, I, I, I, P(ctypes.c_size_t)]
    lib.ctc_b200_fc_workspace_size.restype = I
    lib.ctc_b200_fc_forward.argtypes = [V, I, I, I, I, P(FcLayer), V, V, V, V, V, ctypes.c_size_t, P(CtcOptions)]
    lib.ctc_b200_fc_forward.restype = I
    lib.ctc_b200_create.argtypes = [P(V)]
    lib.ctc_b200_create.restype = I
    lib.ctc_b200_destroy.argtypes = [V]
    lib.ctc_b200_destroy.restype = I
    lib.ctc_b200_compute_host.argtypes = [V, V, V, V, V, V, I, I, I, V, P(ctypes.c_double), P(CtcOptions)]
    lib.ctc_b200_compute_host.restype = I
    lib.ctc_b200_compute_decode.argtypes = [V, V, V, V, V, I, I, I, V, V, ctypes.c_size_t, V, V, V, V, P(CtcOptions)]
    lib.ctc_b200_compute_decode.restype = I
    lib.ctc_b200_error_counts.argtypes = [V, I, V, V, V, I, I, I, I, V, P(CtcOptions)]
    lib.ctc_b200_error_counts.restype = I
    lib.ctc_b200_comm_create.argtypes = [P(V), I, I]
    lib.ctc_b200_comm_create.restype = I
    lib.ctc_b200_comm_handle.argtypes = [V, V]
    lib.ctc_b200_comm_handle.restype = I
    lib.ctc_b200_comm_local.argtypes = [V]
    lib.ctc_b200_comm_local.restype = V
    lib.ctc_b200_comm_open.argtypes = [V, V]
    lib.ctc_b200_comm_open.restype = I
    lib.ctc_b200_comm_set_peers.argtypes = [V, P(V)]
    lib.ctc_b200_comm_set_peers.restype = I
    lib.ctc_b200_comm_destroy.argtypes = [V]
    lib.ctc_b200_comm_destroy.restype = I
    lib.ctc_b200_comm_set_timeout.argtypes = [V, ctypes.c_longlong]
    lib.ctc_b200_comm_set_timeout.restype = I
    lib.ctc_b200_comm_timed_out.argtypes = [V]
    lib.ctc_b200_comm_timed_out.restype = I
    lib.ctc_b200_scale_grads.argtypes = [V, ctypes.c_size_t, V, P(CtcOptions)]
    lib.ctc_b200_scale_grads.restype = I
    lib.ctc_b200_greedy_decode.argtypes = [V, ctypes.c_longlong, ctypes.c_longlong, V, I, I, I, V, V, V, V,
                                           P(CtcOptions)]
    lib.ctc_b200_greedy_decode.restype = I
    lib.ctc_b200_last_launch_count.restype = I
    lib.ctc_b200_last_copy_bytes.argtypes = [P(ctypes.c_ulonglong), P(ctypes.c_ulonglong)]
    lib.ctc_b200_last_copy_bytes.restype = None
    _lib = lib
    return lib


def check(status: int, what: str = "ctc_b200") -> None:
    if status != 0:
        msg = load().ctc_b200_status_string(status).decode()
        raise RuntimeError(f"{what}: {msg} (status {status})")
