"""oracle/ctc_oracle.py -- TEST INFRASTRUCTURE ONLY.

Python face of the CPU oracle for the CTC loss path the reference applies at
/root/reference/train.py:600 (train) and :352/:457 (eval) through `warpctc_pytorch.CTCLoss`
(/root/reference/train.py:12, :904).  PARITY UNPINNED: that third-party package is absent and un-pinned and the
reference has no tests or golden vectors for the path; see ctc_oracle_impl.h.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product (deepspeech.pytorch_b200/) never does.

Contents
  compute(...)        ctypes call into libctc_oracle.so (fp64 golden, or fp32 timed baseline)
  brute_force(...)    independent second opinion: enumerate all C^T alignments (tiny cases only)
"""
from __future__ import annotations

import ctypes
import itertools
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libctc_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C oracle (gcc, OpenMP).  Building the checker is not using it."""
    src = [os.path.join(_HERE, f) for f in ("ctc_oracle.c", "ctc_oracle_impl.h")]
    stale = (not os.path.exists(_LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "libctc_oracle.so"], check=True, capture_output=True)
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        common = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                  ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        for name in ("ctc_oracle_compute_f64", "ctc_oracle_compute_f32"):
            fn = getattr(lib, name)
            fn.argtypes = common
            fn.restype = ctypes.c_int
        lib.ctc_oracle_max_threads.restype = ctypes.c_int
        _lib = lib
    return _lib


def max_threads() -> int:
    return int(_load().ctc_oracle_max_threads())


def compute(acts, labels, act_lens, label_lens, blank: int = 0, zero_infinity: bool = True, want_grads: bool = True,
            precision: str = "f64", num_threads: int = 0):
    """CTC costs [B] (float64) and d(sum costs)/d acts [T,B,C] for fp32 logits `acts` [T,B,C].

    Same argument order as the reference call `criterion(acts, labels, act_lens, label_lens)` (train.py:600).
    precision "f64": golden oracle (grads float64).  "f32": fp32 log-space timed baseline (grads float32).
    """
    lib = _load()
    acts = np.ascontiguousarray(acts, dtype=np.float32)
    T, B, C = acts.shape
    labels = np.ascontiguousarray(labels, dtype=np.int32).reshape(-1)
    act_lens = np.ascontiguousarray(act_lens, dtype=np.int32).reshape(-1)
    label_lens = np.ascontiguousarray(label_lens, dtype=np.int32).reshape(-1)
    assert act_lens.shape[0] == B and label_lens.shape[0] == B
    assert int(label_lens.sum()) == labels.shape[0], "labels must be the flat concatenation of the targets"
    costs = np.zeros(B, dtype=np.float64)
    gdt = np.float64 if precision == "f64" else np.float32
    grads = np.zeros((T, B, C), dtype=gdt) if want_grads else None
    fn = lib.ctc_oracle_compute_f64 if precision == "f64" else lib.ctc_oracle_compute_f32
    rc = fn(acts.ctypes.data, T, B, C, labels.ctypes.data if labels.size else None, label_lens.ctypes.data,
            act_lens.ctypes.data, int(blank), int(bool(zero_infinity)), costs.ctypes.data,
            grads.ctypes.data if want_grads else None, int(num_threads))
    if rc != 0:
        raise ValueError("ctc oracle: invalid arguments (label out of range / equal to blank / bad length)")
    return costs, grads


# ----------------------------------------------------------------------------------------------------------------
# Second opinion: definition-level enumeration (Graves et al. 2006, eq. (3)-(4)): p(l|x) = sum over all
# alignments pi in C^T with collapse(pi) == l of prod_t softmax(x_t)[pi_t].  Exponential: tiny T, C only.
# ----------------------------------------------------------------------------------------------------------------
def _collapse(path, blank):
    out, prev = [], None
    for p in path:
        if p != prev and p != blank:
            out.append(p)
        prev = p
    return tuple(out)


def brute_force_cost(acts_tb, label, blank: int = 0) -> float:
    """-log p(label | acts) for ONE utterance, acts_tb [T,C] (float64 maths), by enumerating alignments."""
    a = np.asarray(acts_tb, dtype=np.float64)
    T, C = a.shape
    a = a - a.max(axis=1, keepdims=True)
    y = np.exp(a) / np.exp(a).sum(axis=1, keepdims=True)
    target = tuple(int(v) for v in label)
    p = 0.0
    for path in itertools.product(range(C), repeat=T):
        if _collapse(path, blank) == target:
            q = 1.0
            for t, c in enumerate(path):
                q *= y[t, c]
            p += q
    return float(-np.log(p)) if p > 0 else float("inf")


def brute_force_grad(acts_tb, label, blank: int = 0, eps: float = 1e-6):
    """Central finite differences of brute_force_cost w.r.t. the logits (tiny cases only)."""
    a = np.asarray(acts_tb, dtype=np.float64)
    g = np.zeros_like(a)
    for idx in np.ndindex(*a.shape):
        ap = a.copy(); ap[idx] += eps
        am = a.copy(); am[idx] -= eps
        g[idx] = (brute_force_cost(ap, label, blank) - brute_force_cost(am, label, blank)) / (2 * eps)
    return g


def count_alignments(T: int, label, blank: int = 0) -> int:
    """#alignments of length T collapsing to `label` (for the uniform-logits KAT: cost = -log(n / C^T))."""
    L = len(label)
    S = 2 * L + 1
    ext = [blank if s % 2 == 0 else int(label[s // 2]) for s in range(S)]
    a = [0] * S
    a[0] = 1
    if S > 1:
        a[1] = 1
    for _ in range(1, T):
        n = [0] * S
        for s in range(S):
            v = a[s]
            if s >= 1:
                v += a[s - 1]
            if s >= 2 and ext[s] != blank and ext[s] != ext[s - 2]:
                v += a[s - 2]
            n[s] = v
        a = n
    return a[S - 1] + (a[S - 2] if S > 1 else 0)
