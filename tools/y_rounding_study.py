"""What would it cost in accuracy to keep only the high word of the fp64 emission probabilities (21 significant bits)
in the lattice kernels?  Plain numpy fp64 CTC on probabilities, exact fp32-softmax y against y rounded to 21 bits.
usage: python tools/y_rounding_study.py [sigma]"""
import sys

import numpy as np


def ctc_occupancy(y, lab, blank=0):
    """y [T,C] probabilities -> (cost, occupancy [T,C]) with per-step rescaling"""
    T, C = y.shape
    L = len(lab)
    S = 2 * L + 1
    ext = np.full(S, blank)
    ext[1::2] = lab
    skip = np.zeros(S, bool)
    skip[3::2] = lab[1:] != lab[:-1]
    ye = y[:, ext]  # [T,S]
    alpha = np.zeros((T, S))
    c = np.zeros(T)
    a = np.zeros(S)
    a[0] = ye[0, 0]
    if S > 1:
        a[1] = ye[0, 1]
    c[0] = a.sum(); a /= c[0]; alpha[0] = a
    for t in range(1, T):
        n = a.copy()
        n[1:] += a[:-1]
        n[2:] += np.where(skip[2:], a[:-2], 0.0)
        a = n * ye[t]
        c[t] = a.sum(); a /= c[t]; alpha[t] = a
    beta = np.zeros((T, S))
    b = np.zeros(S)
    b[S - 1] = 1.0
    if S > 1:
        b[S - 2] = 1.0
    beta[T - 1] = b
    for t in range(T - 2, -1, -1):
        w = b * ye[t + 1]
        n = w.copy()
        n[:-1] += w[1:]
        n[:-2] += np.where(skip[2:], w[2:], 0.0)
        b = n / n.sum()
        beta[t] = b
    post = alpha * beta
    post /= post.sum(axis=1, keepdims=True)
    occ = np.zeros((T, C))
    for s in range(S):
        occ[:, ext[s]] += post[:, s]
    return -np.log(c).sum(), occ


def round21(y):
    bits = y.view(np.uint64)
    hi = ((bits + np.uint64(1 << 31)) >> np.uint64(32)) << np.uint64(32)
    return hi.view(np.float64)


sigma = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
rng = np.random.default_rng(0)
T, C, L = 1500, 29, 400
x = (rng.standard_normal((T, C)) * sigma).astype(np.float32)
m = x.max(1, keepdims=True)
e = np.exp(x - m, dtype=np.float32)
y = np.maximum(e / e.sum(1, keepdims=True, dtype=np.float32), np.float32(2.0 ** -126)).astype(np.float64)
lab = rng.integers(1, C, L)
c0, o0 = ctc_occupancy(y, lab)
c1, o1 = ctc_occupancy(round21(y.copy()), lab)
print(f"sigma {sigma}: cost {c0:.6f} vs {c1:.6f} rel {abs(c1 - c0) / c0:.2e}; occupancy max abs diff {np.abs(o1 - o0).max():.2e}")
