"""tests/numeric_model.py -- numpy model of the NUMERIC SCHEME the CUDA lattice kernels use (test support).

Not an oracle and not the product: it mirrors, array-for-thread, the arithmetic of csrc/ctc_device.cuh and
csrc/ctc_kernels.cuh so the representation can be checked on the CPU against oracle/ before any GPU time is spent:

  * emissions y = softmax(acts) in fp32, clamped to >= 2^-126, widened (exactly) to fp64 (the kernels have since gone
    to the rounded high word of that value, 21 significant bits: DESIGN.md section 3; this model keeps the exact one)
  * alpha / beta in the LINEAR domain, fp64 mantissa, one int exponent per group of K adjacent states
    ("thread"), value = m * 2^E
  * renormalisation every N steps: group max -> [1,2)  (exact power-of-two scaling)
  * neighbour-group inflow scaled by 2^(E_nbr - E_me), clamped to the fp64 exponent range
  * posterior(t,s) = m_alpha * m_betabar * 2^(E_alpha + E_beta - E_P) / m_P
  * engine / gradient split with checkpoints every N steps; checkpoints and parked alpha rows kept as rounded high
    words (21 significant bits), which is where the gradient's extra ~5e-7 comes from
(The exponent lift is modelled as a scan with LIFT = 500; the kernel iterates a neighbour rule with kLift = 250 --
both only act on threads whose content is negligible next to their inflow.)
"""
from __future__ import annotations

import numpy as np


def softmax_f32(acts_tc: np.ndarray) -> np.ndarray:
    a = acts_tc.astype(np.float32)
    m = a.max(axis=1, keepdims=True)
    e = np.exp(a - m, dtype=np.float32)
    y = e / e.sum(axis=1, keepdims=True, dtype=np.float32)
    return np.maximum(y, np.float32(2.0 ** -126)).astype(np.float64)


LIFT = 500


def _pow2(d):
    d = np.clip(d, -1100, 1000)
    return np.ldexp(1.0, d.astype(np.int64))


def _renorm(v, E, inflow_from_below: bool):
    """v [G,K] mantissas, E [G] exponents.  Live groups: scale so the group max lands in [1,2).
    Dead groups (max is zero/denormal): flush, and adopt the exponent of the NEAREST LIVE group on the side
    they receive inflow from (kernel: ballot + clz/ffs + one indexed shuffle), so that a wavefront running at
    full speed (2 states/step, the T == L + repeats case) is never scaled by a stale 2^(E_nbr - E_me)."""
    mx = v.max(axis=1)
    live = mx >= 2.0 ** -1022
    _, ex = np.frexp(np.where(live, mx, 1.0))  # mx = f * 2^ex, f in [0.5,1)
    shift = np.where(live, 1 - ex, 0)
    v = np.where(live[:, None], v * np.ldexp(1.0, shift)[:, None], 0.0)
    E1 = E - shift
    G = E1.shape[0]
    idx = np.arange(G)
    if inflow_from_below:
        src = np.maximum.accumulate(np.where(live, idx, -1))
    else:
        src = np.minimum.accumulate(np.where(live, idx, G)[::-1])[::-1]
    ok = (src >= 0) & (src < G)
    E2 = np.where(ok, E1[np.clip(src, 0, G - 1)], E1)
    # Inflow guard: if the inflow neighbour's scale is more than 2^LIFT above mine, my content is negligible
    # next to what is about to flow in -> lift my exponent (mantissas shrink, possibly to 0).  Keeps the
    # boundary factor 2^(E_nbr - E_me) <= 2^LIFT so nothing overflows between two renormalisations.
    # (a warp scan in the kernel: E = max(E, shfl(E, d) - d*LIFT) for d = 1,2,4,8,16)
    E3 = E2.copy()
    d = 1
    while d < G:
        sh = np.full_like(E3, -(1 << 40))
        if inflow_from_below:
            sh[d:] = E3[:-d]
        else:
            sh[:-d] = E3[d:]
        E3 = np.maximum(E3, sh - d * LIFT)
        d *= 2
    lift = E3 - E2
    v = v * np.ldexp(1.0, -np.minimum(lift, 2000))[:, None]
    return v, E3


class Lattice:
    def __init__(self, label, S_pad_K: int, blank: int = 0):
        K = S_pad_K
        L = len(label)
        self.S = 2 * L + 1
        self.K = K
        self.G = (self.S + K - 1) // K
        Sp = self.G * K
        ext = np.full(Sp, blank, dtype=np.int64)
        ext[1:self.S:2] = np.asarray(label, dtype=np.int64)
        self.ext = ext
        valid = np.arange(Sp) < self.S
        self.valid = valid
        skip_a = np.zeros(Sp)
        skip_b = np.zeros(Sp)
        for s in range(self.S):
            if s % 2 == 1:
                if s == 1 or ext[s] != ext[s - 2]:
                    skip_a[s] = 1.0  # s==1: the virtual alpha_{-1}(-1)=1 inflow
                if s + 2 < self.S and ext[s + 2] != ext[s]:
                    skip_b[s] = 1.0
        self.skip_a = skip_a.reshape(self.G, K)
        self.skip_b = skip_b.reshape(self.G, K)

    def yrow(self, y_t):
        return np.where(self.valid, y_t[self.ext], 0.0).reshape(self.G, self.K)


def alpha_step(lat: Lattice, v, E, y_t, first: bool):
    G, K = lat.G, lat.K
    f = np.ones(G)
    f[1:] = _pow2(E[:-1] - E[1:])
    in1 = np.zeros(G)
    in2 = np.zeros(G)
    in1[1:] = v[:-1, K - 1] * f[1:]
    in2[1:] = v[:-1, K - 2] * f[1:]
    if first:
        in1[0] = 1.0
    p1 = np.concatenate([in1[:, None], v[:, :-1]], axis=1)
    p2 = np.concatenate([in2[:, None], in1[:, None], v[:, :-2]], axis=1)
    return (v + p1 + lat.skip_a * p2) * lat.yrow(y_t)


def betabar(lat: Lattice, v, E):
    """beta-bar_t from standard beta_{t+1} (mantissas v, exps E); same exponents."""
    G, K = lat.G, lat.K
    f = np.ones(G)
    f[:-1] = _pow2(E[1:] - E[:-1])
    in1 = np.zeros(G)
    in2 = np.zeros(G)
    in1[:-1] = v[1:, 0] * f[:-1]
    in2[:-1] = v[1:, 1] * f[:-1]
    n1 = np.concatenate([v[:, 1:], in1[:, None]], axis=1)
    n2 = np.concatenate([v[:, 2:], in1[:, None], in2[:, None]], axis=1)
    return v + n1 + lat.skip_b * n2


def round_hi(v):
    """What csrc/ctc_device.cuh hi_round keeps of a non-negative double: the high word, rounded to nearest --
    the full exponent range and 21 significant bits."""
    m, e = np.frexp(v)
    return np.ldexp(np.round(m * 2.0 ** 21) / 2.0 ** 21, e)


def run(acts_tc, label, K: int = 8, N: int = 8, blank: int = 0, hi_words: bool = True):
    """Full fwd+bwd for one utterance with the scaled-linear scheme.  Returns (cost, grad[T,C]).

    Mirrors the kernels' split: the ENGINE runs the alpha and beta chains in full fp64 and leaves a checkpoint
    (mantissas, exponents) at every group boundary; the GRADIENT pass restarts both chains from the checkpoints of
    each group and multiplies parked alpha rows into live betabar rows.  With hi_words the checkpoints and the parked
    rows keep only the rounded high word of each double, as on the device."""
    T, C = acts_tc.shape
    y = softmax_f32(acts_tc)
    lat = Lattice(label, K, blank)
    G = lat.G
    S = lat.S
    keep = round_hi if hi_words else (lambda x: x)
    ngroups = (T + N - 1) // N
    # ---- engine, alpha: checkpoint g = state BEFORE row g*N (after renormalisation)
    ckA = [None] * ngroups
    v = np.zeros((G, K))
    E = np.zeros(G, dtype=np.int64)
    for t in range(T):
        if t % N == 0:
            ckA[t // N] = (v.copy(), E.copy())
        v = alpha_step(lat, v, E, y[t], t == 0)
        if (t + 1) % N == 0 or t == T - 1:
            v, E = _renorm(v, E, True)
    # ---- P = alpha_{T-1}(S-1) + alpha_{T-1}(S-2)
    flat = v.reshape(-1)
    gl, gs = (S - 1) // K, (S - 2) // K
    eP = max(E[gl], E[gs]) if S > 1 else E[gl]
    mP = flat[S - 1] * _pow2(np.array(E[gl] - eP))
    if S > 1:
        mP = mP + flat[S - 2] * _pow2(np.array(E[gs] - eP))
    if mP == 0:
        return float("inf"), np.zeros((T, C))
    fr, ex = np.frexp(mP)  # normalise P's mantissa (kernel: exponent field on one thread)
    mP, eP = fr * 2.0, eP + int(ex) - 1
    cost = -(np.log(mP) + eP * np.log(2.0))
    # ---- engine, beta: checkpoint g = standard beta AFTER row (g+1)*N, i.e. the state the group g sweep starts from
    ckB = [None] * ngroups
    vb = np.zeros((G, K))
    vb.reshape(-1)[S - 1] = 1.0  # virtual beta_T
    Eb = np.zeros(G, dtype=np.int64)
    for t in range(T - 1, -1, -1):
        if t == T - 1 or (t + 1) % N == 0:
            ckB[t // N] = (vb.copy(), Eb.copy())
        vb = betabar(lat, vb, Eb) * lat.yrow(y[t])
        if t % N == 0:
            vb, Eb = _renorm(vb, Eb, False)
    # ---- gradient pass, one group at a time from the (rounded) checkpoints
    grad = np.zeros((T, C))
    for g in range(ngroups):
        t0, t1 = g * N, min(T, (g + 1) * N)
        va, Ea = ckA[g]
        va = keep(va) if g > 0 else va            # block 0 starts from the exact virtual state
        rows = []
        for t in range(t0, t1):
            va = alpha_step(lat, va, Ea, y[t], t == 0)
            rows.append(keep(va))                 # parked row
        vb, Eb = ckB[g]
        vb = keep(vb) if t1 < T else vb           # the last block starts from the exact virtual beta_T
        for t in range(t1 - 1, t0 - 1, -1):
            bb = betabar(lat, vb, Eb)
            gsc = _pow2(Ea + Eb - eP) / mP
            post = (rows[t - t0] * bb * gsc[:, None]).reshape(-1)
            occ = np.zeros(C)
            np.add.at(occ, lat.ext[:S], post[:S])
            grad[t] = y[t] - occ
            vb = bb * lat.yrow(y[t])
    return float(cost), grad
