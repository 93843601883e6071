"""The confidently-predicted batch of tests/test_gpu_parity.py, per-utterance costs against the oracle (diagnostic)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402
from oracle import ctc_oracle as oracle  # noqa: E402

oracle.build()
C, T = int(sys.argv[1]), int(sys.argv[2])
margin = float(sys.argv[3]) if len(sys.argv) > 3 else 9.0
rng = np.random.default_rng(C)
B = 6
acts = rng.standard_normal((T, B, C)).astype(np.float32) * 0.3
labels, lens = [], []
for b in range(B):
    L = T // 4 - 3 * b
    lab = rng.integers(1, C, size=L).astype(np.int32) if C > 2 else np.ones(L, np.int32)
    t = 0
    for sym in lab:
        acts[t:t + 2, b, 0] += margin
        acts[t + 2:t + 4, b, sym] += margin
        t += 4
    acts[t:, b, 0] += margin
    labels.append(lab)
    lens.append(L)
labels = np.concatenate(labels)
al = np.full(B, T, np.int32)
ll = np.array(lens, np.int32)
oc, og = oracle.compute(acts, labels, al, ll)
c32, _ = oracle.compute(acts, labels, al, ll, precision="f32")
c, g = b200ctc.ctc_forward_backward(torch.from_numpy(acts).cuda(), torch.from_numpy(labels), torch.from_numpy(al),
                                    torch.from_numpy(ll))
c = c.cpu().numpy().astype(np.float64)
for b in range(B):
    print(f"b={b} L={lens[b]} gpu {c[b]:.9g} oracle {oc[b]:.9g} (fp32 port {c32[b]:.9g}) abs {abs(c[b]-oc[b]):.3e} rel {abs(c[b]-oc[b])/abs(oc[b]):.2e}")
print("grad max abs err", np.abs(g.cpu().numpy().astype(np.float64) - og).max())
