"""Re-run one case of tests/test_gpu_parity.py::test_randomised_shapes and print per-utterance costs against the oracle.
usage: python tools/debug_random_shape.py SEED"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import b200ctc  # noqa: E402
from oracle import ctc_oracle as oracle  # noqa: E402

oracle.build()

seed = int(sys.argv[1])
rng = np.random.default_rng(1000 + seed)
C = int(rng.choice([2, 3, 7, 29, 32, 33, 64, 97, 128, 129, 300]))
B = int(rng.choice([1, 3, 16, 21, 40]))
T = int(rng.choice([1, 9, 64, 257, 600]))
blank = int(rng.integers(0, C))
os.environ["CTC_B200_SLICES"] = str(int(rng.choice([1, 2, 5, 8])))
os.environ["CTC_B200_GRAD_CLASSES"] = str(int(rng.integers(0, 2)))
act_lens = rng.integers(max(1, T // 3), T + 1, size=B).astype(np.int32)
act_lens[rng.integers(0, B)] = T
lens, labs = [], []
symbols = [c for c in range(C) if c != blank]
for b in range(B):
    n = int(rng.integers(0, min(int(act_lens[b]), 400) + 1))
    if rng.random() < 0.3:
        n = n // 8
    few = rng.random() < 0.3
    pool = symbols[:2] if few and len(symbols) > 1 else symbols
    lab = rng.choice(pool, size=n).astype(np.int32) if n else np.zeros(0, np.int32)
    lens.append(n)
    labs.append(lab)
acts = (rng.standard_normal((T, B, C)) * rng.choice([0.5, 1.0, 4.0])).astype(np.float32)
labels = np.concatenate(labs) if labs else np.zeros(0, np.int32)
ll = np.array(lens, dtype=np.int32)
print("C", C, "B", B, "T", T, "blank", blank)
oc, og = oracle.compute(acts, labels, act_lens, ll, blank=blank)
c, g = b200ctc.ctc_forward_backward(torch.from_numpy(acts).cuda(), torch.from_numpy(labels), torch.from_numpy(act_lens),
                                    torch.from_numpy(ll), blank=blank)
c = c.cpu().numpy().astype(np.float64)
for b in range(B):
    rel = abs(c[b] - oc[b]) / max(abs(oc[b]), 1e-30) if np.isfinite(oc[b]) else 0.0
    if rel > 1e-6:
        print(f"b={b} T={act_lens[b]} L={lens[b]} gpu {c[b]:.9g} oracle {oc[b]:.9g} rel {rel:.2e} abs {abs(c[b]-oc[b]):.2e}")
print("grad max abs err", np.abs(g.cpu().numpy().astype(np.float64) - og).max())
