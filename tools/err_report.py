"""Print the worst cost / gradient deviation of the CUDA path from the fp64 CPU oracle for every workload config
(a development aid: the pass/fail bars live in tests/test_gpu_parity.py).  Run on the B200 box from the repo root."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402
from oracle import ctc_oracle  # noqa: E402  (checker only)


def main():
    ctc_oracle.build()
    names = sys.argv[1:] or ["cpu_b16_t500", "headline_b64_t1500", "long_b32_t4000"]
    for name in names:
        for sigma in (1.0, 5.0, 20.0):
            d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name], sigma=sigma)
            costs, grads = b200ctc.ctc_forward_backward(d["acts"].cuda(), d["labels"], d["act_lens"], d["label_lens"],
                                                        blank=0, zero_infinity=True, grad_scale=1.0, want_grads=True)
            torch.cuda.synchronize()
            oc, og = ctc_oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(),
                                        d["label_lens"].numpy(), zero_infinity=True)
            c = costs.cpu().numpy().astype(np.float64)
            g = grads.cpu().numpy().astype(np.float64)
            fin = np.isfinite(oc)
            rel = np.abs(c[fin] - oc[fin]) / np.maximum(np.abs(oc[fin]), 1e-30)
            print(f"{name:22s} sigma {sigma:5.1f}  cost rel {rel.max():.2e}  grad abs {np.abs(g - og).max():.2e}  "
                  f"rowsum {np.abs(g.sum(axis=2)).max():.2e}")


if __name__ == "__main__":
    main()
