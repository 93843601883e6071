"""Repeat the B = 1024 step and one shard of it; print the gradient row that differed (diagnostic)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS["sweep_b1024_t1500"])
acts = d["acts"].cuda()
outs = []
for cls in ("1", "1", "0", "0"):
    os.environ["CTC_B200_GRAD_CLASSES"] = cls
    c, g = b200ctc.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"])
    torch.cuda.synchronize()
    outs.append(g.cpu())
    print("classes", cls, "row", outs[-1][1307, 168, :6].tolist())
for i in range(1, 4):
    dg = (outs[i] - outs[0]).abs()
    print("run", i, "vs 0: entries differing", int((dg > 0).sum()))
