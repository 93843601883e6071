"""Step time of shards of the B = 1024 sweep batch with and without the gradient kernel's warp classes (diagnostic).
usage: python tools/grad_classes_sweep.py 64 128 256 512"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

w = b200ctc.workloads.CONFIGS["sweep_b1024_t1500"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for B in [int(a) for a in sys.argv[1:]] or [128, 256]:
    d = b200ctc.workloads.make(w, B=B)
    acts, lab, al, ll = d["acts"].cuda(), d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
    lmax = int(d["label_lens"].max())
    res = {}
    for cls in ("0", "1"):
        os.environ["CTC_B200_GRAD_CLASSES"] = cls
        for _ in range(5):
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
        for a, b in ev:
            flush.zero_()
            a.record()
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
            b.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        res[cls] = ms[len(ms) // 2]
    print(f"B={B}: one launch {1e3 * res['0']:.1f} us, warp classes {1e3 * res['1']:.1f} us")
