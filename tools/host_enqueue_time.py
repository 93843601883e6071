"""Host enqueue time against device time of one criterion step (diagnostic).
usage: python tools/host_enqueue_time.py [workload]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "headline_b64_t1500"
d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name])
acts, lab, al, ll = d["acts"].cuda(), d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
lmax = int(d["label_lens"].max())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for env in sys.argv[2:] or ["CTC_B200_GRAD_CLASSES=1", "CTC_B200_GRAD_CLASSES=0"]:
    k, v = env.split("=")
    os.environ[k] = v
    for _ in range(5):
        b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
    torch.cuda.synchronize()
    N = 100
    t0 = time.perf_counter()
    for _ in range(N):
        b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    # device time per step with the L2 flushed, the host running ahead
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(30)]
    for a, b in ev:
        flush.zero_()
        a.record()
        b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
        b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    print(f"{env}: host enqueue {1e6 * (t1 - t0) / N:.1f} us/call, back-to-back wall {1e6 * (t2 - t0) / N:.1f} us/call, "
          f"flushed device step median {1e3 * ms[len(ms) // 2]:.1f} us (min {1e3 * ms[0]:.1f})")
