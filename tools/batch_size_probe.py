"""Step time of the first B utterances' worth of the sweep batch under environment settings given as arguments
(diagnostic).  usage: python tools/batch_size_probe.py B VAR=value [VAR=value ...]   (each setting is timed separately)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

B = int(sys.argv[1])
d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS["sweep_b1024_t1500"], B=B)
acts, lab, al, ll = d["acts"].cuda(), d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
lmax = int(d["label_lens"].max())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for env in sys.argv[2:]:
    k, v = env.split("=")
    os.environ[k] = v
    out = []
    for mask in (0, 3, 6, 5):  # whole step, prep only, engine only, gradient only (phase_mask bits: skip engine/grad/prep)
        for _ in range(4):
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, phase_mask=mask)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
        for a, b in ev:
            flush.zero_()
            a.record()
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, phase_mask=mask)
            b.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        out.append(1e3 * ms[len(ms) // 2])
    print(f"B={B} {env}: step {out[0]:.1f} us, prep {out[1]:.1f}, engine {out[2]:.1f}, gradient {out[3]:.1f}")
