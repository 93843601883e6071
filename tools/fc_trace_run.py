"""fc kernel under CTC_FC_TRACE=1 (per-role clocks, per-CTA start / end times), with and without an L2 flush in front
of the call.  Run on the B200 box:  CTC_FC_TRACE=1 python tools/fc_trace_run.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

T, B, H, C = 1500, 64, 800, 30
g = torch.Generator().manual_seed(1)
x = torch.randn(T, B, H, generator=g).cuda()
w = (torch.rand(C, H, generator=g) / H ** .5).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for flushed in (False, True, False, True):
    for i in range(3):
        if flushed:
            for _ in range(6):
                flush.zero_()
        print("--- flushed" if flushed else "--- back to back", file=sys.stderr, flush=True)
        b200ctc.fc.fc_forward(x, w)
    torch.cuda.synchronize()
