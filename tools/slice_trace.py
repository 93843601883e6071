"""Device timeline of the sliced launch (CTC_B200_TRACE=1): per length-bucket slice, when its engine and its gradient
kernel finished, in us after the prep kernel.  Run on the B200 box:  python tools/slice_trace.py [workload]"""
import os
import sys

os.environ.setdefault("CTC_B200_TRACE", "1")
import torch  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "headline_b64_t1500"
    d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name])
    acts = d["acts"].cuda().requires_grad_(True)
    lab, al, ll = d["labels"], d["act_lens"], d["label_lens"]  # CPU int32, as the reference passes them
    crit = b200ctc.CTCLoss()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for i in range(6):
        torch.cuda.synchronize()
        for _ in range(8):  # ~0.4 ms of device work ahead of the call: the host enqueues the step while the device
            flush.zero_()   # is still busy, as in a training loop (and the L2 is flushed)
        loss = crit(acts, lab, al, ll)
        torch.cuda.synchronize()
    print("loss", float(loss), file=sys.stderr)


if __name__ == "__main__":
    main()
