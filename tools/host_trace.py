"""Timeline of the host-buffer entry (ctc_b200_compute_host) on the headline workload: runs a few calls with
CTC_B200_HOST_TRACE=1 so the library prints, per utterance slice, when its upload, kernels and download finished.
Run on the B200 box from the repo root:  python tools/host_trace.py [workload] [--sorted]"""
import ctypes
import os
import sys
import time

os.environ["CTC_B200_HOST_TRACE"] = "1"
import torch  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    name = args[0] if args else "headline_b64_t1500"
    d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name])
    if "--sorted" in sys.argv:  # the reference's batch order (data_loader_aug.py:527): longest first
        order = torch.argsort(d["act_lens"], descending=True, stable=True)
        offs = torch.cumsum(d["label_lens"], 0) - d["label_lens"]
        d["labels"] = torch.cat([d["labels"][int(offs[b]):int(offs[b]) + int(d["label_lens"][b])]
                                 for b in order.tolist()])
        d["acts"] = d["acts"][:, order, :].contiguous()
        d["act_lens"] = d["act_lens"][order].contiguous()
        d["label_lens"] = d["label_lens"][order].contiguous()
    T, B, C = d["acts"].shape
    lib = b200ctc.pkg._lib.load()
    h_acts = d["acts"].pin_memory()
    h_grads = torch.empty_like(d["acts"]).pin_memory()
    h_costs = torch.empty(B, dtype=torch.float32).pin_memory()
    ctx = ctypes.c_void_p()
    b200ctc.pkg._lib.check(lib.ctc_b200_create(ctypes.byref(ctx)))
    opts = b200ctc.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(opts))
    opts.stream = torch.cuda.current_stream().cuda_stream
    loss = ctypes.c_double(0.0)
    hl, hal, hll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()
    for i in range(6):
        t0 = time.perf_counter()
        st = lib.ctc_b200_compute_host(ctx, h_acts.data_ptr(), h_grads.data_ptr(), hl.data_ptr(), hll.data_ptr(),
                                       hal.data_ptr(), T, B, C, h_costs.data_ptr(), ctypes.byref(loss),
                                       ctypes.byref(opts))
        b200ctc.pkg._lib.check(st, "ctc_b200_compute_host")
        print(f"call {i}: wall {1e6 * (time.perf_counter() - t0):.0f} us, loss {loss.value:.3f}", file=sys.stderr)
    lib.ctc_b200_destroy(ctx)


if __name__ == "__main__":
    main()
