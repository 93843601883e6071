"""Timing of the logits producer (ctc_b200_fc_forward) on the headline frame count: T=1500, B=64, H=800 (the
reference's rnn_hidden_size default), C=30.  CUDA events on the launching stream, L2 flushed between timed calls.
Beside it: torch's own eval-mode layer (F.batch_norm + F.linear = cuBLAS fp32) on the same tensors, and the chain
"layer -> CTC loss+grad" both ways (logits re-read by the criterion vs emission rows handed over by the fc kernel).
Prints one JSON line.  Run on the B200 box:  python tools/bench_fc.py"""
import json
import os
import statistics
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402


def run():
    """-> dict (one JSON-able record); importable from bench.py"""
    w = b200ctc.workloads.CONFIGS["headline_b64_t1500"]
    d = b200ctc.workloads.make(w, C=30, acts=False)
    T, B, H, C = w.T, w.B, 800, 30
    if os.environ.get("FC_T"):  # experiment: another frame count (and FC_NOFLUSH=1: inputs stay in L2 between calls)
        T = int(os.environ["FC_T"])
        d["act_lens"].clamp_(max=T)
        d["label_lens"].clamp_(max=max(1, T // 4))
        d["labels"] = d["labels"][: int(d["label_lens"].sum())]
    noflush = os.environ.get("FC_NOFLUSH") == "1"
    g = torch.Generator().manual_seed(5)
    x = torch.randn(T, B, H, generator=g).cuda()
    weight = ((torch.rand(C, H, generator=g) * 2 - 1) / H ** 0.5).cuda()
    gamma, beta = (torch.rand(H, generator=g) + 0.5).cuda(), torch.zeros(H).cuda()
    mean, var = torch.zeros(H).cuda(), torch.ones(H).cuda()
    lab, al, ll = d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
    lmax = int(d["label_lens"].max())
    ws = b200ctc.fc.ctc_workspace_for(T, B, C, lmax, "cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    grads = torch.empty(T, B, C, device="cuda")

    flush_i32 = flush.view(torch.int32)

    def timed(fn, n=20, flushed=True, read_flush=False):
        for _ in range(3):
            fn()
        evs = []
        for _ in range(n):
            for _ in range(6 if flushed and not noflush else 0):
                if read_flush:
                    flush_i32.sum()  # evicts the L2 with CLEAN lines (nothing to write back during the timed call)
                else:
                    flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return statistics.mean(a.elapsed_time(b) for a, b in evs)

    fc = lambda **kw: b200ctc.fc.fc_forward(x, weight, gamma, beta, mean, var, 1e-5, **kw)
    ms_logits = timed(lambda: fc())
    # no flush between calls: the 307 MB input is 2.4x the L2, so every call still reads it from HBM, but the call does
    # not start behind 126 MB of dirty flush lines that are written back while x streams in
    ms_logits_back_to_back = timed(lambda: fc(), flushed=False)
    # L2 evicted by READING a 256 MB buffer: cold like the write flush, but no dirty lines, and the host is ahead of the
    # device (the back-to-back figure above contains the host's launch latency, the device being idle between calls)
    ms_logits_read_flush = timed(lambda: fc(), read_flush=True)
    ms_all = timed(lambda: fc(ctc_workspace=ws, act_lens=al, want_path=True))
    ms_stats = timed(lambda: fc(training=True))

    def torch_layer():
        return F.linear(F.batch_norm(x.view(-1, H), mean, var, gamma, beta, False, 0.0, 1e-5), weight).view(T, B, C)

    ms_torch = timed(torch_layer)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    ms_torch_tf32 = timed(torch_layer)
    torch.backends.cuda.matmul.allow_tf32 = prev

    def chain_plain():
        lg = fc()["logits"]
        b200ctc.ctc_forward_backward(lg, lab, al, ll, max_label_len=lmax, out_grads=grads)

    def chain_fused():
        lg = fc(ctc_workspace=ws, act_lens=al)["logits"]
        b200ctc.ctc_forward_backward(lg, lab, al, ll, max_label_len=lmax, out_grads=grads, workspace=ws,
                                     emissions_ready=True)

    ms_plain, ms_fused = timed(chain_plain), timed(chain_fused)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        kind = "measured"
    except Exception:
        peak, kind = 6650.0, "fallback"
    frames = T * B
    alg = frames * (4 * H + 4 * C)  # read x once, write the fp32 logits once
    alg_all = alg + int(al.sum().item()) * 8 * 32 + frames * 4
    return ({
        "kernel": "fc_logits_kernel<32>", "config": {"T": T, "B": B, "H": H, "C": C},
        "ms_logits_only": ms_logits, "ms_logits_only_back_to_back_no_flush": ms_logits_back_to_back,
        "ms_logits_only_after_read_flush": ms_logits_read_flush,
        "ms_logits_emissions_path": ms_all, "ms_training_mode_batch_stats": ms_stats,
        "roofline": {"bound": "hbm", "alg_bytes": alg, "achieved": alg / (ms_logits * 1e-3) / 1e9, "peak": peak,
                     "peak_kind": kind, "unit": "GB/s", "frac": alg / (ms_logits * 1e-3) / 1e9 / peak,
                     "frac_back_to_back_no_flush_input_2p4x_l2": alg / (ms_logits_back_to_back * 1e-3) / 1e9 / peak,
                     "frac_after_read_flush": alg / (ms_logits_read_flush * 1e-3) / 1e9 / peak,
                     "frac_with_emissions": alg_all / (ms_all * 1e-3) / 1e9 / peak},
        "frames_per_s": frames / (ms_logits * 1e-3),
        "comparators": {"torch_batch_norm_plus_linear_fp32_ms": ms_torch, "torch_same_with_allow_tf32_ms": ms_torch_tf32},
        "chain_layer_then_ctc_ms": {"logits_reread_by_criterion": ms_plain, "emission_rows_handed_over": ms_fused}})


if __name__ == "__main__":
    print(json.dumps(run()))
