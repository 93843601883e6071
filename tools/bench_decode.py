"""Measurement of the greedy-decoding row on the headline shape.  Prints ONE JSON line:
device time of the fused by-product (ctc_b200_compute_decode minus ctc_b200_compute), of the stand-alone
ctc_b200_greedy_decode, and the CPU restatement (oracle/greedy_oracle.py: numpy argmax + the reference's per-frame
Python loop, decoder.py:165-180) timed on a bounded sample.  Run on the B200 box from the repo root."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402
from oracle import greedy_oracle as go  # noqa: E402  (checker / CPU baseline only)


def timed(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for _ in range(n):
        ev[0].record()
        fn()
        ev[1].record()
        torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    ts.sort()
    return ts[len(ts) // 2]


def device_time(fn, n=20):
    """GPU time per call with the CPU enqueue cost hidden: the stream is kept busy (torch.cuda._sleep) while all n
    calls are enqueued, then they run back to back between two events."""
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda._sleep(40_000_000)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    w = b200ctc.workloads.CONFIGS[b200ctc.workloads.HEADLINE]
    d = b200ctc.workloads.make(w, seed=1234)
    acts = d["acts"].cuda()
    al, ll, lab = d["act_lens"].cuda(), d["label_lens"].cuda(), d["labels"].cuda()
    lmax = int(d["label_lens"].max())
    grads = torch.empty_like(acts)
    kw = dict(max_label_len=lmax, out_grads=grads)
    t_plain = timed(lambda: b200ctc.ctc_forward_backward(acts, lab, al, ll, **kw))
    t_fused = timed(lambda: b200ctc.ctc_forward_backward(acts, lab, al, ll, decode=True, **kw))
    t_alone = timed(lambda: b200ctc.greedy_decode(acts, al))
    t_alone_dev = device_time(lambda: b200ctc.greedy_decode(acts, al))
    t_plain_dev = device_time(lambda: b200ctc.ctc_forward_backward(acts, lab, al, ll, **kw))
    t_fused_dev = device_time(lambda: b200ctc.ctc_forward_backward(acts, lab, al, ll, decode=True, **kw))
    t0 = time.perf_counter()
    nb = 8  # bounded CPU sample: 8 utterances
    go.decode(d["acts"][:, :nb].numpy(), d["act_lens"][:nb].numpy(), 0)
    t_cpu = (time.perf_counter() - t0) * w.B / nb * 1e3
    frames = w.B * w.T
    # error counts of the best path against the targets: device kernel vs the CPU restatement
    from oracle import wer_oracle as wo  # noqa: E402  (checker / CPU baseline only)
    _, _, best = b200ctc.ctc_forward_backward(acts, lab, al, ll, decode=True, **kw)
    space = b200ctc.ALPHABET.index(" ") if w.C >= 30 else w.C - 1
    t_err_dev = device_time(lambda: b200ctc.error_counts(best["ids"], best["lens"], lab, ll, space, lmax))
    ids_h, lens_h = best["ids"].cpu().numpy(), best["lens"].cpu().numpy()
    t0 = time.perf_counter()
    wo.batch_counts(ids_h[:2], lens_h[:2], d["labels"].numpy(), d["label_lens"][:2].numpy(), space)
    t_err_cpu = (time.perf_counter() - t0) * w.B / 2 * 1e3
    print(json.dumps({"op": "error_counts", "workload": w.name, "device_ms": t_err_dev,
                      "mean_hyp_len": float(lens_h.mean()), "mean_ref_len": float(d["label_lens"].float().mean()),
                      "cpu_port_ms_full_batch": t_err_cpu,
                      "cpu_sample": f"2 of {w.B} utterances, pure-Python DP (the reference uses python-Levenshtein's "
                                    "C implementation, not installed here), scaled"}))
    print(json.dumps({"op": "greedy_decode", "workload": w.name, "frames": frames,
                      "ctc_step_ms": t_plain, "ctc_step_with_decode_ms": t_fused,
                      "fused_extra_ms": t_fused - t_plain, "standalone_ms": t_alone,
                      "device_only": {"ctc_step_ms": t_plain_dev, "ctc_step_with_decode_ms": t_fused_dev,
                                      "standalone_ms": t_alone_dev,
                                      "standalone_frames_per_s": frames / (t_alone_dev * 1e-3),
                                      "standalone_gbs": 4 * w.C * frames / (t_alone_dev * 1e-3) / 1e9},
                      "alg_bytes": 4 * w.C * frames,
                      "cpu_port_ms_full_batch": t_cpu, "cpu_sample": f"{nb} of {w.B} utterances, scaled"}))


if __name__ == "__main__":
    main()
