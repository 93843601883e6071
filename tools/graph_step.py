"""Eager launch vs CUDA-graph replay of the whole loss step (device time, L2 flushed between steps).
Run on the B200 box from the repo root:  python tools/graph_step.py [workload]"""
import json
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "headline_b64_t1500"
    d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name])
    acts = d["acts"].cuda()
    lab, al, ll = d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
    lmax = int(d["label_lens"].max())
    grads = torch.empty_like(acts)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step():
        return b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()

    def timed(fn, n=100):
        evs = []
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return statistics.mean(a.elapsed_time(b) for a, b in evs)

    timed(step, 10)
    timed(graph.replay, 10)
    print(json.dumps({"workload": name, "eager_ms": timed(step), "graph_replay_ms": timed(graph.replay),
                      "eager_ms_2": timed(step), "graph_replay_ms_2": timed(graph.replay)}))


if __name__ == "__main__":
    main()
