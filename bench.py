#!/usr/bin/env python
"""bench.py -- CTC fwd+bwd frames/s (B*T padded frames) on BASELINE.json's headline shape.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]

One "step" = one pass of the hot path (ctc_b200_compute: costs + gradients) over one synthetic batch of the
workload (default: configs[1], B=64 T=1500 C=29, LibriSpeech-shaped variable lengths).  At N>1 (torchrun, one
rank per GPU) every rank owns its own batch of that shape (the reference's DDP flow: each rank has its own batch,
train.py:660-663), i.e. weak scaling, and each step ends with the path's one exchange: an all-reduce of the
summed loss (replaces data/utils.py:40-44).

value      whole-job frames/s, inputs resident in HBM, device-timed (CUDA events), max over ranks
e2e        the same metric through the C ABI's host-buffer entry (ctc_b200_compute_host): pinned HOST acts in,
           HOST grads + costs out, copies inside the timed region
roofline   dominant kernel, algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json
cpu_baseline  the oracle's fp32 port on the box's host cores (bounded sample), a reported baseline

--impl reference: the reference's CPU CTC (warp-ctc) is an absent third-party dependency, so the arm times the
oracle's fp32 OpenMP port of that path on all host threads (the one other place oracle/ may be executed).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ctc_fwd_bwd_frames_per_sec"
UNIT = "frames/s"


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for (ts, r) in self.rows if (t0 is None or ts >= t0 - 0.05) and (t1 is None or ts <= t1 + 0.15)]
        rows = rows or [r for (_, r) in self.rows]
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_port(d, steps=3, warmup=1, threads=0):
    """Time the oracle's fp32 OpenMP port on the host cores over the given batch; returns (frames/s, cores)."""
    from oracle import ctc_oracle
    a, lab, al, ll = d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy()
    # all the host cores this process may run on (torchrun exports OMP_NUM_THREADS=1: do not let that shrink the arm)
    cores = threads or max(len(os.sched_getaffinity(0)), ctc_oracle.max_threads())
    for _ in range(warmup):
        ctc_oracle.compute(a, lab, al, ll, precision="f32", num_threads=cores)
    best = float("inf")
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        ctc_oracle.compute(a, lab, al, ll, precision="f32", num_threads=cores)
        ts.append(time.perf_counter() - t0)
        best = min(best, ts[-1])
    frames = d["T"] * d["B"]
    return frames / best, cores, sum(ts) / len(ts)


def load_traffic(kernel, workload):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` capture of this workload (or None)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if workload in t.get("workloads", {}):
            return t["workloads"][workload]["bytes"][kernel]
        return t["bytes"][kernel] if t.get("workload") == workload else None
    except (OSError, KeyError, ValueError):
        return None


def workload_config(w, d):
    """The `config` object both arms print (same keys, so the driver can match the lines)."""
    return {"workload": w.name, "B_per_gpu": d["B"], "T": d["T"], "C": d["C"],
            "max_label_len": int(d["label_lens"].max()), "valid_frames": int(d["act_lens"].sum())}


def run_reference(args, w):
    """--impl reference: rank 0 only, host cores only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import b200ctc
    d = b200ctc.workloads.make(w, seed=1234)
    fps, cores, mean_s = cpu_port(d, steps=max(1, args.steps), warmup=max(1, min(args.warmup, 2)))
    sample = f"full batch B={d['B']} T={d['T']} C={d['C']} per step, best of {max(1, args.steps)} steps"
    line = {"impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": d["T"] * d["B"] / fps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(w, d),
                           note="reference CTC (warpctc_pytorch) absent from the image: oracle fp32 OpenMP port, "
                                "host buffers only"),
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def run_b200(args, w):
    import ctypes

    import torch
    import torch.distributed as dist

    import b200ctc

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    if args.global_batch and world > 1:
        # BASELINE config 5: ONE global batch of the workload shape, dealt to the ranks by lattice work
        # (sharding.length_bucketed_shards: LPT, equal shard sizes) -- strong scaling, no data-path collective
        dg = b200ctc.workloads.make(w, seed=1234)
        ids = b200ctc.sharding.length_bucketed_shards(dg["act_lens"].tolist(), dg["label_lens"].tolist(), world)[rank]
        a_, l_, al_, ll_ = b200ctc.sharding.take_shard(dg["acts"], dg["labels"], dg["act_lens"], dg["label_lens"], ids)
        d = {"acts": a_, "labels": l_, "act_lens": al_, "label_lens": ll_, "T": dg["T"], "B": len(ids), "C": dg["C"]}
    else:
        # every rank: its own batch of the workload shape (weak scaling); rank r uses seed 1234 + r
        d = b200ctc.workloads.make(w, seed=1234 + rank)
    T, B, C = d["T"], d["B"], d["C"]
    frames = T * B
    acts = d["acts"].to(dev)
    lab, al, ll = d["labels"].to(dev), d["act_lens"].to(dev), d["label_lens"].to(dev)
    lmax = int(d["label_lens"].max())
    grads = torch.empty_like(acts)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
    loss_acc = torch.zeros(1, dtype=torch.float32, device=dev)
    loss_local = torch.zeros(1, dtype=torch.float32, device=dev)

    comm = torch.cuda.Stream(device=dev) if world > 1 else None
    ready = torch.cuda.Event() if world > 1 else None
    # N > 1: the path's one exchange, the all-reduce of the summed loss.  Default: fused into the library's
    # cost-summing kernel over NVLink peer memory (sharding.PeerLossAllReduce, no collective launch);
    # CTC_B200_ALLREDUCE=nccl selects one NCCL all-reduce on a communication stream, overlapped with the gradient pass.
    peer = None
    exchange = "none (1 GPU)"
    if world > 1:
        exchange = "nccl all-reduce on a side stream, overlapped with the gradient kernels"
        if os.environ.get("CTC_B200_ALLREDUCE", "peer") != "nccl":
            try:
                peer = b200ctc.sharding.PeerLossAllReduce()
                exchange = "fused into the cost-summing kernel over NVLink peer memory (no collective launch)"
            except Exception as exc:  # no peer access on this box: say so and use NCCL
                print(f"[bench] peer all-reduce unavailable ({exc}); using NCCL", file=sys.stderr)
                peer = None

    def step(phase_mask=0):
        nonlocal_peer = peer
        if phase_mask or world == 1:
            # the summed loss (the criterion's return value) is written to loss_acc by the library itself, by a
            # one-warp kernel that runs beside the gradient kernels
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads,
                                         phase_mask=phase_mask, loss_sum=loss_acc if phase_mask == 0 else None)
            return
        # N > 1: the path's one exchange -- the summed loss over NVLink (NCCL) -- starts on a communication stream
        # as soon as the costs are final (`costs_ready`: recursions done) and overlaps the gradient kernels; the
        # compute stream joins it at the end of the step, so the step still contains the whole exchange.
        if nonlocal_peer is not None:
            b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads, loss_sum=loss_local,
                                         loss_sum_all=loss_acc, comm=nonlocal_peer)
            return
        costs, _ = b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads,
                                                costs_ready=ready, loss_sum=loss_local)
        b200ctc.sharding.overlapped_loss_all_reduce(costs, ready, comm, loss_acc, local_sum=loss_local)

    def timed(nsteps, phase_mask=0):
        evs = []
        for _ in range(nsteps):
            flush.zero_()  # L2 flush between timed iterations (not timed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step(phase_mask)
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    if world > 1:
        dist.barrier()  # ranks start their first exchange together
    for _ in range(max(3, args.warmup)):
        flush.zero_()
        step()
    torch.cuda.synchronize()
    launches_per_step = b200ctc.last_launch_count()
    if peer is not None:
        # the fused exchange must give what NCCL gives on the same per-rank sums (fp32 rounding of the total apart)
        costs, _ = b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads,
                                                loss_sum=loss_local)
        ref = loss_local.clone()
        dist.all_reduce(ref)
        step()
        torch.cuda.synchronize()
        good = abs(float(loss_acc.item()) - float(ref.item())) <= 1e-6 * abs(float(ref.item()))
        ok = torch.tensor([1 if good else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:  # never time a wrong exchange: fall back to NCCL on every rank and say so
            print(f"[bench] rank {rank}: peer all-reduce {loss_acc.item()} != NCCL {ref.item()}; using NCCL",
                  file=sys.stderr)
            peer = None
            exchange = "nccl all-reduce on a side stream (the fused peer exchange failed its check on this box)"
            for _ in range(3):
                flush.zero_()
                step()
            torch.cuda.synchronize()
            launches_per_step = b200ctc.last_launch_count()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_start = time.time()
    ms = timed(args.steps)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_end = time.time()
    total_ms = torch.tensor([sum(ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(total_ms.item()) / args.steps
    clocks = sampler.stop(t_start, t_end) if rank == 0 else None

    # per-kernel durations (same stream, CUDA events, L2 flushed): engine alone, grad alone
    step()
    nk = max(5, min(args.steps, 20))
    k_prep = timed(nk, phase_mask=1 | 2)
    k_eng = timed(nk, phase_mask=2 | 4)
    k_grad = timed(nk, phase_mask=1 | 4)
    prep_ms, eng_ms, grad_ms = statistics.mean(k_prep), statistics.mean(k_eng), statistics.mean(k_grad)

    # e2e through the C ABI host-buffer entry: pinned host acts in, host grads + costs out
    lib = b200ctc.pkg._lib.load()
    h_acts = d["acts"].pin_memory()
    h_grads = torch.empty_like(d["acts"]).pin_memory()
    h_costs = torch.empty(B, dtype=torch.float32).pin_memory()
    ctx = ctypes.c_void_p()
    b200ctc.pkg._lib.check(lib.ctc_b200_create(ctypes.byref(ctx)))
    opts = b200ctc.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(opts))
    opts.stream = torch.cuda.current_stream(dev).cuda_stream
    loss_sum = ctypes.c_double(0.0)
    hl, hal, hll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()

    def e2e_step():
        st = lib.ctc_b200_compute_host(ctx, h_acts.data_ptr(), h_grads.data_ptr(), hl.data_ptr(), hll.data_ptr(),
                                       hal.data_ptr(), T, B, C, h_costs.data_ptr(), ctypes.byref(loss_sum),
                                       ctypes.byref(opts))
        b200ctc.pkg._lib.check(st, "ctc_b200_compute_host")

    for _ in range(3):
        e2e_step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_fps = frames * world * args.steps / float(e2e_s.item())
    lib.ctc_b200_destroy(ctx)
    h2d = acts.numel() * 4 + (hl.numel() + 2 * B) * 4
    d2h = acts.numel() * 4 + B * 4

    if rank == 0:
        peak, peak_kind = load_peaks()
        alg_grad = 8 * C * frames   # grad kernel: read fp32 acts once + write fp32 grads once
        alg_eng = 4 * C * frames    # engine kernel: read fp32 acts once
        if C <= 128:
            dom = ("ctc_grad_kernel", alg_grad, grad_ms) if grad_ms >= eng_ms else ("ctc_engine_kernel", alg_eng, eng_ms)
            bound_note = "recursion latency (T dependent steps), not HBM -- see DESIGN.md"
        else:
            # large alphabets: the row kernel reads every logit once and writes every gradient element once -- the
            # whole algorithmic traffic of the step; the lattice kernels only touch 1 + L elements per frame
            dom = ("ctc_prep_big_kernel", alg_grad, prep_ms)
            bound_note = "HBM (softmax rows); the recursions overlap the row kernels of later slices"
        achieved = dom[1] / (dom[2] * 1e-3) / 1e9
        step_gbs = 8 * C * frames / (ms_per_step * 1e-3) / 1e9
        cpu = None
        if world == 1:  # reported at N=1 only
            d0 = d
            fps_cpu, cores, mean_s = cpu_port(d0, steps=3, warmup=1)
            cpu = {"value": fps_cpu, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"oracle fp32 OpenMP port, full batch B={B} T={T} C={C}, best of 3 ({mean_s:.2f} s each)"}
        line = {"metric": METRIC, "value": frames * world / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "strong" if (args.global_batch and world > 1) else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": dict(workload_config(w, d), l2="flushed (256 MB memset) between timed steps",
                               lengths="device-resident (no host sync in the step)",
                               launch="8 length-bucket slices on side streams (CTC_B200_SLICES)",
                               loss_exchange=exchange),
                "roofline": {"bound": "hbm", "kernel": dom[0], "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "peak_kind": peak_kind, "traffic": load_traffic(dom[0], w.name),
                             "alg_bytes_per_launch": dom[1], "kernel_ms": dom[2],
                             "prep_kernel_ms": prep_ms, "engine_kernel_ms": eng_ms, "grad_kernel_ms": grad_ms,
                             "step_alg_gbs": step_gbs, "step_frac": step_gbs / peak,
                             "active_bound": bound_note},
                "cpu_baseline": cpu,
                "e2e": {"value": e2e_fps, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches_per_step * args.steps,
                "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None)
    ap.add_argument("--global-batch", action="store_true",
                    help="N>1: shard ONE global batch of the workload over the ranks (strong scaling, BASELINE config 5)")
    args = ap.parse_args()
    import b200ctc
    w = b200ctc.workloads.CONFIGS[args.workload or b200ctc.workloads.HEADLINE]
    if args.impl == "reference":
        if args.steps > 20:
            args.steps = 20  # each step is ~0.5 s of host work; keep the run to a few minutes
        run_reference(args, w)
    else:
        run_b200(args, w)


if __name__ == "__main__":
    main()
