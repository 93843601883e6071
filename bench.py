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
           HOST grads + costs out, copies inside the timed region; the byte counts are what the entry really moved
roofline   dominant kernel, algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json
cpu_baseline  the oracle's fp32 port on the box's host cores (bounded sample), a reported baseline
workloads  the other BASELINE.json configs timed in the same process with fewer steps (long, bigc, sweep at N=1;
           at N>1 config 5's ONE global batch of 1024 sharded over the ranks = strong scaling), same timing rules
comparators  torch's own CUDA CTC (log_softmax + F.ctc_loss(sum) + backward, fp32) on the same device tensors --
           the only other GPU CTC on the box; outside the product path, never a fallback
module     the drop-in nn.Module as train.py uses it: criterion(...) ; loss / B ; loss.backward()

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
SECONDARY = ("long_b32_t4000", "bigc_b64_t1000_c5000", "sweep_b1024_t1500")


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
    """DRAM bytes per launch of `kernel` from the committed `ncu` capture of this workload (or None)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        if workload in t.get("workloads", {}):
            return t["workloads"][workload]["bytes"][kernel]
        return t["bytes"][kernel] if t.get("workload") == workload else None
    except (OSError, KeyError, ValueError):
        return None


def workload_config(w, d):
    """The `config` object BOTH arms print: exactly these keys, so the driver can match the two lines.  Everything
    else an arm wants to say goes into its sibling `notes` object."""
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
            "config": workload_config(w, d),
            "notes": {"what": "reference CTC (warpctc_pytorch) absent from the image: oracle fp32 OpenMP port, "
                              "host buffers only"},
            "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def bind_to_gpu_numa_node(torch, local_rank):
    """Run this process (and first-touch its pinned buffers) on the CPU cores of the NUMA node the GPU hangs off,
    as a production launcher would: with every rank on node 0 the host copies of the far-socket GPUs cross the
    inter-socket link.  Best effort -- returns a note for the bench line."""
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (int(pr.pci_domain_id), int(pr.pci_bus_id), int(pr.pci_device_id))
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "gpu numa node unknown (single node?)"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            return f"gpu on numa node {node}, none of its cores in this process's cpuset"
        os.sched_setaffinity(0, allowed)
        return f"bound to numa node {node} ({len(allowed)} cores)"
    except Exception as exc:
        return f"not bound ({type(exc).__name__})"


class Bench:
    """One workload on this rank's GPU: the device-resident step, the per-kernel legs, the host-buffer (e2e) leg,
    the module leg and the torch comparator.  All timing: CUDA events on the launching stream, L2 flushed (256 MB
    memset) between timed steps, >= 3 warm-up steps."""

    def __init__(self, b200ctc, torch, dist, dev, rank, world, peer, flush):
        self.b, self.torch, self.dist, self.dev = b200ctc, torch, dist, dev
        self.rank, self.world, self.peer, self.flush = rank, world, peer, flush
        self.comm_stream = torch.cuda.Stream(device=dev) if world > 1 else None
        self.ready = torch.cuda.Event() if world > 1 else None

    def load(self, d):
        torch, dev = self.torch, self.dev
        self.d = d
        self.T, self.B, self.C = d["T"], d["B"], d["C"]
        self.frames = self.T * self.B
        self.acts = d["acts"].to(dev)
        self.lab, self.al, self.ll = d["labels"].to(dev), d["act_lens"].to(dev), d["label_lens"].to(dev)
        self.lmax = int(d["label_lens"].max())
        self.grads = torch.empty_like(self.acts)
        self.loss_acc = torch.zeros(1, dtype=torch.float32, device=dev)
        self.loss_local = torch.zeros(1, dtype=torch.float32, device=dev)

    def unload(self):
        for k in ("acts", "grads", "lab", "al", "ll", "d"):
            setattr(self, k, None)
        self.b.pkg.ctc_loss.clear_workspace_cache()
        self.torch.cuda.empty_cache()

    def step(self, phase_mask=0):
        b = self.b
        if phase_mask or self.world == 1:
            # the summed loss (the criterion's return value) is written to loss_acc by the library itself, by a
            # one-warp kernel that runs beside the gradient kernels
            b.ctc_forward_backward(self.acts, self.lab, self.al, self.ll, max_label_len=self.lmax,
                                   out_grads=self.grads, phase_mask=phase_mask,
                                   loss_sum=self.loss_acc if phase_mask == 0 else None)
            return
        # N > 1: the path's one exchange -- the summed loss over NVLink.  Fused into the library's cost-summing kernel
        # (peer memory, no collective launch), or one NCCL all-reduce on a communication stream that starts as soon
        # as the costs are final (`costs_ready`) and overlaps the gradient kernels; either way the step contains it.
        if self.peer is not None:
            b.ctc_forward_backward(self.acts, self.lab, self.al, self.ll, max_label_len=self.lmax,
                                   out_grads=self.grads, loss_sum=self.loss_local, loss_sum_all=self.loss_acc,
                                   comm=self.peer)
            return
        costs, _ = b.ctc_forward_backward(self.acts, self.lab, self.al, self.ll, max_label_len=self.lmax,
                                          out_grads=self.grads, costs_ready=self.ready, loss_sum=self.loss_local)
        b.sharding.overlapped_loss_all_reduce(costs, self.ready, self.comm_stream, self.loss_acc,
                                              local_sum=self.loss_local)

    def timed(self, nsteps, fn=None, run_ahead=1, **kw):
        """CUDA-event time of each of `nsteps` calls of fn.  run_ahead > 1 queues that many L2-flush memsets (~60 us
        of device work each) in front of every timed call, so that the host enqueues the call while the device is
        still busy -- the situation inside a training loop, where the model's kernels are in flight when the loss is
        called -- and the events bracket device time only (used for the Python-heavy module / comparator legs)."""
        torch = self.torch
        fn = fn or self.step
        evs = []
        for _ in range(nsteps):
            for _ in range(run_ahead):
                self.flush.zero_()  # L2 flush between timed iterations (not timed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn(**kw)
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return [a.elapsed_time(b) for a, b in evs]

    def warm(self, n):
        for _ in range(max(3, n)):
            self.flush.zero_()
            self.step()
        self.torch.cuda.synchronize()
        return self.b.last_launch_count()

    def graph_replay_ms(self, nsteps):
        """the same step captured once in a CUDA graph and replayed (one launch per step instead of ~25 API calls)"""
        torch = self.torch
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(3):
                self.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            self.step()
        self.timed(5, fn=graph.replay)
        ms = self.timed(nsteps, fn=graph.replay)
        del graph
        return sum(ms) / len(ms)

    def device_step_ms(self, nsteps):
        """max over ranks of the summed device time of `nsteps` steps, per step (barrier + sync on both sides)"""
        torch, dist = self.torch, self.dist
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = self.timed(nsteps)
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        tot = torch.tensor([sum(ms)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        return float(tot.item()) / nsteps

    def kernel_legs(self, nk):
        """per-kernel durations (same stream, CUDA events, L2 flushed): prep alone, engine alone, grad alone"""
        self.step()
        k_prep = self.timed(nk, phase_mask=1 | 2)
        k_eng = self.timed(nk, phase_mask=2 | 4)
        k_grad = self.timed(nk, phase_mask=1 | 4)
        return statistics.mean(k_prep), statistics.mean(k_eng), statistics.mean(k_grad)

    def e2e(self, nsteps):
        """through the C ABI host-buffer entry: pinned host acts in, host grads + costs out, wall clock around
        synchronous calls; returns (frames/s over all ranks, h2d bytes, d2h bytes per step as really copied)"""
        import ctypes

        torch, dist, b, d = self.torch, self.dist, self.b, self.d
        lib = b.pkg._lib.load()
        h_acts = d["acts"].pin_memory()
        h_grads = torch.empty_like(d["acts"]).pin_memory()
        h_costs = torch.empty(self.B, dtype=torch.float32).pin_memory()
        ctx = ctypes.c_void_p()
        b.pkg._lib.check(lib.ctc_b200_create(ctypes.byref(ctx)))
        opts = b.pkg._lib.CtcOptions()
        lib.ctc_b200_default_options(ctypes.byref(opts))
        opts.stream = torch.cuda.current_stream(self.dev).cuda_stream
        loss_sum = ctypes.c_double(0.0)
        hl, hal, hll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()

        def one():
            st = lib.ctc_b200_compute_host(ctx, h_acts.data_ptr(), h_grads.data_ptr(), hl.data_ptr(), hll.data_ptr(),
                                           hal.data_ptr(), self.T, self.B, self.C, h_costs.data_ptr(),
                                           ctypes.byref(loss_sum), ctypes.byref(opts))
            b.pkg._lib.check(st, "ctc_b200_compute_host")

        for _ in range(3):
            one()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(nsteps):
            one()
        torch.cuda.synchronize()
        sec = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(sec, op=dist.ReduceOp.MAX)
        h2d, d2h = ctypes.c_ulonglong(0), ctypes.c_ulonglong(0)
        lib.ctc_b200_last_copy_bytes(ctypes.byref(h2d), ctypes.byref(d2h))
        lib.ctc_b200_destroy(ctx)
        del h_acts, h_grads
        return self.frames * self.world * nsteps / float(sec.item()), int(h2d.value), int(d2h.value)

    def module_ms(self, nsteps):
        """the drop-in nn.Module as the reference's train step uses it (train.py:600-602,619-620):
        loss = criterion(logits, targets, sizes, target_sizes); loss = loss / B; loss.backward() -- and the same with
        the 1/B folded into the criterion (size_average=True), where backward multiplies by 1."""
        torch, b = self.torch, self.b
        x = self.acts.detach().clone().requires_grad_(True)
        out = {}
        for name, crit, div in (("criterion_div_B_backward", b.CTCLoss(), float(self.B)),
                                ("size_average_backward", b.CTCLoss(size_average=True), 1.0)):
            def one():
                x.grad = None
                loss = crit(x, self.lab, self.al, self.ll, max_label_len=self.lmax)
                if div != 1.0:
                    loss = loss / div
                loss.backward()
            for _ in range(3):
                one()
            out[name] = statistics.mean(self.timed(nsteps, fn=one, run_ahead=12))
        x.grad = None
        return out

    def torch_cuda_ctc_ms(self, nsteps):
        """torch's own CUDA CTC on the same device tensors: fp32 log_softmax + F.ctc_loss(sum) + backward"""
        torch = self.torch
        import torch.nn.functional as F

        x = self.acts.detach().clone().requires_grad_(True)
        lab_l, al_l, ll_l = self.lab.long(), self.al.long(), self.ll.long()

        def one():
            x.grad = None
            lp = F.log_softmax(x, dim=-1)
            loss = F.ctc_loss(lp, lab_l, al_l, ll_l, blank=0, reduction="sum", zero_infinity=True)
            loss.backward()

        # torch may pick cuDNN's CTC for some shapes; its native kernel is the comparator (deterministic off is fine)
        with torch.backends.cudnn.flags(enabled=False):
            for _ in range(3):
                one()
            ms = statistics.mean(self.timed(nsteps, fn=one, run_ahead=12))
        x.grad = None
        return ms


def roofline_block(w, d, ms_per_step, prep_ms, eng_ms, grad_ms, peak, peak_kind):
    T, B, C = d["T"], d["B"], d["C"]
    frames = T * B
    valid = int(d["act_lens"].sum())
    alg_step = 8 * C * frames      # read fp32 acts once + write fp32 grads once, per padded frame
    if C <= 128:
        alg_grad = 8 * C * frames  # grad kernel: fp32 acts read once (as emission rows) + fp32 grads written once
        alg_eng = 4 * C * frames   # engine kernel: the emission rows read once
        dom = ("ctc_grad_kernel", alg_grad, grad_ms) if grad_ms >= eng_ms else ("ctc_engine_kernel", alg_eng, eng_ms)
        bound_note = "recursion latency (T dependent steps), not HBM -- see DESIGN.md"
        alg_note = "8*C*B*T (grad kernel) or 4*C*B*T (engine kernel)"
    else:
        # large alphabets: the row kernel reads every VALID logit row once and writes every gradient row once
        # (padding rows are written as zeros, never read): its real algorithmic bytes, not the step's 8*C*B*T
        nv = 5 if C <= 5120 else (8 if C <= 8192 else 16)
        name = (f"ctc_prep_big_reg_kernel<256,{nv},0>" if (C % 4 == 0 and C <= 16384) else "ctc_prep_big_kernel<256>")
        dom = (name, 4 * C * valid + 4 * C * frames, prep_ms)
        bound_note = "HBM (softmax rows); recursion + scattered occupancy reductions follow with the DRAM mostly idle"
        alg_note = "4*C*valid_frames read + 4*C*B*T written (row kernel)"
    achieved = dom[1] / (dom[2] * 1e-3) / 1e9
    step_gbs = alg_step / (ms_per_step * 1e-3) / 1e9
    tkey = "ctc_prep_big_kernel" if dom[0].startswith("ctc_prep_big") else dom[0]
    return {"bound": "hbm", "kernel": dom[0], "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "peak_kind": peak_kind, "traffic": load_traffic(tkey, w.name),
            "alg_bytes_per_launch": dom[1], "alg_bytes_formula": alg_note, "kernel_ms": dom[2],
            "prep_kernel_ms": prep_ms, "engine_kernel_ms": eng_ms, "grad_kernel_ms": grad_ms,
            "step_alg_gbs": step_gbs, "step_frac": step_gbs / peak, "active_bound": bound_note}


def run_b200(args, w):
    import torch
    import torch.distributed as dist

    import b200ctc

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    numa_note = bind_to_gpu_numa_node(torch, local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def make_batch(wl, global_batch, same_seed=False):
        if global_batch and world > 1:
            # BASELINE config 5: ONE global batch of the workload shape, dealt to the ranks by lattice work
            # (sharding.length_bucketed_shards: LPT, equal shard sizes) -- strong scaling, no data-path collective
            dg = b200ctc.workloads.make(wl, seed=1234)
            ids = b200ctc.sharding.length_bucketed_shards(dg["act_lens"].tolist(), dg["label_lens"].tolist(),
                                                          world)[rank]
            a_, l_, al_, ll_ = b200ctc.sharding.take_shard(dg["acts"], dg["labels"], dg["act_lens"], dg["label_lens"],
                                                           ids)
            return {"acts": a_, "labels": l_, "act_lens": al_, "label_lens": ll_, "T": dg["T"], "B": len(ids),
                    "C": dg["C"]}
        # every rank: its own batch of the workload shape (weak scaling); rank r uses seed 1234 + r
        return b200ctc.workloads.make(wl, seed=1234 + (0 if same_seed else rank))

    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2
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

    bn = Bench(b200ctc, torch, dist, dev, rank, world, peer, flush)
    d = make_batch(w, args.global_batch)
    bn.load(d)
    if world > 1:
        dist.barrier()  # ranks start their first exchange together
    warm = max(3, args.warmup)
    launches_per_step = bn.warm(warm)
    if peer is not None:
        # the fused exchange must give what NCCL gives on the same per-rank sums (fp32 rounding of the total apart)
        b200ctc.ctc_forward_backward(bn.acts, bn.lab, bn.al, bn.ll, max_label_len=bn.lmax, out_grads=bn.grads,
                                     loss_sum=bn.loss_local)
        ref = bn.loss_local.clone()
        dist.all_reduce(ref)
        bn.step()
        torch.cuda.synchronize()
        good = abs(float(bn.loss_acc.item()) - float(ref.item())) <= 1e-6 * abs(float(ref.item()))
        ok = torch.tensor([1 if good else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:  # never time a wrong exchange: fall back to NCCL on every rank and say so
            print(f"[bench] rank {rank}: peer all-reduce {bn.loss_acc.item()} != NCCL {ref.item()}; using NCCL",
                  file=sys.stderr)
            bn.peer = peer = None
            exchange = "nccl all-reduce on a side stream (the fused peer exchange failed its check on this box)"
            launches_per_step = bn.warm(3)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    t_start = time.time()
    ms_per_step = bn.device_step_ms(args.steps)
    t_end = time.time()
    clocks = sampler.stop(t_start, t_end) if rank == 0 else None

    nk = max(5, min(args.steps, 20))
    prep_ms, eng_ms, grad_ms = bn.kernel_legs(nk)
    e2e_fps, h2d, d2h = bn.e2e(args.steps)
    T, B, C = bn.T, bn.B, bn.C
    frames = bn.frames

    comparators, module = None, None
    control, replay = None, None
    if not args.quick:
        if world == 1:
            rms = bn.graph_replay_ms(nk)
            replay = {"ms_per_step": rms, "value": frames / (rms * 1e-3), "unit": UNIT,
                      "what": "the same step captured in a CUDA graph and replayed; `value` above is the eager launch"}
        module = bn.module_ms(nk)
        tc_ms = bn.torch_cuda_ctc_ms(min(nk, 10))
        comparators = {"torch_cuda_ctc": {"ms_per_step": tc_ms, "value": frames * world / (tc_ms * 1e-3), "unit": UNIT,
                                          "what": "fp32 log_softmax + F.ctc_loss(sum, native CUDA kernel) + backward "
                                                  "on the same device tensors, this rank; not an oracle"}}
        if world > 1 and not args.global_batch:
            # weak-scaling control: every rank times the SAME batch (seed 1234), so the max over ranks carries no
            # batch skew; what is left against N=1 is the exchange and launch jitter
            bn.unload()
            bn.load(make_batch(w, False, same_seed=True))
            bn.warm(3)
            cms = bn.device_step_ms(max(10, args.steps // 2))
            control = {"same_batch_on_every_rank_ms_per_step": cms, "value": frames * world / (cms * 1e-3)}
            bn.unload()
            bn.load(d)

    # ---- the other BASELINE.json configurations, same process, fewer steps
    others = {}
    if not args.quick and w.name == b200ctc.workloads.HEADLINE:
        peak, peak_kind = load_peaks()
        bn.unload()
        names = SECONDARY if world == 1 else ("sweep_b1024_t1500",)
        for name in names:
            wl = b200ctc.workloads.CONFIGS[name]
            dd = make_batch(wl, global_batch=(world > 1))
            bn.load(dd)
            ns = max(5, min(args.steps, 10))
            lps = bn.warm(3)
            ms = bn.device_step_ms(ns)
            p_ms, e_ms, g_ms = bn.kernel_legs(5)
            ef, eh, ed = bn.e2e(3)
            entry = {"config": workload_config(wl, dd), "steps": ns, "ms_per_step": ms,
                     "value": bn.frames * world / (ms * 1e-3), "unit": UNIT,
                     "scaling": "strong (one global batch of 1024 sharded by lattice work)" if world > 1 else "n/a",
                     "gpu_launches_per_step": lps,
                     "e2e": {"value": ef, "unit": UNIT, "h2d_bytes_per_step": eh, "d2h_bytes_per_step": ed},
                     "roofline": roofline_block(wl, dd, ms, p_ms, e_ms, g_ms, peak, peak_kind)}
            entry["step_frac"] = entry["roofline"]["step_frac"]
            if world == 1:
                entry["module_ms"] = bn.module_ms(5)
                tc = bn.torch_cuda_ctc_ms(3)
                entry["torch_cuda_ctc_ms"] = tc
            others[name] = entry
            bn.unload()

    fc_layer = None
    if not args.quick and world == 1 and w.name == b200ctc.workloads.HEADLINE:
        # the next row of the scope table: the logits producer feeding this loss, same frame count
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_fc
            fc_layer = bench_fc.run()
        except Exception as exc:  # never lose the headline line to the extra block
            fc_layer = {"error": repr(exc)}
    if rank == 0:
        peak, peak_kind = load_peaks()
        cpu = None
        if world == 1:  # reported at N=1 only
            fps_cpu, cores, mean_s = cpu_port(d, steps=3, warmup=1)
            cpu = {"value": fps_cpu, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"oracle fp32 OpenMP port, full batch B={B} T={T} C={C}, best of 3 ({mean_s:.2f} s each)"}
        line = {"metric": METRIC, "value": frames * world / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": warm, "warmup_requested": args.warmup, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "strong" if (args.global_batch and world > 1) else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(w, d),
                "notes": {"l2": "flushed (256 MB memset) between timed steps",
                          "lengths": "device-resident (no host sync in the step)",
                          "launch": "length-bucket slices on side streams (CTC_B200_SLICES)",
                          "loss_exchange": exchange, "host_numa": numa_note,
                          "batch_order": "longest first, as the reference's collate function emits batches "
                                         "(data_loader_aug.py:527)"},
                "roofline": roofline_block(w, d, ms_per_step, prep_ms, eng_ms, grad_ms, peak, peak_kind),
                "cpu_baseline": cpu,
                "e2e": {"value": e2e_fps, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches_per_step * args.steps,
                "clocks": clocks}
        if comparators:
            line["comparators"] = comparators
        if module:
            line["module_step_ms"] = module
        if replay:
            line["cuda_graph_replay"] = replay
        if control:
            line["weak_scaling_control"] = control
        if others:
            line["workloads"] = others
        if fc_layer:
            line["fc_layer"] = fc_layer
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
    ap.add_argument("--quick", action="store_true",
                    help="headline line only: skip the module / comparator legs and the other configurations")
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
