"""Batch sharding across the GPUs of one box, and the single loss all-reduce.

Utterances are independent, so the CTC path shards on the batch axis with no data-path collective: each rank keeps
acts[:, R_r, :], its label slice and lengths, and produces grads[:, R_r, :] locally.  The only exchange step is a
scalar: the summed loss.  This generalises the reference's DistributedBucketingSampler
(/root/reference/data/data_loader_aug.py:582-617: rank-strided bins of duration-sorted ids) and replaces
`reduce_tensor` (/root/reference/data/utils.py:40-44: clone -> all_reduce(SUM) -> /world_size, gloo by default,
train.py:99) with one all-reduce on the compute stream (NCCL over NVLink/NVSwitch on the GPU box).
"""
from __future__ import annotations

from typing import List, Sequence

import torch


def lattice_work(act_lens: Sequence[int], label_lens: Sequence[int]) -> List[int]:
    """Per-utterance lattice work T_b * S_b (S_b = 2 L_b + 1): what the recursion kernels' time follows."""
    return [int(t) * (2 * int(l) + 1) for t, l in zip(act_lens, label_lens)]


def length_bucketed_shards(act_lens: Sequence[int], label_lens: Sequence[int], world_size: int) -> List[List[int]]:
    """Longest-processing-time-first deal of utterance ids to ranks, balanced on lattice work, with equal shard
    sizes when B divides evenly (so every rank launches the same grid).  Within a shard ids are ordered
    longest-first so the longest recursion starts first.  Deterministic."""
    B = len(act_lens)
    w = lattice_work(act_lens, label_lens)
    order = sorted(range(B), key=lambda i: (-w[i], i))
    cap = (B + world_size - 1) // world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    load = [0] * world_size
    for i in order:
        r = min((r for r in range(world_size) if len(shards[r]) < cap), key=lambda r: (load[r], r))
        shards[r].append(i)
        load[r] += w[i]
    return shards


def take_shard(acts: torch.Tensor, labels: torch.Tensor, act_lens: torch.Tensor, label_lens: torch.Tensor,
               ids: Sequence[int]):
    """Slice one rank's utterances out of a packed batch (acts [T,B,C], flat labels)."""
    offs = torch.zeros(label_lens.numel() + 1, dtype=torch.int64)
    offs[1:] = torch.cumsum(label_lens.to(torch.int64), 0)
    idx = torch.as_tensor(list(ids), dtype=torch.int64)
    lab = [labels[int(offs[i]):int(offs[i + 1])] for i in ids]
    flat = torch.cat(lab) if lab else labels[:0]
    return acts[:, idx.to(acts.device), :].contiguous(), flat.contiguous(), act_lens[idx].contiguous(), \
        label_lens[idx].contiguous()


def all_reduce_loss(loss_sum: torch.Tensor, mean_of_means: bool = False, local_batch: int = 1) -> torch.Tensor:
    """The path's one collective.  Default: global SUM of the per-rank summed losses (north-star).
    mean_of_means=True reproduces the reference's logging value: each rank's batch-mean, summed, / world_size
    (data/utils.py:40-44 applied to train.py:601's `loss / B`)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return loss_sum / local_batch if mean_of_means else loss_sum
    rt = (loss_sum / local_batch).clone() if mean_of_means else loss_sum.clone()
    dist.all_reduce(rt, op=dist.ReduceOp.SUM)
    if mean_of_means:
        rt /= dist.get_world_size()
    return rt


def overlapped_loss_all_reduce(costs: torch.Tensor, ready: "torch.cuda.Event", comm: "torch.cuda.Stream",
                               out: torch.Tensor, local_sum: "torch.Tensor | None" = None) -> None:
    """The same collective, off the critical path: `ready` is the event handed to the CTC call as
    `costs_ready` (recorded when the per-utterance costs are final, while the gradient kernels still run).  The sum and
    the NCCL all-reduce run on `comm`; the current stream joins `comm` before returning, so work enqueued after this
    call sees `out` (1-element fp32, global SUM of the costs).  With `local_sum` (the `loss_sum` tensor handed to the
    same CTC call) no local reduction is launched at all."""
    import torch.distributed as dist

    cur = torch.cuda.current_stream(costs.device)
    with torch.cuda.stream(comm):
        comm.wait_event(ready)
        # local_sum: the library's own loss_sum output of the same call (already final when `ready` fires)
        s = local_sum.reshape(1).clone() if local_sum is not None else costs.sum().reshape(1)
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(s)
        out.copy_(s)
    costs.record_stream(comm)
    cur.wait_stream(comm)


class PeerLossAllReduce:
    """Communicator of the fused loss all-reduce (include/ctc_b200.h: ctc_b200_comm_*): the kernel that sums the
    per-utterance costs also exchanges the per-rank sums over NVLink peer memory, so the path's one collective
    (reduce_tensor, data/utils.py:40-44) costs no launch of its own.  One process per GPU; the CUDA IPC handles of the
    ranks' slot buffers are exchanged once through torch.distributed (any backend).

        comm = PeerLossAllReduce()                                  # after init_process_group, device set
        ctc_forward_backward(..., loss_sum=loss, comm=comm)         # loss: sum over all ranks, same bits everywhere

    `PeerLossAllReduce.local_group(n)` builds n communicators inside ONE process (one device) for tests."""

    def __init__(self, rank: int | None = None, world: int | None = None, _exchange: bool = True):
        import ctypes

        import torch.distributed as dist

        from . import _lib

        self._lib = _lib.load()
        if rank is None or world is None:
            if not (dist.is_available() and dist.is_initialized()):
                raise RuntimeError("PeerLossAllReduce needs an initialised torch.distributed process group")
            rank, world = dist.get_rank(), dist.get_world_size()
        self.rank, self.world = int(rank), int(world)
        h = ctypes.c_void_p()
        _lib.check(self._lib.ctc_b200_comm_create(ctypes.byref(h), self.rank, self.world), "ctc_b200_comm_create")
        self.handle = h.value
        if not _exchange:
            return
        buf = ctypes.create_string_buffer(64)
        _lib.check(self._lib.ctc_b200_comm_handle(self.handle, buf), "ctc_b200_comm_handle")
        gathered: list = [None] * self.world
        dist.all_gather_object(gathered, bytes(buf.raw))
        blob = ctypes.create_string_buffer(b"".join(gathered), 64 * self.world)
        st = self._lib.ctc_b200_comm_open(self.handle, blob)
        # one collective doubles as the barrier (every rank has mapped every buffer before anyone's first call stores
        # into them) and as the agreement that the mapping worked everywhere
        ok = torch.tensor([1 if st == 0 else 0], dtype=torch.int32,
                          device="cuda" if dist.get_backend() == "nccl" else "cpu")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            self.close()
            raise RuntimeError("CUDA IPC peer mapping of the loss all-reduce buffers failed on at least one rank")

    @classmethod
    def local_group(cls, n: int):
        import ctypes

        from . import _lib

        comms = [cls(r, n, _exchange=False) for r in range(n)]
        lib = _lib.load()
        locals_ = (ctypes.c_void_p * n)(*[lib.ctc_b200_comm_local(c.handle) for c in comms])
        for c in comms:
            _lib.check(lib.ctc_b200_comm_set_peers(c.handle, locals_), "ctc_b200_comm_set_peers")
        return comms

    def set_timeout(self, milliseconds: int) -> None:
        """Wall-clock bound on one exchange (default 120 s): a late peer is waited for that long."""
        from . import _lib

        _lib.check(self._lib.ctc_b200_comm_set_timeout(self.handle, int(milliseconds)), "ctc_b200_comm_set_timeout")

    @property
    def timed_out(self) -> bool:
        """True once an exchange was given up (a peer skipped a call or stalled past the timeout).  Sticky: every
        later CTC call that names this communicator raises; build a new one on every rank or use NCCL."""
        return bool(self.handle) and bool(self._lib.ctc_b200_comm_timed_out(self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self._lib.ctc_b200_comm_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
