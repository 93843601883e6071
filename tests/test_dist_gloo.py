"""world_size-2 gloo test (CPU) of the multi-GPU host logic: length-bucketed sharding of one global batch, per-rank
costs, and the path's single collective (all-reduce of the summed loss).  The per-rank costs come from the CPU
oracle here (there is no GPU in this container); on the GPU box the same code path runs with the CUDA op."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200ctc
    from oracle import ctc_oracle

    w = b200ctc.workloads.Workload("g", 12, 50, 30, 12, 0.05, 0.25)
    d = b200ctc.workloads.make(w, seed=77)  # every rank draws the same global batch
    shards = b200ctc.sharding.length_bucketed_shards(d["act_lens"].tolist(), d["label_lens"].tolist(), world)
    a, lab, al, ll = b200ctc.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], shards[rank])
    costs, _ = ctc_oracle.compute(a.numpy(), lab.numpy(), al.numpy(), ll.numpy(), num_threads=1)
    local = torch.tensor([costs.sum()], dtype=torch.float64)
    total = b200ctc.sharding.all_reduce_loss(local)
    mom = b200ctc.sharding.all_reduce_loss(local.clone(), mean_of_means=True, local_batch=len(shards[rank]))
    full, _ = ctc_oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(),
                                 d["label_lens"].numpy(), num_threads=1)
    ok = abs(float(total) - full.sum()) <= 1e-9 * full.sum()
    ok = ok and abs(float(mom) - full.sum() / w.B) <= 1e-9 * full.sum()  # equal shard sizes: mean of means == mean
    ok = ok and float(local) < float(total)
    out[rank] = 1 if ok else 0
    dist.destroy_process_group()


def test_sharded_loss_allreduce_world2():
    world = 2
    out = mp.get_context("spawn").Array("i", [0] * world)
    port = _free_port()
    procs = [mp.get_context("spawn").Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert list(out) == [1] * world


def test_all_reduce_loss_without_process_group():
    sys.path.insert(0, ROOT)
    import b200ctc

    x = torch.tensor([6.0])
    assert float(b200ctc.sharding.all_reduce_loss(x)) == 6.0
    assert float(b200ctc.sharding.all_reduce_loss(x, mean_of_means=True, local_batch=3)) == 2.0
