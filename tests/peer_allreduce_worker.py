"""Worker of tests/test_gpu_parity.py::test_peer_loss_all_reduce_across_processes (launched by torchrun, one process
per GPU): the loss all-reduce fused into the cost-summing kernel over CUDA-IPC peer memory must give, call after
call, the bits NCCL gives on the same per-rank sums -- replaces reduce_tensor (/root/reference/data/utils.py:40-44).
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import b200ctc

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    w = b200ctc.workloads.Workload("peer", 24, 300, 29, 60, 0.05, 0.25)
    dg = b200ctc.workloads.make(w, seed=77)
    ids = b200ctc.sharding.length_bucketed_shards(dg["act_lens"].tolist(), dg["label_lens"].tolist(), world)[rank]
    a, lab, al, ll = b200ctc.sharding.take_shard(dg["acts"], dg["labels"], dg["act_lens"], dg["label_lens"], ids)
    comm = b200ctc.sharding.PeerLossAllReduce()
    mine = torch.zeros(1, device=dev)
    tot = torch.zeros(1, device=dev)
    for it in range(6):
        x = (a * (1.0 + 0.05 * it)).to(dev)
        if rank == 1 and it == 3:
            torch.cuda._sleep(200_000_000)  # a late rank (~0.1 s) is waited for, not timed out
        b200ctc.ctc_forward_backward(x, lab, al, ll, loss_sum=mine, loss_sum_all=tot, comm=comm)
        ref = mine.clone()
        dist.all_reduce(ref)
        torch.cuda.synchronize()
        parts = [torch.zeros(1, device=dev) for _ in range(world)]
        dist.all_gather(parts, mine)
        want = np.float32(sum(np.float64(p.item()) for p in parts))  # rank-order fp64 sum, rounded once
        assert tot.item() == want, (rank, it, tot.item(), want)
        assert abs(tot.item() - ref.item()) <= 1e-6 * abs(ref.item()), (rank, it, tot.item(), ref.item())
    # the drop-in module: returned loss rank-local, criterion.loss_all_ranks the reduce_tensor value
    crit = b200ctc.CTCLoss(comm=comm)
    x = a.to(dev).requires_grad_(True)
    loss = crit(x, lab, al, ll)
    (loss / len(ids)).backward()
    ref = loss.detach().clone()
    dist.all_reduce(ref)
    torch.cuda.synchronize()
    assert abs(crit.loss_all_ranks.item() - ref.item()) <= 1e-6 * abs(ref.item())
    assert not comm.timed_out
    comm.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("peer all-reduce worker ok")


if __name__ == "__main__":
    main()
