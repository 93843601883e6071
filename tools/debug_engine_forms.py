"""Save costs / gradients of shard SHARD (of 8) of the B = 1024 batch under the current CTC_B200_ENGINE setting, or
compare two saved files.  usage: python tools/debug_engine_forms.py save OUT.pt SHARD | cmp A.pt B.pt"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if sys.argv[1] == "save":
    import b200ctc
    d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS["sweep_b1024_t1500"])
    shards = b200ctc.sharding.length_bucketed_shards(d["act_lens"].tolist(), d["label_lens"].tolist(), 8)
    ids = shards[int(sys.argv[3])]
    a, lab, al, ll = b200ctc.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], ids)
    c, g = b200ctc.ctc_forward_backward(a.cuda(), lab, al, ll)
    torch.cuda.synchronize()
    torch.save({"c": c.cpu(), "g": g.cpu()}, sys.argv[2])
else:
    a, b = torch.load(sys.argv[2]), torch.load(sys.argv[3])
    dg = (a["g"] - b["g"]).abs()
    print("costs equal", torch.equal(a["c"], b["c"]), "grads equal", torch.equal(a["g"], b["g"]), "n entries differing",
          int((dg > 0).sum()), "max diff", float(dg.max()))
