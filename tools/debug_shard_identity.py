"""Which utterances of the B = 1024 batch differ between the whole-batch run and a shard run (diagnostic)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ctc  # noqa: E402

d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS["sweep_b1024_t1500"])
c0, g0 = b200ctc.ctc_forward_backward(d["acts"].cuda(), d["labels"], d["act_lens"], d["label_lens"])
torch.cuda.synchronize()
c0, g0 = c0.cpu(), g0.cpu()
shards = b200ctc.sharding.length_bucketed_shards(d["act_lens"].tolist(), d["label_lens"].tolist(), 8)
for si, ids in enumerate(shards):
    a, lab, al, ll = b200ctc.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], ids)
    for env in (("X", "0"), ("X", "1")):
        os.environ[env[0]] = env[1]
        c, g = b200ctc.ctc_forward_backward(a.cuda(), lab, al, ll)
        torch.cuda.synchronize()
        c, g = c.cpu(), g.cpu()
        bad = [(j, int(ids[j])) for j in range(len(ids)) if not torch.equal(g[:, j], g0[:, ids[j]])]
        cb = [(j, int(ids[j])) for j in range(len(ids)) if c[j] != c0[ids[j]]]
        print("shard", si, env, "grad differs for", len(bad), "utterances; cost differs for", len(cb))
        for j, b in bad[:3]:
            diff = (g[:, j] - g0[:, b]).abs()
            t = int(diff.max(1).values.argmax())
            print("   utt", b, "T", int(d["act_lens"][b]), "L", int(d["label_lens"][b]), "max diff", float(diff.max()), "at t", t,
                  "n frames differing", int((diff.max(1).values > 0).sum()))
