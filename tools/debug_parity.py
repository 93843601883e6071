"""GPU debugging aid: per-utterance parity report for one synthetic workload."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200ctc
from oracle import ctc_oracle

def report(name, dev_lens, **kw):
    d = b200ctc.workloads.make(b200ctc.workloads.CONFIGS[name], **kw)
    acts = d["acts"].cuda()
    lab, al, ll = d["labels"], d["act_lens"], d["label_lens"]
    extra = {}
    if dev_lens:
        extra["max_label_len"] = int(ll.max()); lab, al, ll = lab.cuda(), al.cuda(), ll.cuda()
    costs, grads = b200ctc.ctc_forward_backward(acts, lab, al, ll, **extra)
    torch.cuda.synchronize()
    oc, og = ctc_oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    g = grads.cpu().numpy().astype(np.float64)
    err = np.abs(g - og)
    print(f"== {name} dev_lens={dev_lens} {kw}: Lmax={int(d['label_lens'].max())} max err {err.max():.3e}")
    per_b = err.max(axis=(0, 2))
    bad = np.where(per_b > 1e-5)[0]
    for b in bad[:12]:
        t, c = np.unravel_index(err[:, b].argmax(), err[:, b].shape)
        tb = np.where(err[:, b].max(axis=1) > 1e-5)[0]
        print(f"  b={b} T_b={int(d['act_lens'][b])} L={int(d['label_lens'][b])} err={per_b[b]:.3e} at t={t} c={c} "
              f"bad rows: {len(tb)} first {tb[:6]} last {tb[-3:]} cost rel {abs(costs[b].item()-oc[b])/oc[b]:.1e}")
    print(f"  bad utterances: {len(bad)} / {len(per_b)}")

if __name__ == "__main__":
    report("headline_b64_t1500", False, no_repeats=True, seed=99)
    report("headline_b64_t1500", True, no_repeats=True, seed=99)
    report("headline_b64_t1500", False, no_repeats=False, seed=99)
    report("headline_b64_t1500", False, no_repeats=True, seed=1234)
