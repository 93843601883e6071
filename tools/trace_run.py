import os, sys, torch
sys.path.insert(0, "/root/repo")
import b200ctc
w = b200ctc.workloads.CONFIGS[sys.argv[1] if len(sys.argv) > 1 else "headline_b64_t1500"]
d = b200ctc.workloads.make(w)
acts = d["acts"].cuda(); lab = d["labels"].cuda(); al = d["act_lens"].cuda(); ll = d["label_lens"].cuda()
lmax = int(d["label_lens"].max())
print("Lmax", lmax, "mean L", float(d["label_lens"].float().mean()), "mean T", float(d["act_lens"].float().mean()))
for i in range(6):
    c, g = b200ctc.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
    torch.cuda.synchronize()
