"""Small invocations of every kernel (all row-kernel forms, decode, error counts, summed loss) -- sized for
compute-sanitizer (`compute-sanitizer --tool memcheck python tools/sanitize_small.py`); the tool is closed on this
round's GPU pool, so here it only serves as a quick smoke run of every launch path."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402


def run(T, B, C, lmax, seed, **kw):
    g = torch.Generator().manual_seed(seed)
    acts = torch.randn(T, B, C, generator=g)
    act_lens = torch.randint(max(1, T // 2), T + 1, (B,), generator=g).int()
    act_lens[0] = T
    label_lens = torch.minimum(torch.randint(0, lmax + 1, (B,), generator=g).int(), act_lens // 2)
    labels = torch.randint(1, C, (int(label_lens.sum()),), generator=g).int()
    loss = torch.zeros(1, device="cuda")
    out = b200ctc.ctc_forward_backward(acts.cuda(), labels, act_lens, label_lens, loss_sum=loss, **kw)
    torch.cuda.synchronize()
    return out, labels, label_lens, float(loss)


def main():
    for (T, B, C, lmax) in ((40, 20, 29, 15), (70, 9, 30, 30), (33, 17, 140, 12), (21, 16, 5000, 9), (12, 16, 6000, 5),
                            (12, 3, 16388, 5)):
        (costs, grads), _, _, loss = run(T, B, C, lmax, seed=T)
        assert torch.isfinite(grads).all() and np.isfinite(loss)
        (costs, grads, best), labels, label_lens, _ = run(T, B, C, lmax, seed=T + 1, decode=True)
        counts = b200ctc.error_counts(best["ids"], best["lens"], labels.cuda(), label_lens.cuda(), C - 1,
                                      int(label_lens.max()))
        torch.cuda.synchronize()
        assert counts.shape == (B, 4)
        print("ok", T, B, C, lmax, float(costs.sum()), counts.sum(0).tolist())


if __name__ == "__main__":
    main()
