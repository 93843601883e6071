"""tests/golden/make_golden.py -- generator of the committed golden fixtures (run once, in the build container).

The reference's own CTC arithmetic (`warpctc_pytorch`, /root/reference/train.py:12) is absent from the image and
un-pinned, and the reference holds no fixtures, so the golden vectors come from an independent
implementation that IS importable here: torch 2.11 CPU, fp64,
    F.ctc_loss(F.log_softmax(acts.double(), -1), labels, act_lens, label_lens, blank=0, reduction='none',
               zero_infinity=True)  + autograd through log_softmax
i.e. exactly the quantities of the reference call `criterion(acts, labels, act_lens, label_lens)` (train.py:600).
Each fixture stores the fp32 inputs and the fp64 costs/grads; oracle/ and the CUDA path are both tested against
them.  Usage:  python tests/golden/make_golden.py   (writes tests/golden/*.npz; a few hundred KB in total)
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))


def torch_ctc_f64(acts, labels, act_lens, label_lens, blank=0):
    a = acts.double().requires_grad_(True)
    lp = F.log_softmax(a, dim=-1)
    costs = F.ctc_loss(lp, labels.long(), act_lens.long(), label_lens.long(), blank=blank, reduction="none",
                       zero_infinity=True)
    costs.sum().backward()
    return costs.detach().numpy(), a.grad.numpy()


def case(name, T, B, C, lens, llens, seed, sigma=1.0, labels=None, blank=0):
    g = torch.Generator().manual_seed(seed)
    acts = (torch.randn(T, B, C, generator=g) * sigma).float()
    act_lens = torch.tensor(lens, dtype=torch.int32)
    label_lens = torch.tensor(llens, dtype=torch.int32)
    if labels is None:
        labels = torch.randint(1, C, (int(label_lens.sum()),), generator=g, dtype=torch.int32)
        if blank != 0:
            labels = torch.where(labels == blank, torch.zeros_like(labels), labels)
    else:
        labels = torch.tensor(labels, dtype=torch.int32)
    costs, grads = torch_ctc_f64(acts, labels, act_lens, label_lens, blank)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), acts=acts.numpy(), labels=labels.numpy(),
                        act_lens=act_lens.numpy(), label_lens=label_lens.numpy(), costs=costs, grads=grads,
                        blank=np.int32(blank))
    print(name, "costs", costs[:4], "grad absmax", np.abs(grads).max())


if __name__ == "__main__":
    torch.set_num_threads(4)
    # the fork's 30-symbol alphabet (labels.json) and the upstream 29; ragged lengths; repeats; infeasible; L=0
    case("g_c30_ragged", 40, 6, 30, [40, 33, 25, 17, 9, 1], [9, 7, 5, 3, 2, 1], seed=1)
    case("g_c29_sigma5", 60, 4, 29, [60, 50, 41, 30], [14, 11, 9, 6], seed=2, sigma=5.0)
    case("g_repeats", 24, 3, 6, [24, 20, 12], [6, 5, 4], seed=3, labels=[1, 1, 2, 2, 3, 3, 4, 4, 4, 5, 5, 2, 2, 2, 2])
    case("g_infeasible_and_empty", 12, 4, 5, [12, 3, 12, 5], [3, 3, 0, 5], seed=4,
         labels=[1, 2, 3, 2, 2, 2, 1, 2, 3, 4, 1])  # utt1: L=3 + 2 repeats > T=3; utt2: L=0; utt3: L=5 == T, no repeats
    case("g_tight", 9, 2, 4, [9, 7], [4, 3], seed=5, labels=[1, 1, 2, 3, 3, 3, 1])  # utt0: 4+1 <= 9; utt1: 3+1 <= 7
    case("g_blank_last", 30, 3, 8, [30, 22, 15], [6, 4, 2], seed=6, blank=7)
    case("g_long_single", 300, 1, 30, [300], [80], seed=7)
