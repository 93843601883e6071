"""Synthetic CTC workloads of BASELINE.json's configs (generator of SURVEY.md section 8(d)).

All tensors are drawn on the CPU from torch.Generator().manual_seed(seed) so that the oracle and the CUDA path
see identical bits; there is no dataset on the box (no network), so this stands in for LibriSpeech-shaped
batches: act_lens ~ U(0.35,1.0)*T with act_lens[0] == T (mirrors `input_percentages`,
data/data_loader_aug.py:544), label_lens = act_lens * r with r ~ U(lo,hi), labels uniform in 1..C-1.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import torch


@dataclass(frozen=True)
class Workload:
    name: str
    B: int
    T: int
    C: int
    L_max: int
    r_lo: float
    r_hi: float


CONFIGS: Dict[str, Workload] = {
    # BASELINE.json configs[0..4]
    "cpu_b16_t500": Workload("cpu_b16_t500", 16, 500, 29, 100, 0.05, 0.20),
    "headline_b64_t1500": Workload("headline_b64_t1500", 64, 1500, 29, 450, 0.15, 0.30),
    "long_b32_t4000": Workload("long_b32_t4000", 32, 4000, 29, 600, 0.05, 0.15),
    "bigc_b64_t1000_c5000": Workload("bigc_b64_t1000_c5000", 64, 1000, 5000, 200, 0.05, 0.20),
    "sweep_b1024_t1500": Workload("sweep_b1024_t1500", 1024, 1500, 29, 450, 0.15, 0.30),
}
HEADLINE = "headline_b64_t1500"


def make(w: Workload, seed: int = 1234, sigma: float = 1.0, B: Optional[int] = None, C: Optional[int] = None,
         no_repeats: bool = False, acts: bool = True, sort_by_length: bool = True):
    """-> dict(acts fp32 [T,B,C] | None, labels int32 [sum L], act_lens int32 [B], label_lens int32 [B]).

    sort_by_length (default): utterances are ordered longest first, which is the order every batch of the reference
    has -- its collate function sorts by spectrogram length (data/data_loader_aug.py:527
    `batch = sorted(batch, key=lambda sample: sample[0].size(1), reverse=True)`).  The draw itself is unchanged (the
    same utterances as with sort_by_length=False, permuted; the logits are i.i.d., so they need no permutation)."""
    g = torch.Generator().manual_seed(seed)
    B = w.B if B is None else B
    C = w.C if C is None else C
    T = w.T
    u = torch.rand(B, generator=g) * 0.65 + 0.35
    act_lens = torch.clamp((u * T).round().int(), min=1, max=T)
    act_lens[0] = T
    r = torch.rand(B, generator=g) * (w.r_hi - w.r_lo) + w.r_lo
    label_lens = torch.clamp((act_lens.float() * r).round().int(), min=1, max=w.L_max)
    total = int(label_lens.sum())
    labels = torch.randint(1, C, (total,), generator=g, dtype=torch.int32)
    if no_repeats and C > 2:
        # the fork's real targets almost never repeat a label (doubled letters become '2'): remove repeats
        off = 0
        for n in label_lens.tolist():
            for i in range(off + 1, off + n):
                if labels[i] == labels[i - 1]:
                    labels[i] = 1 + (int(labels[i]) % (C - 1))  # next class, wraps within 1..C-1
            off += n
    if sort_by_length:
        order = torch.argsort(act_lens, descending=True, stable=True)
        offs = torch.cumsum(label_lens, 0) - label_lens
        labels = torch.cat([labels[int(offs[b]):int(offs[b]) + int(label_lens[b])] for b in order.tolist()])
        act_lens, label_lens = act_lens[order].contiguous(), label_lens[order].contiguous()
    a = torch.randn(T, B, C, generator=g) * sigma if acts else None
    return {"acts": a, "labels": labels, "act_lens": act_lens, "label_lens": label_lens, "T": T, "B": B, "C": C}


def algorithmic_bytes(T: int, B: int, C: int, total_labels: int = 0, score_only: bool = False) -> int:
    """read fp32 acts once + write fp32 grads once (+ labels, lengths, costs)."""
    per_frame = (4 if score_only else 8) * C
    return per_frame * B * T + 4 * (total_labels + 2 * B) + 4 * B
