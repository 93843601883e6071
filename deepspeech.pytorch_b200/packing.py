"""Host-side input layout of the CTC call: alphabet, target packing, length producers.

Mirrors (re-stated, not copied) the pieces of the reference that define what the loss call receives:
  * alphabet            /root/reference/labels.json (30 symbols, blank "_" at 0) + data/labels.py:13-58
  * target packing      data/data_loader_aug.py:523-548  (_collate_fn: flat int32 targets + target_sizes)
  * act_lens producer   model.py:382-393 (get_seq_lens over the conv stack) and train.py:557 (input_sizes)
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch

# The fork's alphabet: blank, apostrophe, A-Z, '2' (doubled-letter marker, data/labels.py:53-54), space.
ALPHABET: Tuple[str, ...] = ("_", "'") + tuple("ABCDEFGHIJKLMNOPQRSTUVWXYZ") + ("2", " ")
BLANK_INDEX = 0
assert len(ALPHABET) == 30


class Labels:
    """char <-> id with the fork's doubled-letter rule: a letter repeated immediately is written as '2'
    ("HELLO" -> H E L 2 O), so consecutive repeats are rare in real targets (data/labels.py:53-54)."""

    def __init__(self, alphabet: Sequence[str] = ALPHABET):
        self.alphabet = tuple(alphabet)
        self.index = {ch: i for i, ch in enumerate(self.alphabet)}
        self.blank = self.index["_"]

    def encode(self, text: str) -> List[int]:
        out: List[int] = []
        for ch in text.upper():
            if ch not in self.index or ch == "_":
                continue
            code = self.index[ch]
            if out and out[-1] == code and "2" in self.index:
                code = self.index["2"]  # double char (data/labels.py:52-54); "AAA" -> A 2 A
            out.append(code)
        return out

    def decode(self, ids: Iterable[int]) -> str:
        out: List[str] = []
        for i in ids:
            ch = self.alphabet[int(i)]
            if ch == "_":
                continue
            if ch == "2" and out:
                out.append(out[-1])
            else:
                out.append(ch)
        return "".join(out)


def pack_targets(targets: Sequence[Sequence[int]]) -> Tuple[torch.Tensor, torch.Tensor]:
    """List of per-utterance label-id sequences -> (flat int32 [sum L], target_sizes int32 [B]), batch order
    preserved -- the layout `_collate_fn` hands to the criterion (data_loader_aug.py:534,545-547)."""
    sizes = torch.tensor([len(t) for t in targets], dtype=torch.int32)
    flat: List[int] = []
    for t in targets:
        flat.extend(int(v) for v in t)
    return torch.tensor(flat, dtype=torch.int32), sizes


def unpack_targets(flat: torch.Tensor, sizes: torch.Tensor) -> List[List[int]]:
    """Inverse of pack_targets (the 'unflatten targets' loop at train.py:341-345)."""
    out, off = [], 0
    for n in sizes.tolist():
        out.append([int(v) for v in flat[off:off + n].tolist()])
        off += n
    return out


def input_sizes_from_percentages(input_percentages: torch.Tensor, max_seqlength: int) -> torch.Tensor:
    """train.py:557: input_sizes = input_percentages.mul_(T_in).int()"""
    return (input_percentages.float() * int(max_seqlength)).int()


def conv_out_lens(input_lens: torch.Tensor, convs: Sequence[Tuple[int, int, int, int]] = ((11, 2, 5, 1), (11, 1, 5, 1)),
                  true_division: bool = False) -> torch.Tensor:
    """act_lens from input lengths through a stack of time-axis convolutions given as (kernel, stride, padding,
    dilation) -- the formula of model.py:382-393.  The default is DeepSpeech2's two Conv2d time axes
    (k=11,s=2,p=5 then k=11,s=1,p=5), i.e. floor((T_in - 1) / 2) + 1.

    The reference writes `(n + 2p - d(k-1) - 1) / s + 1` on an IntTensor and truncates once at the end
    (`seq_len.int()`).  On the torch versions it was written for `/` on integer tensors floors, i.e. every layer
    floors -- which is also what the convolutions themselves produce and is the default here.  On torch >= 1.5 the
    same expression is a TRUE division carried through the stack in floating point; `true_division=True` reproduces
    that reading.  Both agree for the default stack (tests/test_host.py pins them against the reference's own
    DeepSpeech.get_seq_lens for T_in = 1..4000); they can differ for stacks with more than one strided layer."""
    if true_division:
        x = input_lens.to(torch.int32)  # the reference starts from `lengths.cpu().int()` (model.py:345)
        for k, s, p, d in convs:
            x = (x + 2 * p - d * (k - 1) - 1) / s + 1
        return x.int()
    n = input_lens.to(torch.int64)
    for k, s, p, d in convs:
        n = torch.div(n + 2 * p - d * (k - 1) - 1, s, rounding_mode="floor") + 1
    return n.to(torch.int32)
