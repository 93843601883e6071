"""oracle/greedy_oracle.py -- CPU restatement of the reference's greedy CTC decoding (TEST INFRASTRUCTURE ONLY:
imported by tests/ and by nothing on the product path).

Follows /root/reference/decoder.py:
  * GreedyDecoder.decode (decoder.py:182-197): `_, max_probs = torch.max(probs, 2)` over batch-first [B,T,C]
    probabilities -- per frame the class of the maximum, the FIRST one on ties;
  * GreedyDecoder.process_string (decoder.py:165-180), called with remove_repetitions=True: walk frames
    i < size; skip the blank; skip a frame whose class equals the class of frame i-1 (the previous FRAME, blank or
    not); otherwise emit the character and record the frame index i as its offset.
Pinned by tests/golden/greedy_*.npz, generated from the reference's own decoder.py (tests/golden/make_golden_greedy.py).
"""
import numpy as np


def best_path(acts_tbc: np.ndarray) -> np.ndarray:
    """acts [T,B,C] -> int32 [B,T]: argmax over classes, first maximum wins (np.argmax, like torch.max)."""
    return np.argmax(acts_tbc, axis=2).T.astype(np.int32)


def collapse(path_bt: np.ndarray, act_lens, blank: int = 0):
    """-> (ids list of int lists, offsets list of int lists), one per utterance."""
    B, T = path_bt.shape
    ids, offs = [], []
    for b in range(B):
        n = T if act_lens is None else int(act_lens[b])
        out, off = [], []
        for i in range(n):
            c = int(path_bt[b, i])
            if c == blank:
                continue
            if i != 0 and c == int(path_bt[b, i - 1]):
                continue
            out.append(c)
            off.append(i)
        ids.append(out)
        offs.append(off)
    return ids, offs


def decode(acts_tbc: np.ndarray, act_lens=None, blank: int = 0):
    path = best_path(acts_tbc)
    ids, offs = collapse(path, act_lens, blank)
    return path, ids, offs


def to_strings(ids, labels):
    return ["".join(labels[i] for i in seq) for seq in ids]
