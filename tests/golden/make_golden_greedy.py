"""Generates tests/golden/greedy_*.npz from the REFERENCE's own greedy decoder (run in the build container, where
/root/reference exists; the GPU box only sees the committed .npz files).

/root/reference/decoder.py is imported unmodified.  Its top-level `import Levenshtein` (used only by wer/cer, which
are not called here) is satisfied by an empty stub module, because python-Levenshtein is not installed in the image.

    python tests/golden/make_golden_greedy.py
"""
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def main():
    sys.modules.setdefault("Levenshtein", types.ModuleType("Levenshtein"))
    sys.path.insert(0, REF)
    import decoder as ref_decoder  # noqa: E402  (the reference's module)

    with open(os.path.join(REF, "labels.json")) as f:
        labels = "".join(json.load(f))
    dec = ref_decoder.GreedyDecoder(labels, blank_index=labels.index("_"))
    g = torch.Generator().manual_seed(2024)
    cases = {"small": (5, 40, len(labels), 1.0), "peaky": (7, 120, len(labels), 6.0), "ties": (3, 50, 5, 0.0)}
    for name, (B, T, C, sigma) in cases.items():
        if sigma > 0:
            logits = torch.randn(B, T, C, generator=g) * sigma
            # long runs of one class and of blanks, as a trained model would produce
            runs = torch.randint(0, C, (B, T // 4 + 1), generator=g).repeat_interleave(4, dim=1)[:, :T]
            logits.scatter_add_(2, runs.unsqueeze(2), torch.full((B, T, 1), 3.0 * sigma))
        else:
            logits = torch.randint(0, 2, (B, T, C), generator=g).float()  # many exact ties
        sizes = torch.randint(T // 2, T + 1, (B,), generator=g).int()
        sizes[0] = T
        probs = torch.softmax(logits, dim=-1) if sigma > 0 else logits
        strings, offsets = dec.decode(probs, sizes)
        _, path = torch.max(probs, 2)
        flat_off = np.concatenate([o[0].numpy().astype(np.int32) for o in offsets]) if B else np.zeros(0, np.int32)
        lens = np.array([len(o[0]) for o in offsets], dtype=np.int32)
        np.savez_compressed(os.path.join(HERE, f"greedy_{name}.npz"), probs_btc=probs.numpy().astype(np.float32),
                            sizes=sizes.numpy().astype(np.int32), path=path.numpy().astype(np.int32),
                            strings=np.array([s[0] for s in strings]), offsets=flat_off, lens=lens,
                            labels=np.array(labels), blank=np.int32(labels.index("_")))
        print(name, [s[0][:30] for s in strings][:2], lens.tolist())


if __name__ == "__main__":
    main()
