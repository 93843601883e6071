"""Generates tests/golden/wer_cases.npz from the REFERENCE's own error-count code (run in the build container, where
/root/reference exists; the GPU box only sees the committed .npz).

/root/reference/decoder.py (Decoder.wer / Decoder.cer) is imported unmodified and get_cer_wer is executed from the
source text of /root/reference/data/utils.py:47-57 (the module itself imports torch.distributed helpers and data
loaders that are irrelevant here).  The one thing the reference cannot supply is `Levenshtein.distance`
(python-Levenshtein is not installed and not vendored): a module with the textbook unit-cost edit distance
(oracle/wer_oracle.levenshtein) is registered under that name.  So these vectors pin the reference's TOKENISATION
(split / replace / strip / word->symbol mapping / the `or 1` denominators), not the third-party distance function.

    python tests/golden/make_golden_wer.py
"""
import json
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def main():
    sys.path.insert(0, ROOT)
    from oracle import wer_oracle

    lev = types.ModuleType("Levenshtein")
    lev.distance = lambda a, b: wer_oracle.levenshtein(a, b)
    sys.modules["Levenshtein"] = lev
    sys.path.insert(0, REF)
    import decoder as ref_decoder  # noqa: E402  (the reference's module)

    src = open(os.path.join(REF, "data", "utils.py")).read()
    m = re.search(r"^def get_cer_wer\(.*?(?=^def |\Z)", src, re.S | re.M)
    ns = {}
    exec(m.group(0), ns)  # the reference's function text, verbatim
    get_cer_wer = ns["get_cer_wer"]

    with open(os.path.join(REF, "labels.json")) as f:
        labels = "".join(json.load(f))
    dec = ref_decoder.Decoder(labels, blank_index=labels.index("_"))
    space = labels.index(" ")
    rng = np.random.default_rng(7)
    letters = [i for i in range(len(labels)) if i not in (labels.index("_"), space)]

    def sentence(nwords, maxlen):
        ids = []
        for w in range(nwords):
            ids += list(rng.choice(letters[:6], size=int(rng.integers(1, maxlen + 1))))  # small alphabet: collisions
            ids += [space] * int(rng.integers(1, 3))  # single and double spaces
        return ids[:-1] if ids and rng.random() < 0.7 else ids

    def corrupt(ids):
        out = []
        for v in ids:
            r = rng.random()
            if r < 0.08:
                continue
            if r < 0.16:
                out.append(int(rng.choice(letters[:6])))
                continue
            out.append(v)
            if r > 0.94:
                out.append(space if rng.random() < 0.5 else int(rng.choice(letters[:6])))
        return out

    hyps, refs = [], []
    for _ in range(40):
        r = sentence(int(rng.integers(1, 12)), 6)
        refs.append(r)
        hyps.append(corrupt(r))
    # edge cases: identical, empty hypothesis, empty reference, only spaces, leading / trailing spaces
    refs += [[2, 3, space, 4], [2, 3, space, 4], [], [space, space], [space, 2, 3, space], [2, 3]]
    hyps += [[2, 3, space, 4], [], [2, 3], [2], [2, 3], [space, space, 2, 3, space]]
    to_s = lambda ids: "".join(labels[i] for i in ids)
    out = []
    for h, r in zip(hyps, refs):
        wer, cer, wer_ref, cer_ref = get_cer_wer(dec, to_s(h), to_s(r))
        out.append((int(wer), int(cer), int(wer_ref), int(cer_ref)))
    P = max(len(h) for h in hyps)
    hyp_ids = np.zeros((len(hyps), P), np.int32)
    for b, h in enumerate(hyps):
        hyp_ids[b, :len(h)] = h
    np.savez_compressed(os.path.join(HERE, "wer_cases.npz"), hyp_ids=hyp_ids,
                        hyp_lens=np.array([len(h) for h in hyps], np.int32),
                        ref_flat=np.array([v for r in refs for v in r], np.int32),
                        ref_lens=np.array([len(r) for r in refs], np.int32), space=np.int32(space),
                        counts=np.array(out, np.int32))
    print(len(out), "cases; first:", out[:5], "edge:", out[-6:])


if __name__ == "__main__":
    main()
