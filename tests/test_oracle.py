"""CPU tests of the oracle itself (no GPU): pin it against the committed golden fixtures (torch-CPU fp64, generated
by tests/golden/make_golden.py), against definition-level enumeration, and against closed forms."""
import glob
import os

import numpy as np
import pytest

GOLDEN = sorted(f for f in glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz"))
                if os.path.basename(f).startswith("g_"))


def test_golden_present():
    assert len(GOLDEN) >= 7


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_golden(oracle, path):
    z = np.load(path)
    costs, grads = oracle.compute(z["acts"], z["labels"], z["act_lens"], z["label_lens"], blank=int(z["blank"]))
    assert np.allclose(costs, z["costs"], rtol=1e-12, atol=1e-12)
    assert np.abs(grads - z["grads"]).max() <= 1e-12


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_fp32_port_is_reference_class(oracle, path):
    # the timed CPU baseline (fp32 log-space) agrees with the golden values to fp32-class accuracy
    z = np.load(path)
    costs, grads = oracle.compute(z["acts"], z["labels"], z["act_lens"], z["label_lens"], blank=int(z["blank"]),
                                  precision="f32")
    assert np.allclose(costs, z["costs"], rtol=1e-4, atol=1e-4)
    assert np.abs(grads - z["grads"]).max() <= 5e-3


@pytest.mark.parametrize("label", [[1], [1, 2], [2, 2], [3, 1, 3], []])
def test_oracle_vs_enumeration(oracle, label):
    rng = np.random.default_rng(len(label) + 10)
    T, C = 4, 4
    a = rng.standard_normal((T, 1, C)).astype(np.float32)
    costs, grads = oracle.compute(a, np.array(label, np.int32), np.array([T], np.int32),
                                  np.array([len(label)], np.int32))
    bf = oracle.brute_force_cost(a[:, 0], label)
    assert abs(costs[0] - bf) <= 1e-10 * max(1.0, abs(bf))
    g = oracle.brute_force_grad(a[:, 0], label)
    assert np.abs(grads[:, 0] - g).max() <= 1e-7  # finite differences


def test_uniform_logits_closed_form(oracle):
    T, C, label = 12, 5, [1, 2, 2, 3]
    costs, _ = oracle.compute(np.zeros((T, 1, C), np.float32), np.array(label, np.int32), np.array([T], np.int32),
                              np.array([len(label)], np.int32))
    want = -(np.log(oracle.count_alignments(T, label)) - T * np.log(C))
    assert abs(costs[0] - want) <= 1e-12 * want


def test_zero_infinity_modes_and_padding(oracle):
    rng = np.random.default_rng(5)
    a = rng.standard_normal((10, 3, 5)).astype(np.float32)
    labels = np.array([1, 1, 1, 1, 1, 1, 2, 3], np.int32)  # utt0: L=6 + 5 repeats = 11 > 10 -> infeasible
    al, ll = np.array([10, 6, 4], np.int32), np.array([6, 1, 1], np.int32)
    c1, g1 = oracle.compute(a, labels, al, ll, zero_infinity=True)
    c0, g0 = oracle.compute(a, labels, al, ll, zero_infinity=False)
    assert c1[0] == 0.0 and np.isinf(c0[0]) and np.all(g1[:, 0] == 0) and np.all(g0[:, 0] == 0)
    assert np.all(g1[6:, 1] == 0) and np.all(g1[4:, 2] == 0)  # padded frames: exactly zero
    assert np.abs(g1.sum(axis=2)).max() <= 1e-12            # softmax - posterior: rows sum to zero
    assert np.array_equal(c1[1:], c0[1:])


def test_oracle_rejects_bad_labels(oracle):
    a = np.zeros((4, 1, 3), np.float32)
    one = np.array([1], np.int32)
    with pytest.raises(ValueError):
        oracle.compute(a, np.array([0], np.int32), np.array([4], np.int32), one)   # label == blank
    with pytest.raises(ValueError):
        oracle.compute(a, np.array([3], np.int32), np.array([4], np.int32), one)   # label >= C
    with pytest.raises(ValueError):
        oracle.compute(a, one, np.array([5], np.int32), one)                        # act_len > T


def test_numeric_model_of_kernel_scheme(oracle):
    # the block-exponent linear-domain fp64 scheme the CUDA kernels use (tests/numeric_model.py), long and peaky
    import numeric_model as nm
    rng = np.random.default_rng(3)
    for T, L, C, sigma in ((200, 40, 29, 1.0), (300, 90, 30, 5.0), (64, 32, 5, 8.0), (203, 40, 29, 1.0)):
        a = (rng.standard_normal((T, 1, C)) * sigma).astype(np.float32)
        lab = rng.integers(1, C, L).astype(np.int32)
        costs, grads = oracle.compute(a, lab, np.array([T], np.int32), np.array([L], np.int32))
        for hi_words, bound in ((False, 5e-7), (True, 2e-6)):
            # hi_words: checkpoints and parked rows keep 21 significant bits, as on the device -- the bound documents
            # what that costs next to the 1e-5 tolerance of the GPU parity tests
            cost, grad = nm.run(a[:, 0], lab.tolist(), K=8, N=8, hi_words=hi_words)
            if np.isfinite(cost):
                assert abs(cost - costs[0]) <= 1e-6 * abs(costs[0])
                assert np.abs(grad - grads[:, 0]).max() <= bound
            else:
                assert costs[0] == 0.0


def _greedy_golden():
    import glob
    import os
    files = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "greedy_*.npz")))
    assert len(files) == 3
    return [np.load(f) for f in files]


def test_greedy_oracle_against_reference_decoder_goldens():
    # tests/golden/greedy_*.npz were produced by the reference's own decoder.py (make_golden_greedy.py)
    from oracle import greedy_oracle as go
    for z in _greedy_golden():
        probs = z["probs_btc"]
        path, ids, offs = go.decode(np.ascontiguousarray(probs.transpose(1, 0, 2)), z["sizes"], int(z["blank"]))
        assert np.array_equal(path, z["path"])
        assert go.to_strings(ids, str(z["labels"])) == [str(s) for s in z["strings"]]
        assert [len(i) for i in ids] == z["lens"].tolist()
        assert np.array_equal(np.concatenate([np.asarray(o, np.int32) for o in offs]), z["offsets"])


def test_wer_oracle_against_reference_goldens():
    # tests/golden/wer_cases.npz: the reference's own get_cer_wer / Decoder.wer / Decoder.cer (make_golden_wer.py)
    from oracle import wer_oracle

    # the distance function itself: definition-level known answers
    lev = wer_oracle.levenshtein
    assert lev("kitten", "sitting") == 3 and lev("flaw", "lawn") == 2 and lev("", "abc") == 3 and lev("abc", "") == 3
    assert lev("abc", "abc") == 0 and lev("ab", "ba") == 2 and lev("intention", "execution") == 5
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "wer_cases.npz"))
    got = wer_oracle.batch_counts(g["hyp_ids"], g["hyp_lens"], g["ref_flat"], g["ref_lens"], int(g["space"]))
    assert np.array_equal(np.array(got, np.int32), g["counts"])
