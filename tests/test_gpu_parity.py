"""Parity of the CUDA path (through the C ABI / public op) against the CPU oracle.  Run with -m gpu on a B200.

Tolerance (BASELINE.json north_star): loss rel <= 1e-5, grad abs <= 1e-5 against the fp64 oracle.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5
GRAD_ATOL = 1e-5


def _run(b200, d, zero_infinity=True, device_lens=False, want_grads=True, grad_scale=1.0, blank=0, as_numpy=True):
    acts = d["acts"].cuda()
    lab, al, ll = d["labels"], d["act_lens"], d["label_lens"]
    kw = {}
    if device_lens:
        kw["max_label_len"] = int(ll.max())
        lab, al, ll = lab.cuda(), al.cuda(), ll.cuda()
    costs, grads = b200.ctc_forward_backward(acts, lab, al, ll, blank=blank, zero_infinity=zero_infinity,
                                             grad_scale=grad_scale, want_grads=want_grads, **kw)
    torch.cuda.synchronize()
    if not as_numpy:  # large tensors: stay in fp32 torch, compare in chunks (_max_abs_diff)
        return costs.cpu().numpy().astype(np.float64), (grads.cpu() if want_grads else None)
    return costs.cpu().numpy().astype(np.float64), (grads.cpu().numpy().astype(np.float64) if want_grads else None)


def _max_abs_diff(grads_f32, og, scale=1.0):
    """max |grads - scale * og| over a [T,B,C] pair, one time chunk at a time (the fp64 copies of a 1.3 GB gradient
    would not be a good use of the box's memory)."""
    g = grads_f32.numpy() if isinstance(grads_f32, torch.Tensor) else grads_f32
    worst = 0.0
    step = max(1, (1 << 24) // max(1, g.shape[1] * g.shape[2]))
    for t0 in range(0, g.shape[0], step):
        worst = max(worst, float(np.abs(g[t0:t0 + step].astype(np.float64) - og[t0:t0 + step] * scale).max(initial=0.0)))
    return worst


def _check(oracle, b200, d, cost_floor=1e-30, **kw):
    """cost_floor (scalar or per utterance): costs below it are compared with the absolute error LOSS_RTOL * cost_floor.
    The lattice reads the emission probabilities with 21 significant bits (the reference's are fp32): a frame's log
    probability is off by up to 1.2e-7, so a cost of 1e-5 -- every frame's probability within 1e-5 of 1 -- cannot be
    right to 5 digits."""
    big = d["acts"].numel() > (1 << 26)
    costs, grads = _run(b200, d, as_numpy=not big, **kw)
    oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy(),
                            blank=kw.get("blank", 0), zero_infinity=kw.get("zero_infinity", True))
    fin = np.isfinite(oc)
    assert np.array_equal(np.isfinite(costs), fin)
    floor = np.broadcast_to(np.asarray(cost_floor, dtype=np.float64), oc.shape)[fin]
    rel = np.abs(costs[fin] - oc[fin]) / np.maximum(np.abs(oc[fin]), floor)
    # per-utterance costs are fp32 on the device: allow fp32 rounding of the value itself
    assert rel.max(initial=0.0) <= LOSS_RTOL, f"cost rel err {rel.max()}"
    tot_rel = abs(costs[fin].sum() - oc[fin].sum()) / max(abs(oc[fin].sum()), float(floor.sum()), 1e-30)
    assert tot_rel <= LOSS_RTOL
    if grads is not None:
        err = _max_abs_diff(grads, og, kw.get("grad_scale", 1.0))
        assert err <= GRAD_ATOL, f"grad max-abs err {err}"
    return costs, grads


def _wl(b200, name, **kw):
    return b200.workloads.make(b200.workloads.CONFIGS[name], **kw)


def test_tiny_closed_form(oracle, b200):
    # T=3, L=2 with and without a repeated label: enumeration over all C^T alignments
    rng = np.random.default_rng(0)
    for label in ([1, 2], [2, 2], [1], []):
        T = 3 if len(label) else 2
        a = rng.standard_normal((T, 1, 4)).astype(np.float32)
        d = {"acts": torch.from_numpy(a), "labels": torch.tensor(label, dtype=torch.int32),
             "act_lens": torch.tensor([T], dtype=torch.int32), "label_lens": torch.tensor([len(label)], dtype=torch.int32)}
        costs, grads = _run(b200, d)
        bf = oracle.brute_force_cost(a[:, 0], label)
        if np.isfinite(bf):
            assert abs(costs[0] - bf) <= 1e-5 * abs(bf)
            g = oracle.brute_force_grad(a[:, 0], label)
            assert np.abs(grads[:, 0] - g).max() <= 1e-5
        else:
            assert costs[0] == 0.0 and np.all(grads == 0)


@pytest.mark.parametrize("C", [29, 30])
def test_config1_cpu_shape(oracle, b200, C):
    _check(oracle, b200, _wl(b200, "cpu_b16_t500", C=C))


@pytest.mark.parametrize("sigma", [1.0, 5.0])
def test_config2_headline(oracle, b200, sigma):
    d = _wl(b200, "headline_b64_t1500", sigma=sigma)
    costs, grads = _check(oracle, b200, d)
    # padded frames: exactly zero gradient rows; valid rows sum to zero over classes
    al = d["act_lens"].numpy()
    for b in range(0, 64, 7):
        assert np.all(grads[al[b]:, b, :] == 0.0)
    assert np.abs(grads.sum(axis=2)).max() <= 2e-5


def test_config2_no_repeats_and_device_lengths(oracle, b200):
    d = _wl(b200, "headline_b64_t1500", no_repeats=True, seed=99)
    _check(oracle, b200, d, device_lens=True)


def test_config3_long_small_batch(oracle, b200):
    _check(oracle, b200, _wl(b200, "long_b32_t4000", B=5, seed=5))  # odd batch: no slicing, nsplit at its cap


def test_config4_large_alphabet_small_batch(oracle, b200):
    # BPE-style vocabulary: the position-indexed emission path (C > 128)
    d = _wl(b200, "bigc_b64_t1000_c5000", B=4)
    costs, grads = _check(oracle, b200, d)
    al = d["act_lens"].numpy()
    for b in range(4):
        assert np.all(grads[al[b]:, b, :] == 0.0)
    assert np.abs(grads.sum(axis=2)).max() <= 5e-5  # 5000 fp32 terms per row


# ---- the BASELINE.json configurations at their REAL batch sizes (grid sizes, slice plans, nsplit and workspace
# offsets all depend on B; the reference call being matched is /root/reference/train.py:600)
@pytest.mark.parametrize("sigma", [1.0, 5.0])
def test_config3_long_full_size(oracle, b200, sigma):
    d = _wl(b200, "long_b32_t4000", sigma=sigma)
    assert d["B"] == 32 and d["T"] == 4000
    _check(oracle, b200, d)


@pytest.mark.parametrize("sigma", [1.0, 5.0])
def test_config4_large_alphabet_full_size(oracle, b200, sigma):
    d = _wl(b200, "bigc_b64_t1000_c5000", sigma=sigma)
    assert d["B"] == 64 and d["C"] == 5000
    costs, grads = _check(oracle, b200, d)
    al = d["act_lens"].numpy()
    for b in range(0, 64, 9):
        assert bool((grads[int(al[b]):, b, :] == 0).all())


def test_config5_sweep_b1024_single_and_sharded(oracle, b200, monkeypatch):
    # config 5 on one GPU at B = 1024, then the same batch dealt into 8 length-bucketed shards (what 8 ranks would
    # each hold, sharding.length_bucketed_shards) run one after another: per-utterance costs and the gradients
    # scattered back must equal the unsharded run BIT FOR BIT (an utterance's result may not depend on its batch).
    # The library picks the form of its recursion kernel by batch size (first form for B > 224); the two forms renormalise
    # in different orders and their checkpoints can differ in the last bits, which -- rarely -- moves one fp32 gradient
    # entry by one ulp (1 entry in 45 million on this batch).  So: bit for bit with the form pinned, and with the
    # library's own choice the costs bit for bit, the gradients within one ulp of a value <= 1.
    d = _wl(b200, "sweep_b1024_t1500")
    assert d["B"] == 1024
    costs, grads = _check(oracle, b200, d)
    shards = b200.sharding.length_bucketed_shards(d["act_lens"].tolist(), d["label_lens"].tolist(), 8)
    for form in ("1", "0"):
        monkeypatch.setenv("CTC_B200_ENGINE", form)
        for ids in shards:
            a, lab, al, ll = b200.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], ids)
            c, g = b200.ctc_forward_backward(a.cuda(), lab, al, ll)
            torch.cuda.synchronize()
            assert np.array_equal(c.cpu().numpy().astype(np.float64), costs[ids])
            want = torch.from_numpy(grads[:, ids, :]).float()
            if form == "1":
                assert torch.equal(g.cpu(), want)
            else:
                diff = (g.cpu() - want).abs()
                assert float(diff.max()) <= 6e-8 and int((diff > 0).sum()) <= 8


# ---- blank != 0 and the 32 < C <= 128 instantiations through the CUDA path
def _relabel_for_blank(d, blank):
    """labels are drawn from 1..C-1: a label equal to the new blank index takes the freed index 0 instead"""
    out = dict(d)
    lab = d["labels"].clone()
    lab[lab == blank] = 0
    out["labels"] = lab
    return out


def test_blank_last_golden_through_cuda(b200):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "g_blank_last.npz"))
    d = {"acts": torch.from_numpy(g["acts"]), "labels": torch.from_numpy(g["labels"]),
         "act_lens": torch.from_numpy(g["act_lens"]), "label_lens": torch.from_numpy(g["label_lens"])}
    costs, grads = _run(b200, d, blank=int(g["blank"]))
    assert int(g["blank"]) == 7
    assert (np.abs(costs - g["costs"]) / np.abs(g["costs"])).max() <= LOSS_RTOL
    assert np.abs(grads - g["grads"]).max() <= GRAD_ATOL


@pytest.mark.parametrize("C,blank", [(29, 28), (30, 13), (48, 47), (100, 50), (700, 699), (700, 123)])
def test_blank_index_other_than_zero(oracle, b200, C, blank):
    w = b200.workloads.Workload("blank", 12, 200, C, 40, 0.05, 0.2)
    d = _relabel_for_blank(b200.workloads.make(w, seed=C + blank), blank)
    _check(oracle, b200, d, blank=blank)
    _check(oracle, b200, d, blank=blank, device_lens=True, grad_scale=0.5)
    c_eval, _ = _run(b200, d, blank=blank, want_grads=False)
    c_train, _ = _run(b200, d, blank=blank)
    assert np.array_equal(c_eval, c_train)


@pytest.mark.parametrize("C", [33, 48, 64, 65, 100, 128])
def test_mid_sized_alphabets(oracle, b200, C):
    # the Cp = 64 and Cp = 128 instantiations of the prep / engine / gradient kernels (16 and 32 lanes per softmax
    # row): random batch, then a hand-made one with repeated labels, an infeasible utterance and L = 0
    w = b200.workloads.Workload("mid", 20, 300, C, 70, 0.05, 0.25)
    _check(oracle, b200, b200.workloads.make(w, seed=C))
    T = 36
    g = torch.Generator().manual_seed(C)
    acts = torch.randn(T, 5, C, generator=g) * 2.0
    targets = [[C - 1, C - 1, 7, 7, 9, C - 1, 3, 3], [], [C // 2] * 20, [1, 2, 3, 1, 2, 3, 1, 2, 3], [C - 2]]
    act_lens = torch.tensor([36, 12, 36, 9, 1], dtype=torch.int32)  # utt2: 20 + 19 repeats > 36 -> infeasible
    labels, ll = b200.pack_targets(targets)
    d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
    costs, grads = _check(oracle, b200, d)
    assert costs[2] == 0.0 and np.all(grads[:, 2] == 0)
    _check(oracle, b200, d, device_lens=True, grad_scale=0.25, zero_infinity=False)


# ---- non-finite logits without the guard: they propagate, they are never turned into a finite wrong answer
@pytest.mark.parametrize("C", [29, 100, 700, 5000])
@pytest.mark.parametrize("bad", [float("nan"), float("inf")])
def test_non_finite_logits_propagate(oracle, b200, C, bad):
    w = b200.workloads.Workload("nf", 6, 80, C, 16, 0.05, 0.2)
    d = b200.workloads.make(w, seed=C)
    acts = d["acts"].clone()
    acts[5, 1, 3] = bad      # valid frame of utterance 1
    acts[79, 2, 0] = bad     # utterance 2: only if that frame is valid (act_lens[2] may be shorter)
    al = d["act_lens"].numpy()
    hit = {1} | ({2} if al[2] > 79 else set())
    d2 = dict(d, acts=acts)
    costs, grads = _run(b200, d2)  # nan_to_zero off (the default)
    clean = dict(d, acts=torch.nan_to_num(acts, nan=0.0, posinf=0.0))
    oc, og = oracle.compute(clean["acts"].numpy(), d["labels"].numpy(), al, d["label_lens"].numpy())
    for b in range(w.B):
        if b in hit:
            assert np.isnan(costs[b]), (b, costs[b])
            assert np.isnan(grads[:int(al[b]), b]).any()
        else:  # the other utterances of the batch are untouched
            assert abs(costs[b] - oc[b]) <= LOSS_RTOL * abs(oc[b])
            assert np.abs(grads[:, b] - og[:, b]).max() <= GRAD_ATOL
    c_eval, _ = _run(b200, d2, want_grads=False)
    assert np.array_equal(np.isnan(c_eval), np.isnan(costs))
    # the summed loss the criterion returns is NaN, which is what train.py:604-611 / 625-627 look for
    loss = b200.CTCLoss()(acts.cuda(), d["labels"], d["act_lens"], d["label_lens"])
    assert torch.isnan(loss).all()


def test_large_alphabet_edges(oracle, b200):
    # C just above the class-row limit, C not a multiple of 4 (unaligned rows), repeated labels, infeasible, L = 0
    for C in (129, 203):
        T = 30
        g = torch.Generator().manual_seed(C)
        acts = torch.randn(T, 5, C, generator=g) * 2.0
        targets = [[5, 5, 5, 7, 7, 128], [], [100] * 20, [1, 2, 3, 1, 2, 3, 1, 2, 3], [C - 1]]
        act_lens = torch.tensor([30, 12, 30, 9, 1], dtype=torch.int32)  # utt2: 20 + 19 repeats > 30 -> infeasible
        labels, ll = b200.pack_targets(targets)
        d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
        costs, grads = _check(oracle, b200, d, zero_infinity=True)
        assert costs[2] == 0.0 and np.all(grads[:, 2] == 0)
        _check(oracle, b200, d, device_lens=True, grad_scale=0.25)


@pytest.mark.parametrize("C", [132, 5120, 6000, 10000, 16388, 20000])
def test_large_alphabet_row_kernels(oracle, b200, C):
    # every form of the large-alphabet row kernel: rows kept in registers (5 / 8 / 16 float4 per thread), rows staged
    # in shared memory (C % 4 != 0 or > 16384 classes), rows too long to stage (re-read through L2); repeated
    # classes within an utterance, padding frames, an infeasible utterance and gradient scaling
    T, B = 24, 6
    g = torch.Generator().manual_seed(C)
    acts = torch.randn(T, B, C, generator=g) * 3.0
    targets = [[C - 1, 7, 7, 9, C - 1, 3], [1], [5] * 10, [2, 4, 6, 8, 10, 12, 14, 16], [], [11, 11, 11]]
    act_lens = torch.tensor([24, 5, 24, 17, 9, 4], dtype=torch.int32)  # utt 5: 3 + 2 repeats > 4 -> infeasible
    labels, ll = b200.pack_targets(targets)
    d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
    costs, grads = _check(oracle, b200, d, zero_infinity=True)
    assert costs[5] == 0.0 and np.all(grads[:, 5] == 0)
    for b in range(B):
        assert np.all(grads[int(act_lens[b]):, b, :] == 0.0)
    _check(oracle, b200, d, device_lens=True, grad_scale=0.125)
    c_eval, _ = _run(b200, d, want_grads=False)  # eval call sites (train.py:352,457): no gradient buffer at all
    assert np.array_equal(c_eval, costs)
    # a view whose rows are 16-byte aligned but whose base is not: falls back to the generic kernel (scalar loads, another summation order)
    if C % 4 == 0:
        pad = torch.zeros(T * B * C + 1, device="cuda")
        pad[1:] = acts.reshape(-1).cuda()
        d2 = dict(d, acts=pad[1:].view(T, B, C))
        assert d2["acts"].data_ptr() % 16 == 4
        c2, g2 = _run(b200, d2)
        assert np.abs(c2 - costs).max() <= 1e-5 * np.abs(costs).max() and np.abs(g2 - grads).max() <= 2e-6


def test_grad_scale_and_score_only(oracle, b200):
    d = _wl(b200, "cpu_b16_t500")
    _check(oracle, b200, d, grad_scale=1.0 / 16)
    c1, _ = _run(b200, d, want_grads=False)
    c2, _ = _run(b200, d, want_grads=True)
    assert np.array_equal(c1, c2)


def test_deterministic(b200):
    d = _wl(b200, "cpu_b16_t500", seed=3)
    c1, g1 = _run(b200, d)
    c2, g2 = _run(b200, d)
    assert np.array_equal(c1, c2) and np.array_equal(g1, g2)


def test_edge_cases(oracle, b200):
    # ragged: infeasible (L + repeats > T), L = 0, T_b = 1, tight T == L + repeats, repeated labels
    C, T = 6, 40
    g = torch.Generator().manual_seed(11)
    acts = torch.randn(T, 6, C, generator=g)
    targets = [[1, 1, 1, 1], [], [2], [3, 3, 4, 4, 5], [1, 2, 3, 4, 5, 1, 2, 3], [5] * 30]
    act_lens = torch.tensor([6, 40, 1, 7, 8, 40], dtype=torch.int32)  # utt0: needs 7 > 6 -> infeasible; utt5: 59 > 40
    labels, ll = b200.pack_targets(targets)
    d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
    costs, grads = _check(oracle, b200, d, zero_infinity=True)
    assert costs[0] == 0.0 and costs[5] == 0.0 and np.all(grads[:, 0] == 0) and np.all(grads[:, 5] == 0)
    costs, grads = _run(b200, d, zero_infinity=False)
    assert np.isinf(costs[0]) and np.isinf(costs[5]) and np.all(np.isfinite(grads)) and np.all(grads[:, 0] == 0)


@pytest.mark.parametrize("seed", range(int(os.environ.get("CTC_TEST_RANDOM_SHAPES", "40"))))
def test_randomised_shapes(oracle, b200, seed, monkeypatch):
    # random small problems over everything that changes a launch plan: alphabet size (all three small-alphabet
    # instantiations and the large-alphabet path), blank index, batch size (slicing starts at 16), slice count, one
    # gradient launch or two warp classes, label lengths from 0 up to several gradient warps, tight and infeasible
    # utterances, heavy label repetition
    rng = np.random.default_rng(1000 + seed)
    C = int(rng.choice([2, 3, 7, 29, 32, 33, 64, 97, 128, 129, 300]))
    B = int(rng.choice([1, 3, 16, 21, 40]))
    T = int(rng.choice([1, 9, 64, 257, 600]))
    blank = int(rng.integers(0, C))
    monkeypatch.setenv("CTC_B200_SLICES", str(int(rng.choice([1, 2, 5, 8]))))
    monkeypatch.setenv("CTC_B200_GRAD_CLASSES", str(int(rng.integers(0, 2))))
    act_lens = rng.integers(max(1, T // 3), T + 1, size=B).astype(np.int32)
    act_lens[rng.integers(0, B)] = T
    lens, labs = [], []
    symbols = [c for c in range(C) if c != blank]
    for b in range(B):
        n = int(rng.integers(0, min(int(act_lens[b]), 400) + 1))
        if rng.random() < 0.3:
            n = n // 8
        few = rng.random() < 0.3  # few distinct symbols: many repeats (each needs a blank between)
        pool = symbols[:2] if few and len(symbols) > 1 else symbols
        lab = rng.choice(pool, size=n).astype(np.int32) if n else np.zeros(0, np.int32)
        lens.append(n)
        labs.append(lab)
    d = {"acts": torch.from_numpy((rng.standard_normal((T, B, C)) * rng.choice([0.5, 1.0, 4.0])).astype(np.float32)),
         "labels": torch.from_numpy(np.concatenate(labs) if labs else np.zeros(0, np.int32)),
         "act_lens": torch.from_numpy(act_lens), "label_lens": torch.tensor(lens, dtype=torch.int32)}
    # (cost_floor: the 21-bit emission probabilities put up to 4.8e-7 (rms 2e-7) on every frame's log probability, a
    # random walk over the frames; costs below 0.1 * sqrt(T) -- utterances predicted almost with certainty -- are
    # therefore compared absolutely, with 1e-6 * sqrt(T), about 4 sigma of that walk)
    _check(oracle, b200, d, cost_floor=0.1 * np.sqrt(np.maximum(act_lens, 1).astype(np.float64)), blank=blank,
           device_lens=bool(rng.integers(0, 2)))


@pytest.mark.parametrize("C,T", [(2, 600), (29, 1500)])
def test_confidently_predicted_utterances(oracle, b200, C, T):
    # logits that put ~99.5 % on one alignment: the cost is small against the number of frames, so a BIAS in how the
    # emission probabilities are rounded shows at once (rounding their high words half-up instead of to even put 1e-5
    # relative on exactly this kind of batch); unbiased rounding leaves a random walk of ~2e-7 per frame
    rng = np.random.default_rng(C)
    B = 6
    acts = rng.standard_normal((T, B, C)).astype(np.float32) * 0.3
    labels, lens = [], []
    for b in range(B):
        L = T // 4 - 3 * b
        lab = rng.integers(1, C, size=L).astype(np.int32) if C > 2 else np.ones(L, np.int32)
        t = 0
        for sym in lab:  # blank, blank, symbol, symbol, ...
            acts[t:t + 2, b, 0] += 9.0
            acts[t + 2:t + 4, b, sym] += 9.0
            t += 4
        acts[t:, b, 0] += 9.0
        labels.append(lab)
        lens.append(L)
    d = {"acts": torch.from_numpy(acts), "labels": torch.from_numpy(np.concatenate(labels)),
         "act_lens": torch.full((B,), T, dtype=torch.int32), "label_lens": torch.tensor(lens, dtype=torch.int32)}
    costs, grads = _check(oracle, b200, d, cost_floor=0.1 * np.sqrt(T))  # (see test_randomised_shapes)
    oc, _ = oracle.compute(acts, d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    assert oc.max() < 0.02 * T  # the point of the case: far less than a nat per frame
    assert np.abs(costs - oc).max() <= 1.5e-6 * np.sqrt(T)  # ~5 sigma of the rounding walk; the biased rounding gave 3e-5


def test_uniform_logits_count(oracle, b200):
    # uniform logits: cost = -log(#alignments / C^T)
    C, T = 5, 12
    label = [1, 2, 2, 3]
    d = {"acts": torch.zeros(T, 1, C), "labels": torch.tensor(label, dtype=torch.int32),
         "act_lens": torch.tensor([T], dtype=torch.int32), "label_lens": torch.tensor([len(label)], dtype=torch.int32)}
    costs, _ = _run(b200, d)
    want = -(np.log(oracle.count_alignments(T, label)) - T * np.log(C))
    assert abs(costs[0] - want) <= 1e-5 * want


def test_module_matches_reference_call(oracle, b200):
    # the reference's train-step block (train.py:593-620): transpose view, criterion, /B, backward
    w = b200.workloads.Workload("m", 8, 120, 30, 30, 0.05, 0.2)
    d = b200.workloads.make(w, seed=21)
    logits_btc = d["acts"].transpose(0, 1).contiguous().cuda().requires_grad_(True)  # model output [B,T,C]
    crit = b200.CTCLoss()
    loss = crit(logits_btc.transpose(0, 1), d["labels"], d["act_lens"], d["label_lens"])
    loss = loss / w.B
    loss = loss.to("cuda")
    loss.backward()
    oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    assert loss.shape == (1,)
    assert abs(loss.item() * w.B - oc.sum()) <= 1e-5 * oc.sum()
    g = logits_btc.grad.transpose(0, 1).cpu().numpy().astype(np.float64) * w.B
    assert np.abs(g - og).max() <= 1e-5
    with torch.no_grad():  # eval call sites train.py:352,457: score only
        l2 = crit(logits_btc.transpose(0, 1), d["labels"], d["act_lens"], d["label_lens"])
    assert abs(l2.item() - oc.sum()) <= 1e-5 * oc.sum()
    crit_avg = b200.CTCLoss(size_average=True, loss_on_cpu=True)
    l3 = crit_avg(logits_btc.detach().transpose(0, 1), d["labels"], d["act_lens"], d["label_lens"])
    assert not l3.is_cuda and abs(l3.item() * w.B - oc.sum()) <= 1e-5 * oc.sum()


def test_errors(b200):
    a = torch.randn(10, 2, 5)
    with pytest.raises(ValueError):
        b200.ctc_forward_backward(a, torch.tensor([1], dtype=torch.int32), torch.tensor([10, 10], dtype=torch.int32),
                                  torch.tensor([1, 0], dtype=torch.int32))  # CPU acts: no fallback
    with pytest.raises(RuntimeError):  # label == blank
        b200.ctc_forward_backward(a.cuda(), torch.tensor([0], dtype=torch.int32),
                                  torch.tensor([10, 10], dtype=torch.int32), torch.tensor([1, 0], dtype=torch.int32))
    with pytest.raises(RuntimeError):  # act_len > T
        b200.ctc_forward_backward(a.cuda(), torch.tensor([1], dtype=torch.int32),
                                  torch.tensor([11, 10], dtype=torch.int32), torch.tensor([1, 0], dtype=torch.int32))


def test_length_bucket_slices_are_invisible(oracle, b200, monkeypatch):
    # The engine/grad launches are cut into length-bucket slices on side streams; the result of an utterance must
    # not depend on how the batch was sliced (bitwise), including batch sizes that do not divide evenly.
    for B in (64, 37, 9):
        d = _wl(b200, "cpu_b16_t500", seed=5, B=B)
        monkeypatch.setenv("CTC_B200_SLICES", "1")
        c1, g1 = _run(b200, d)
        for ns in ("2", "5", "8"):
            monkeypatch.setenv("CTC_B200_SLICES", ns)
            c2, g2 = _run(b200, d)
            assert np.array_equal(c1, c2) and np.array_equal(g1, g2), (B, ns)
        monkeypatch.delenv("CTC_B200_SLICES")
        if B == 37:
            _check(oracle, b200, d)


def test_sliced_launch_orders_with_caller_stream(b200):
    # work enqueued on the caller's stream after the call must see the finished gradients (fork/join events)
    d = _wl(b200, "cpu_b16_t500", seed=6, B=32)
    acts = d["acts"].cuda()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        costs, grads = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"])
        total = grads.abs().sum() + costs.sum()  # consumer on the same (non-default) stream
    s.synchronize()
    c2, g2 = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"])
    torch.cuda.synchronize()
    assert torch.equal(grads, g2) and torch.equal(costs, c2)
    assert abs(float(total) - float(g2.abs().sum() + c2.sum())) <= 1e-3 * abs(float(total))


def _sorted_by_length(d):
    """The batch order the reference's collate function produces (data_loader_aug.py:527): longest utterance first."""
    order = torch.argsort(d["act_lens"], descending=True, stable=True)
    offs = torch.cumsum(d["label_lens"], 0) - d["label_lens"]
    labels = torch.cat([d["labels"][int(offs[b]):int(offs[b]) + int(d["label_lens"][b])] for b in order.tolist()])
    out = dict(d)
    out.update(acts=d["acts"][:, order, :].contiguous(), labels=labels, act_lens=d["act_lens"][order].contiguous(),
               label_lens=d["label_lens"][order].contiguous())
    return out


@pytest.mark.parametrize("B,by_length", [(3, False), (16, False), (33, False), (64, False), (64, True), (40, True)])
def test_host_buffer_entry(oracle, b200, B, by_length):
    # ctc_b200_compute_host (the C-ABI entry bench.py's e2e leg times): pinned host buffers in and out, the batch
    # pipelined in utterance slices; must equal the device-resident entry bit for bit.  by_length: the reference's
    # batch order, where later slices are shorter than T and their padding frames never cross the bus.
    import ctypes

    lib = b200.pkg._lib.load()
    d = _wl(b200, "cpu_b16_t500", seed=8, B=B)
    if by_length:
        d = _sorted_by_length(d)
    T, C = d["T"], d["C"]
    h_acts = d["acts"].contiguous().pin_memory()
    h_grads = torch.full_like(d["acts"], float("nan")).pin_memory()
    h_costs = torch.empty(B, dtype=torch.float32).pin_memory()
    ctx = ctypes.c_void_p()
    b200.pkg._lib.check(lib.ctc_b200_create(ctypes.byref(ctx)))
    opts = b200.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(opts))
    loss_sum = ctypes.c_double(0.0)
    hl, hal, hll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()
    for _ in range(2):  # second call reuses the context's buffers and streams
        st = lib.ctc_b200_compute_host(ctx, h_acts.data_ptr(), h_grads.data_ptr(), hl.data_ptr(), hll.data_ptr(),
                                       hal.data_ptr(), T, B, C, h_costs.data_ptr(), ctypes.byref(loss_sum),
                                       ctypes.byref(opts))
        b200.pkg._lib.check(st, "ctc_b200_compute_host")
    lib.ctc_b200_destroy(ctx)
    costs, grads = _run(b200, d)
    assert np.array_equal(h_costs.numpy().astype(np.float64), costs)
    assert np.array_equal(h_grads.numpy().astype(np.float64), grads)
    assert abs(loss_sum.value - costs.sum()) <= 1e-6 * abs(costs.sum())
    if B == 33:
        oc, og = oracle.compute(d["acts"].numpy(), hl.numpy(), hal.numpy(), hll.numpy())
        assert np.abs(h_grads.numpy().astype(np.float64) - og).max() <= GRAD_ATOL


def test_deterministic_headline_and_long(b200, monkeypatch):
    # a race in shared memory (parked rows, posterior rows, in-place unit sums, slot rings) would show up as run-to-run
    # or slicing-dependent bit differences on the shapes with the most warps and the most concurrency
    for name in ("headline_b64_t1500", "long_b32_t4000"):
        d = _wl(b200, name, seed=17)
        c0, g0 = _run(b200, d)
        for _ in range(2):
            c1, g1 = _run(b200, d)
            assert np.array_equal(c0, c1) and np.array_equal(g0, g1), name
        monkeypatch.setenv("CTC_B200_SLICES", "1")
        c2, g2 = _run(b200, d)
        monkeypatch.delenv("CTC_B200_SLICES")
        assert np.array_equal(c0, c2) and np.array_equal(g0, g2), name


def _greedy_check(best, path, ids, offs, act_lens):
    lens = best["lens"].cpu().numpy()
    assert lens.tolist() == [len(i) for i in ids]
    gp, gi, go_ = best["path"].cpu().numpy(), best["ids"].cpu().numpy(), best["offsets"].cpu().numpy()
    for b in range(len(ids)):
        n = gp.shape[1] if act_lens is None else int(act_lens[b])
        assert np.array_equal(gp[b, :n], path[b, :n]), b
        assert gi[b, :lens[b]].tolist() == ids[b] and go_[b, :lens[b]].tolist() == offs[b], b


def test_greedy_decode_matches_reference_goldens(b200):
    # stand-alone entry on the reference decoder's own batch-first input; bit-exact integers
    import glob
    import os
    for f in sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "greedy_*.npz"))):
        z = np.load(f)
        labels = str(z["labels"])
        dec = b200.GreedyDecoder(labels, blank_index=int(z["blank"]))
        probs = torch.from_numpy(z["probs_btc"]).cuda()
        strings, offsets = dec.decode(probs, torch.from_numpy(z["sizes"]))
        assert [s[0] for s in strings] == [str(s) for s in z["strings"]], f
        assert np.array_equal(np.concatenate([o[0].numpy() for o in offsets]), z["offsets"]), f
        # same through the loss layout [T,B,C] (a transposed, non-contiguous view is read in place)
        best = b200.greedy_decode(probs.transpose(0, 1), torch.from_numpy(z["sizes"]), blank=int(z["blank"]))
        assert best["lens"].cpu().numpy().tolist() == z["lens"].tolist()


@pytest.mark.parametrize("name", ["cpu_b16_t500", "headline_b64_t1500"])
def test_fused_decode_is_a_by_product(oracle, b200, name):
    # ctc_b200_compute_decode: identical costs / gradients, plus the greedy decoding of the same logits
    from oracle import greedy_oracle as go
    d = _wl(b200, name, seed=23)
    acts = d["acts"].cuda()
    c0, g0 = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"])
    c1, g1, best = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"], decode=True)
    torch.cuda.synchronize()
    assert torch.equal(c0, c1) and torch.equal(g0, g1)
    path, ids, offs = go.decode(d["acts"].numpy(), d["act_lens"].numpy(), 0)
    _greedy_check(best, path, ids, offs, d["act_lens"].numpy())
    alone = b200.greedy_decode(acts, d["act_lens"], blank=0)
    _greedy_check(alone, path, ids, offs, d["act_lens"].numpy())


def test_fused_decode_large_alphabet(b200):
    from oracle import greedy_oracle as go
    w = b200.workloads.Workload("bigdec", 6, 90, 700, 20, 0.05, 0.2)
    d = b200.workloads.make(w, seed=29)
    _, _, best = b200.ctc_forward_backward(d["acts"].cuda(), d["labels"], d["act_lens"], d["label_lens"], decode=True)
    path, ids, offs = go.decode(d["acts"].numpy(), d["act_lens"].numpy(), 0)
    _greedy_check(best, path, ids, offs, d["act_lens"].numpy())


def test_costs_ready_event(oracle, b200):
    # the event fires when costs[] is final; a consumer on another stream that waits on it reads the right costs
    d = _wl(b200, "cpu_b16_t500", seed=31, B=24)
    acts = d["acts"].cuda()
    ev = torch.cuda.Event()
    side = torch.cuda.Stream()
    costs, grads = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"], costs_ready=ev)
    with torch.cuda.stream(side):
        side.wait_event(ev)
        total = costs.sum()
    costs.record_stream(side)
    side.synchronize()
    oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    assert abs(float(total) - oc.sum()) <= 1e-5 * abs(oc.sum())
    torch.cuda.synchronize()
    assert np.abs(grads.cpu().numpy().astype(np.float64) - og).max() <= GRAD_ATOL


@pytest.mark.parametrize("name,B", [("cpu_b16_t500", 16), ("headline_b64_t1500", 64), ("bigc_b64_t1000_c5000", 3)])
def test_nan_guard_matches_reference_workaround(oracle, b200, name, B):
    # train.py:595-598: `logits[torch.isnan(logits)] = 0` ahead of the criterion.  nan_to_zero folds that into the
    # softmax pass: same costs / gradients as the oracle run on the zeroed logits, NaNs of valid frames counted,
    # the caller's logits left as they were.
    d = _wl(b200, name, seed=41, B=B)
    acts = d["acts"].clone()
    T, _, C = acts.shape
    rng = np.random.default_rng(5)
    al = d["act_lens"].numpy()
    expected = 0
    for _ in range(200):
        t, b, c = int(rng.integers(T)), int(rng.integers(B)), int(rng.integers(C))
        if not torch.isnan(acts[t, b, c]):
            acts[t, b, c] = float("nan")
            expected += int(t < al[b])
    acts[0, 0, :] = float("nan")  # a whole valid frame of NaNs -> uniform row
    expected = int((torch.isnan(acts) & (torch.arange(T)[:, None, None] < d["act_lens"][None, :, None])).sum())
    zeroed = torch.nan_to_num(acts, nan=0.0).numpy()
    dev = acts.cuda()
    count = torch.zeros(1, dtype=torch.int32, device="cuda")
    costs, grads = b200.ctc_forward_backward(dev, d["labels"], d["act_lens"], d["label_lens"], nan_to_zero=True,
                                             nan_count=count)
    torch.cuda.synchronize()
    oc, og = oracle.compute(zeroed, d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    c = costs.cpu().numpy().astype(np.float64)
    assert np.all(np.isfinite(c))
    assert (np.abs(c - oc) / np.maximum(np.abs(oc), 1e-30)).max() <= LOSS_RTOL
    g = grads.cpu().numpy().astype(np.float64)
    assert np.all(np.isfinite(g)) and np.abs(g - og).max() <= GRAD_ATOL
    assert int(count.item()) == expected
    assert torch.equal(torch.isnan(dev).cpu(), torch.isnan(acts))  # input untouched


def test_nan_guard_module(oracle, b200):
    w = b200.workloads.Workload("m", 8, 120, 30, 30, 0.05, 0.2)
    d = b200.workloads.make(w, seed=22)
    acts = d["acts"].clone()
    acts[3, 1, 7] = float("nan")
    acts[10, 5, 0] = float("nan")
    logits = acts.cuda().requires_grad_(True)
    crit = b200.CTCLoss(nan_to_zero=True)
    loss = crit(logits, d["labels"], d["act_lens"], d["label_lens"])
    loss.backward()
    oc, og = oracle.compute(torch.nan_to_num(acts, nan=0.0).numpy(), d["labels"].numpy(), d["act_lens"].numpy(),
                            d["label_lens"].numpy())
    assert abs(loss.item() - oc.sum()) <= 1e-5 * oc.sum()
    assert np.abs(logits.grad.cpu().numpy().astype(np.float64) - og).max() <= GRAD_ATOL
    assert int(crit.nan_count.item()) == 2
    # clean data: nothing counted, same result as the plain criterion
    clean = d["acts"].cuda()
    l1 = crit(clean, d["labels"], d["act_lens"], d["label_lens"])
    l0 = b200.CTCLoss()(clean, d["labels"], d["act_lens"], d["label_lens"])
    assert int(crit.nan_count.item()) == 0 and l1.item() == l0.item()


def test_cuda_graph_capture_of_the_loss_step(oracle, b200):
    # with lengths and labels resident on the device the whole loss step (prep, sliced engine and
    # gradient launches on the library's side streams, forked from / joined to the caller's stream by events) is
    # capturable in a CUDA graph; replaying it on new logits gives the bits of an eager call.
    for name, B in (("cpu_b16_t500", 32), ("bigc_b64_t1000_c5000", 16)):
        d = _wl(b200, name, seed=9, B=B)
        acts = d["acts"].cuda()
        lab, al, ll = d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
        lmax = int(d["label_lens"].max())
        grads = torch.empty_like(acts)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up outside the capture: pools, attributes, workspace
            for _ in range(2):
                b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            costs_g, _ = b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads)
        new_acts = torch.randn(acts.shape, generator=torch.Generator().manual_seed(77)) * 2.0
        acts.copy_(new_acts.cuda())
        grads.zero_()
        graph.replay()
        torch.cuda.synchronize()
        c_graph, g_graph = costs_g.cpu().numpy().copy(), grads.cpu().numpy().copy()
        c_eager, g_eager = b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax)
        torch.cuda.synchronize()
        assert np.array_equal(c_graph, c_eager.cpu().numpy()) and np.array_equal(g_graph, g_eager.cpu().numpy())
        oc, og = oracle.compute(new_acts.numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
        assert np.abs(c_graph - oc).max() <= LOSS_RTOL * np.abs(oc).max()
        assert np.abs(g_graph - og).max() <= GRAD_ATOL


def test_error_counts_match_reference_goldens(b200):
    # WER / CER counts on the device, bit-exact against the reference's get_cer_wer goldens
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "wer_cases.npz"))
    hyp = torch.from_numpy(g["hyp_ids"]).cuda()
    got = b200.error_counts(hyp, torch.from_numpy(g["hyp_lens"]).cuda(), torch.from_numpy(g["ref_flat"]).cuda(),
                            torch.from_numpy(g["ref_lens"]).cuda(), int(g["space"]), int(g["ref_lens"].max()))
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), g["counts"])


def test_error_counts_of_the_fused_decode(b200):
    # the training-loop use: loss + gradient + best path in one call, then error counts of the best path against the
    # same flat targets, everything on the device; checked against the CPU restatement at the headline shape
    # (random logits: hypotheses of ~1400 symbols against references of up to 415)
    from oracle import greedy_oracle, wer_oracle

    d = _wl(b200, "headline_b64_t1500", C=30)
    labels = "".join(b200.ALPHABET)
    dec = b200.GreedyDecoder(labels, blank_index=labels.index("_"))
    costs, grads, best = b200.ctc_forward_backward(d["acts"].cuda(), d["labels"], d["act_lens"], d["label_lens"],
                                                   decode=True)
    counts = dec.error_counts(best, d["labels"], d["label_lens"], int(d["label_lens"].max()))
    torch.cuda.synchronize()
    _, ids, _ = greedy_oracle.decode(d["acts"].numpy(), d["act_lens"].numpy(), blank=0)
    P = max(len(i) for i in ids)
    hyp = np.zeros((len(ids), P), np.int32)
    for b, i in enumerate(ids):
        hyp[b, :len(i)] = i
    want = wer_oracle.batch_counts(hyp, [len(i) for i in ids], d["labels"].numpy(), d["label_lens"].numpy(),
                                   dec.space_index)
    assert np.array_equal(counts.cpu().numpy(), np.array(want, np.int32))
    with pytest.raises(ValueError):
        b200.error_counts(best["ids"].cpu(), best["lens"], d["labels"].cuda(), d["label_lens"].cuda(), 29, 10)


def test_loss_sum_output(oracle, b200):
    # the criterion's one-element return value produced by the library (fixed order, fp64 accumulation), sliced and
    # unsliced, small and large alphabet; the module's loss is that value
    for name, B in (("headline_b64_t1500", 64), ("cpu_b16_t500", 5), ("bigc_b64_t1000_c5000", 8)):
        d = _wl(b200, name, B=B)
        acts = d["acts"].cuda()
        loss = torch.full((1,), float("nan"), device="cuda")
        costs, _ = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"], loss_sum=loss)
        torch.cuda.synchronize()
        want = np.float32(costs.cpu().numpy().astype(np.float64).sum())
        assert loss.item() == want
        a = acts.clone().requires_grad_(True)
        out = b200.CTCLoss()(a, d["labels"], d["act_lens"], d["label_lens"])
        assert out.shape == (1,) and out.item() == want
        out.backward()
        assert a.grad is not None and torch.isfinite(a.grad).all()


def test_fused_loss_all_reduce_two_ranks_one_device(b200):
    # the path's one exchange (reduce_tensor, data/utils.py:40-44) fused into the cost-summing kernel: two logical
    # ranks in this process, each with its own shard, stream and communicator; both must obtain the same global sum
    # (their kernels poll each other's stores, so they have to run concurrently), call after call
    d = _wl(b200, "cpu_b16_t500", seed=11, B=24)
    shards = b200.sharding.length_bucketed_shards(d["act_lens"].tolist(), d["label_lens"].tolist(), 2)
    comms = b200.sharding.PeerLossAllReduce.local_group(2)
    for c in comms:
        c.set_timeout(5000)  # if the two kernels were ever not co-resident, fail in seconds rather than minutes
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    parts = [b200.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], ids) for ids in shards]
    outs = [torch.zeros(1, device="cuda") for _ in range(2)]
    mine = [torch.zeros(1, device="cuda") for _ in range(2)]
    for it in range(3):
        costs = []
        for r in range(2):
            a, l, al, ll = parts[r]
            with torch.cuda.stream(streams[r]):
                c, _ = b200.ctc_forward_backward(a.cuda() * (1.0 + 0.1 * it), l, al, ll, loss_sum=mine[r],
                                                 loss_sum_all=outs[r], comm=comms[r])
                costs.append(c)
        torch.cuda.synchronize()
        if any(c.timed_out for c in comms):
            # two kernels that wait for each other on ONE device are only a stand-in for two GPUs: nothing guarantees
            # that they are resident at the same time
            pytest.skip("the two ranks' kernels were not co-resident on this device")
        local = [np.float32(c.cpu().numpy().astype(np.float64).sum()) for c in costs]
        want = np.float32(np.float64(local[0]) + np.float64(local[1]))
        assert outs[0].item() == outs[1].item() == want, (it, outs[0].item(), outs[1].item(), want)
        assert [m.item() for m in mine] == local
    # (the drop-in module with a communicator -- criterion.loss_all_ranks -- is exercised by the multi-process test
    # below, where every rank has a GPU of its own)
    for c in comms:
        c.close()


def test_fused_loss_all_reduce_timeout_is_reported(b200):
    # a peer that never makes its call (a rank that skipped a step): after the communicator's wall-clock timeout the
    # exchange is given up -- NaN in loss_sum_all, the rank's own results untouched, the sticky flag set -- and the
    # NEXT call that names the communicator fails with a status instead of running the protocol out of step
    d = _wl(b200, "cpu_b16_t500", seed=12, B=8)
    comms = b200.sharding.PeerLossAllReduce.local_group(2)
    comms[0].set_timeout(30)  # milliseconds
    acts = d["acts"].cuda()
    mine = torch.zeros(1, device="cuda")
    tot = torch.zeros(1, device="cuda")
    assert not comms[0].timed_out
    costs, grads = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"], loss_sum=mine,
                                             loss_sum_all=tot, comm=comms[0])  # rank 1 never calls
    torch.cuda.synchronize()
    c0, g0 = b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"])
    torch.cuda.synchronize()
    assert torch.isnan(tot).all() and torch.equal(costs, c0) and torch.equal(grads, g0)
    assert mine.item() == np.float32(c0.cpu().numpy().astype(np.float64).sum())
    assert comms[0].timed_out and not comms[1].timed_out
    with pytest.raises(RuntimeError):
        b200.ctc_forward_backward(acts, d["labels"], d["act_lens"], d["label_lens"], loss_sum_all=tot, comm=comms[0])
    for c in comms:
        c.close()


def test_peer_loss_all_reduce_across_processes():
    # the real thing: one process per GPU, CUDA-IPC slot buffers, the exchange fused into the cost-summing kernel,
    # checked against NCCL on the same per-rank sums (tests/peer_allreduce_worker.py); needs two GPUs
    import subprocess
    import sys

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29731", os.path.join(root, "tests", "peer_allreduce_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert "peer all-reduce worker ok" in res.stdout


def test_graph_replay_survives_workspace_regrowth(oracle, b200):
    # a graph captured at a small shape must not point into the shared eager workspace, which a later, larger eager
    # call regrows (and frees): the capture allocates its own scratch inside the graph's pool
    d = _wl(b200, "cpu_b16_t500", seed=13, B=16)
    acts = d["acts"].cuda()
    lab, al, ll = d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
    lmax = int(d["label_lens"].max())
    grads = torch.empty_like(acts)
    b200.pkg.ctc_loss.clear_workspace_cache()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        costs_g, _ = b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=lmax, out_grads=grads)
    big = _wl(b200, "headline_b64_t1500", seed=14)  # regrows every cached eager workspace ...
    for s in (torch.cuda.current_stream(), side):
        with torch.cuda.stream(s):
            b200.ctc_forward_backward(big["acts"].cuda(), big["labels"], big["act_lens"], big["label_lens"])
    torch.cuda.synchronize()
    junk = torch.full((64 << 20,), 0x7F, dtype=torch.uint8, device="cuda")  # ... and whatever was freed gets reused
    grads.zero_()
    graph.replay()
    torch.cuda.synchronize()
    del junk
    oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    assert np.abs(costs_g.cpu().numpy() - oc).max() <= LOSS_RTOL * np.abs(oc).max()
    assert np.abs(grads.cpu().numpy().astype(np.float64) - og).max() <= GRAD_ATOL


def test_module_backward_scales_in_place(oracle, b200):
    # train.py:600-602,619-620: loss = criterion(...); loss = loss / B; loss.backward().  The binding's backward is
    # cached_grads * grad_output; here the cached gradients are scaled in place by the device scalar (no second
    # [T,B,C] tensor), and a criterion that already carries the 1/B (size_average) leaves them untouched
    w = b200.workloads.Workload("m", 8, 150, 30, 30, 0.05, 0.2)
    d = b200.workloads.make(w, seed=31)
    oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    x = d["acts"].cuda().requires_grad_(True)
    loss = b200.CTCLoss()(x, d["labels"], d["act_lens"], d["label_lens"])
    (loss / w.B).backward(retain_graph=True)
    g1 = x.grad.clone()
    assert np.abs(g1.cpu().numpy().astype(np.float64) * w.B - og).max() <= GRAD_ATOL
    x.grad = None
    (loss * 3.0).backward()  # second trip through the same graph: the first factor is undone
    assert np.abs(x.grad.cpu().numpy().astype(np.float64) / 3.0 - og).max() <= GRAD_ATOL
    x2 = d["acts"].cuda().requires_grad_(True)
    l2 = b200.CTCLoss(size_average=True)(x2, d["labels"], d["act_lens"], d["label_lens"])
    l2.backward()
    assert abs(l2.item() * w.B - oc.sum()) <= LOSS_RTOL * oc.sum()
    assert torch.equal(x2.grad, g1) or np.abs((x2.grad - g1).cpu().numpy()).max() <= 1e-9
    # the entry itself: a factor of exactly 1 leaves the bits alone, anything else multiplies
    g = torch.randn(1000003, device="cuda")
    ref = g.clone()
    b200.pkg.ctc_loss.scale_grads_(g, torch.ones(1, device="cuda"))
    assert torch.equal(g, ref)
    b200.pkg.ctc_loss.scale_grads_(g[1:], torch.full((1,), 0.25, device="cuda"))  # unaligned base
    torch.cuda.synchronize()
    assert torch.equal(g[1:], ref[1:] * 0.25) and g[0] == ref[0]


@pytest.mark.parametrize("C", [30, 200])
def test_module_predicts_grad_output_across_steps(oracle, b200, C):
    # the criterion folds the grad_output it saw at the previous backward() (train.py:602: loss / B, every step) into the
    # kernels' gradient scale; backward() then multiplies by grad_output / predicted on the device: exact whether the
    # prediction holds (steps 2, 3: nothing is touched), is absent (step 1) or is wrong (step 4: another divisor)
    w = b200.workloads.Workload("m", 6, 120, C, 25, 0.05, 0.2)
    crit = b200.CTCLoss()
    for step, div in enumerate((6.0, 6.0, 6.0, 4.0, 4.0)):
        d = b200.workloads.make(w, seed=60 + step)
        oc, og = oracle.compute(d["acts"].numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
        x = d["acts"].cuda().requires_grad_(True)
        loss = crit(x, d["labels"], d["act_lens"], d["label_lens"])
        (loss / div).backward()
        torch.cuda.synchronize()
        assert abs(loss.item() - oc.sum()) <= LOSS_RTOL * oc.sum()  # the returned loss never carries the prediction
        assert np.abs(x.grad.cpu().numpy().astype(np.float64) * div - og).max() <= GRAD_ATOL, (step, div)
    assert crit._go_pred == 0.25  # learnt from the observations, without a host sync in the step itself
    plain = b200.CTCLoss(predict_grad_output=False)
    x = d["acts"].cuda().requires_grad_(True)
    (plain(x, d["labels"], d["act_lens"], d["label_lens"]) / 4.0).backward()
    assert plain._go_pred == 1.0 and np.abs(x.grad.cpu().numpy().astype(np.float64) * 4.0 - og).max() <= GRAD_ATOL


def test_gradient_warp_classes_match_single_launch(oracle, b200, monkeypatch):
    # the gradient kernel's two warp classes (own launch with half-size CTAs for utterances that need at most half of
    # the warps; default for B >= 256) against one launch for all, on label lengths at the class boundaries
    # (112 labels = 225 states = the first that need a second warp, 223 / 224 the third), incl. empty and infeasible
    B, T, C = 12, 500, 29
    lens = [0, 1, 111, 112, 113, 200, 223, 224, 225, 300, 335, 336]
    g = torch.Generator().manual_seed(808)
    acts = torch.randn(T, B, C, generator=g)
    act_lens = torch.tensor([T - 7 * i for i in range(B)], dtype=torch.int32)
    act_lens[9] = 200  # 300 labels in 200 frames: infeasible
    labels = torch.cat([torch.randint(1, C, (n,), generator=g, dtype=torch.int32) for n in lens])
    label_lens = torch.tensor(lens, dtype=torch.int32)
    out = {}
    for cls in ("0", "1"):
        monkeypatch.setenv("CTC_B200_GRAD_CLASSES", cls)
        for slices in ("1", "4"):
            monkeypatch.setenv("CTC_B200_SLICES", slices)
            c, gr = b200.ctc_forward_backward(acts.cuda(), labels, act_lens, label_lens)
            torch.cuda.synchronize()
            out[cls, slices] = (c.cpu(), gr.cpu())
    oc, og = oracle.compute(acts.numpy(), labels.numpy(), act_lens.numpy(), label_lens.numpy())
    for key, (c, gr) in out.items():
        assert torch.equal(c, out["0", "1"][0]) and torch.equal(gr, out["0", "1"][1]), key  # bit-identical
    c, gr = out["1", "4"]
    fin = np.isfinite(oc)
    assert (np.abs(c.numpy()[fin] - oc[fin]) / np.maximum(np.abs(oc[fin]), 1e-30)).max() <= LOSS_RTOL
    assert np.abs(gr.numpy().astype(np.float64) - og).max() <= GRAD_ATOL


def test_zero_frames_with_a_sliceable_batch(b200):
    # T == 0 skips the prep launch (which writes the length ranking of a sliced launch): the engines must then run
    # unsliced.  Straight through the C ABI, since an empty torch tensor has no data pointer.
    import ctypes

    lib = b200.pkg._lib.load()
    B, C = 32, 5
    opts = b200.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(opts))
    opts.stream = torch.cuda.current_stream().cuda_stream
    n = ctypes.c_size_t(0)
    b200.pkg._lib.check(lib.ctc_b200_workspace_size(0, B, C, 3, ctypes.byref(opts), ctypes.byref(n)))
    ws = torch.empty(n.value + 256, dtype=torch.uint8, device="cuda")
    ws_ptr = (ws.data_ptr() + 255) & ~255
    dummy = torch.zeros(16, device="cuda")
    costs = torch.full((B,), -1.0, device="cuda")
    ll = torch.tensor([0, 2] * (B // 2), dtype=torch.int32)
    labels = torch.tensor([1, 2] * (B // 2), dtype=torch.int32)
    al = torch.zeros(B, dtype=torch.int32)
    for zi, want in ((1, 0.0), (0, float("inf"))):
        opts.zero_infinity = zi
        st = lib.ctc_b200_compute(dummy.data_ptr(), dummy.data_ptr(), labels.data_ptr(), ll.data_ptr(), al.data_ptr(),
                                  0, B, C, costs.data_ptr(), ws_ptr, n.value, ctypes.byref(opts))
        b200.pkg._lib.check(st)
        torch.cuda.synchronize()
        c = costs.cpu().numpy()
        assert np.all(c[0::2] == 0.0) and np.all(c[1::2] == want)


def test_cluster_engine_matches_single_cta_engine(oracle, b200, monkeypatch):
    # more than 4 lattice warps (L > 511): the recursion runs as 2-CTA clusters whose boundary crosses through
    # distributed shared memory (ctc_engine2c_kernel); it must reproduce the one-CTA engine bit for bit, in both
    # directions, for utterances that need 5..8 lattice warps as well as short ones that leave the second CTA idle,
    # ragged last groups included
    T, C = 2300, 29
    g = torch.Generator().manual_seed(101)
    lens = [(2300, 1023), (2299, 900), (2041, 770), (1800, 640), (1700, 513), (1500, 512), (900, 300), (77, 20),
            (1300, 600), (8, 2), (9, 3), (1, 1)]
    B = len(lens)
    acts = torch.randn(T, B, C, generator=g)
    act_lens = torch.tensor([t for t, _ in lens], dtype=torch.int32)
    ll = torch.tensor([l for _, l in lens], dtype=torch.int32)
    labels = torch.randint(1, C, (int(ll.sum()),), generator=g, dtype=torch.int32)
    # no repeats, so that every utterance is feasible (L + repeats <= T)
    for i in range(1, labels.numel()):
        if labels[i] == labels[i - 1]:
            labels[i] = 1 + (int(labels[i]) % (C - 1))
    d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
    monkeypatch.setenv("CTC_B200_CLUSTER", "0")
    c0, g0 = _run(b200, d)
    monkeypatch.setenv("CTC_B200_CLUSTER", "1")
    c1, g1 = _check(oracle, b200, d)
    assert np.array_equal(c0, c1) and np.array_equal(g0, g1)
    ce, _ = _run(b200, d, want_grads=False)  # score-only launch: alpha clusters only
    assert np.array_equal(ce, c1)


def test_two_part_recursion_is_bit_identical(oracle, b200, monkeypatch):
    # the recursion kernels can run in two launches (3/4 + 1/4 of each direction's groups, per-thread state handed over
    # in full fp64; an off-by-default switch, see make_plan): costs and gradients must not change by a bit
    for name, kw in (("headline_b64_t1500", {}), ("cpu_b16_t500", {"B": 40, "seed": 3})):
        d = _wl(b200, name, **kw)
        d["act_lens"][-1] = 57   # fewer than 8 groups: everything in part 1
        d["act_lens"][-2] = 64   # exactly 8 groups
        d["act_lens"][-3] = 129  # ragged last group
        d["label_lens"][-3:] = torch.tensor([20, 9, 7], dtype=torch.int32)
        d["labels"] = d["labels"][: int(d["label_lens"].sum())]
        monkeypatch.setenv("CTC_B200_TWO_PART", "0")
        c0, g0 = _run(b200, d)
        monkeypatch.setenv("CTC_B200_TWO_PART", "2")
        c2, g2 = _run(b200, d)
        assert np.array_equal(c0, c2) and np.array_equal(g0, g2), name
        monkeypatch.delenv("CTC_B200_TWO_PART")
        _check(oracle, b200, d)


@pytest.mark.parametrize("C,Lmax", [(700, 300), (5000, 400), (1000, 511)])
def test_large_alphabet_long_labels(oracle, b200, C, Lmax):
    # C > 128 with label sequences beyond 255 symbols (a BPE target at T = 1000 easily is): emission rows indexed by
    # label position at pitch 512, positions derived from the lattice thread index.  Ragged batch: one utterance at the
    # limit, short ones that stay below 256, repeats, an infeasible one.
    T, B = 1100, 6
    g = torch.Generator().manual_seed(C + Lmax)
    acts = torch.randn(T, B, C, generator=g) * 1.5
    ll = torch.tensor([Lmax, 256, 255, 12, Lmax - 37, 600 if Lmax > 500 else 300], dtype=torch.int32).clamp(max=Lmax)
    act_lens = torch.tensor([T, T - 3, 900, 40, 2 * (Lmax - 37) + 5, int(ll[5]) - 1], dtype=torch.int32)  # last: L > T
    labels = torch.randint(1, C, (int(ll.sum()),), generator=g, dtype=torch.int32)
    labels[5:9] = labels[4]  # a run of repeats inside the long utterance
    d = {"acts": acts, "labels": labels, "act_lens": act_lens, "label_lens": ll}
    costs, grads = _check(oracle, b200, d)
    assert costs[5] == 0.0 and np.all(grads[:, 5] == 0)
    _check(oracle, b200, d, device_lens=True, grad_scale=0.5)
    ce, _ = _run(b200, d, want_grads=False)
    assert np.array_equal(ce, costs)
