"""CPU tests of the host side: C-ABI exports and argument checking that needs no device, alphabet and target
packing (rows a7-a9 of SURVEY.md 8(a)), sharding."""
import ctypes
import json
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_exports_every_declared_symbol(b200):
    b200.build()
    lib = b200.pkg._lib.load()
    hdr = open(os.path.join(ROOT, "include", "ctc_b200.h")).read()
    declared = set(re.findall(r"\b(ctc_b200_[a-z_]+)\s*\(", hdr))
    assert declared == set(b200.pkg._lib.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.ctc_b200_get_version() >= 100
    assert lib.ctc_b200_status_string(2) == b"invalid value"
    assert lib.ctc_b200_status_string(0) == b"no error"


def test_options_struct_layout_and_defaults(b200):
    lib = b200.pkg._lib.load()
    o = b200.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(o))
    assert (o.blank, o.zero_infinity, o.grad_scale, o.lengths_on_device) == (0, 1, 1.0, 0)
    assert ctypes.sizeof(o) == 88  # 8 + 7 * 4 (+4 pad) + 5 pointers + emissions_ready (+4 pad)


def test_workspace_query_is_shape_only(b200):
    lib = b200.pkg._lib.load()
    n = ctypes.c_size_t(0)
    assert lib.ctc_b200_workspace_size(1500, 64, 29, 450, None, ctypes.byref(n)) == 0 and n.value > 0
    small = n.value
    assert lib.ctc_b200_workspace_size(1500, 128, 29, 450, None, ctypes.byref(n)) == 0 and n.value > small
    assert lib.ctc_b200_workspace_size(4000, 32, 29, 600, None, ctypes.byref(n)) == 0    # long targets
    assert lib.ctc_b200_workspace_size(1000, 64, 5000, 200, None, ctypes.byref(n)) == 0   # large alphabet
    assert lib.ctc_b200_workspace_size(1000, 64, 5000, 300, None, ctypes.byref(n)) == 0   # pitch-512 emission rows
    assert lib.ctc_b200_workspace_size(1000, 64, 5000, 512, None, ctypes.byref(n)) == 2   # ... needs L <= 511
    assert lib.ctc_b200_workspace_size(10, 0, 29, 5, None, ctypes.byref(n)) == 2          # B < 1
    assert lib.ctc_b200_workspace_size(10, 1, 1, 5, None, ctypes.byref(n)) == 2           # C < 2
    assert lib.ctc_b200_workspace_size(10, 1, 29, 5, None, None) == 2
    assert lib.ctc_b200_workspace_size(0, 16, 29, 5, None, ctypes.byref(n)) == 0          # T == 0 is a valid shape


def test_compute_rejects_bad_arguments_before_touching_the_device(b200):
    lib = b200.pkg._lib.load()
    assert lib.ctc_b200_compute(None, None, None, None, None, 10, 1, 5, None, None, 0, None) == 2


def test_no_cpu_fallback(b200):
    a = torch.randn(10, 2, 5)
    with pytest.raises(ValueError):
        b200.ctc_forward_backward(a, torch.tensor([1], dtype=torch.int32), torch.tensor([10, 10], dtype=torch.int32),
                                  torch.tensor([1, 0], dtype=torch.int32))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "deepspeech.pytorch_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"import\s+oracle|from\s+oracle|ctc_oracle|libctc_oracle|oracle/_ref", src), f


def test_alphabet_matches_labels_json(b200):
    # the fork's labels.json (30 symbols, blank "_" at 0), restated in packing.ALPHABET
    assert len(b200.ALPHABET) == 30 and b200.ALPHABET[0] == "_" and b200.ALPHABET[28] == "2" and b200.ALPHABET[29] == " "
    ref = "/root/reference/labels.json"
    if os.path.exists(ref):  # only in the build container
        assert tuple(json.load(open(ref))) == b200.ALPHABET


def test_labels_doubled_letter_rule_and_round_trip(b200):
    lab = b200.Labels()
    ids = lab.encode("hello all")
    assert lab.blank == 0 and 0 not in ids
    assert [b200.ALPHABET[i] for i in ids] == list("HEL2O AL2")
    assert lab.decode(ids) == "HELLO ALL"
    assert all(a != b for a, b in zip(ids, ids[1:]))  # no consecutive repeats survive the '2' rule


def test_pack_targets_layout(b200):
    flat, sizes = b200.pack_targets([[1, 2, 3], [], [4, 4]])
    assert flat.dtype == torch.int32 and sizes.dtype == torch.int32
    assert flat.tolist() == [1, 2, 3, 4, 4] and sizes.tolist() == [3, 0, 2]
    assert b200.packing.unpack_targets(flat, sizes) == [[1, 2, 3], [], [4, 4]]


def test_conv_out_lens_formula(b200):
    t_in = torch.tensor([1, 2, 3, 100, 101, 3001], dtype=torch.int32)
    assert b200.packing.conv_out_lens(t_in).tolist() == [(int(t) - 1) // 2 + 1 for t in t_in]
    assert b200.packing.input_sizes_from_percentages(torch.tensor([1.0, 0.5]), 301).tolist() == [301, 150]


def test_conv_out_lens_matches_the_reference_get_seq_lens(b200):
    # row a7 pinned to the reference itself: DeepSpeech.get_seq_lens (/root/reference/model.py:382-393) imports in
    # the build container (pure torch); it only walks self.conv, so a small model is enough
    if not os.path.exists("/root/reference/model.py"):
        pytest.skip("reference tree only exists in the build container")
    import importlib.util
    import sys

    spec = importlib.util.spec_from_file_location("ref_model", "/root/reference/model.py")
    ref = importlib.util.module_from_spec(spec)
    sys.path.insert(0, "/root/reference")
    try:
        spec.loader.exec_module(ref)
    finally:
        sys.path.remove("/root/reference")
    model = ref.DeepSpeech(rnn_type="gru", labels="".join(b200.ALPHABET), rnn_hidden_size=16, nb_layers=1)
    t_in = torch.arange(1, 4001, dtype=torch.int32)
    want = model.get_seq_lens(t_in)
    assert torch.equal(b200.packing.conv_out_lens(t_in), want.to(torch.int32))
    assert torch.equal(b200.packing.conv_out_lens(t_in, true_division=True), want.to(torch.int32))
    # a stack with two strided time axes: the reference's expression as torch evaluates it today (true division)
    model.conv.seq_module[3] = torch.nn.Conv2d(32, 32, kernel_size=(21, 11), stride=(2, 2), padding=(10, 5))
    want2 = model.get_seq_lens(t_in)
    got2 = b200.packing.conv_out_lens(t_in, convs=((11, 2, 5, 1), (11, 2, 5, 1)), true_division=True)
    assert torch.equal(got2, want2.to(torch.int32))


def test_length_bucketed_shards(b200):
    d = b200.workloads.make(b200.workloads.CONFIGS["sweep_b1024_t1500"], acts=False)
    al, ll = d["act_lens"].tolist(), d["label_lens"].tolist()
    for world in (2, 4, 8):
        shards = b200.sharding.length_bucketed_shards(al, ll, world)
        assert sorted(i for s in shards for i in s) == list(range(1024))
        assert {len(s) for s in shards} == {1024 // world}
        w = b200.sharding.lattice_work(al, ll)
        loads = [sum(w[i] for i in s) for s in shards]
        assert max(loads) / min(loads) < 1.01
        assert all(w[s[i]] >= w[s[i + 1]] for s in shards for i in range(len(s) - 1))  # longest first


def test_take_shard_slices_labels(b200):
    w = b200.workloads.Workload("t", 6, 20, 5, 6, 0.1, 0.3)
    d = b200.workloads.make(w, seed=2)
    ids = [4, 1]
    a, lab, al, ll = b200.sharding.take_shard(d["acts"], d["labels"], d["act_lens"], d["label_lens"], ids)
    tg = b200.packing.unpack_targets(d["labels"], d["label_lens"])
    assert a.shape == (20, 2, 5) and torch.equal(a[:, 0], d["acts"][:, 4])
    assert b200.packing.unpack_targets(lab, ll) == [tg[4], tg[1]]
    assert al.tolist() == [int(d["act_lens"][4]), int(d["act_lens"][1])]


def test_workload_generator_shapes(b200):
    for name, w in b200.workloads.CONFIGS.items():
        d = b200.workloads.make(w, acts=False)
        assert int(d["act_lens"][0]) == w.T and int(d["act_lens"].max()) == w.T and int(d["act_lens"].min()) >= 1
        assert int(d["label_lens"].max()) <= w.L_max and int(d["labels"].min()) >= 1 and int(d["labels"].max()) < w.C
    assert b200.workloads.algorithmic_bytes(1500, 64, 29) == 8 * 29 * 96000 + 4 * 128 + 4 * 64
