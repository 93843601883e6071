"""ctc_b200_compute_static -- the allocation-free entry point (no counterpart in the reference):
same results as ctc_b200_compute with the scratch in a file-scope static __device__ array.

Collected last on purpose.  Everything the two timed calls need is allocated BEFORE
them; the calls themselves go straight through the C ABI with raw pointers.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5  # tests/test_gpu_parity.py
GRAD_ATOL = 1e-5


def _opts(b200, stream):
    lib = b200.pkg._lib.load()
    o = b200.pkg._lib.CtcOptions()
    lib.ctc_b200_default_options(ctypes.byref(o))
    o.stream = stream.cuda_stream
    return o


def test_static_scratch_entry_matches_compute_twice_in_a_row(oracle, b200):
    lib = b200.pkg._lib.load()
    stream = torch.cuda.current_stream()
    # two different batches of BASELINE config 1's shape (B=16, T=500, C=29), back to back: the second call finds the
    # first call's contents in the shared array and must not be affected by them
    batches = [b200.workloads.make(b200.workloads.CONFIGS["cpu_b16_t500"], seed=s, sigma=sg)
               for s, sg in ((51, 1.0), (52, 3.0))]
    prepared = []
    for d in batches:  # every allocation happens here, ahead of the calls under test
        acts = d["acts"].cuda()
        lab, al, ll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()
        prepared.append((d, acts, torch.full_like(acts, float("nan")), torch.full((d["B"],), float("nan"), device="cuda"),
                         lab, al, ll))
    torch.cuda.synchronize()
    opts = _opts(b200, stream)
    for d, acts, grads, costs, lab, al, ll in prepared:
        st = lib.ctc_b200_compute_static(acts.data_ptr(), grads.data_ptr(), lab.data_ptr(), ll.data_ptr(),
                                         al.data_ptr(), d["T"], d["B"], d["C"], costs.data_ptr(), ctypes.byref(opts))
        assert st == 0, lib.ctc_b200_status_string(st)
    torch.cuda.synchronize()
    for d, acts, grads, costs, lab, al, ll in prepared:
        c_ref, g_ref = b200.ctc_forward_backward(acts, lab, al, ll)  # the existing entry (caller-owned workspace)
        torch.cuda.synchronize()
        c, cr = costs.cpu().numpy().astype(np.float64), c_ref.cpu().numpy().astype(np.float64)
        assert np.all(np.isfinite(c))
        assert (np.abs(c - cr) / np.abs(cr)).max() <= LOSS_RTOL
        assert np.abs(grads.cpu().numpy().astype(np.float64) - g_ref.cpu().numpy().astype(np.float64)).max() <= GRAD_ATOL
        oc, og = oracle.compute(d["acts"].numpy(), lab.numpy(), al.numpy(), ll.numpy())
        assert (np.abs(c - oc) / np.abs(oc)).max() <= LOSS_RTOL
        assert np.abs(grads.cpu().numpy().astype(np.float64) - og).max() <= GRAD_ATOL


def test_static_scratch_entry_score_only_and_device_lengths(b200):
    lib = b200.pkg._lib.load()
    d = b200.workloads.make(b200.workloads.Workload("s", 8, 200, 700, 40, 0.05, 0.2), seed=53)  # large-alphabet path
    acts = d["acts"].cuda()
    lab, al, ll = d["labels"].cuda(), d["act_lens"].cuda(), d["label_lens"].cuda()
    costs = torch.full((d["B"],), float("nan"), device="cuda")
    opts = _opts(b200, torch.cuda.current_stream())
    opts.lengths_on_device = 1
    opts.max_label_len = int(d["label_lens"].max())
    st = lib.ctc_b200_compute_static(acts.data_ptr(), None, lab.data_ptr(), ll.data_ptr(), al.data_ptr(), d["T"],
                                     d["B"], d["C"], costs.data_ptr(), ctypes.byref(opts))
    assert st == 0
    c_ref, _ = b200.ctc_forward_backward(acts, lab, al, ll, max_label_len=opts.max_label_len, want_grads=False)
    torch.cuda.synchronize()
    assert (torch.abs(costs - c_ref) / torch.abs(c_ref)).max().item() <= LOSS_RTOL


def test_static_scratch_entry_rejects_shapes_beyond_the_limit(b200, capfd):
    lib = b200.pkg._lib.load()
    d = b200.workloads.make(b200.workloads.CONFIGS["headline_b64_t1500"], acts=False)  # needs 129 MiB of scratch
    dummy = torch.zeros(64, device="cuda")
    lab, al, ll = d["labels"].contiguous(), d["act_lens"].contiguous(), d["label_lens"].contiguous()
    opts = _opts(b200, torch.cuda.current_stream())
    st = lib.ctc_b200_compute_static(dummy.data_ptr(), dummy.data_ptr(), lab.data_ptr(), ll.data_ptr(), al.data_ptr(),
                                     d["T"], d["B"], d["C"], dummy.data_ptr(), ctypes.byref(opts))
    assert st == 2  # CTC_B200_STATUS_INVALID_VALUE, nothing enqueued
    assert "CTC_B200_STATIC_WORKSPACE_BYTES" in capfd.readouterr().err
    torch.cuda.synchronize()
