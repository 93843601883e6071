"""GPU parity of the logits producer (BatchNorm1d + Linear of /root/reference/model.py:335-341
as one TMA + tcgen05 kernel) against goldens generated from the reference's own SequenceWise / nn.Sequential
(tests/golden/make_golden_fc.py), and of the fused hand-over of its softmax rows to the CTC kernels.

Tolerance: logits abs <= 3e-5 against the reference evaluated in float64.  (The reference runs the layer in fp32; its
own fp32 result is ~1.5e-6 away from fp64 at H = 800.  A single TF32 pass would be 2e-3 away; the split-TF32 product
used here measures 2e-6 .. 1e-5, the remainder being the tensor core's fp32 accumulation.)
"""
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LOGITS_ATOL = 3e-5
LOSS_RTOL = 1e-5
GRAD_ATOL = 1e-5
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "fc_*.npz")))


def _layer(z):
    return {k: torch.from_numpy(z[k]).cuda() for k in ("weight", "bn_weight", "bn_bias", "running_mean", "running_var")}


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_fc_matches_reference_goldens(b200, path):
    z = np.load(path)
    p = _layer(z)
    x = torch.from_numpy(z["x"]).cuda()
    train = bool(z["train"])
    out = b200.fc.fc_forward(x, p["weight"], p["bn_weight"], p["bn_bias"], p["running_mean"], p["running_var"],
                             float(z["eps"]), training=train, want_path=True, return_batch_stats=True)
    torch.cuda.synchronize()
    got = out["logits"].cpu().numpy().astype(np.float64)
    assert got.shape == z["logits"].shape
    assert np.abs(got - z["logits"]).max() <= LOGITS_ATOL
    assert np.array_equal(out["path"].cpu().numpy().T, got.argmax(-1))  # first maximum, like torch.max
    if train:
        assert np.abs(out["batch_mean"].cpu().numpy() - z["batch_mean"]).max() <= 1e-6
        assert np.abs(out["batch_var"].cpu().numpy() - z["batch_var"]).max() <= 1e-5
    # second call, same inputs: bit-identical (fixed summation orders everywhere)
    again = b200.fc.fc_forward(x, p["weight"], p["bn_weight"], p["bn_bias"], p["running_mean"], p["running_var"],
                               float(z["eps"]), training=train)
    assert torch.equal(again["logits"], out["logits"])


def test_fc_module_wraps_the_reference_layer(b200):
    # the drop-in module holds the reference layer's own BatchNorm1d / Linear modules (model.py:335-341 builds them so)
    H, C = 160, 30
    torch.manual_seed(3)
    fully_connected = torch.nn.Sequential(torch.nn.BatchNorm1d(H), torch.nn.Linear(H, C, bias=False)).cuda()
    with torch.no_grad():
        fully_connected[0].running_mean.normal_()
        fully_connected[0].running_var.uniform_(0.2, 2.0)
        fully_connected[0].weight.uniform_(0.5, 1.5)
        fully_connected[0].bias.normal_()
    fused = b200.fc.SequenceWiseBNLinear(fully_connected[0], fully_connected[1])
    x = torch.randn(50, 7, H, device="cuda")
    for mode in (False, True):
        fully_connected.train(mode)
        with torch.no_grad():
            stats = (fully_connected[0].running_mean.clone(), fully_connected[0].running_var.clone())
            want = fully_connected.double()(x.double().view(-1, H)).view(50, 7, C)
            fully_connected.float()
            fully_connected[0].running_mean.copy_(stats[0])  # the training forward updated them; put them back
            fully_connected[0].running_var.copy_(stats[1])
            got = fused(x)
        assert (got.double() - want).abs().max().item() <= LOGITS_ATOL
    with pytest.raises(RuntimeError):
        fused(x.clone().requires_grad_(True))  # forward only
    with pytest.raises(ValueError):
        b200.fc.fc_forward(x.cpu(), fully_connected[1].weight)  # no CPU fallback


def test_fc_full_size_and_fused_emission_rows(oracle, b200):
    # BASELINE's headline frame count through the layer's real size (T=1500, B=64, H=800, C=30), then the CTC loss
    # computed (a) the reference way, from the logits, and (b) from the emission rows the fc kernel wrote into the CTC
    # workspace, with the criterion's softmax pass skipped
    w = b200.workloads.CONFIGS["headline_b64_t1500"]
    d = b200.workloads.make(w, C=30, acts=False)
    T, B, H, C = w.T, w.B, 800, 30
    g = torch.Generator().manual_seed(77)
    x = torch.randn(T, B, H, generator=g).cuda()
    weight = ((torch.rand(C, H, generator=g) * 2 - 1) * 3.0 / H ** 0.5).cuda()
    gamma, beta = (torch.rand(H, generator=g) + 0.5).cuda(), (torch.randn(H, generator=g) * 0.2).cuda()
    mean, var = (torch.randn(H, generator=g) * 0.3).cuda(), (torch.rand(H, generator=g) + 0.5).cuda()
    lmax = int(d["label_lens"].max())
    ws = b200.fc.ctc_workspace_for(T, B, C, lmax, "cuda")
    al_dev = d["act_lens"].cuda()
    out = b200.fc.fc_forward(x, weight, gamma, beta, mean, var, 1e-5, ctc_workspace=ws, act_lens=al_dev, want_path=True)
    logits = out["logits"]
    with torch.no_grad():  # the reference's arithmetic in float64 on the device
        xn = (x.double() - mean.double()) / torch.sqrt(var.double() + 1e-5) * gamma.double() + beta.double()
        want = (xn.view(-1, H) @ weight.double().t()).view(T, B, C)
    assert (logits.double() - want).abs().max().item() <= LOGITS_ATOL
    del xn, want
    lab, ll = d["labels"].cuda(), d["label_lens"].cuda()
    c_fused, g_fused = b200.ctc_forward_backward(logits, lab, al_dev, ll, max_label_len=lmax, workspace=ws,
                                                 emissions_ready=True)
    c_plain, g_plain = b200.ctc_forward_backward(logits, lab, al_dev, ll, max_label_len=lmax)
    torch.cuda.synchronize()
    oc, og = oracle.compute(logits.cpu().numpy(), d["labels"].numpy(), d["act_lens"].numpy(), d["label_lens"].numpy())
    for c, gr in ((c_fused, g_fused), (c_plain, g_plain)):
        cc = c.cpu().numpy().astype(np.float64)
        assert (np.abs(cc - oc) / np.abs(oc)).max() <= LOSS_RTOL
        assert np.abs(gr.cpu().numpy().astype(np.float64) - og).max() <= GRAD_ATOL
    al = d["act_lens"].numpy()
    for b in range(0, B, 9):  # padding frames: exactly zero gradient on the fused path too
        assert bool((g_fused[int(al[b]):, b, :] == 0).all())
    # best path by-product == argmax of the logits on valid frames
    pth = out["path"].cpu().numpy()
    am = logits.argmax(-1).cpu().numpy()
    for b in range(0, B, 5):
        assert np.array_equal(pth[b, :al[b]], am[:al[b], b])


def test_fc_rejects_unsupported_shapes(b200):
    x = torch.randn(4, 2, 30, device="cuda")  # H not a multiple of 4
    with pytest.raises(RuntimeError):
        b200.fc.fc_forward(x, torch.randn(5, 30, device="cuda"))
    x = torch.randn(4, 2, 32, device="cuda")
    with pytest.raises(RuntimeError):
        b200.fc.fc_forward(x, torch.randn(65, 32, device="cuda"))  # C > 64


def test_fc_ts_form_matches_goldens():
    # the alternative form of the kernel (A operands through TMEM: tcgen05.st by the converters, TS-form tcgen05.mma),
    # selected per process by CTC_FC_TS=1: same goldens; one accumulator per tile half, so its error is ~2x the default's
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, CTC_FC_TS="1")
    res = subprocess.run([sys.executable, os.path.join(root, "tools", "fc_check.py")], capture_output=True, text=True,
                         env=env, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    errs = [float(line.split("max abs err")[1].split()[0]) for line in res.stdout.splitlines() if "max abs err" in line]
    assert len(errs) == len(GOLD) and max(errs) <= LOGITS_ATOL, res.stdout
    assert "BAD" not in res.stdout
