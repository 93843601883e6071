"""tests/golden/make_golden_fc.py -- golden vectors of the logits producer, generated FROM THE REFERENCE ITSELF.

The reference's final layer (/root/reference/model.py:335-341)
    fully_connected = nn.Sequential(nn.BatchNorm1d(rnn_hidden_size), nn.Linear(rnn_hidden_size, num_classes, bias=False))
    self.fc = nn.Sequential(SequenceWise(fully_connected))
is pure torch and imports in the build container, so the fixtures below are outputs of the reference's own
`SequenceWise` class wrapped around exactly that Sequential, evaluated on the CPU in float64 (parameters and input
are float32 values, widened exactly), in both BatchNorm modes:
    eval   running statistics (what check_model_quality / test.py run under model.eval())
    train  statistics of the batch itself (what the training forward uses, model.py:373 under model.train())
Each .npz holds the fp32 input [T,B,H], the fp32 parameters, and the fp64 logits [T,B,C] (+ the batch mean / biased
variance for the train case).  Usage:  python tests/golden/make_golden_fc.py   (run once; writes tests/golden/fc_*.npz)
"""
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_model():
    spec = importlib.util.spec_from_file_location("ref_model", "/root/reference/model.py")
    ref = importlib.util.module_from_spec(spec)
    sys.path.insert(0, "/root/reference")
    try:
        spec.loader.exec_module(ref)
    finally:
        sys.path.remove("/root/reference")
    return ref


def case(ref, name, T, B, H, C, seed, train, x_scale=1.0):
    g = torch.Generator().manual_seed(seed)
    fully_connected = nn.Sequential(nn.BatchNorm1d(H), nn.Linear(H, C, bias=False))  # model.py:335-338
    fc = nn.Sequential(ref.SequenceWise(fully_connected))                            # model.py:339-341
    bn, lin = fully_connected[0], fully_connected[1]
    with torch.no_grad():  # non-trivial affine parameters and running statistics (a trained model's are)
        bn.weight.copy_(torch.rand(H, generator=g) * 1.5 + 0.25)
        bn.bias.copy_(torch.randn(H, generator=g) * 0.3)
        bn.running_mean.copy_(torch.randn(H, generator=g) * 0.5)
        bn.running_var.copy_(torch.rand(H, generator=g) * 2.0 + 0.1)
        lin.weight.copy_((torch.rand(C, H, generator=g) * 2 - 1) / H ** 0.5)
    x = (torch.randn(T, B, H, generator=g) * x_scale + torch.randn(H, generator=g) * 0.4).float()
    params = {k: v.detach().clone().numpy() for k, v in (("bn_weight", bn.weight), ("bn_bias", bn.bias),
                                                          ("running_mean", bn.running_mean),
                                                          ("running_var", bn.running_var), ("weight", lin.weight))}
    fc = fc.double()
    fc.train(train)
    with torch.no_grad():
        logits = fc(x.double())
    assert logits.shape == (T, B, C)
    extra = {}
    if train:
        flat = x.double().view(T * B, H)
        extra = {"batch_mean": flat.mean(0).numpy(), "batch_var": flat.var(0, unbiased=False).numpy()}
    np.savez_compressed(os.path.join(HERE, name + ".npz"), x=x.numpy(), logits=logits.numpy(), eps=np.float64(bn.eps),
                        train=np.int32(train), **params, **extra)
    print(name, "logits absmax", float(logits.abs().max()), "shape", tuple(logits.shape))


if __name__ == "__main__":
    torch.set_num_threads(4)
    ref = load_reference_model()
    # the reference's own layer sizes: H = 800 (rnn_hidden_size default, train.py:45), C = 30 (labels.json)
    case(ref, "fc_eval_h800_c30", 13, 5, 800, 30, seed=11, train=False)
    case(ref, "fc_train_h800_c30", 11, 6, 800, 30, seed=12, train=True, x_scale=2.0)
    # ragged sizes: rows not a multiple of 128, H not a multiple of 32, the upstream 29-symbol alphabet, C > 32
    case(ref, "fc_eval_h100_c29", 37, 9, 100, 29, seed=13, train=False)
    case(ref, "fc_eval_h256_c50", 40, 4, 256, 50, seed=14, train=False)
