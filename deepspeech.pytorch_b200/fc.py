"""The logits producer of the CTC path on the B200 -- forward of the reference's final layer

    fully_connected = nn.Sequential(nn.BatchNorm1d(H), nn.Linear(H, C, bias=False))
    self.fc = nn.Sequential(SequenceWise(fully_connected))          # /root/reference/model.py:335-341
    x = self.fc(x)                                                   # model.py:373, x [T,B,H] -> [T,B,C]

through ctc_b200_fc_forward (include/ctc_b200.h): one TMA + tcgen05 kernel that reads x once and writes the fp32
logits and, optionally, the CTC workspace's emission rows (the softmax the criterion would otherwise recompute from
the logits) and the per-frame argmax.  Forward only -- evaluation / inference (train.py:329-512 run the model under
no_grad in eval mode); there is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib

_fc_ws_cache = {}
_fold_sig = {}  # fc workspace data_ptr -> signature of the layer whose folded weights it holds


def _aligned(t: torch.Tensor) -> int:
    return (t.data_ptr() + 255) & ~255


def ctc_workspace_for(T: int, B: int, C: int, max_label_len: int, device) -> torch.Tensor:
    """A CTC workspace (uint8 tensor) sized by ctc_b200_workspace_size, to hand to BOTH fc_forward(ctc_workspace=...)
    and ctc_forward_backward(workspace=..., emissions_ready=True)."""
    lib = _lib.load()
    n = ctypes.c_size_t(0)
    _lib.check(lib.ctc_b200_workspace_size(T, B, C, int(max_label_len), None, ctypes.byref(n)), "ctc_b200_workspace_size")
    return torch.empty(n.value + 256, dtype=torch.uint8, device=device)


def fc_forward(x: torch.Tensor, weight: torch.Tensor, bn_weight: Optional[torch.Tensor] = None,
               bn_bias: Optional[torch.Tensor] = None, running_mean: Optional[torch.Tensor] = None,
               running_var: Optional[torch.Tensor] = None, eps: float = 1e-5, training: bool = False,
               want_logits: bool = True, ctc_workspace: Optional[torch.Tensor] = None,
               act_lens: Optional[torch.Tensor] = None, want_path: bool = False, return_batch_stats: bool = False):
    """logits [T,B,C] = Linear(BatchNorm1d(x)) for x [T,B,H] CUDA fp32 (SequenceWise: rows are the T*B frames).

    training=False: BatchNorm uses running_mean / running_var (module.eval()).  training=True: the statistics of x itself
    (biased variance), as the reference's training forward does; with return_batch_stats they are returned so the caller
    can update its running statistics.  ctc_workspace: a tensor from `ctc_workspace_for` -- its emission rows are
    written by the same kernel (act_lens: CUDA int32 [B], frames t >= act_lens[b] are skipped).  want_path: also return
    the per-frame argmax [B,T] int32 (the greedy decoder's input).
    Returns a dict: {"logits"?, "path"?, "batch_mean"?, "batch_var"?}.
    """
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.float32 and x.dim() == 3):
        raise ValueError("x must be a CUDA float32 tensor of shape [T,B,H] (no CPU fallback exists)")
    if x.requires_grad and torch.is_grad_enabled():
        raise RuntimeError("fc_forward is forward-only (evaluation / inference); run it under torch.no_grad()")
    lib = _lib.load()
    x = x.detach().contiguous()
    T, B, H = x.shape
    dev = x.device

    def f32(t, n, name):
        if t is None:
            return None
        t = t.detach()
        if not (t.is_cuda and t.dtype == torch.float32 and t.numel() == n):
            raise ValueError(f"{name} must be a CUDA float32 tensor with {n} elements")
        return t.contiguous()

    if weight.dim() != 2 or weight.shape[1] != H:
        raise ValueError("weight must be [C,H]")
    C = int(weight.shape[0])
    w = f32(weight, C * H, "weight")
    keep = [w] + [f32(t, H, n) for t, n in ((bn_weight, "bn_weight"), (bn_bias, "bn_bias"),
                                             (running_mean, "running_mean"), (running_var, "running_var"))]
    layer = _lib.FcLayer()
    layer.weight = w.data_ptr()
    layer.bn_weight = keep[1].data_ptr() if keep[1] is not None else None
    layer.bn_bias = keep[2].data_ptr() if keep[2] is not None else None
    layer.bn_mean = keep[3].data_ptr() if (keep[3] is not None and not training) else None
    layer.bn_var = keep[4].data_ptr() if (keep[4] is not None and not training) else None
    layer.bn_eps = float(eps)
    layer.bn_batch_stats = int(bool(training))
    out = {}
    with torch.cuda.device(dev):
        if training and return_batch_stats:
            out["batch_mean"] = torch.empty(H, dtype=torch.float32, device=dev)
            out["batch_var"] = torch.empty(H, dtype=torch.float32, device=dev)
            layer.batch_mean_out = out["batch_mean"].data_ptr()
            layer.batch_var_out = out["batch_var"].data_ptr()
        n = ctypes.c_size_t(0)
        _lib.check(lib.ctc_b200_fc_workspace_size(T, B, H, C, ctypes.byref(n)), "ctc_b200_fc_workspace_size")
        key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
        ws = _fc_ws_cache.get(key)
        if torch.cuda.is_current_stream_capturing():
            ws = torch.empty(n.value + 256, dtype=torch.uint8, device=dev)
        elif ws is None or ws.numel() < n.value + 256:
            ws = torch.empty(n.value + 256, dtype=torch.uint8, device=dev)
            _fc_ws_cache[key] = ws
        # evaluation: the folded weights in the workspace stay valid while the layer's tensors are untouched
        # (same storage, same in-place version counters): skip the fold kernel then
        sig = (T * 0 + H, C, float(eps)) + tuple((t.data_ptr(), t._version) if t is not None else None for t in keep)
        reuse = (not training) and not torch.cuda.is_current_stream_capturing() and _fold_sig.get(ws.data_ptr()) == sig
        layer.reuse_folded = int(reuse)
        _fold_sig[ws.data_ptr()] = None if training else sig
        logits = torch.empty(T, B, C, dtype=torch.float32, device=dev) if want_logits else None
        path = torch.empty(B, T, dtype=torch.int32, device=dev) if want_path else None
        lens = None
        if act_lens is not None:
            lens = act_lens.to(device=dev, dtype=torch.int32).contiguous()
        opts = _lib.CtcOptions()
        lib.ctc_b200_default_options(ctypes.byref(opts))
        opts.stream = torch.cuda.current_stream(dev).cuda_stream
        st = lib.ctc_b200_fc_forward(x.data_ptr(), T, B, H, C, ctypes.byref(layer),
                                     logits.data_ptr() if logits is not None else None,
                                     _aligned(ctc_workspace) if ctc_workspace is not None else None,
                                     lens.data_ptr() if lens is not None else None,
                                     path.data_ptr() if path is not None else None, _aligned(ws),
                                     ws.numel() - (_aligned(ws) - ws.data_ptr()), ctypes.byref(opts))
        _lib.check(st, "ctc_b200_fc_forward")
    if logits is not None:
        out["logits"] = logits
    if path is not None:
        out["path"] = path
    return out


class SequenceWiseBNLinear(torch.nn.Module):
    """Drop-in for the reference's `self.fc` of the RNN models (model.py:335-341) in evaluation: holds the SAME
    parameters / buffers (pass the reference's BatchNorm1d and Linear modules; nothing is copied) and maps
    x [T,B,H] -> logits [T,B,C] through the fused kernel.

        fused = SequenceWiseBNLinear.from_reference_fc(model.fc)       # model.fc = Sequential(SequenceWise(Sequential(BN, Linear)))
        logits = fused(x)                                              # replaces `x = self.fc(x)` at model.py:373
    """

    def __init__(self, bn: torch.nn.BatchNorm1d, linear: torch.nn.Linear):
        super().__init__()
        if linear.bias is not None:
            raise ValueError("the reference's output layer has bias=False (model.py:337)")
        self.bn, self.linear = bn, linear

    @classmethod
    def from_reference_fc(cls, fc: torch.nn.Module):
        inner = fc[0].module  # Sequential(SequenceWise(Sequential(BatchNorm1d, Linear)))
        return cls(inner[0], inner[1])

    def forward(self, x: torch.Tensor, **kw):
        bn = self.bn
        training = bool(bn.training or not bn.track_running_stats)
        out = fc_forward(x, self.linear.weight, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps,
                         training=training, **kw)
        return out["logits"] if set(out) == {"logits"} else out
