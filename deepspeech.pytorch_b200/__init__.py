"""deepspeech.pytorch_b200 -- B200-native CTC loss for the DeepSpeech2 training path.

Only the one hot path of vadimkantorov/deepspeech.pytorch is here: the CTC loss forward+backward behind
`criterion(acts, labels, act_lens, label_lens)` (/root/reference/train.py:600, :352, :457), as hand-written
sm_100a CUDA (csrc/) behind a C ABI (include/ctc_b200.h), plus the host-side mirror of the reference interface.

The directory name contains a dot, so it is imported through the loader in /root/repo/b200ctc.py
(`import b200ctc`), which registers this package as `deepspeech_pytorch_b200`.
"""
from . import _lib, fc, packing, sharding, workloads  # noqa: F401
from ._lib import build  # noqa: F401
from .ctc_loss import (CTCLoss, GreedyDecoder, ctc_forward_backward, error_counts, greedy_decode,  # noqa: F401
                       last_launch_count)
from .packing import ALPHABET, BLANK_INDEX, Labels, pack_targets  # noqa: F401

__all__ = ["CTCLoss", "GreedyDecoder", "ctc_forward_backward", "error_counts", "greedy_decode", "last_launch_count", "build", "ALPHABET", "BLANK_INDEX", "Labels",
           "pack_targets", "packing", "sharding", "workloads"]
