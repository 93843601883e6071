"""Loader shim: the product package lives in the directory `deepspeech.pytorch_b200/` (a dot in the name, as the
task lays it out), which the import system cannot address by dotted name.  `import b200ctc` loads it from that
path, registers it as `deepspeech_pytorch_b200`, and re-exports its public names.

    from b200ctc import CTCLoss           # drop-in for `from warpctc_pytorch import CTCLoss` (train.py:12)
"""
import importlib.util
import os
import sys

_NAME = "deepspeech_pytorch_b200"
_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "deepspeech.pytorch_b200")


def _load():
    if _NAME in sys.modules:
        return sys.modules[_NAME]
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(_DIR, "__init__.py"),
                                                  submodule_search_locations=[_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod


pkg = _load()
CTCLoss = pkg.CTCLoss
ctc_forward_backward = pkg.ctc_forward_backward
GreedyDecoder = pkg.GreedyDecoder
greedy_decode = pkg.greedy_decode
error_counts = pkg.error_counts
last_launch_count = pkg.last_launch_count
build = pkg.build
Labels = pkg.Labels
pack_targets = pkg.pack_targets
ALPHABET = pkg.ALPHABET
workloads = pkg.workloads
sharding = pkg.sharding
packing = pkg.packing
fc = pkg.fc
