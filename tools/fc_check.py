"""Quick check of the fc kernel against the reference-generated goldens (tests/golden/fc_*.npz) on the GPU box."""
import glob
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200ctc  # noqa: E402

for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "fc_*.npz"))):
    z = np.load(f)
    cu = lambda k: torch.from_numpy(z[k]).cuda()
    train = bool(z["train"])
    out = b200ctc.fc.fc_forward(cu("x"), cu("weight"), cu("bn_weight"), cu("bn_bias"), cu("running_mean"),
                                cu("running_var"), float(z["eps"]), training=train, want_path=True,
                                return_batch_stats=True)
    torch.cuda.synchronize()
    got = out["logits"].cpu().numpy().astype(np.float64)
    err = np.abs(got - z["logits"]).max()
    path_ok = np.array_equal(out["path"].cpu().numpy().T, got.argmax(-1))
    msg = f"{os.path.basename(f)}: max abs err {err:.3e} path {'ok' if path_ok else 'BAD'}"
    if train:
        msg += f" mean err {np.abs(out['batch_mean'].cpu().numpy() - z['batch_mean']).max():.2e}"
        msg += f" var err {np.abs(out['batch_var'].cpu().numpy() - z['batch_var']).max():.2e}"
    print(msg, flush=True)
