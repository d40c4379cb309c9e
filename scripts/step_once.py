"""A few train steps of one ensemble configuration - the short command line used for ncu captures.
    python scripts/step_once.py [--M 8 --d 512 --N 4096 --B 2048 --steps 6 --tied 0 --fmt auto]
"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=8)
ap.add_argument("--d", type=int, default=512)
ap.add_argument("--N", type=int, default=4096)
ap.add_argument("--B", type=int, default=2048)
ap.add_argument("--steps", type=int, default=6)
ap.add_argument("--tied", type=int, default=0)
ap.add_argument("--fmt", default="auto")
a = ap.parse_args()
torch.manual_seed(0)
sig = sc.FunctionalTiedSAE if a.tied else sc.FunctionalSAE
models = [sig.init(a.d, a.N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, a.M)]
ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
ens.operand_format = a.fmt
data = torch.randn(4 * a.B, a.d, device="cuda")
for i in range(a.steps):
    losses, _ = ens.step_batch(data[(i % 4) * a.B:(i % 4 + 1) * a.B])
torch.cuda.synchronize()
print({k: [round(float(x), 5) for x in v] for k, v in losses.items()})
