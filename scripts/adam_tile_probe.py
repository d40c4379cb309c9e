"""The weight-gradient + Adam kernel with almost no GEMM work (B = 128 rows: two k-blocks per tile), i.e. its tile-wise
Adam epilogue alone, timed with the library's profiling marks - against the plain streaming pass of
scripts/adam_stream_probe.py this says what the tile access pattern itself costs."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402
from sparse_coding__b200 import _native as nat  # noqa: E402

for M, d, N, B in ((8, 512, 4096, 128), (8, 512, 4096, 2048), (8, 128, 16384, 128)):
    torch.manual_seed(1)
    models = [sc.FunctionalSAE.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
    ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    x = torch.randn(4 * B, d, device="cuda")
    for i in range(10):
        ens.step_batch(x[(i % 4) * B:(i % 4 + 1) * B])
    nat.profile_begin()
    for i in range(20):
        ens.step_batch(x[(i % 4) * B:(i % 4 + 1) * B])
    per = {}
    for n, ms in nat.profile_end():
        per.setdefault(n, []).append(ms)
    us = {k: round(float(np.mean(v)) * 1e3, 1) for k, v in per.items()}
    state = 6 * 2 * M * N * d * 4          # p, mu, nu in and out, both dictionaries
    halves = 3 * M * N * d * 2             # fp16 copies left for the next step (hand-over up to 3072 rows)
    codes = 2 * M * N * ((B + 7) & ~7) * 2  # c^T and dz^T read once
    t = us["G45_dW_adam"] * 1e-6
    print({"d": d, "N": N, "B": B, "kernel_us": us, "G45_algorithmic_GB_per_s": round((state + halves + codes) / t / 1e9, 1)})
