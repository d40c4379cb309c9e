"""FISTA at config 4 (B = 8192, d = 512, N = 4096): ms per iteration over 300 iterations (device-timed)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402

torch.manual_seed(0)
B, d, N = 8192, 512, 4096
D = torch.nn.functional.normalize(torch.randn(N, d, device="cuda"), dim=1)
x = torch.randn(B, d, device="cuda")
a0 = torch.relu(x @ D.T - 1.0)
for _ in range(2):
    sc.FunctionalFista.fista(x, D, 1e-3, a0, 20, "cuda")
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
a, res = sc.FunctionalFista.fista(x, D, 1e-3, a0, 300, "cuda")
e1.record()
torch.cuda.synchronize()
print({"ms_per_iteration": round(e0.elapsed_time(e1) / 300, 4), "nnz_frac": round(float((a > 0).float().mean()), 4),
       "res_rms": round(float(res.pow(2).mean().sqrt()), 5)})
