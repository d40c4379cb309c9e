"""One pass over the inference-side entry points (encode with statistics, predict / FVU, FISTA, chunk gather) - the
short command line used for ncu launch lists of those kernels."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402
from sparse_coding__b200 import standard_metrics as sm  # noqa: E402
from sparse_coding__b200.chunks import ResidentChunk  # noqa: E402

torch.manual_seed(0)
n, d, N = 65536, 512, 4096
x = torch.randn(n, d, device="cuda")
We, Wd, b = torch.randn(N, d, device="cuda") / d ** 0.5, torch.randn(N, d, device="cuda"), torch.zeros(N, device="cuda")
m = sc.UntiedSAE(We, Wd, b)
for _ in range(2):
    c = m.encode(x[:8192])
    xh = m.predict(x[:8192])
    fvu = sm.fraction_variance_unexplained(m, x)
    l0 = sm.mean_nonzero_activations(m, x).sum()
    D = m.get_learned_dict()
    a, res = sc.FunctionalFista.fista(x[:8192], D, 1e-3, torch.relu(x[:8192] @ D.T - 1.0), 3, "cuda")
    rc = ResidentChunk(x.half())
    batch = rc[torch.randint(0, n, (2048,), device="cuda")]
torch.cuda.synchronize()
print("fvu", float(fvu), "L0", float(l0), "fista", float(a.sum()), "gather", tuple(batch.shape))
