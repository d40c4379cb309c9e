"""HBM rate of the plain elementwise Adam pass (sc_adam_apply: reads p, mu, nu, g, writes p, mu, nu - seven linear
streams) on tensors of the headline size: the yardstick for the tile-wise Adam epilogue of the weight-gradient kernel."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sparse_coding__b200 import _native as nat  # noqa: E402

L = nat.lib()
n = 8 * 4096 * 512
p, mu, nu, g = (torch.randn(2 * n, device="cuda") for _ in range(4))   # 2 dictionaries
nu.abs_()
adam = nat.ScAdam(1e-3, 0.9, 0.999, 1e-8, 0.0, 0.1, 0.001)
st = nat.stream_ptr(p.device)


def run():
    nat.check(L.sc_adam_apply(nat.ptr(p), nat.ptr(mu), nat.ptr(nu), nat.ptr(g), p.numel(), C.byref(adam), st), "adam")


for _ in range(5):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
print({"ms": round(ms, 4), "GB_per_s": round(7 * 4 * p.numel() / ms / 1e6, 1), "bytes_MB": 7 * 4 * p.numel() / 1e6})
# a copy of the same size for reference
a, b = torch.empty(7 * n, device="cuda"), torch.empty(7 * n, device="cuda")
for _ in range(5):
    b.copy_(a)
e0.record()
for _ in range(50):
    b.copy_(a)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 50
print({"copy_ms": round(ms, 4), "GB_per_s": round(2 * 4 * a.numel() / ms / 1e6, 1)})
