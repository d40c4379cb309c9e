"""Host time per step_batch call (enqueue only) against the device time per step, at the toy and the headline shape.
    python scripts/host_overhead.py
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402

for name, sig, M, d, N, B in (("c1_toy_tied", sc.FunctionalTiedSAE, 4, 256, 512, 4096),
                              ("c2_headline", sc.FunctionalSAE, 8, 512, 4096, 2048)):
    torch.manual_seed(1)
    models = [sig.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
    ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    x = torch.randn(4 * B, d, device="cuda")
    for i in range(20):
        ens.step_batch(x[(i % 4) * B:(i % 4 + 1) * B])
    torch.cuda.synchronize()
    n = 100  # fewer launches than the driver queues (1024), so the host is never blocked by a full queue
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(n):
        ens.step_batch(x[(i % 4) * B:(i % 4 + 1) * B])
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(name, {"host_enqueue_us_per_step": round((t1 - t0) / n * 1e6, 1), "device_us_per_step": round(e0.elapsed_time(e1) / n * 1e3, 1),
                 "wall_us_per_step": round((t2 - t0) / n * 1e6, 1)})
