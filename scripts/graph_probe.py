"""What would CUDA-graph replay of a whole step buy?  Captures k steps of step_batch on fixed batch buffers with
torch.cuda.graph and compares replay time with ordinary stream launches (device-timed).
    python scripts/graph_probe.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402


def timed(fn, n):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for name, sig, M, d, N, B in (("c1_toy_tied", sc.FunctionalTiedSAE, 4, 256, 512, 4096),
                              ("c2_headline", sc.FunctionalSAE, 8, 512, 4096, 2048)):
    torch.manual_seed(1)
    models = [sig.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
    ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    x = torch.randn(4 * B, d, device="cuda")
    xs = [x[i * B:(i + 1) * B] for i in range(4)]
    for i in range(24):
        ens.step_batch(xs[i % 4])
    plain = timed(lambda i: ens.step_batch(xs[i % 4]), 100)
    ens._poll_range_flag = lambda device: None  # its event query / pinned copy are not capturable
    k = 4
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for i in range(k):
            ens.step_batch(xs[i % 4])
    torch.cuda.current_stream().wait_stream(s)
    keep = []
    with torch.cuda.graph(g):
        for i in range(k):
            keep.append(ens.step_batch(xs[i % 4]))
    replay = timed(lambda i: g.replay(), 50) / k
    print(name, {"stream_launch_ms_per_step": round(plain, 4), "graph_replay_ms_per_step": round(replay, 4)})
