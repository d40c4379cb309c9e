"""PROBE: can the HBM-bound weight-gradient + Adam phase of one member group run under the tensor-bound encode /
decode / dL/dz phase of another group on a disjoint set of SMs?  Two 8-member ensembles A and B, two streams; the
library's debug hooks restrict a launch to its phase (sc_debug_set_phase) and to a number of CTAs
(sc_debug_set_sm_limit).  Baseline: both groups stepped one after the other on one stream with all SMs.
    python scripts/overlap_probe.py
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402
from sparse_coding__b200 import _native as nat  # noqa: E402

L = nat.lib()
L.sc_debug_set_sm_limit.argtypes = [C.c_int]
L.sc_debug_set_phase.argtypes = [C.c_int]
M, d, N, B = 8, 512, 4096, int(os.environ.get("PROBE_B", "2048"))


def group(seed):
    torch.manual_seed(seed)
    models = [sc.FunctionalSAE.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
    ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    ens._ws = nat.Workspace()  # own workspace per group (the class default is shared per ensemble object anyway)
    x = torch.randn(B, d, device="cuda")
    keep = [ens.step_batch(x) for _ in range(3)]
    args = ens._step_ctx.args
    args.operands_valid = 1
    return ens, x, keep, args


A, Bg = group(1), group(2)
torch.cuda.synchronize()


def launch(g, mask, limit, stream):
    L.sc_debug_set_phase(mask)
    L.sc_debug_set_sm_limit(limit)
    g[3].stream = stream.cuda_stream
    nat.check(L.sc_sae_step(C.byref(g[3])), "sc_sae_step")


def timed(fn, n=60, warm=10):
    for i in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


main = torch.cuda.current_stream()


def sequential():
    launch(A, 3, 0, main)
    launch(Bg, 3, 0, main)


print({"rows": B, "sequential_ms_per_16_members": round(timed(sequential), 4)})
for lim in (0, 100, 74, 48):
    print({"phase1_only_limit": lim, "ms": round(timed(lambda: launch(A, 1, lim, main)), 4)})
for lim in (0, 100, 74, 60, 48):
    print({"phase2_only_limit": lim, "ms": round(timed(lambda: launch(A, 2, lim, main)), 4)})

sA, sB = torch.cuda.Stream(), torch.cuda.Stream()
eA1, eB1 = torch.cuda.Event(), torch.cuda.Event()
for l1, l2 in ((100, 48), (88, 60), (74, 74), (60, 88)):
    def pipelined():
        sA.wait_event(eB1)
        launch(A, 1, l1, sA)
        eA1.record(sA)
        launch(A, 2, l2, sA)
        sB.wait_event(eA1)
        launch(Bg, 1, l1, sB)
        eB1.record(sB)
        launch(Bg, 2, l2, sB)

    def run():
        pipelined()
        main.wait_stream(sA)
        main.wait_stream(sB)

    sA.wait_stream(main)
    sB.wait_stream(main)
    eB1.record(sB)
    # time a batch of macro steps between two main-stream events that enclose both streams
    for i in range(10):
        pipelined()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    sA.wait_event(e0)
    sB.wait_event(e0)
    n = 60
    for i in range(n):
        pipelined()
    main.wait_stream(sA)
    main.wait_stream(sB)
    e1.record(main)
    torch.cuda.synchronize()
    print({"pipelined_limits": (l1, l2), "ms_per_16_members": round(e0.elapsed_time(e1) / n, 4)})
L.sc_debug_set_phase(3)
L.sc_debug_set_sm_limit(0)
