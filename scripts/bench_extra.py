"""Secondary measurements for profiles/: batch sweep of the headline config, the other BASELINE.json
configs on one GPU (C1 toy tied, C3 per-GPU share, C4 FISTA), all with CUDA events after warm-up.
    python scripts/bench_extra.py > gpurun_out/extra.json
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402


def timed(fn, warm=5, steps=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def ensemble_case(name, sig, M, d, N, B, out):
    torch.manual_seed(0)
    models = [sig.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
    ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    data = torch.randn(8 * B, d, device="cuda")
    i = [0]

    def step():
        s = i[0] % 8
        i[0] += 1
        ens.step_batch(data[s * B:(s + 1) * B])
    ms = timed(step)
    out[name] = {"M": M, "d": d, "N": N, "B": B, "ms_per_step": ms, "member_act_per_s": B * M / ms * 1e3,
                 "step_tflops": 10.0 * B * N * d * M / ms * 1e-9}
    del ens, data
    torch.cuda.empty_cache()


def input_pipeline_case(M=8, d=512, N=4096, B=2048, rows=1 << 20):
    """ensemble_train_loop fed (a) the reference way - fp32 chunk in host memory, torch BatchSampler, per-step host
    gather + H2D (big_sweep.py:170) - and (b) from a ResidentChunk (fp16 in HBM, sc_gather_rows, device sampler)."""
    class Cfg:
        use_wandb = False

    class Counter:
        value = 0

    torch.manual_seed(0)
    chunk_h = torch.randn(rows, d).half()
    res = {}
    for mode in ("host_fp32_chunk", "resident_fp16_chunk"):
        models = [sc.FunctionalSAE.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-4, -2, M)]
        ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
        if mode == "host_fp32_chunk":
            ds = chunk_h.to(torch.float32)
            n_batches = 40
            sampler = [b for _, b in zip(range(n_batches), torch.utils.data.BatchSampler(
                torch.utils.data.RandomSampler(range(rows)), batch_size=B, drop_last=False))]
        else:
            ds = sc.ResidentChunk(chunk_h, device="cuda")
            n_batches = 400
            sampler = [b for _, b in zip(range(n_batches), sc.DeviceBatchSampler(rows, B, device="cuda", seed=0))]
        sc.ensemble_train_loop(ens, Cfg(), {"device": "cuda"}, "e", sampler[:5], ds, Counter(), torch.float32,
                               fista_dictionary_update=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sc.ensemble_train_loop(ens, Cfg(), {"device": "cuda"}, "e", sampler, ds, Counter(), torch.float32,
                               fista_dictionary_update=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        res[mode] = {"batches": n_batches, "ms_per_step": dt / n_batches * 1e3, "member_act_per_s": M * B * n_batches / dt}
        del ens, ds
        torch.cuda.empty_cache()
    return res


def main():
    out = {}
    for B in (1024, 2048, 4096, 8192):
        ensemble_case(f"C2_untied_B{B}", sc.FunctionalSAE, 8, 512, 4096, B, out)
    ensemble_case("C1_toy_tied", sc.FunctionalTiedSAE, 4, 256, 512, 4096, out)
    ensemble_case("C3_gpt2_share_untied", sc.FunctionalSAE, 4, 768, 6144, 2048, out)
    ensemble_case("C3_gpt2_share_tied", sc.FunctionalTiedSAE, 4, 768, 6144, 2048, out)
    ensemble_case("C2_tied_B2048", sc.FunctionalTiedSAE, 8, 512, 4096, 2048, out)
    # C4 FISTA
    B, d, N, iters = 8192, 512, 4096, 300
    torch.manual_seed(1)
    D = torch.randn(N, d, device="cuda")
    D = D / D.norm(dim=1, keepdim=True)
    x = torch.randn(B, d, device="cuda")
    a0 = torch.relu(x @ D.T)
    sc.FunctionalFista.fista(x, D, 1e-3, a0, 5, "cuda")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    a, res = sc.FunctionalFista.fista(x, D, 1e-3, a0, iters, "cuda")
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    sc.FunctionalFista.fista(x, D, 1e-3, a0, iters, "cuda", eta=0.05)
    torch.cuda.synchronize()
    dt2 = time.perf_counter() - t0
    out["C4_fista"] = {"B": B, "d": d, "N": N, "iters": iters, "seconds_incl_eta": dt, "seconds_given_eta": dt2,
                       "ms_per_iter": dt2 / iters * 1e3, "algorithmic_tflops": 4.0 * B * N * d * iters / dt2 * 1e-12,
                       "tensor_tflops_3xtf32": 12.0 * B * N * d * iters / dt2 * 1e-12,
                       "algorithmic_GBps": 16.0 * B * N * iters / dt2 * 1e-9}
    # metrics on 1M rows
    n = 1 << 18
    ld = sc.UntiedSAE(torch.randn(N, d, device="cuda") / 22, D, torch.zeros(N, device="cuda"))
    xs = torch.randn(n, d, device="cuda")
    from sparse_coding__b200 import standard_metrics as sm
    out["metrics_n262144"] = {
        "fvu_ms": timed(lambda: sm.fraction_variance_unexplained(ld, xs), 2, 5),
        "mean_nonzero_ms": timed(lambda: sm.mean_nonzero_activations(ld, xs), 2, 5),
        "encode_ms": timed(lambda: ld.encode(xs[:65536]), 2, 5),
    }
    out["train_loop_input_pipeline"] = input_pipeline_case()
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
