#!/usr/bin/env python
"""Headline benchmark: ensemble-SAE train throughput (BASELINE.json configs[1]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

Workload (config.workload "pythia70m_l2_untied_sae"): M = 8 untied SAEs (L1 sweep
np.logspace(-4,-2,8)), d = 512, 4096 features, Adam lr 1e-3, one FunctionalEnsemble.step_batch per
step on a synthetic activation batch of B rows.  Multi-GPU: the ensemble shards naturally, one
8-member group per GPU, no data-path collective ("weak" scaling), same per-GPU work.

metric = member-activations per second = B * M_total / step_time.
  value : inputs resident in HBM (a chunk larger than L2, a different batch slice each step)
  e2e   : through the public API with the batch in pinned HOST memory, H2D copy of the batch and
          D2H read of the loss scalars inside the timed region
  roofline    : dominant kernel's algorithmic flops / its CUDA-event time vs the measured TF32 dense peak
  cpu_baseline: the reference's own FunctionalEnsemble.step_batch (oracle/_ref, unmodified files) on the host cores,
                on the same B rows per step; the CPU port (oracle/restated.py) beside it as port_value
--impl reference times the reference's own code with all host threads on the same workload and prints the same line shape.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D_ACT, N_FEATS, M_PER_GPU = 512, 4096, 8
METRIC = "sae_train_member_activations_per_sec"
UNIT = "member-activations/s"


def l1_values(m):
    return np.logspace(-4, -2, m)


def make_chunk(rows, d, seed, device):
    """Synthetic sparse-feature activations of the named shape (sc_datasets/random_dataset.py style)."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    n_gt = 1024
    feats = torch.randn(n_gt, d, generator=g)
    feats /= feats.norm(dim=1, keepdim=True)
    out = torch.empty(rows, d)
    for i in range(0, rows, 16384):
        n = min(16384, rows - i)
        codes = torch.rand(n, n_gt, generator=g)
        codes = codes * (torch.rand(n, n_gt, generator=g) < 10.0 / n_gt)
        out[i:i + n] = codes @ feats
    return out.to(device)


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_port_steps(batch, steps, warmup, threads):
    """Time the CPU port of the reference path (oracle/restated.EnsembleOracle) on `batch` rows."""
    from oracle import restated as R
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    params = {"encoder": torch.stack([R.xavier_uniform(N_FEATS, D_ACT) for _ in range(M_PER_GPU)]),
              "encoder_bias": torch.zeros(M_PER_GPU, N_FEATS),
              "decoder": torch.stack([R.xavier_uniform(N_FEATS, D_ACT) for _ in range(M_PER_GPU)])}
    buffers = {"l1_alpha": torch.tensor(l1_values(M_PER_GPU), dtype=torch.float32),
               "bias_decay": torch.zeros(M_PER_GPU)}
    orc = R.EnsembleOracle(params, buffers, tied=False, lr=1e-3)
    data = make_chunk(batch * 4, D_ACT, 1, "cpu")
    for i in range(warmup):
        orc.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
    t0 = time.perf_counter()
    for i in range(steps):
        orc.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
    dt = time.perf_counter() - t0
    return batch * M_PER_GPU * steps / dt, dt / steps


def reference_available():
    """The reference's own modules under oracle/_ref (copied there by oracle/build_ref.py; they travel with the
    snapshot - /root/reference is never read at run time)."""
    from oracle import build_ref
    return build_ref.available()


def cpu_reference_steps(batch, steps, warmup, threads, budget_s=None):
    """Time the UNMODIFIED reference - its FunctionalEnsemble.step_batch (autoencoders/ensemble.py:175-193) over
    FunctionalSAE (sae_ensemble.py:13-77) with torchopt-style Adam - on the host cores: the same workload as the GPU
    arm (M members, d, N, `batch` rows per step).  With `budget_s` the timed steps stop early once the budget is spent
    (at least 2 are run).  Returns (member-activations / s, seconds per step, steps timed)."""
    os.environ["SC_REFERENCE_ROOT"] = os.path.join(ROOT, "oracle", "_ref")
    from oracle import ref_loader
    import warnings
    warnings.filterwarnings("ignore", message=".*batching rule.*")
    ns = ref_loader.load()
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    models = [ns.FunctionalSAE.init(D_ACT, N_FEATS, float(l1)) for l1 in l1_values(M_PER_GPU)]
    ens = ns.FunctionalEnsemble(models, ns.FunctionalSAE, ns.torchopt.adam, {"lr": 1e-3}, device="cpu")
    data = make_chunk(batch * 4, D_ACT, 1, "cpu")
    for i in range(warmup):
        ens.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
    done = 0
    t0 = time.perf_counter()
    for i in range(steps):
        ens.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
        done += 1
        if budget_s is not None and done >= 2 and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return batch * M_PER_GPU * done / dt, dt / done, done


def gpu_reference_steps(batch, steps, warmup, allow_tf32):
    """The same UNMODIFIED reference code (oracle/_ref) with device="cuda": the reference as its users run it - torch
    eager + functorch vmap/grad + torchopt-style Adam, cuBLAS SGEMM (true fp32, or TF32 when `allow_tf32`).  Not the
    contract's reference arm (that one is the CPU run): `bench.py --impl reference --ref-device cuda` prints it as an
    extra comparison line.  Device-timed with CUDA events.  Returns (member-activations / s, seconds per step)."""
    os.environ["SC_REFERENCE_ROOT"] = os.path.join(ROOT, "oracle", "_ref")
    from oracle import ref_loader
    import warnings
    warnings.filterwarnings("ignore", message=".*batching rule.*")
    ns = ref_loader.load()
    torch.backends.cuda.matmul.allow_tf32 = bool(allow_tf32)
    torch.backends.cudnn.allow_tf32 = bool(allow_tf32)
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    models = [ns.FunctionalSAE.init(D_ACT, N_FEATS, float(l1)) for l1 in l1_values(M_PER_GPU)]
    ens = ns.FunctionalEnsemble(models, ns.FunctionalSAE, ns.torchopt.adam, {"lr": 1e-3}, device=dev)
    data = make_chunk(batch * 4, D_ACT, 1, dev)
    for i in range(warmup):
        ens.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        ens.step_batch(data[(i % 4) * batch:(i % 4 + 1) * batch])
    e1.record()
    torch.cuda.synchronize(dev)
    sec = e0.elapsed_time(e1) * 1e-3 / steps
    torch.backends.cuda.matmul.allow_tf32 = False
    return batch * M_PER_GPU / sec, sec


def run_reference_on_gpu(args):
    """Extra comparison (not the contract's arm): the reference's own code on the B200 through torch."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if not reference_available():
        print(json.dumps({"impl": "reference", "ref_device": "cuda", "unavailable": "oracle/_ref is missing"}))
        return
    B, steps, warmup = args.batch, max(2, min(args.steps, 100)), max(3, min(args.warmup, 10))
    out = {}
    for name, tf32 in (("fp32", False), ("tf32_allowed", True)):
        val, sec = gpu_reference_steps(B, steps, warmup, tf32)
        out[name] = {"value": val, "ms_per_step": sec * 1e3}
    print(json.dumps({
        "impl": "reference", "ref_device": "cuda", "metric": METRIC, "unit": UNIT, "value": out["fp32"]["value"],
        "ms_per_step": out["fp32"]["ms_per_step"], "tf32_allowed": out["tf32_allowed"], "steps": steps, "warmup": warmup,
        "n_gpus": 1, "higher_is_better": True, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "pythia70m_l2_untied_sae", "d": D_ACT, "n_feats": N_FEATS, "members_per_gpu": M_PER_GPU,
                   "batch": B},
        "note": "the reference's FunctionalEnsemble.step_batch (oracle/_ref, unmodified; optree / torchopt / torchtyping "
                "from oracle/shims) on cuda:0 through torch eager + functorch: cuBLAS SGEMM, ~100 ATen launches per step",
    }))


def run_reference(args):
    """CPU arm: the reference's own code on the box's host cores, same workload (B rows per step, M members), the
    driver's --steps / --warmup (the timed steps stop after ~150 s if that many do not fit; `steps` says how many ran)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    B = args.batch
    warmup = max(1, args.warmup)
    if reference_available():
        kind = "reference"
        val, sec, steps = cpu_reference_steps(B, args.steps, warmup, threads, budget_s=150.0)
        what = "the reference's FunctionalEnsemble.step_batch (oracle/_ref/autoencoders/*.py, unmodified; optree / torchopt / torchtyping from oracle/shims)"
        port_val, _ = cpu_port_steps(B, 2, 1, threads)
    else:
        kind = "port"
        steps = max(2, min(args.steps, 40))
        val, sec = cpu_port_steps(B, steps, warmup, threads)
        what = "oracle/restated.py (CPU port of the reference path; oracle/_ref is missing)"
        port_val = val
    sample = f"{steps} steps of {B} rows x {M_PER_GPU} members after {warmup} warm-up, {what}, torch CPU fp32, {threads} threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "pythia70m_l2_untied_sae", "d": D_ACT, "n_feats": N_FEATS, "members_per_gpu": M_PER_GPU,
                   "batch": B},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample,
                         "port_value": port_val},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    def __init__(self, index):
        self.index, self.samples, self.reasons, self.stop = index, [], set(), False

        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                 0x80: "hw_power_brake"}
        while not self.stop and self.nv is not None:
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                    self.nv, "nvmlDeviceGetCurrentClocksEventReasons") else self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.001)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=1)

    def summary(self):
        # the first sample is taken as the timed region starts, before the load shows in the clock / power readings
        if len(self.samples) > 2:
            self.samples = self.samples[1:]
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def measure_tf32_peak(device):
    """Dense TF32 cuBLAS peak, same method as MEASURED_PEAKS.json's bf16 figure (8192^3, best of 10)."""
    n = 8192
    a = torch.randn(n, n, device=device)
    b = torch.randn(n, n, device=device)
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    best = 1e9
    for i in range(12):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            best = min(best, e0.elapsed_time(e1))
    torch.backends.cuda.matmul.allow_tf32 = old
    del a, b
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12



# ------------------------------------------------------------------------------------------ secondary measurements
def timed_ms(fn, warm, steps, device):
    """Mean milliseconds per call of fn(i): CUDA events on the current stream after `warm` untimed calls."""
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize(device)
    return e0.elapsed_time(e1) / steps


def ensemble_ms(sc, sig, M, d, N, B, device, fmt="auto", read_c=False, warm=6, steps=30):
    """ms per FunctionalEnsemble.step_batch of M members (L1 sweep) on fresh synthetic batches of B rows."""
    torch.manual_seed(1)
    models = [sig.init(d, N, float(l1), device=device) for l1 in l1_values(M)]
    ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device=device)
    ens.operand_format = fmt
    data = torch.randn(4 * B, d, device=device)

    def step(i):
        _, aux = ens.step_batch(data[(i % 4) * B:(i % 4 + 1) * B])
        if read_c:
            aux["c"]
    ms = timed_ms(step, warm, steps, device)
    used = ens.last_operand_format
    del ens, data, models
    torch.cuda.empty_cache()
    return ms, used


def extras_one_gpu(sc, device, B, hbm_peak, tensor_peak, tf32_peak):
    """The other BASELINE.json configurations and the variants of the headline step a reader would ask for, each a
    short CUDA-event run on this GPU (rank 0, single-GPU runs only)."""
    from oracle import restated as R
    ex = {}
    flops = lambda M, d, N, b: 10.0 * b * N * d * M
    # headline step with aux["c"] read as float32 [M,B,N] every step (the reference returns it, sae_ensemble.py:73-75)
    ms, _ = ensemble_ms(sc, sc.FunctionalSAE, M_PER_GPU, D_ACT, N_FEATS, B, device, read_c=True)
    ex["with_c"] = {"ms_per_step": ms, "member_act_per_s": B * M_PER_GPU / ms * 1e3,
                    "note": "aux['c'] materialised every step: + one fp16 -> fp32 transposing pass, 4*M*B*N bytes written"}
    ms, used = ensemble_ms(sc, sc.FunctionalSAE, M_PER_GPU, D_ACT, N_FEATS, B, device, fmt="tf32")
    ex["tf32"] = {"ms_per_step": ms, "member_act_per_s": B * M_PER_GPU / ms * 1e3, "operand_format": used,
                  "step_tflops": flops(M_PER_GPU, D_ACT, N_FEATS, B) / ms * 1e-9,
                  "frac_of_tf32_peak": flops(M_PER_GPU, D_ACT, N_FEATS, B) / ms * 1e-9 / tf32_peak}
    # config 1: toy tied ensemble (generate_test_data shape), config 3: one GPU's share of the 32-member GPT-2 sweep
    for name, sig, M, d, N, b in (("c1_toy_tied", sc.FunctionalTiedSAE, 4, 256, 512, 4096),
                                  ("c3_share_tied", sc.FunctionalTiedSAE, 4, 768, 6144, 2048),
                                  ("c3_share_untied", sc.FunctionalSAE, 4, 768, 6144, 2048)):
        ms, used = ensemble_ms(sc, sig, M, d, N, b, device)
        tf = flops(M, d, N, b) / ms * 1e-9
        ex[name] = {"M": M, "d": d, "N": N, "B": b, "ms_per_step": ms, "member_act_per_s": b * M / ms * 1e3,
                    "step_tflops": tf, "frac_of_tensor_peak": tf / tensor_peak, "operand_format": used}
    # config 5 on one GPU: the single large dictionary, 8192 rows per step
    ms, used = ensemble_ms(sc, sc.FunctionalSAE, 1, 2048, 65536, 8192, device, warm=3, steps=8)
    tf = flops(1, 2048, 65536, 8192) / ms * 1e-9
    ex["c5_one_gpu"] = {"d": 2048, "N": 65536, "B": 8192, "ms_per_step": ms, "activations_per_s": 8192 / ms * 1e3,
                        "step_tflops": tf, "frac_of_tensor_peak": tf / tensor_peak, "operand_format": used}
    # config 4: FISTA, 300 iterations on 8192 rows
    Bf, d, N, iters = 8192, 512, 4096, 300
    torch.manual_seed(2)
    D = torch.randn(N, d, device=device)
    D /= D.norm(dim=1, keepdim=True)
    x = torch.randn(Bf, d, device=device)
    a0 = torch.relu(x @ D.T - 1.0)
    l1 = torch.tensor(1e-3, device=device)
    run = lambda i: sc.FunctionalFista.fista(x, D, l1, a0, iters, device)
    ms = timed_ms(run, 1, 2, device)
    per_iter = ms / iters
    t0 = time.perf_counter()
    R.fista(x.cpu(), D.cpu(), 1e-3, a0.cpu(), 10)
    cpu_iter = (time.perf_counter() - t0) / 10
    ex["fista_c4"] = {"B": Bf, "d": d, "N": N, "iterations": iters, "seconds": ms * 1e-3, "ms_per_iteration": per_iter,
                      "products": "fp16 hi/lo pairs, three kind::f16 passes per product (fp32-accurate; the TF32 pair is the "
                                  "fallback for shapes the fp16 kernels do not take)",
                      "tflops_fp32_equivalent": 4.0 * Bf * N * d / per_iter * 1e-9,
                      "frac_of_tensor_peak_3_passes": 3 * 4.0 * Bf * N * d / per_iter * 1e-9 / tensor_peak,
                      "alg_GBps": 16.0 * Bf * N / per_iter * 1e-6, "frac_of_hbm_peak": 16.0 * Bf * N / per_iter * 1e-6 / hbm_peak,
                      "cpu_oracle_ms_per_iteration": cpu_iter * 1e3, "cpu_threads": torch.get_num_threads()}
    return ex


def dp_config5(sc, device, rank, world):
    """BASELINE.json configs[4]: ONE dictionary d = 2048 x 65536 trained data-parallel, 8192 rows per GPU (weak
    scaling), through both multi-GPU paths of sparse_coding__b200.dist - NCCL all-reduce of the gradients
    (DataParallelEnsemble) and the reduce-scatter fused into the weight-gradient kernel's epilogue over peer memory
    (FusedDataParallel).  Every rank calls this; rank 0 returns the record.  Times are CUDA events, max over ranks."""
    import torch.distributed as dist
    from sparse_coding__b200 import dist as scd
    d, N, rows = 2048, 65536, 8192
    flops = 10.0 * rows * N * d  # per GPU and step

    def fresh():
        torch.manual_seed(3)  # same initial dictionary on every rank
        p, b = sc.FunctionalSAE.init(d, N, 1e-3, device=device)
        return sc.FunctionalEnsemble([(p, b)], sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device=device)

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    torch.manual_seed(100 + rank)
    local = torch.randn(4 * rows, d, device=device)
    rec = {"d": d, "N": N, "rows_per_gpu": rows, "world": world, "scaling": "weak"}
    # the same build's single-GPU step on the same rows (every rank runs it at the same time: same clocks, no traffic)
    ens = fresh()
    t1 = max_over_ranks(timed_ms(lambda i: ens.step_batch(local[(i % 4) * rows:(i % 4 + 1) * rows]), 3, 8, device))
    rec["one_gpu_ms_per_step"] = t1
    # loss of one global step computed by ONE GPU on all ranks' rows (the parity oracle of SURVEY 8e); the global batch
    # of this check is kept at <= 16384 rows so that the one-GPU side stays a shape the tests cover
    check_rows = min(rows, 16384 // world)
    rec["loss_check_rows_per_gpu"] = check_rows
    gathered = [torch.empty(check_rows, d, device=device) for _ in range(world)]
    dist.all_gather(gathered, local[:check_rows].contiguous())
    ref_loss = None
    if rank == 0:
        ens0 = fresh()
        _, l0, _ = ens0.compute_grads(torch.cat(gathered))
        ref_loss = float(l0["loss"][0])
        del ens0
    del ens, gathered
    torch.cuda.empty_cache()
    for name, cls in (("nccl_allreduce", scd.DataParallelEnsemble), ("fused_reduce_scatter", scd.FusedDataParallel)):
        try:
            dp = cls(fresh())
            first, _ = dp.step_batch(local[:check_rows])
            first_loss = float(first["loss"][0])
            dist.barrier()
            ms = max_over_ranks(timed_ms(lambda i: dp.step_batch(local[(i % 4) * rows:(i % 4 + 1) * rows]), 2, 8, device))
            rec[name] = {"ms_per_step": ms, "activations_per_s": rows * world / ms * 1e3,
                         "tflops_per_gpu": flops / ms * 1e-9, "efficiency_vs_one_gpu": t1 / ms,
                         "first_step_loss": first_loss}
            del dp
        except Exception as e:  # a fabric that cannot be set up must not take the headline line down
            rec[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
        dist.barrier()
    # what the NCCL path pays for its collective: the all-reduce of the flat gradient buffer, alone
    flat = torch.zeros(2 * N * d + N + 3, device=device)
    t_ar = max_over_ranks(timed_ms(lambda i: dist.all_reduce(flat), 2, 5, device))
    grad_bytes = 4.0 * (2 * N * d + N)
    rec["allreduce_alone_ms"] = t_ar
    rec["allreduce_busbw_GBps"] = 2.0 * (world - 1) / world * grad_bytes / t_ar * 1e-6
    # the fused path moves (W-1)/W of the gradient out (reduce-add) and (W-1)/W of the new parameters out (broadcast)
    out_bytes = 2.0 * (world - 1) / world * grad_bytes
    rec["fused_nvlink_bytes_out_per_gpu_per_step"] = out_bytes
    if "ms_per_step" in rec.get("fused_reduce_scatter", {}):
        f = rec["fused_reduce_scatter"]
        f["nvlink_out_GBps_if_spread_over_the_step"] = out_bytes / f["ms_per_step"] * 1e-6
        f["limiter"] = ("the step's own kernels: link traffic is %.0f GB/s of the 770 GB/s a peer copy reaches; the gap "
                        "to the one-GPU step is two cross-rank barriers + the owner's Adam pass + broadcast stores"
                        % f["nvlink_out_GBps_if_spread_over_the_step"])
    if "ms_per_step" in rec.get("nccl_allreduce", {}):
        n = rec["nccl_allreduce"]
        n["limiter"] = ("exposed collective: all-reduce alone %.2f ms + sc_adam_apply on the full dictionary, after a "
                        "gradient-only step" % t_ar)
    if rank == 0 and ref_loss is not None:
        rec["one_gpu_full_batch_loss"] = ref_loss
        for name in ("nccl_allreduce", "fused_reduce_scatter"):
            if "first_step_loss" in rec.get(name, {}):
                rec[name]["loss_rel_dev_vs_one_gpu_full_batch"] = abs(rec[name]["first_step_loss"] - ref_loss) / abs(ref_loss)
    return rec


# ------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch.distributed as dist
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import _native as nat

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    B, M = args.batch, M_PER_GPU
    torch.manual_seed(rank)  # every GPU trains its own group of members
    models = [sc.FunctionalSAE.init(D_ACT, N_FEATS, float(l1), device=device) for l1 in l1_values(M)]
    ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device=device)
    n_slices = max(8, (160 << 20) // (B * D_ACT * 4) + 1)  # chunk > L2 (126 MB)
    chunk = make_chunk(n_slices * B, D_ACT, 1, device)
    host_chunk = chunk.cpu().pin_memory()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step(i):
        s = i % n_slices
        return ens.step_batch(chunk[s * B:(s + 1) * B])

    # End-to-end loop as a user would write it: the batch lives in pinned host memory, is copied to the
    # device on a side stream one step ahead (double buffered), and every step's loss vector is copied back
    # to pinned host memory and read on the host one step late (so the host never stalls the GPU).
    copy_stream = torch.cuda.Stream(device=device)
    dev_batch = [torch.empty(B, D_ACT, device=device) for _ in range(2)]
    host_loss = [torch.empty(M, dtype=torch.float32).pin_memory() for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    seen = {"sum": 0.0}

    def e2e_prefetch(i):
        s, b = i % n_slices, i % 2
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(done[b])  # the step that last used this buffer has finished
            dev_batch[b].copy_(host_chunk[s * B:(s + 1) * B], non_blocking=True)
            ready[b].record(copy_stream)

    def e2e_run(n):
        cur = torch.cuda.current_stream(device)
        for b in range(2):
            done[b].record(cur)
        e2e_prefetch(0)
        for i in range(n):
            b = i % 2
            cur.wait_event(ready[b])
            losses, _ = ens.step_batch(dev_batch[b])
            host_loss[b].copy_(losses["loss"], non_blocking=True)
            done[b].record(cur)
            e2e_prefetch(i + 1)
            if i > 0:
                done[1 - b].synchronize()  # previous step's loss is on the host now
                seen["sum"] += float(host_loss[1 - b].sum())
        done[(n - 1) % 2].synchronize()
        seen["sum"] += float(host_loss[(n - 1) % 2].sum())

    for i in range(args.warmup):
        device_step(i)
    barrier()
    with ClockSampler(local) as clocks:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            device_step(i)
        e1.record()
        barrier()
        ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    value = B * M * world / (ms_step * 1e-3)

    if args.quick:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "ms_per_step": ms_step, "quick": True}))
        if world > 1:
            dist.destroy_process_group()
        return

    # end to end: pinned host batch -> device, loss scalars -> host, every step
    e2e_run(min(args.warmup, 5))
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    barrier()
    t = torch.tensor([time.perf_counter() - t0], device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = B * M * world / (float(t.item()) / args.steps)

    dp5 = None
    if world > 1 and not args.no_extras:
        del chunk
        torch.cuda.empty_cache()
        chunk = None
        dp5 = dp_config5(sc, device, rank, world)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if chunk is None:
        chunk = make_chunk(n_slices * B, D_ACT, 1, device)

    # per-kernel CUDA-event times over profiled steps of the same workload
    prof_steps = min(args.steps, 20)
    nat.profile_begin()
    for i in range(prof_steps):
        device_step(i)
    marks = nat.profile_end()
    per = {}
    for nm, ms in marks:
        per.setdefault(nm, []).append(ms)
    kernel_ms = {k: float(np.mean(v)) for k, v in per.items()}
    gemm_flops = 2.0 * B * N_FEATS * D_ACT * M  # per GEMM launch (all members)
    gemms = {k: v for k, v in kernel_ms.items() if k.startswith("G")}
    # per launch: G1/G2/G3 are one GEMM of 2*B*N*d*M flops, G45 batches the two weight-gradient GEMMs
    launch_flops = {k: gemm_flops * (2.0 if k.startswith("G45") else 1.0) for k in gemms}
    # algorithmic HBM bytes per launch: codes / dz / Adam state dominate
    fmt = getattr(ens, "last_operand_format", "tf32")
    eb = 2.0 if fmt == "f16" else 4.0  # bytes per element of the GEMM operands / intermediates c, r, dz
    MBN, MNd, MBd, Bd = eb * M * B * N_FEATS, eb * M * N_FEATS * D_ACT, eb * M * B * D_ACT, eb * B * D_ACT
    adam_bytes = 2 * 6 * 4.0 * M * N_FEATS * D_ACT  # encoder + decoder: read and write p, mu, nu in fp32
    # fp16 mode: the Adam epilogue also writes the fp16 copies of the updated dictionaries for the next step (encoder,
    # decoder and the decoder transposed: 3 * MNd), which replaces the separate pass that re-read the fp32 weights
    handover = 3 * MNd if fmt == "f16" else 0.0
    launch_bytes = {"G1_encode": MBN + MNd + Bd, "G2_decode_mse": MBN + MNd + MBd + 4.0 * B * D_ACT,
                    "G3_back_code": 2 * MBN + MNd + MBd, "G45_dW_adam": 2 * MBN + adam_bytes + MBd + Bd + handover}
    tf32_peak = measure_tf32_peak(device)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "6650 GB/s (of fallback)"
    traffic = {}
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if traffic.get("batch") != B:
            traffic = {}
    except Exception:
        pass
    if fmt == "f16":
        tensor_peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1400.0)))
        tensor_src = ("MEASURED_PEAKS.json bf16_tflops_sustained (16-bit dense peak; kernels timed inside a long step, "
                      "kind::f16)") if peaks else "1400 TFLOP/s 16-bit sustained (of fallback)"
    else:
        tensor_peak = tf32_peak
        tensor_src = "measured here: torch.matmul TF32 8192^3 best of 10 (MEASURED_PEAKS.json has bf16 only; kind::tf32)"
    per_kernel = {}
    for k in gemms:
        t = gemms[k] * 1e-3
        t_tensor = launch_flops[k] / (tensor_peak * 1e12)
        t_hbm = launch_bytes.get(k, 0.0) / (hbm_peak * 1e9)
        bound = "tensor" if t_tensor >= t_hbm else "hbm"
        per_kernel[k] = {"ms": gemms[k], "tflops": launch_flops[k] / t / 1e12, "alg_GBps": launch_bytes.get(k, 0.0) / t / 1e9,
                         "bound": bound, "frac": max(t_tensor, t_hbm) / t,
                         "ncu_dram_bytes": traffic.get("kernels", {}).get(k)}
    dom = max(gemms, key=gemms.get)
    dk = per_kernel[dom]
    step_flops = 10.0 * B * N_FEATS * D_ACT * M
    roofline = {
        "kernel": dom, "bound": dk["bound"],
        "achieved": dk["tflops"] if dk["bound"] == "tensor" else dk["alg_GBps"],
        "peak": tensor_peak if dk["bound"] == "tensor" else hbm_peak,
        "unit": "TFLOP/s" if dk["bound"] == "tensor" else "GB/s",
        "frac": dk["frac"], "traffic": dk["ncu_dram_bytes"],
        "peak_source": "tensor peak: " + tensor_src + "; hbm peak: " + hbm_src,
        "operand_format": fmt, "tensor_peak_tflops": tensor_peak, "tf32_peak_tflops": tf32_peak, "hbm_peak_GBps": hbm_peak, "bf16_peak_measured": peaks.get("bf16_tflops"),
        "step_tflops": step_flops / (ms_step * 1e-3) / 1e12,
        "step_frac_of_tensor_peak": step_flops / (ms_step * 1e-3) / 1e12 / tensor_peak,
        "kernels": per_kernel, "kernel_ms": kernel_ms,
    }

    # the same step at larger batches (the Adam traffic of the weight-gradient kernel is per step, not per row, so
    # tensor-core utilisation rises with the batch): short CUDA-event runs, reported beside the headline batch
    by_batch = {str(B): {"ms_per_step": ms_step, "member_act_per_s": value, "step_tflops": roofline["step_tflops"],
                         "frac_of_tensor_peak": roofline["step_frac_of_tensor_peak"]}}
    if world == 1:
        for b2 in (1024, 4096, 8192):
            if b2 == B or b2 * 4 > chunk.shape[0]:
                continue
            n2 = chunk.shape[0] // b2
            for i in range(10):
                ens.step_batch(chunk[(i % n2) * b2:(i % n2 + 1) * b2])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(100):
                ens.step_batch(chunk[(i % n2) * b2:(i % n2 + 1) * b2])
            e1.record()
            torch.cuda.synchronize()
            ms2 = e0.elapsed_time(e1) / 100
            tf2 = 10.0 * b2 * N_FEATS * D_ACT * M / (ms2 * 1e-3) / 1e12
            by_batch[str(b2)] = {"ms_per_step": ms2, "member_act_per_s": b2 * M / (ms2 * 1e-3), "step_tflops": tf2,
                                 "frac_of_tensor_peak": tf2 / tensor_peak}
    roofline["by_batch"] = by_batch

    threads = os.cpu_count() or 1
    if reference_available():
        cpu_val, cpu_sec, cpu_n = cpu_reference_steps(B, 3, 1, threads)
        port_val, _ = cpu_port_steps(B, 2, 1, threads)
        cpu_baseline = {"value": cpu_val, "unit": UNIT, "cores": threads, "kind": "reference", "port_value": port_val,
                        "sample": f"{cpu_n} steps of {B} rows x {M} members after 1 warm-up: the reference's own "
                                  "FunctionalEnsemble.step_batch (oracle/_ref, unmodified) on torch CPU fp32; port_value = "
                                  "oracle/restated.py on the same batch"}
    else:
        cpu_val, cpu_sec = cpu_port_steps(B, 3, 1, threads)
        cpu_baseline = {"value": cpu_val, "unit": UNIT, "cores": threads, "kind": "port",
                        "sample": f"3 steps of {B} rows x {M} members after 1 warm-up, oracle/restated.py torch-CPU fp32 "
                                  "(oracle/_ref missing)"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": fmt,
        "data": "synthetic",
        "config": {"workload": "pythia70m_l2_untied_sae", "d": D_ACT, "n_feats": N_FEATS, "members_per_gpu": M,
                   "batch": B, "optimizer": "adam lr=1e-3 (reference-faithful t=1)",
                   "precision": ("fp16 GEMM operands (11-bit significand, same as TF32), fp32 accumulate / weights / Adam"
                                 if fmt == "f16" else "TF32 GEMM operands, fp32 accumulate / weights / Adam"), "parallelism": f"ensemble-shard x{world}",
                   "l2_policy": f"inputs larger than L2: {n_slices} batch slices = {n_slices * B * D_ACT * 4 >> 20} MiB chunk, "
                                "working set per step > 1 GB"},
        "activations_per_sec": value / (M * world) ,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * D_ACT * 4, "d2h_bytes_per_step": M * 4},
        # our kernels per step: one launch mark each in the library (fp16 mode: prologue, G1, G2, G3, G4+G5 = 5);
        # in TF32 mode one mark covers two launches (operand preparation + bias norm)
        "gpu_launches": (len(marks) // prof_steps + (0 if fmt == "f16" else 1)) * args.steps,
        "roofline": roofline, "cpu_baseline": cpu_baseline, "clocks": clocks.summary(),
    }
    if dp5 is not None:
        line["dp_config5"] = dp5
    if world == 1 and not args.no_extras:
        del ens, chunk
        torch.cuda.empty_cache()
        line["extra"] = extras_one_gpu(sc, device, B, hbm_peak, tensor_peak, tf32_peak)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--batch", type=int, default=2048)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ref-device", default="cpu", choices=["cpu", "cuda"],
                    help="with --impl reference: cpu (the contract's arm) or cuda (extra: the reference through torch on the GPU)")
    ap.add_argument("--quick", action="store_true", help="profiling runs: skip e2e, peak measurement and the CPU baseline")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary records (other configs, config-5 data-parallel)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        if args.ref_device == "cuda":
            run_reference_on_gpu(args)
        else:
            run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
