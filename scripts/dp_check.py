"""Multi-GPU check of the data-parallel single-dictionary path (BASELINE.json configs[4] shape family):
torchrun --nproc-per-node G scripts/dp_check.py [--width 2048 --feats 65536 --batch 8192 --steps 5]
Every rank trains the same dictionary on its slice of the rows (gradient-only fused step + one NCCL
all-reduce + sc_adam_apply); rank 0 also steps a single-GPU copy on the full batch and reports the
difference, then times the data-parallel step.  --fused 1 uses dist.FusedDataParallel instead (gradient tiles
reduce-scattered into the owner's buffer by the dW epilogue over NVLink, sharded Adam with parameter broadcast; no
NCCL call on the data path)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402
from sparse_coding__b200.dist import DataParallelEnsemble, FusedDataParallel, shard_rows  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=512)
    ap.add_argument("--feats", type=int, default=8192)
    ap.add_argument("--batch", type=int, default=2048, help="rows per GPU")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--check", type=int, default=1)
    ap.add_argument("--fused", type=int, default=0)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    model = sc.FunctionalSAE.init(a.width, a.feats, 1e-3, device=dev)
    ens = sc.FunctionalEnsemble([model], sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device=dev)
    if a.fused:
        if world == 1:
            dist.init_process_group("nccl", device_id=dev, init_method="tcp://127.0.0.1:29577", rank=0, world_size=1)
        dist.broadcast(ens.params["encoder"], src=0)  # also brings NCCL up before symmetric memory is set up
        for t in ens.params.values():
            dist.broadcast(t, src=0)
        dp = FusedDataParallel(ens)
    else:
        dp = DataParallelEnsemble(ens)
        dp.broadcast_parameters(0)
    B = a.batch * world
    g = torch.Generator(device="cpu").manual_seed(1)
    full = torch.randn(B, a.width, generator=g).to(dev)
    lo, hi = shard_rows(B, rank, world)
    out = {"world": world, "d": a.width, "n": a.feats, "global_batch": B, "fused": bool(a.fused)}
    if a.check:
        ref = None
        if rank == 0:
            torch.manual_seed(0)
            ref = sc.FunctionalEnsemble([sc.FunctionalSAE.init(a.width, a.feats, 1e-3, device=dev)], sc.FunctionalSAE,
                                        sc.torchopt_compat.adam, {"lr": 1e-3}, device=dev)
        worst = 0.0
        for s in range(3):
            l, _ = dp.step_batch(full[lo:hi])
            if rank == 0:
                lr, _ = ref.step_batch(full)
                worst = max(worst, float(((l["loss"] - lr["loss"]).abs() / lr["loss"].abs()).max()))
        if rank == 0:
            out["max_rel_loss_diff_vs_single_gpu_full_batch"] = worst
            out["mean_abs_param_diff"] = {k: float((ens.params[k] - ref.params[k]).abs().mean()) for k in ens.params}
            del ref
    for s in range(3):  # warm-up (allocator, lazy module loading, NCCL channels)
        dp.step_batch(full[lo:hi])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(a.steps):
        dp.step_batch(full[lo:hi])
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        out["ms_per_step"] = float(t)
        out["activations_per_s"] = B / (float(t) * 1e-3)
        out["step_tflops"] = 10.0 * B * a.feats * a.width / (float(t) * 1e-3) / 1e12
        print(json.dumps(out))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
