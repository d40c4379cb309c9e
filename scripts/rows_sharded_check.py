"""Row-sharded FISTA and metrics on N GPUs, one process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/rows_sharded_check.py
Every rank holds its own rows (weak scaling: 8192 FISTA rows / 65536 metric rows per GPU); rank 0 prints one JSON line
with the times (CUDA events, max over ranks) and the agreement of the all-reduced metrics with one GPU on all rows."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparse_coding__b200 as sc  # noqa: E402
from sparse_coding__b200 import dist as scd, standard_metrics as sm  # noqa: E402


def timed(fn, dev, reps=3):
    fn()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), out


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    d, N, rows_f, rows_m, iters = 512, 4096, 8192, 65536, 100
    torch.manual_seed(0)  # same dictionary everywhere
    We, Wd, b = torch.randn(N, d, device=dev) / d ** 0.5, torch.randn(N, d, device=dev), -0.5 * torch.ones(N, device=dev)
    model = sc.UntiedSAE(We, Wd, b)
    D = model.get_learned_dict()
    torch.manual_seed(1 + rank)  # different rows on every rank
    x = torch.randn(rows_m, d, device=dev)
    xf = x[:rows_f]
    a0 = torch.relu(xf @ D.T - 1.0)
    rec = {"world": world, "d": d, "N": N, "fista_rows_per_gpu": rows_f, "metric_rows_per_gpu": rows_m, "fista_iterations": iters}
    ms, (a, res) = timed(lambda: scd.fista_rows(xf, D, 1e-3, a0, iters, device=dev), dev, reps=2)
    rec["fista_ms"] = ms
    rec["fista_rows_per_s"] = rows_f * world / ms * 1e3
    ms, l0 = timed(lambda: scd.mean_nonzero_activations_sharded(model, x).sum(), dev)
    rec["mean_l0_ms"], rec["mean_l0"] = ms, float(l0)
    ms, fvu = timed(lambda: scd.fraction_variance_unexplained_sharded(model, x), dev)
    rec["fvu_ms"], rec["fvu"] = ms, float(fvu)
    ms, n_act = timed(lambda: scd.n_ever_active_sharded(model, x, threshold=10), dev)
    rec["n_ever_active_ms"], rec["n_ever_active"] = ms, n_act
    # one GPU on ALL rows: the numbers the all-reduced metrics must reproduce
    gathered = [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(gathered, x)
    if rank == 0:
        full = torch.cat(gathered)
        rec["one_gpu_all_rows"] = {"mean_l0": float(sm.mean_nonzero_activations(model, full).sum()),
                                   "fvu": float(sm.fraction_variance_unexplained(model, full)),
                                   "n_ever_active": sm.batched_calc_feature_n_ever_active(model, full, threshold=10)}
        print(json.dumps(rec))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
