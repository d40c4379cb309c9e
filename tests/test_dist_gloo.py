"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: member sharding, row sharding and the
data-parallel gradient all-reduce.  The per-rank "engine" here is the CPU oracle evaluated with the
same global-batch normalisation the CUDA step uses in gradient-only mode, so the test checks the
algebra the GPU path relies on: shard gradients, sum them, apply Adam == full-batch step."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import restated as R
from sparse_coding__b200 import dist as scd
from sparse_coding__b200.dist import DataParallelEnsemble, shard_members, shard_rows


def test_sharding_helpers():
    assert [len(shard_members(list(range(32)), r, 8)) for r in range(8)] == [4] * 8
    parts = [shard_members(list(range(10)), r, 4) for r in range(4)]
    assert sum(parts, []) == list(range(10)) and [len(p) for p in parts] == [3, 3, 2, 2]
    spans = [shard_rows(1001, r, 4) for r in range(4)]
    assert spans[0][0] == 0 and spans[-1][1] == 1001 and all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


class OracleGradEngine:
    """CPU stand-in with FunctionalEnsemble's gradient-only interface (untied SAE, M members)."""

    def __init__(self, params, buffers, lr=1e-3):
        self.params = {k: v.clone() for k, v in params.items()}
        self.buffers = {k: v.clone() for k, v in buffers.items()}
        self.mu = {k: torch.zeros_like(v) for k, v in params.items()}
        self.nu = {k: torch.zeros_like(v) for k, v in params.items()}
        self.lr = lr

    def compute_grads(self, x, expand_dims=True, batch_total=0, bias_decay=None):
        P = self.params
        M = P["encoder"].shape[0]
        B = x.shape[0]
        Bn = batch_total or B
        grads = {k: torch.zeros_like(v) for k, v in P.items()}
        out = {k: [] for k in ("l_reconstruction", "l_l1", "l_bias_decay")}
        cs = []
        for m in range(M):
            lam = self.buffers["l1_alpha"][m]
            beta = (self.buffers["bias_decay"] if bias_decay is None else bias_decay)[m]
            ls, fw = R.sae_forward(P["encoder"][m], P["encoder_bias"][m], P["decoder"][m], lam, beta, x)
            gWe, gb, gWd = R.sae_backward(P["encoder"][m], P["encoder_bias"][m], P["decoder"][m], lam, beta, x, fw)
            # sae_backward normalises by the local B; the engine's gradient-only mode by batch_total
            scale = B / Bn
            db_decay = beta * P["encoder_bias"][m] / fw["bn"] if float(fw["bn"]) > 0 else 0.0
            grads["encoder"][m] = gWe * scale
            grads["decoder"][m] = gWd * scale
            grads["encoder_bias"][m] = (gb - db_decay) * scale + db_decay
            out["l_reconstruction"].append(ls["l_reconstruction"] * scale)
            out["l_l1"].append(ls["l_l1"] * scale)
            out["l_bias_decay"].append(ls["l_bias_decay"])
            cs.append(fw["c"])
        return grads, {k: torch.stack(v) for k, v in out.items()}, {"c": torch.stack(cs)}

    def apply_grads(self, grads):
        for k, g in grads.items():
            for m in range(g.shape[0]):
                R.adam_step_(self.params[k][m], g[m], self.mu[k][m], self.nu[k][m], self.lr)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    d, N, B, M = 16, 24, 50, 2
    params = {"encoder": torch.stack([R.xavier_uniform(N, d) for _ in range(M)]),
              "encoder_bias": 0.01 * torch.randn(M, N),
              "decoder": torch.stack([R.xavier_uniform(N, d) for _ in range(M)])}
    buffers = {"l1_alpha": torch.tensor([1e-3, 1e-2]), "bias_decay": torch.tensor([0.0, 0.05])}
    if rank != 0:  # wrong start on purpose: broadcast_parameters must fix it
        params = {k: v + 1.0 for k, v in params.items()}
    dp = DataParallelEnsemble(OracleGradEngine(params, buffers))
    dp.broadcast_parameters(0)
    data = torch.randn(3, B, d)
    losses = []
    for step in range(3):
        lo, hi = shard_rows(B, rank, world)
        l, aux = dp.step_batch(data[step, lo:hi])
        assert aux["c"].shape == (M, hi - lo, N)
        losses.append(l)
    torch.save({"params": dp.engine.params, "losses": losses}, os.path.join(tmp, f"rank{rank}.pt"))
    dist.destroy_process_group()


def test_data_parallel_equals_full_batch_step(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "rank0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "rank1.pt", weights_only=False)
    for k in r0["params"]:
        assert torch.equal(r0["params"][k], r1["params"][k]), k  # ranks stay in lock step
    # single-process full-batch oracle
    torch.manual_seed(0)
    d, N, B, M = 16, 24, 50, 2
    params = {"encoder": torch.stack([R.xavier_uniform(N, d) for _ in range(M)]),
              "encoder_bias": 0.01 * torch.randn(M, N),
              "decoder": torch.stack([R.xavier_uniform(N, d) for _ in range(M)])}
    buffers = {"l1_alpha": torch.tensor([1e-3, 1e-2]), "bias_decay": torch.tensor([0.0, 0.05])}
    orc = R.EnsembleOracle(params, buffers, tied=False, lr=1e-3)
    data = torch.randn(3, B, d)
    for step in range(3):
        l, _ = orc.step_batch(data[step])
        for k in ("loss", "l_reconstruction", "l_l1", "l_bias_decay"):
            assert torch.allclose(r0["losses"][step][k], l[k], rtol=1e-4, atol=1e-7), (step, k)
    for k in orc.params:
        assert float((orc.params[k] - r0["params"][k]).abs().mean()) < 5e-6, k


# ---------------------------------------------------------------------------------------------- row-sharded FISTA / metrics
class _OracleModel:
    """UntiedSAE stand-in on the CPU oracle (identity centring)."""

    def __init__(self, We, b, Wd):
        self.We, self.b, self.Wd = We, b, Wd

    def center(self, x):
        return x

    def predict(self, x):
        return R.untied_predict(self.We, self.b, self.Wd, x)


def _rows_worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.load(os.path.join(tmp, "inputs.pt"), weights_only=False)
    x, D, a0, We, b, Wd = (g[k] for k in ("x", "D", "a0", "We", "b", "Wd"))
    lo, hi = shard_rows(x.shape[0], rank, world)
    model = _OracleModel(We, b, Wd)
    counts_fn = lambda m, xc: (R.untied_encode(m.We, m.b, xc) != 0).float().sum(0)
    rss_fn = lambda m, xs: (xs - m.predict(xs)).pow(2).sum()
    out = {}
    # eta deliberately different per rank before the broadcast: rank 0's must win
    eta_fn = lambda xs, Dd, l1, a: R.fista_eta(Dd).reshape(1) * (1.0 if rank == 0 else 0.5)
    fista_fn = lambda xs, Dd, l1, a, it, eta: R.fista(xs, Dd, l1, a, it, eta=eta)
    out["a"], out["res"] = scd.fista_rows(x[lo:hi], D, 1e-2, a0[lo:hi], 25, fista_fn=fista_fn, eta_fn=eta_fn)
    out["mnz"] = scd.mean_nonzero_activations_sharded(model, x[lo:hi], counts_fn=counts_fn)
    out["ever"] = scd.n_ever_active_sharded(model, x[lo:hi], threshold=3, counts_fn=counts_fn)
    out["fvu"] = scd.fraction_variance_unexplained_sharded(model, x[lo:hi], rss_fn=rss_fn)
    out["rows"] = (lo, hi)
    torch.save(out, os.path.join(tmp, f"rows{rank}.pt"))
    dist.destroy_process_group()


def test_row_sharded_fista_and_metrics_equal_the_unsharded_ones(tmp_path):
    """FISTA rows are independent given D and eta; the metrics need the [N] counts and the two
    FVU sums added across ranks.  World 2 over gloo with the CPU oracle as the local engine, ragged split (101 rows)."""
    torch.manual_seed(3)
    n, d, N = 101, 12, 20
    D = R.row_normalise(torch.randn(N, d))[0]
    x = torch.randn(n, d) + 0.3
    We, b, Wd = torch.randn(N, d) / d ** 0.5, 0.05 * torch.randn(N) - 0.2, torch.randn(N, d)
    a0 = torch.relu(x @ D.T)
    torch.save({"x": x, "D": D, "a0": a0, "We": We, "b": b, "Wd": Wd}, tmp_path / "inputs.pt")
    world = 2
    mp.spawn(_rows_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(tmp_path / f"rows{k}.pt", weights_only=False) for k in range(world)]
    a_ref, res_ref = R.fista(x, D, 1e-2, a0, 25)
    assert torch.allclose(torch.cat([r[0]["a"], r[1]["a"]]), a_ref, atol=1e-6)
    assert torch.allclose(torch.cat([r[0]["res"], r[1]["res"]]), res_ref, atol=1e-6)
    c = R.untied_encode(We, b, x)
    for k in range(world):
        assert torch.allclose(r[k]["mnz"], R.mean_nonzero_activations(c), atol=1e-6)
        assert r[k]["ever"] == int(((c != 0).float().sum(0) > 3).sum())
        want = R.fraction_variance_unexplained(x, R.untied_predict(We, b, Wd, x))
        assert abs(float(r[k]["fvu"]) - float(want)) <= 1e-5 * abs(float(want))
