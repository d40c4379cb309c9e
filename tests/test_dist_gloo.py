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
