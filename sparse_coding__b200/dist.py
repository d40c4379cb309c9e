"""Multi-GPU use of the engine, one process per GPU over torch.distributed.

* Ensembles shard with NO data-path collective: every rank builds a FunctionalEnsemble from its own
  slice of the members (`shard_members`) and steps it on the same batch - what the reference does
  with one process per device (cluster_runs.py:100-157).
* A single large dictionary trains data-parallel (reference analogue: DDP over gloo in
  experiments/huge_batch_size.py:259-335): `DataParallelEnsemble` shards the batch rows, runs the
  fused step in gradient-only mode with the GLOBAL batch size in the loss normalisation, sums the
  flat gradient buffer with ONE all-reduce (NCCL over NVLink on GPUs), then applies Adam with
  `sc_adam_apply`.  The gradient through the decoder row-normalisation is linear in the batch, so
  all-reducing the projected per-rank gradients is exact.
* FISTA and the metrics shard by ROWS: codes of different rows are independent given the dictionary, so
  `fista_rows` runs each rank's rows with no data collective (only the step size is agreed on once), and the
  `*_sharded` metrics add per-rank partial results - the [N] per-feature counts and the FVU sums - with one small
  all-reduce each (standard_metrics.py:303-312,439-452).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_members(items: Sequence, rank: int, world: int) -> List:
    """Contiguous, balanced slice of the ensemble members for this rank (first ranks get the extra)."""
    n = len(items)
    base, extra = divmod(n, world)
    start = rank * base + min(rank, extra)
    return list(items[start:start + base + (1 if rank < extra else 0)])


def shard_rows(n_rows: int, rank: int, world: int):
    """[start, stop) of the batch rows this rank owns (balanced, contiguous)."""
    base, extra = divmod(n_rows, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class DataParallelEnsemble:
    """Data-parallel wrapper: same parameters on every rank, different batch rows.

    `engine` is a FunctionalEnsemble (or anything with `compute_grads(x, batch_total=, bias_decay=)`,
    `apply_grads(grads)`, `.buffers`, `.params`).  `step_batch(local_rows)` returns the same
    (loss_dict, aux) as FunctionalEnsemble.step_batch, with global-batch losses and this rank's codes.
    """

    def __init__(self, engine, group=None):
        self.engine = engine
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._flat = None

    def broadcast_parameters(self, src: int = 0):
        """Make every rank start from rank `src`'s parameters (DDP does this at construction)."""
        if self.world == 1:
            return
        for t in self.engine.params.values():
            dist.broadcast(t, src=src, group=self.group)

    def _grad_views(self, n_extra: int):
        """One flat buffer holding every gradient (+ n_extra loss scalars): the fused step writes into views of it,
        so a single in-place all-reduce covers everything and nothing is copied."""
        params = self.engine.params
        n = sum(p.numel() for p in params.values()) + n_extra
        ref = next(iter(params.values()))
        if self._flat is None or self._flat.numel() != n or self._flat.device != ref.device:
            self._flat = torch.empty(n, dtype=torch.float32, device=ref.device)
            self._views, off = {}, 0
            for k, p in params.items():
                self._views[k] = self._flat[off:off + p.numel()].view(p.shape)
                off += p.numel()
            self._extra = self._flat[off:]
        return self._views, self._extra

    def step_batch(self, local_rows: torch.Tensor, expand_dims=True):
        eng, world = self.engine, self.world
        # global batch size: one small all-reduce + host read, repeated only when this rank's row count changes
        # (ranks are expected to change their row counts together, as a sharded sampler does)
        n_local = int(local_rows.shape[-2])
        if getattr(self, "_b_cache", (None, None))[0] != n_local:
            b_local = torch.tensor([n_local], device=local_rows.device, dtype=torch.int64)
            if world > 1:
                dist.all_reduce(b_local, group=self.group)
            self._b_cache = (n_local, int(b_local.item()))
        b_total = self._b_cache[1]
        # the bias-decay term does not depend on the batch: give each rank 1/world of it so the sum is right
        bd = eng.buffers["bias_decay"].to(torch.float32) / world
        M = eng.params["encoder"].shape[0]
        can_view = all(hasattr(p, "is_cuda") and p.is_cuda and p.is_contiguous() for p in eng.params.values())
        if can_view:
            views, extra = self._grad_views(3 * M)
            grads, losses, aux = eng.compute_grads(local_rows, expand_dims=expand_dims, batch_total=b_total, bias_decay=bd,
                                                   out=views)
            parts = extra.view(3, M)
            parts.copy_(torch.stack([losses["l_reconstruction"], losses["l_l1"], losses["l_bias_decay"]]))
            if world > 1:
                dist.all_reduce(self._flat, group=self.group)
            parts = parts.clone()
        else:  # engines without preallocated-output support (the CPU stand-ins of the gloo tests)
            grads, losses, aux = eng.compute_grads(local_rows, expand_dims=expand_dims, batch_total=b_total, bias_decay=bd)
            parts = torch.stack([losses["l_reconstruction"], losses["l_l1"], losses["l_bias_decay"]]).to(torch.float32)
            if world > 1:
                flat = torch.cat([g.reshape(-1) for g in grads.values()] + [parts.reshape(-1)])
                dist.all_reduce(flat, group=self.group)
                off = 0
                for k, g in grads.items():
                    g.copy_(flat[off:off + g.numel()].view_as(g))
                    off += g.numel()
                parts = flat[off:].view_as(parts).clone()
        eng.apply_grads(grads)
        out = {"loss": parts.sum(0), "l_reconstruction": parts[0], "l_l1": parts[1]}
        if "l_bias_decay" in losses and not getattr(eng, "_kind", "") == "tied":
            out["l_bias_decay"] = parts[2]
        return out, aux


class _SymmetricFabric:
    """Peer-mapped buffers and cross-rank barriers from torch.distributed._symmetric_memory (plumbing only: the
    data moves in this library's kernels, through the peer addresses handed out here)."""

    def __init__(self, group=None):
        import torch.distributed._symmetric_memory as symm
        self._symm = symm
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self._handles = []

    def empty(self, shape, device):
        """-> (local tensor, [device address of every rank's tensor, usable from THIS rank's kernels])"""
        t = self._symm.empty(*shape, dtype=torch.float32, device=device)
        h = self._symm.rendezvous(t, self.group)
        self._handles.append(h)
        return t, [int(p) for p in h.buffer_ptrs]

    def barrier(self):
        self._handles[0].barrier()  # on the current stream: signals exchanged with system-scope release / acquire


class FusedDataParallel:
    """Data-parallel training of ONE dictionary with the gradient collective fused into the kernels (SURVEY 8e:
    reduce-scatter + sharded Adam + parameter broadcast, overlapped with the weight-gradient GEMMs) - no NCCL call
    on the data path.

    Rank r owns dictionary rows [r N/W, (r+1) N/W).  Per step every rank runs the fused step on its batch rows in
    scatter mode: the weight-gradient kernel's epilogue ADDS each 32 x 32 gradient tile straight into the owner's
    accumulation buffer over NVLink (TMA reduce-add on peer memory), so the reduction traffic is spread under the
    GEMMs; after one barrier the owner runs Adam on its slice and stores the new values into every rank's parameter
    copy (`sc_dp_adam_broadcast`), after a second barrier all copies agree.  Summation order across ranks is not
    fixed, so results agree with the all-reduce path to fp32 rounding, not bitwise.

    `engine`: a FunctionalEnsemble with one member; its parameters are moved into peer-mapped memory.  Every rank must
    pass the same number of batch rows.  `fabric` (tests): anything with world / rank / empty() / barrier().
    """

    def __init__(self, engine, group=None, fabric=None):
        if engine.n_models != 1:
            raise ValueError("FusedDataParallel trains one dictionary (n_models == 1); shard ensembles by member instead")
        self.engine = engine
        self.fabric = fabric if fabric is not None else _SymmetricFabric(group)
        self.world, self.rank = self.fabric.world, self.fabric.rank
        enc = engine.params["encoder"]
        _, N, d = enc.shape
        if N % (32 * self.world):
            raise ValueError(f"dictionary size {N} must be a multiple of 32 * world ({32 * self.world})")
        self.rows, self.d, self.N = N // self.world, d, N
        dev = enc.device
        self.keys = [k for k in ("encoder", "decoder") if k in engine.params]
        self._peer_params = {}
        for k in list(engine.params):  # parameters live in peer-mapped memory from here on
            t, ptrs = self.fabric.empty(tuple(engine.params[k].shape), dev)
            t.copy_(engine.params[k])
            engine.params[k] = t
            self._peer_params[k] = ptrs
        # this rank's accumulation buffers: summed gradient of its rows of each matrix, of its bias slice, loss terms
        self._off = {}
        n = 0
        for k in self.keys:
            self._off[k] = n
            n += self.rows * d
        self._off["encoder_bias"] = n
        n += self.rows
        self._off["losses"] = n
        n += 4
        self.acc, self._peer_acc = self.fabric.empty((n,), dev)
        self.acc.zero_()
        torch.cuda.synchronize(dev)
        self.fabric.barrier()

    def _rs_args(self):
        a = lambda key: [p + 4 * self._off[key] for p in self._peer_acc]
        rs = {"world": self.world, "rows": self.rows, "grad_enc": a("encoder"), "grad_bias": a("encoder_bias"),
              "losses": a("losses")}
        if "decoder" in self._off:
            rs["grad_dec"] = a("decoder")
        return rs

    def scatter(self, local_rows, expand_dims=True):
        """Phase 1: forward, backward, gradients added into their owners' buffers.  Returns aux."""
        eng = self.engine
        bd = eng.buffers["bias_decay"].to(torch.float32) / self.world  # each rank contributes 1/W of the bias-decay term
        _, aux = eng.scatter_grads(local_rows, self._rs_args(), expand_dims=expand_dims,
                                   batch_total=local_rows.shape[-2] * self.world, bias_decay=bd)
        return aux

    def update(self):
        """Phase 2 (after every rank finished phase 1): Adam on the owned slices, broadcast into all copies."""
        import ctypes as C
        from . import _native as nat
        eng, L = self.engine, nat.lib()
        adam = eng.adam_constants(advance=True)
        st = eng.optim_states[0]
        parts = self.acc[self._off["losses"]:self._off["losses"] + 3].clone()
        self.acc[self._off["losses"]:].zero_()
        lo, hi = self.rank * self.rows, (self.rank + 1) * self.rows
        for k in self.keys + ["encoder_bias"]:
            per_row = self.d if k != "encoder_bias" else 1
            n = self.rows * per_row
            g = self.acc[self._off[k]:self._off[k] + n]
            p, mu, nu = eng.params[k][0, lo:hi], st.mu[k][0, lo:hi], st.nu[k][0, lo:hi]
            dst = (C.c_void_p * self.world)(*[ptr + 4 * lo * per_row for ptr in self._peer_params[k]])
            with torch.cuda.device(p.device):
                nat.check(L.sc_dp_adam_broadcast(nat.ptr(g), nat.ptr(mu), nat.ptr(nu), nat.ptr(p), dst, self.world, n,
                                                 C.byref(adam), nat.stream_ptr(p.device)), "sc_dp_adam_broadcast")
        out = {"loss": parts.sum(0, keepdim=True), "l_reconstruction": parts[0:1], "l_l1": parts[1:2]}
        if getattr(eng, "_kind", "") != "tied":
            out["l_bias_decay"] = parts[2:3]
        return out

    def step_batch(self, local_rows, expand_dims=True):
        aux = self.scatter(local_rows, expand_dims)
        self.fabric.barrier()   # every contribution has landed in its owner's buffer
        out = self.update()
        self.fabric.barrier()   # every copy of the parameters is complete before anyone's next step reads it
        return out, aux


# --------------------------------------------------------------------------------------------------------------------
# Row-sharded FISTA and metrics.  Every function takes THIS rank's rows; results that describe rows stay local, results
# that describe the dictionary (per-feature counts, FVU) are identical on every rank after one all-reduce.  The local
# computation can be swapped (`*_fn` arguments): the gloo tests run the CPU oracle there, the default is the CUDA engine.

def _world(group):
    return dist.get_world_size(group) if dist.is_initialized() else 1


def _all_reduce_(t, group):
    if _world(group) > 1:
        dist.all_reduce(t, group=group)
    return t


def fista_rows(local_batch, learned_dict, l1_coef, local_coefficients, num_iter, device=None, eta=None, group=None,
               fista_fn=None, eta_fn=None):
    """FunctionalFista.fista (fista.py:99-128) on this rank's rows of a row-sharded batch: (ahat_local, Res_local).
    The step size eta = 1/lambda_max(D D^T) depends on the dictionary only; when it is not given, rank 0's value is
    broadcast so that every rank iterates with bit-identical eta (no other communication: rows are independent)."""
    if fista_fn is None:
        from .autoencoders.fista import FunctionalFista, fista_native
        fista_fn = lambda x, D, l1, a0, it, eta: FunctionalFista.fista(x, D, l1, a0, it, device, eta=eta)
        if eta_fn is None:
            eta_fn = lambda x, D, l1, a0: fista_native(x[:1], D, l1, a0[:1], 0)[2].reshape(1)
    if eta is None:
        if eta_fn is None:
            raise ValueError("fista_rows: give eta or eta_fn together with a custom fista_fn")
        eta_t = eta_fn(local_batch, learned_dict, l1_coef, local_coefficients).to(torch.float32).reshape(1).clone()
        if _world(group) > 1:
            dist.broadcast(eta_t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        eta = float(eta_t.item())
    return fista_fn(local_batch, learned_dict, l1_coef, local_coefficients, num_iter, eta)


def mean_nonzero_activations_sharded(model, local_batch, group=None, counts_fn=None):
    """standard_metrics.py:303-306 over a row-sharded batch: [N] fraction of ALL rows on which each feature fires
    (callers sum it for the mean L0).  One all-reduce of N + 1 floats."""
    if counts_fn is None:
        from . import standard_metrics as sm
        counts_fn = sm.feature_nonzero_counts
    xc = model.center(local_batch)
    counts = counts_fn(model, xc).to(torch.float64)
    pack = torch.cat([counts, counts.new_tensor([float(xc.shape[0])])])
    _all_reduce_(pack, group)
    return (pack[:-1] / pack[-1]).to(torch.float32)


def n_ever_active_sharded(model, local_activations, threshold=10, group=None, counts_fn=None):
    """batched_calc_feature_n_ever_active (standard_metrics.py:444-452) over row-sharded (already centred)
    activations: per-feature non-zero counts are added across ranks before the threshold."""
    if counts_fn is None:
        from . import standard_metrics as sm
        counts_fn = sm.feature_nonzero_counts
    counts = counts_fn(model, local_activations).to(torch.float64)
    _all_reduce_(counts, group)
    return int((counts > threshold).sum().item())


def fraction_variance_unexplained_sharded(model, local_batch, group=None, rss_fn=None):
    """standard_metrics.py:308-312 over a row-sharded batch.  The denominator needs the column means of ALL rows: first
    all-reduce (d + 1 numbers: column sums and the row count), then the two sums of squares (second all-reduce of 2)."""
    if rss_fn is None:
        from . import standard_metrics as sm
        rss_fn = sm.residual_sum_of_squares
    x = local_batch
    first = torch.cat([x.sum(0, dtype=torch.float64), torch.tensor([float(x.shape[0])], dtype=torch.float64, device=x.device)])
    _all_reduce_(first, group)
    mean = (first[:-1] / first[-1]).to(x.dtype)
    sums = torch.stack([rss_fn(model, x).to(torch.float64).reshape(()), (x - mean).pow(2).sum(dtype=torch.float64)])
    _all_reduce_(sums, group)
    return (sums[0] / sums[1]).to(torch.float32)
