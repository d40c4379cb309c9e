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

    def _flatten(self, grads: Dict[str, torch.Tensor], extra: torch.Tensor) -> torch.Tensor:
        n = sum(g.numel() for g in grads.values()) + extra.numel()
        ref = next(iter(grads.values()))
        if self._flat is None or self._flat.numel() != n or self._flat.device != ref.device:
            self._flat = torch.empty(n, dtype=ref.dtype, device=ref.device)
        off = 0
        for g in grads.values():
            self._flat[off:off + g.numel()].copy_(g.reshape(-1))
            off += g.numel()
        self._flat[off:].copy_(extra.reshape(-1))
        return self._flat

    def step_batch(self, local_rows: torch.Tensor, expand_dims=True):
        eng, world = self.engine, self.world
        b_local = torch.tensor([local_rows.shape[-2]], device=local_rows.device, dtype=torch.int64)
        if world > 1:
            dist.all_reduce(b_local, group=self.group)
        b_total = int(b_local.item())
        # the bias-decay term does not depend on the batch: give each rank 1/world of it so the sum is right
        bd = eng.buffers["bias_decay"].to(torch.float32) / world
        grads, losses, aux = eng.compute_grads(local_rows, expand_dims=expand_dims, batch_total=b_total, bias_decay=bd)
        parts = torch.stack([losses["l_reconstruction"], losses["l_l1"], losses["l_bias_decay"]]).to(torch.float32)
        if world > 1:
            flat = self._flatten(grads, parts)
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
