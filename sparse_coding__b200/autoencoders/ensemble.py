"""FunctionalEnsemble on the B200 engine - host-side mirror of the reference's
autoencoders/ensemble.py:15-193 (same constructor, attributes, state_dict keys and
`step_batch` contract).  The vmap(grad(loss)) -> vmap(adam.update) -> apply_updates chain
(ensemble.py:175-193) is one call into libsc_b200 (`sc_sae_step`).
"""
from __future__ import annotations

import ctypes as C
import weakref
from collections.abc import Mapping
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _native as nat
from ..torchopt_compat import EmptyState, ScaleByAdamState, spec_from


class DictSignature:
    """ensemble.py:15-22."""

    @staticmethod
    def to_learned_dict(params, buffers):
        pass

    @staticmethod
    def loss(params, buffers, batch):
        pass


def optim_str_to_func(optim_str):
    """ensemble.py:25-31."""
    from .. import torchopt_compat
    if optim_str == "adam":
        return torchopt_compat.adam
    raise ValueError("Unknown optimizer string: {}".format(optim_str))


def stack_dict(models: list, device=None) -> Dict[str, torch.Tensor]:
    """ensemble.py:50-56: stack the same leaf of every member on a new leading dim."""
    keys = list(models[0].keys())
    out = {}
    for k in keys:
        ts = [m[k] for m in models]
        out[k] = torch.stack([t.detach() for t in ts]).to(device=device)
    return out


def unstack_dict(params: Dict[str, torch.Tensor], n_models: int, device=None) -> List[Dict[str, torch.Tensor]]:
    """ensemble.py:59-65."""
    return [{k: t[i].to(device=device) for k, t in params.items()} for i in range(n_models)]


def _sig_kind(sig) -> str:
    kind = getattr(sig, "_sc_kind", None)
    if kind is None:
        kind = {"FunctionalSAE": "untied", "FunctionalFista": "untied", "FunctionalTiedSAE": "tied",
                "FunctionalMaskedSAE": "masked_untied", "FunctionalMaskedTiedSAE": "masked_tied"}.get(
            getattr(sig, "__name__", ""), None)
    if kind is None:
        raise NotImplementedError(
            f"signature {sig!r}: the B200 engine implements FunctionalSAE, FunctionalFista, FunctionalTiedSAE and the "
            "two Masked variants; "
            "other signatures have no fused path (and there is no generic fallback)")
    return kind


class _StepCtx:
    """Everything of a plain train step that stays the same while the same tensors sit in the same places."""
    __slots__ = ("args", "xshape", "device", "dev_index", "expand_dims", "fmt", "f16", "tied", "masked", "dims", "st",
                 "tensors", "buffers", "keep", "ws_need", "enc", "dec")

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)


class LazyCodes(Mapping):
    """aux of step_batch in fp16-operand mode: behaves like the reference's {"c": codes} (sae_ensemble.py:73-75).
    The engine keeps the codes as an fp16 TRANSPOSED tensor c^T [M, N, ldb] (that is what its GEMMs read,
    `codes_t`); the float32 [M, B, N] tensor the reference returns is produced by `sc_codes_to_float` on first
    access, so callers that ignore `aux` - most do - never pay for the 4*M*B*N-byte write.  Every access path
    (indexing, iteration, dict(aux), {**aux}, copy, pickle) goes through __getitem__ and sees the real tensor."""

    def __init__(self, codes_t, batch_rows, code_scale=None):
        self.codes_t = codes_t
        self.code_scale = code_scale  # [M, N] factor per feature (the untied step stores c / ||W_d||), or None
        self.batch_rows = int(batch_rows)
        self._c = None

    def _materialise(self):
        if self._c is None:
            M, N, _ = self.codes_t.shape
            out = torch.empty((M, self.batch_rows, N), dtype=torch.float32, device=self.codes_t.device)
            with torch.cuda.device(out.device):
                nat.check(nat.lib().sc_codes_to_float(nat.ptr(self.codes_t), nat.ptr(self.code_scale), nat.ptr(out), M,
                                                      self.batch_rows, N, nat.stream_ptr(out.device)), "sc_codes_to_float")
            self._c = out
        return self._c

    def __getitem__(self, key):
        if key == "c":
            return self._materialise()
        raise KeyError(key)

    def __iter__(self):
        return iter(("c",))

    def __len__(self):
        return 1

    def __reduce__(self):
        return (dict, ({"c": self._materialise()},))

    def __copy__(self):
        return {"c": self._materialise()}

    def copy(self):
        return self.__copy__()

    def __repr__(self):
        return f"LazyCodes(M={self.codes_t.shape[0]}, B={self.batch_rows}, N={self.codes_t.shape[1]})"


class FunctionalEnsemble:
    """Drop-in for reference ensemble.py:68-193.

    `operand_format`: "auto" (default: fp16 operands when the shape allows it, else TF32), "tf32" or "f16"
    (env SC_B200_OPERANDS overrides the default).  Both keep an 11-bit significand on every GEMM operand
    and fp32 accumulation / master weights; fp16 halves the operand bytes and doubles the MMA rate but
    overflows for |activation| > 65504.

    `reference_faithful` (default True) reproduces the reference's Adam as it actually runs:
    the stored step count is never advanced (ensemble.py:185-189 writes the new optimizer state
    into a clone), so bias correction stays at t = 1.  Set it to False for textbook Adam.
    """

    def __init__(self, models, sig, optimizer_func, optimizer_kwargs, device=None, no_stacking=False):
        if device is None:
            self.device = models[0][0]["encoder"].device
        else:
            self.device = device
        if torch.device(self.device).type != "cuda":
            raise NotImplementedError(f"device {self.device}: this engine runs on CUDA sm_100a only (no CPU path)")
        self.n_models = len(models)
        params, buffers = tuple(zip(*models))
        self.params = stack_dict(list(params), device=self.device)
        self.buffers = stack_dict(list(buffers), device=self.device)
        self.sig = sig
        self.no_stacking = no_stacking
        self.optimizer_func = optimizer_func
        self.optimizer_kwargs = optimizer_kwargs
        self.optim_states = self._init_optim_states()
        self.reference_faithful = True
        self.init_functions()

    # ------------------------------------------------------------------ set-up
    def _init_optim_states(self):
        mu = {k: torch.zeros_like(v) for k, v in self.params.items()}
        nu = {k: torch.zeros_like(v) for k, v in self.params.items()}
        count = {k: torch.zeros(self.n_models, dtype=torch.int64, device=self.device) for k in self.params}
        return (ScaleByAdamState(mu=mu, nu=nu, count=count), EmptyState())

    def init_functions(self):
        """ensemble.py:99-123 builds the vmapped closures; here: resolve the fused path."""
        self._kind = _sig_kind(self.sig)
        self._adam = spec_from(self.optimizer_func, self.optimizer_kwargs)
        self._ws = nat.Workspace()
        self._step = 0
        if not hasattr(self, "operand_format"):
            import os
            self.operand_format = os.environ.get("SC_B200_OPERANDS", "auto")
        if not hasattr(self, "reference_faithful"):
            self.reference_faithful = True
        nat.lib()  # fail now, loudly, if the native library is missing

    @staticmethod
    def from_state(state_dict):
        """ensemble.py:125-143."""
        self = FunctionalEnsemble.__new__(FunctionalEnsemble)
        self.device = state_dict["device"]
        self.n_models = state_dict["n_models"]
        self.params = state_dict["params"]
        self.buffers = state_dict["buffers"]
        self.sig = state_dict["sig"]
        self.no_stacking = state_dict["no_stacking"]
        self.optimizer_func = state_dict["optimizer_func"]
        self.optimizer_kwargs = state_dict["optimizer_kwargs"]
        self.optim_states = state_dict["optim_states"]
        self.reference_faithful = state_dict.get("reference_faithful", True)
        self.init_functions()
        if not self.reference_faithful:  # textbook Adam continues from the stored step count
            self._step = int(self.optim_states[0].count["encoder"].max())
        return self

    def unstack(self, device=None):
        """ensemble.py:145-148."""
        params = unstack_dict(self.params, self.n_models, device=device)
        buffers = unstack_dict(self.buffers, self.n_models, device=device)
        return list(zip(params, buffers))

    def state_dict(self):
        """ensemble.py:150-161 (same nine keys)."""
        sd = {
            "device": self.device,
            "n_models": self.n_models,
            "params": self.params,
            "buffers": self.buffers,
            "sig": self.sig,
            "no_stacking": self.no_stacking,
            "optimizer_func": self.optimizer_func,
            "optimizer_kwargs": self.optimizer_kwargs,
            "optim_states": self.optim_states,
        }
        if not self.reference_faithful:  # a tenth key only for the non-default (textbook Adam) mode
            sd["reference_faithful"] = False
        return sd

    def _map_state(self, fn):
        st = self.optim_states[0]
        self.optim_states = (ScaleByAdamState(mu={k: fn(v) for k, v in st.mu.items()},
                                              nu={k: fn(v) for k, v in st.nu.items()},
                                              count={k: fn(v) for k, v in st.count.items()}), EmptyState())

    def to_device(self, device):
        """ensemble.py:163-168."""
        self.device = device
        self.params = {k: v.to(device) for k, v in self.params.items()}
        self.buffers = {k: v.to(device) for k, v in self.buffers.items()}
        self._map_state(lambda t: t.to(device))

    def to_shared_memory(self):
        """ensemble.py:170-173 (CPU tensors only, as in torch)."""
        for v in self.params.values():
            v.share_memory_()
        for v in self.buffers.values():
            v.share_memory_()
        st = self.optim_states[0]
        for group in (st.mu, st.nu, st.count):
            for v in group.values():
                v.share_memory_()

    def reinit_dead_features(self, member, c_totals, worst_vectors, encoder_norm_ratio=0.2):
        """Dead-feature re-initialisation of experiments/huge_batch_size.py:224-254 for one ensemble member (that
        script keeps its encoder as [d, N]; here dictionaries are [N, d], so columns there are rows here): features
        whose summed activation `c_totals` [N] is zero get the worst-reconstructed data vectors `worst_vectors`
        [n_dead, d] as encoder rows, scaled to `encoder_norm_ratio` / (mean encoder row norm), and the Adam moments
        of every parameter of those features are reset.  Returns the indices that were replaced."""
        dead = torch.where(c_totals.to(self.params["encoder"].device) == 0)[0]
        n = int(dead.numel())
        if n == 0:
            return dead
        wv = torch.as_tensor(worst_vectors, dtype=torch.float32, device=dead.device)
        if wv.dim() == 1:  # huge_batch_size.py:234-235, n_replace == 1
            wv = wv.unsqueeze(0)
        if wv.shape[0] < n:
            raise ValueError(f"{n} dead features but only {wv.shape[0]} replacement vectors")
        enc = self.params["encoder"][member]
        av_norm = torch.norm(enc, dim=1).mean()
        enc[dead] = wv[:n] * encoder_norm_ratio / av_norm
        st = self.optim_states[0]
        for group in (st.mu, st.nu):
            for v in group.values():
                v[member][dead] = 0
        return dead

    # ------------------------------------------------------------------ the hot path
    def _contig(self, store: dict, key: str, shape=None) -> torch.Tensor:
        t = store[key]
        if not t.is_cuda or t.dtype != torch.float32:
            nat.require_cuda(t, key)
        if not t.is_contiguous() or t.data_ptr() % 16:
            t = t.contiguous().clone() if t.data_ptr() % 16 else t.contiguous()
            store[key] = t  # callers assign into .params (big_sweep.py:198); keep the rebinding visible
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError(f"{key}: expected shape {tuple(shape)}, got {tuple(t.shape)}")
        return t

    def _scalar_buffer(self, key: str) -> torch.Tensor:
        """Per-member scalars as a contiguous float32 [M] device vector.  Recipes often build them from
        numpy float64 values (basic_l1_sweep.py:71); the stored buffer keeps its dtype, the kernel gets a
        cached float32 copy."""
        t = self.buffers[key]
        if t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and tuple(t.shape) == (self.n_models,):
            return t
        cache = self.__dict__.setdefault("_scalar_cache", {})
        tag = (t.data_ptr(), t._version, t.dtype)
        hit = cache.get(key)
        if hit is None or hit[0] != tag:
            v = t.detach().to(device=self.device, dtype=torch.float32).reshape(self.n_models).contiguous()
            cache[key] = (tag, v)
            hit = cache[key]
        return hit[1]

    def _check_prefix_mask(self, dict_size):
        """The fused step masks feature n of member m iff n >= dict_size[m]; the reference masks with
        buffers["coef_mask"] (sae_ensemble.py:356, :425), which `init` writes as exactly that prefix mask.  A caller
        who edits coef_mask to anything else would be silently ignored here, so check (once per buffer version)."""
        cm = self.buffers.get("coef_mask")
        if cm is None:
            return
        key = (cm.data_ptr(), cm._version, dict_size.data_ptr(), self.buffers["dict_size"]._version)
        if getattr(self, "_mask_key", None) == key:
            return
        N = cm.shape[-1]
        want = torch.arange(N, device=cm.device)[None, :] >= dict_size.to(cm.device)[:, None]
        if not torch.equal(cm.to(torch.bool).reshape(self.n_models, N), want):
            raise NotImplementedError("buffers['coef_mask'] is not the prefix mask arange(N) >= dict_size that init writes; "
                                      "the fused step only implements that form")
        self._mask_key = key

    def _centered_batch(self, x: torch.Tensor, expand: bool):
        """FunctionalTiedSAE.center (sae_ensemble.py:126-127) with per-member buffers.  Identity
        buffers (the default, :97-108) skip it; otherwise the batch is centred once per step with one batched
        [B x d x d] product on the tcgen05 core at fp32 accuracy (sc_gemm_f32) - a constant pre-transform of the
        input, not part of the trained graph."""
        b = self.buffers
        rot, tr, sc = b["center_rot"], b["center_trans"], b["center_scale"]
        key = (rot.data_ptr(), rot._version, tr.data_ptr(), tr._version, sc.data_ptr(), sc._version)
        if getattr(self, "_center_key", None) != key:
            d = rot.shape[-1]
            eye = torch.eye(d, device=rot.device, dtype=rot.dtype)
            self._center_identity = bool((rot == eye).all() and (tr == 0).all() and (sc == 1).all())
            self._center_key = key
        if self._center_identity:
            return x, not expand
        from .learned_dict import center_native
        xm = x.unsqueeze(0) if expand else x
        xc = center_native(xm, rot.to(torch.float32), tr.to(torch.float32), sc.to(torch.float32))
        return xc.contiguous(), True

    # ------------------------------------------------------------------ fp16 range guard
    def _poll_range_flag(self, device):
        """fp16 operands hold |x| <= 65504.  The step saturates instead of overflowing and raises a device flag
        (ScStepArgs.range_flag) when a batch leaves that range; the flag travels to pinned host memory every few steps
        without a sync, and is looked at here before the next launch: "auto" then continues with TF32 operands (which
        have fp32's range), an explicit "f16" keeps going and warns."""
        if getattr(self, "_range_flag", None) is None or self._range_flag.device != torch.device(device):
            self._range_flag = torch.zeros(1, dtype=torch.int32, device=device)
            self._range_host = torch.zeros(1, dtype=torch.int32).pin_memory()
            self._range_tag, self._range_seen, self._range_event = 0, 0, None
        self._range_tag += 1
        if self._range_event is not None and self._range_event.query():
            self._range_event = None
            self._handle_range_flag(int(self._range_host[0]))
        if self._range_event is None and self._range_tag % 8 == 0:
            self._range_host.copy_(self._range_flag, non_blocking=True)
            self._range_event = torch.cuda.Event()
            self._range_event.record(torch.cuda.current_stream(device))

    def _handle_range_flag(self, tag):
        if tag <= self._range_seen:
            return
        first, self._range_seen = tag, tag
        import warnings
        if self.operand_format == "auto":
            self.operand_format = "tf32"
            warnings.warn(f"a batch (step {first} of this ensemble) held activations beyond the fp16 range (|x| > 65504 or "
                          "non-finite); they were saturated in that step.  operand_format 'auto' continues with TF32 operands.")
        else:
            warnings.warn(f"a batch (step {first}) held activations beyond the fp16 range (|x| > 65504 or non-finite); they "
                          "were saturated.  Use operand_format 'tf32' for such data.")

    def check_operand_range(self):
        """One sync: look at the fp16 range flag now (what `_poll_range_flag` does lazily).  Returns True when every
        batch so far was inside the fp16 range."""
        if getattr(self, "_range_flag", None) is None:
            return True
        self._handle_range_flag(int(self._range_flag.item()))
        return self._range_seen == 0

    def step_batch(self, minibatches, expand_dims=True):
        """ensemble.py:175-193.  Returns (loss_dict of [M] tensors, {"c": [M,B,N]})."""
        return self._run(minibatches, expand_dims, grad_only=False)

    def compute_grads(self, minibatches, expand_dims=True, batch_total=0, bias_decay=None, out=None):
        """Gradient-only twin of step_batch: what vmap(grad(sig.loss)) returns at ensemble.py:180,
        without the Adam update.  `batch_total` (> 0) replaces the local batch size in the loss
        normalisation (data-parallel shards); `bias_decay` overrides buffers["bias_decay"].
        `out`: optional dict of preallocated contiguous tensors shaped like .params to receive the gradients (e.g.
        views into one flat buffer that is all-reduced in place).
        Returns (grads dict like .params, loss_dict, {"c": ...})."""
        return self._run(minibatches, expand_dims, grad_only=True, batch_total=batch_total, bias_decay=bias_decay,
                         grad_out=out)

    def adam_constants(self, advance=True):
        """(ScAdam for this step)."""
        if advance:
            self._step += 1
        t = 1 if self.reference_faithful else max(self._step, 1)
        a = self._adam
        # torchopt: mu.div(1 - b1 ** count_inc) with a float32 tensor power
        bc1 = float(np.float32(1.0) - np.float32(a.b1) ** np.float32(t))
        bc2 = float(np.float32(1.0) - np.float32(a.b2) ** np.float32(t))
        if advance and not self.reference_faithful:
            for cnt in self.optim_states[0].count.values():
                cnt += 1
        return nat.ScAdam(a.lr, a.b1, a.b2, a.eps, a.eps_root, bc1, bc2)

    def apply_grads(self, grads):
        """vmap(optimizer.update) + apply_updates (ensemble.py:182-191) from explicit gradients."""
        L = nat.lib()
        adam = self.adam_constants(advance=True)
        self._ops_key = None  # the parameters change outside the fused step: its operand hand-over is void
        st = self.optim_states[0]
        for k, g in grads.items():
            p = self._contig(self.params, k)
            mu, nu = self._contig(st.mu, k, p.shape), self._contig(st.nu, k, p.shape)
            g = g.contiguous()
            with torch.cuda.device(p.device):
                nat.check(L.sc_adam_apply(nat.ptr(p), nat.ptr(mu), nat.ptr(nu), nat.ptr(g), p.numel(), C.byref(adam),
                                          nat.stream_ptr(p.device)), "sc_adam_apply")

    def scatter_grads(self, minibatches, reduce_scatter, expand_dims=True, batch_total=0, bias_decay=None):
        """Fused reduce-scatter twin of compute_grads for ONE dictionary trained data-parallel (dist.FusedDataParallel):
        the gradients are not returned but ADDED, tile by tile, into their owner rank's accumulation buffers
        (`reduce_scatter`: {"world", "rows", "grad_enc", "grad_dec", "grad_bias", "losses"} with one device address
        per rank each - ScStepArgs.rs_*).  Returns (loss_dict of this rank's share, aux)."""
        _, losses, aux = self._run(minibatches, expand_dims, grad_only=True, batch_total=batch_total, bias_decay=bias_decay,
                                   reduce_scatter=reduce_scatter)
        return losses, aux

    def _run_cached(self, x, expand_dims):
        """step_batch when nothing but the batch changed since the last plain step: the argument record of that step is
        reused (only the batch, the fresh output tensors, the hand-over flag and the stream are written into it), which
        halves the host time per step - at the toy shapes the step is bound by it.  Returns None when anything differs
        (another shape / device / format setting, a tensor of .params / .buffers / the optimiser state replaced or a
        buffer edited): the caller then takes the full path, which rebuilds the record."""
        ctx = self._step_ctx
        if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype is torch.float32 and tuple(x.shape) == ctx.xshape
                and x.device == ctx.device and x.is_contiguous() and bool(expand_dims) == ctx.expand_dims
                and self.operand_format == ctx.fmt and self.optim_states[0] is ctx.st):
            return None
        for store, k, t, p in ctx.tensors:
            if store.get(k) is not t or t.data_ptr() != p:
                return None
        buffers = self.buffers
        for k, t, ver in ctx.buffers:
            if buffers.get(k) is not t or t._version != ver:
                return None
        f16 = ctx.f16
        if f16:
            self._poll_range_flag(x.device)
            if self.operand_format != ctx.fmt:  # a batch left the fp16 range: "auto" continues with TF32 operands
                return None
        M, B, d, N = ctx.dims
        args = ctx.args
        args.adam = self.adam_constants(advance=True)
        dev = x.device
        losses = torch.empty((4, M), dtype=torch.float32, device=dev)
        if f16:
            c = None
            c_half = torch.empty((M, N, (B + 7) & ~7), dtype=torch.float16, device=dev)
            code_scale = torch.empty((M, N), dtype=torch.float32, device=dev)
            args.out_c_half, args.out_code_scale = c_half.data_ptr(), code_scale.data_ptr()
            args.step_tag = self._range_tag
        else:
            c = torch.empty((M, B, N), dtype=torch.float32, device=dev)
            args.out_c = c.data_ptr()
        ws_ptr, ws_bytes = self._ws.aligned(ctx.ws_need, dev)
        args.workspace, args.workspace_bytes = ws_ptr, ws_bytes
        ops_key = None
        if f16 and not ctx.tied:
            enc, dec = ctx.enc, ctx.dec
            ops_key = (ws_ptr.value, M, d, N, enc.data_ptr(), enc._version, dec.data_ptr(), dec._version)
        args.operands_valid = int(ops_key is not None and self._ops_key == ops_key)
        self._ops_key = None
        args.x, args.out_losses = x.data_ptr(), losses.data_ptr()
        args.stream = torch.cuda.current_stream(dev).cuda_stream
        L = nat.lib()
        if torch.cuda.current_device() == ctx.dev_index:
            nat.check(L.sc_sae_step(C.byref(args)), "sc_sae_step")
        else:
            with torch.cuda.device(dev):
                nat.check(L.sc_sae_step(C.byref(args)), "sc_sae_step")
        self._ops_key = ops_key
        rows = losses.unbind(0)
        loss_dict = {"loss": rows[0], "l_reconstruction": rows[1], "l_l1": rows[2]}
        if not ctx.tied and not ctx.masked:  # see the key rules at the end of _run
            loss_dict["l_bias_decay"] = rows[3]
        return loss_dict, (LazyCodes(c_half, B, code_scale) if f16 else {"c": c})

    def _run(self, minibatches, expand_dims, grad_only, batch_total=0, bias_decay=None, reduce_scatter=None, grad_out=None):
        plain_step = not grad_only and reduce_scatter is None and bias_decay is None and batch_total == 0
        if plain_step and getattr(self, "_step_ctx", None) is not None:
            out = self._run_cached(minibatches, expand_dims)
            if out is not None:
                return out
        self._step_ctx = None
        L = nat.lib()
        M = self.n_models
        x = minibatches
        nat.require_cuda(x, "minibatches")
        if self.no_stacking and expand_dims and x.dim() == 3:
            # reference ensemble.py:100-116 indexes batch[i] per member in this mode; a 2-D batch is expanded to
            # [M,B,d] first (ensemble.py:177-178), which is the shared-batch path here
            expand_dims = False
        if expand_dims:
            if x.dim() != 2:
                raise ValueError(f"expected a [B, d] batch, got {tuple(x.shape)}")
        else:
            if x.dim() != 3 or x.shape[0] != M:
                raise ValueError(f"expected an [M={M}, B, d] batch, got {tuple(x.shape)}")
        x = x.contiguous()
        masked = self._kind.startswith("masked")
        tied = self._kind in ("tied", "masked_tied")
        per_member = not expand_dims
        if tied and not masked:  # the masked tied signature has no centering buffers (sae_ensemble.py:307-371)
            x, per_member = self._centered_batch(x, expand_dims)
        B, d = x.shape[-2], x.shape[-1]
        enc = self._contig(self.params, "encoder")
        N = enc.shape[1]
        if enc.shape != (M, N, d):
            raise ValueError(f"encoder shape {tuple(enc.shape)} does not match batch width {d}")
        bias = self._contig(self.params, "encoder_bias", (M, N))
        st = self.optim_states[0]
        enc_mu, enc_nu = self._contig(st.mu, "encoder", enc.shape), self._contig(st.nu, "encoder", enc.shape)
        b_mu, b_nu = self._contig(st.mu, "encoder_bias", (M, N)), self._contig(st.nu, "encoder_bias", (M, N))
        if tied:
            dec = dec_mu = dec_nu = None
        else:
            dec = self._contig(self.params, "decoder", enc.shape)
            dec_mu, dec_nu = self._contig(st.mu, "decoder", enc.shape), self._contig(st.nu, "decoder", enc.shape)
        l1 = self._scalar_buffer("l1_alpha")
        bd = self._scalar_buffer("bias_decay") if bias_decay is None else bias_decay.to(torch.float32).contiguous()
        dict_size = None
        if masked:
            # the masked losses carry no bias-decay term (sae_ensemble.py:360-361, :429-430)
            bd = torch.zeros(M, dtype=torch.float32, device=x.device)
            dict_size = self.buffers["dict_size"].to(device=x.device, dtype=torch.int32).contiguous()
            self._check_prefix_mask(dict_size)

        adam = self.adam_constants(advance=not grad_only)
        grads = None
        if grad_only and reduce_scatter is None:
            if grad_out is not None:
                grads = {k: grad_out[k] for k in (("encoder", "encoder_bias") if tied else ("encoder", "encoder_bias", "decoder"))}
                for k, g in grads.items():
                    if not (g.is_contiguous() and g.shape == self.params[k].shape and g.dtype == torch.float32 and g.is_cuda):
                        raise ValueError(f"out[{k!r}] must be a contiguous float32 CUDA tensor of shape {tuple(self.params[k].shape)}")
            else:
                grads = {"encoder": torch.empty_like(enc), "encoder_bias": torch.empty_like(bias)}
                if not tied:
                    grads["decoder"] = torch.empty_like(dec)

        fmt = self.operand_format
        if fmt not in ("auto", "tf32", "f16"):
            raise ValueError(f"operand_format {fmt!r}: expected 'auto', 'tf32' or 'f16'")
        if fmt != "tf32":
            self._poll_range_flag(x.device)
            fmt = self.operand_format  # "auto" falls back to "tf32" once a batch has left the fp16 range
        f16 = fmt != "tf32" and bool(L.sc_sae_step_f16_supported(B, d, N))
        if fmt == "f16" and not f16:
            raise ValueError(f"operand_format 'f16' needs d, N multiples of 64 and >= 128 and B > 64 (got B={B} d={d} N={N})")
        self.last_operand_format = "f16" if f16 else "tf32"
        losses = torch.empty((4, M), dtype=torch.float32, device=x.device)
        c = None if f16 else torch.empty((M, B, N), dtype=torch.float32, device=x.device)
        # fp16 mode keeps the codes transposed, c^T [M][N][ldb], ldb = B rounded up to 8 (include/sc_b200.h)
        c_half = torch.empty((M, N, (B + 7) & ~7), dtype=torch.float16, device=x.device) if f16 else None
        code_scale = torch.empty((M, N), dtype=torch.float32, device=x.device) if f16 else None
        need = L.sc_sae_step_workspace_bytes(M, B, d, N, int(tied), int(per_member))
        ws_ptr, ws_bytes = self._ws.aligned(need, x.device)
        # Hand-over of the fp16 operand copies from the previous step's Adam epilogue (ScStepArgs.operands_valid): intact
        # while the workspace, the shapes and the parameter tensors (identity and torch version counter - the engine's own
        # writes do not bump it) are what the previous full step left.
        ops_key = None
        if f16 and not tied and not grad_only:
            ops_key = (ws_ptr.value, M, d, N, enc.data_ptr(), enc._version, dec.data_ptr(), dec._version)
        # the tensor OBJECTS must be the same as well: a caller who assigns a new tensor into .params (big_sweep.py:198)
        # can get the old one's address back from the caching allocator, with an equal version counter
        last_refs = getattr(self, "_ops_refs", (None, None))
        same_objects = last_refs[0] is not None and last_refs[0]() is enc and last_refs[1]() is dec
        ops_valid = ops_key is not None and same_objects and getattr(self, "_ops_key", None) == ops_key
        self._ops_key = None  # re-armed below once the step has been enqueued
        args = nat.ScStepArgs(
            M=M, B=B, d=d, N=N, tied=int(tied), x_per_member=int(per_member),
            x=nat.ptr(x), enc=nat.ptr(enc), enc_mu=nat.ptr(enc_mu), enc_nu=nat.ptr(enc_nu),
            bias=nat.ptr(bias), bias_mu=nat.ptr(b_mu), bias_nu=nat.ptr(b_nu),
            dec=nat.ptr(dec), dec_mu=nat.ptr(dec_mu), dec_nu=nat.ptr(dec_nu),
            l1_alpha=nat.ptr(l1), bias_decay=nat.ptr(bd), adam=adam,
            out_losses=nat.ptr(losses), out_c=nat.ptr(c),
            workspace=ws_ptr, workspace_bytes=ws_bytes, stream=nat.stream_ptr(x.device),
            grad_enc=nat.ptr(grads["encoder"]) if grads else None,
            grad_bias=nat.ptr(grads["encoder_bias"]) if grads else None,
            grad_dec=nat.ptr(grads.get("decoder")) if grads else None,
            B_total=int(batch_total),
            operand_format=int(f16), out_c_half=nat.ptr(c_half), dict_size=nat.ptr(dict_size),
            operands_valid=int(ops_valid), out_code_scale=nat.ptr(code_scale),
            range_flag=nat.ptr(self._range_flag) if f16 else None, step_tag=self._range_tag if f16 else 0,
        )
        if reduce_scatter is not None:
            rs = reduce_scatter
            args.rs_world, args.rs_rows = int(rs["world"]), int(rs["rows"])
            for r in range(args.rs_world):
                args.rs_grad_enc[r] = rs["grad_enc"][r]
                args.rs_grad_bias[r] = rs["grad_bias"][r]
                args.rs_losses[r] = rs["losses"][r]
                if not tied:
                    args.rs_grad_dec[r] = rs["grad_dec"][r]
        with torch.cuda.device(x.device):
            nat.check(L.sc_sae_step(C.byref(args)), "sc_sae_step")
        if ops_key is not None:
            self._ops_key = ops_key
            self._ops_refs = (weakref.ref(enc), weakref.ref(dec))
        if plain_step and x is minibatches and (not tied or masked or self._center_identity):
            # what the next plain step can reuse while the same tensors sit in the same places (_run_cached)
            pst = self.params
            tensors = [(pst, "encoder", enc), (pst, "encoder_bias", bias), (st.mu, "encoder", enc_mu), (st.nu, "encoder", enc_nu),
                       (st.mu, "encoder_bias", b_mu), (st.nu, "encoder_bias", b_nu)]
            if not tied:
                tensors += [(pst, "decoder", dec), (st.mu, "decoder", dec_mu), (st.nu, "decoder", dec_nu)]
            watched = ["l1_alpha"] if masked else ["l1_alpha", "bias_decay"]
            if masked:
                watched += ["dict_size", "coef_mask"]
            elif tied:
                watched += ["center_rot", "center_trans", "center_scale"]
            self._step_ctx = _StepCtx(
                args=args, xshape=tuple(x.shape), device=x.device, dev_index=x.device.index, expand_dims=bool(expand_dims),
                fmt=self.operand_format, f16=f16, tied=tied, masked=masked, dims=(M, B, d, N), st=st,
                tensors=[(store, k, t, t.data_ptr()) for store, k, t in tensors],
                buffers=[(k, self.buffers[k], self.buffers[k]._version) for k in watched if k in self.buffers],
                keep=(l1, bd, dict_size), ws_need=need, enc=enc, dec=dec)
        loss_dict = {"loss": losses[0], "l_reconstruction": losses[1], "l_l1": losses[2]}
        if masked:
            pass  # loss, l_reconstruction, l_l1 only
        elif not tied:  # FunctionalTiedSAE's loss_data has no l_bias_decay key (sae_ensemble.py:150-154)
            loss_dict["l_bias_decay"] = losses[3]
        elif grad_only:
            loss_dict["l_bias_decay"] = losses[3]
        aux = LazyCodes(c_half, B, code_scale) if f16 else {"c": c}
        if grad_only:
            return grads, loss_dict, aux
        return loss_dict, aux
