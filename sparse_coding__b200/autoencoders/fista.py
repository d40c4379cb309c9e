"""FISTA - mirror of the reference's autoencoders/fista.py (FunctionalFista :18-138, Fista :201-301).

`FunctionalFista.loss` is the same loss as FunctionalSAE (fista.py:60-84), so the ensemble trains it
through the same fused step.  `fista()` runs the batched proximal-gradient solver in libsc_b200
(`sc_fista`, split 3xTF32 tensor-core products).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.nn as nn

from .. import _native as nat
from .ensemble import DictSignature
from .learned_dict import LearnedDict, TiedSAE, UntiedSAE, _encode_native, _prep, _row_normalise_native
from .sae_ensemble import FunctionalSAE

_ws = nat.Workspace()


def momentum_sequence(num_iter: int) -> np.ndarray:
    """fista.py:108-118: t_{k+1} = (1 + sqrt(1 + 4 t_k^2)) / 2 in float64; returns (t_k - 1) / t_{k+1}."""
    tk_n = 1.0
    out = np.empty(num_iter, dtype=np.float32)
    for i in range(num_iter):
        tk = tk_n
        tk_n = (1 + np.sqrt(1 + 4 * tk ** 2)) / 2
        out[i] = (tk - 1) / tk_n
    return out


def _pair_format(x, a0):
    """0 (fp16 operand pairs, twice the tensor rate) unless the data could leave fp16's range: every FISTA iterate is
    bounded by a small multiple of the largest row norm of x for a unit-row dictionary, the warm start is checked
    directly.  One tiny reduction + sync per call (a call is hundreds of GEMMs)."""
    bound = max(float(torch.linalg.vector_norm(x, dim=1).max()), float(a0.abs().max()))
    return 0 if bound < 8192.0 else 1


def fista_native(batch, learned_dict, l1_coef, coefficients, num_iter, eta=None):
    """(ahat, Res) of fista.py:99-128 on the GPU."""
    L = nat.lib()
    x = _prep(batch, "batch")
    D = _prep(learned_dict, "learned_dict")
    a0 = _prep(coefficients, "coefficients")
    B, d = x.shape
    N = D.shape[0]
    if D.shape[1] != d or tuple(a0.shape) != (B, N):
        raise ValueError(f"shape mismatch: batch {tuple(x.shape)}, dict {tuple(D.shape)}, coefficients {tuple(a0.shape)}")
    dev = x.device
    mom = momentum_sequence(int(num_iter))
    out_a = torch.empty((B, N), dtype=torch.float32, device=dev)
    out_res = torch.empty((B, d), dtype=torch.float32, device=dev)
    out_eta = torch.empty((1,), dtype=torch.float32, device=dev)
    if eta is None:
        eta_f = -1.0
    else:
        eta_f = float(eta.item() if isinstance(eta, torch.Tensor) else eta)
    # a coefficient that already lives on the device (buffers["l1_alpha"] of an ensemble member, big_sweep.py:186) is
    # read there by the kernel: no .item() sync per member and step
    l1_dev = None
    if isinstance(l1_coef, torch.Tensor) and l1_coef.is_cuda:
        l1_dev = l1_coef.detach().to(device=dev, dtype=torch.float32).reshape(1)
        l1_f = 0.0
    else:
        l1_f = float(l1_coef.item() if isinstance(l1_coef, torch.Tensor) else l1_coef)
    need = L.sc_fista_workspace_bytes(B, d, N)
    ws_ptr, ws_bytes = _ws.aligned(need, dev)
    args = nat.ScFistaArgs(B=B, d=d, N=N, num_iter=int(num_iter), x=nat.ptr(x), D=nat.ptr(D), a0=nat.ptr(a0),
                           l1_coef=l1_f, eta=eta_f, momentum=mom.ctypes.data_as(C.POINTER(C.c_float)),
                           out_a=nat.ptr(out_a), out_res=nat.ptr(out_res), out_eta=nat.ptr(out_eta),
                           workspace=ws_ptr, workspace_bytes=ws_bytes, stream=nat.stream_ptr(dev),
                           l1_coef_dev=nat.ptr(l1_dev), pair_format=_pair_format(x, a0))
    with torch.cuda.device(dev):
        nat.check(L.sc_fista(C.byref(args)), "sc_fista")
    return out_a, out_res, out_eta


class FunctionalFista(DictSignature):
    _sc_kind = "untied"

    @staticmethod
    def init(activation_size, n_dict_components, l1_alpha, bias_decay=0.0, device=None, dtype=None):
        """fista.py:20-45 (FunctionalSAE.init plus the hessian_diag buffer)."""
        params, buffers = FunctionalSAE.init(activation_size, n_dict_components, l1_alpha, bias_decay, device, dtype)
        buffers["hessian_diag"] = torch.zeros(n_dict_components).to(device)
        return params, buffers

    to_learned_dict = staticmethod(FunctionalSAE.to_learned_dict)   # fista.py:47-49
    encode = staticmethod(FunctionalSAE.encode)                     # fista.py:51-57
    loss = staticmethod(FunctionalSAE.loss)                         # fista.py:59-84

    @staticmethod
    def fista(batch, learned_dict, l1_coef, coefficients, num_iter, device, eta=None, threshold=1e-4):
        """fista.py:99-128.  `device` and `threshold` are unused there too."""
        a, res, _ = fista_native(batch, learned_dict, l1_coef, coefficients, num_iter, eta)
        return a.to(batch.dtype), res

    @staticmethod
    def quadraticBasisUpdate(learned_dict, Res, ahat, lowestActivation, HessianDiag, stepSize=0.001, Noneg=False):
        """fista.py:131-138, kept verbatim in meaning: only shape-valid when N == d, and the final
        normalisation is over dim 0 (column norms)."""
        dBasis = stepSize * nat.gemm_f32(Res, ahat, a_mn=True, b_mn=True) / ahat.size(0)   # Res^T ahat, [d, N]
        dBasis = dBasis / (HessianDiag + lowestActivation)
        learned_dict = learned_dict + (dBasis.t())
        if Noneg:
            learned_dict = learned_dict.clamp(min=0.)
        return learned_dict / (learned_dict.norm(2, 0))

    @staticmethod
    def quadratic_basis_update_rows(learned_dict, Res, ahat, lowestActivation, HessianDiag, stepSize=0.001, Noneg=False):
        """What fista.py:131-138 is after, valid for any N and d: the same preconditioned gradient
        step on the reconstruction error, dBasis[n] = stepSize * mean_b(a_bn Res_b) / (H_n + lowestActivation), with
        every ATOM (row of the [N, d] dictionary) renormalised to unit length afterwards."""
        dBasis = stepSize * nat.gemm_f32(ahat, Res, a_mn=True, b_mn=True) / ahat.size(0)   # ahat^T Res, [N, d]
        dBasis = dBasis / (HessianDiag + lowestActivation)[:, None]
        learned_dict = learned_dict + dBasis
        if Noneg:
            learned_dict = learned_dict.clamp(min=0.)
        return learned_dict / learned_dict.norm(2, 1, keepdim=True).clamp_min(1e-8)

    @staticmethod
    def dictionary_update_corrected(params, buffers, batch_centered, coeffs, learned_dict, num_iter=500):
        """dictionary_update without the reference's defects: works for N != d (row-wise update and
        normalisation) and the running Hessian diagonal PERSISTS (updated in place in buffers["hessian_diag"])."""
        device = params["encoder"].device
        coeffs_fista, res = FunctionalFista.fista(batch_centered, learned_dict, buffers["l1_alpha"], coeffs, num_iter, device)
        ACT_HISTORY_LEN = 300
        buffers["hessian_diag"].mul_((ACT_HISTORY_LEN - 1.0) / ACT_HISTORY_LEN).add_(
            torch.pow(coeffs_fista, 2).mean(0) / ACT_HISTORY_LEN)
        new_dict = FunctionalFista.quadratic_basis_update_rows(learned_dict, res, coeffs_fista, 0.001, buffers["hessian_diag"])
        return new_dict, res

    @staticmethod
    def dictionary_update(params, buffers, batch_centered, coeffs, learned_dict):
        """fista.py:88-96: 500 FISTA iterations from the SAE codes, running Hessian diagonal (kept in
        the dict the caller hands in), quadratic basis update."""
        device = params["encoder"].device
        coeffs_fista, res = FunctionalFista.fista(batch_centered, learned_dict, buffers["l1_alpha"], coeffs, 500, device, eta=None)
        ACT_HISTORY_LEN = 300
        buffers["hessian_diag"] = buffers["hessian_diag"].mul((ACT_HISTORY_LEN - 1.0) / ACT_HISTORY_LEN) \
            + torch.pow(coeffs_fista, 2).mean(0) / ACT_HISTORY_LEN
        new_dict = FunctionalFista.quadraticBasisUpdate(learned_dict, res, coeffs_fista, 0.001, buffers["hessian_diag"])
        return new_dict, res


class Fista(TiedSAE):
    """fista.py:201-301: a TiedSAE-shaped LearnedDict with a `fista` method."""

    def fista(self, batch, coefficients, device, l1_coef, num_iter, eta=None):
        """fista.py:274-301."""
        a, res, _ = fista_native(batch, self.get_learned_dict(), l1_coef, coefficients, num_iter, eta)
        return a, res


Fista.__module__ = "autoencoders.fista"
