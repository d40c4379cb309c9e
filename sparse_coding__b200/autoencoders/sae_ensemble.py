"""Model signatures - mirror of the reference's autoencoders/sae_ensemble.py:13-160.

`init` is plain torch (same parameter order and RNG consumption as the reference, so equal seeds
give equal tensors); `loss` / `encode` evaluate one member through the native kernels; training goes
through FunctionalEnsemble.step_batch, which never calls `loss` (the fused step computes it).
"""
import torch
import torch.nn as nn

from .ensemble import DictSignature
from .learned_dict import TiedSAE, UntiedSAE, _encode_native, _decode_native


def _loss_terms(encoder, bias, dictionary_src, buffers, batch, center=None):
    """Forward of sae_ensemble.py:53-77 / :134-160 for ONE member, no update."""
    xb = batch if center is None else center(buffers, batch)
    c, stats = _encode_native(xb, encoder, bias, want_moments=True)
    _, sq = _decode_native(c, dictionary_src, normalise=True, target=xb, want_x_hat=False)
    B, d = xb.shape
    l_rec = sq / (B * d)
    l_l1 = buffers["l1_alpha"] * (stats["sum"].sum() / B)
    l_bd = buffers["bias_decay"] * torch.sqrt((bias * bias).sum())
    return c, l_rec, l_l1, l_bd


class FunctionalSAE(DictSignature):
    _sc_kind = "untied"

    @staticmethod
    def init(activation_size, n_dict_components, l1_alpha, bias_decay=0.0, device=None, dtype=None):
        """sae_ensemble.py:15-38 (encoder, bias, decoder drawn in that order)."""
        params, buffers = {}, {}
        params["encoder"] = torch.empty((n_dict_components, activation_size), device=device, dtype=dtype)
        nn.init.xavier_uniform_(params["encoder"])
        params["encoder_bias"] = torch.empty((n_dict_components,), device=device, dtype=dtype)
        nn.init.zeros_(params["encoder_bias"])
        params["decoder"] = torch.empty((n_dict_components, activation_size), device=device, dtype=dtype)
        nn.init.xavier_uniform_(params["decoder"])
        buffers["l1_alpha"] = torch.tensor(l1_alpha, device=device, dtype=dtype)
        buffers["bias_decay"] = torch.tensor(bias_decay, device=device, dtype=dtype)
        return params, buffers

    @staticmethod
    def to_learned_dict(params, buffers):
        """sae_ensemble.py:40-42."""
        return UntiedSAE(params["encoder"], params["decoder"], params["encoder_bias"])

    @staticmethod
    def encode(params, buffers, batch):
        """sae_ensemble.py:45-50."""
        return _encode_native(batch, params["encoder"], params["encoder_bias"])[0]

    @staticmethod
    def loss(params, buffers, batch):
        """sae_ensemble.py:53-77."""
        c, l_rec, l_l1, l_bd = _loss_terms(params["encoder"], params["encoder_bias"], params["decoder"], buffers, batch)
        total = l_rec + l_l1 + l_bd
        loss_data = {"loss": total, "l_reconstruction": l_rec, "l_l1": l_l1, "l_bias_decay": l_bd}
        return total, (loss_data, {"c": c})


class FunctionalTiedSAE(DictSignature):
    _sc_kind = "tied"

    @staticmethod
    def init(activation_size, n_dict_components, l1_alpha, device=None, dtype=None, bias_decay=0.0,
             translation=None, rotation=None, scaling=None):
        """sae_ensemble.py:82-118."""
        params, buffers = {}, {}
        if rotation is None:
            rotation = torch.eye(activation_size, device=device, dtype=dtype)
        if translation is None:
            translation = torch.zeros(activation_size, device=device, dtype=dtype)
        if scaling is None:
            scaling = torch.ones(activation_size, device=device, dtype=dtype)
        buffers["center_rot"] = rotation
        buffers["center_trans"] = translation
        buffers["center_scale"] = scaling
        params["encoder"] = torch.empty((n_dict_components, activation_size), device=device, dtype=dtype)
        nn.init.xavier_uniform_(params["encoder"])
        params["encoder_bias"] = torch.empty((n_dict_components,), device=device, dtype=dtype)
        nn.init.zeros_(params["encoder_bias"])
        buffers["l1_alpha"] = torch.tensor(l1_alpha, device=device, dtype=dtype)
        buffers["bias_decay"] = torch.tensor(bias_decay, device=device, dtype=dtype)
        return params, buffers

    @staticmethod
    def to_learned_dict(params, buffers):
        """sae_ensemble.py:120-122."""
        return TiedSAE(params["encoder"], params["encoder_bias"],
                       centering=(buffers["center_trans"], buffers["center_rot"], buffers["center_scale"]),
                       norm_encoder=True)

    @staticmethod
    def center(buffers, batch):
        """sae_ensemble.py:125-127."""
        return torch.einsum("cu,bu->bc", buffers["center_rot"], batch - buffers["center_trans"][None, :]) \
            * buffers["center_scale"][None, :]

    @staticmethod
    def uncenter(buffers, batch):
        """sae_ensemble.py:129-131."""
        return torch.einsum("cu,bc->bu", buffers["center_rot"], batch / buffers["center_scale"][None, :]) \
            + buffers["center_trans"][None, :]

    @staticmethod
    def loss(params, buffers, batch):
        """sae_ensemble.py:134-160 (loss_data carries no l_bias_decay key, as in the reference)."""
        from .learned_dict import _row_normalise_native
        D = _row_normalise_native(params["encoder"])
        c, l_rec, l_l1, l_bd = _loss_terms(D, params["encoder_bias"], D, buffers, batch, FunctionalTiedSAE.center)
        total = l_rec + l_l1 + l_bd
        loss_data = {"loss": total, "l_reconstruction": l_rec, "l_l1": l_l1}
        return total, (loss_data, {"c": c})
