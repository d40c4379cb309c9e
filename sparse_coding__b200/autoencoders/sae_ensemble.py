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
        from .learned_dict import center_native
        return center_native(batch, buffers["center_rot"], buffers["center_trans"], buffers["center_scale"])

    @staticmethod
    def uncenter(buffers, batch):
        """sae_ensemble.py:129-131."""
        from .learned_dict import uncenter_native
        return uncenter_native(batch, buffers["center_rot"], buffers["center_trans"], buffers["center_scale"])

    @staticmethod
    def loss(params, buffers, batch):
        """sae_ensemble.py:134-160 (loss_data carries no l_bias_decay key, as in the reference)."""
        from .learned_dict import _row_normalise_native
        D = _row_normalise_native(params["encoder"])
        c, l_rec, l_l1, l_bd = _loss_terms(D, params["encoder_bias"], D, buffers, batch, FunctionalTiedSAE.center)
        total = l_rec + l_l1 + l_bd
        loss_data = {"loss": total, "l_reconstruction": l_rec, "l_l1": l_l1}
        return total, (loss_data, {"c": c})


def _masked_loss(encoder, bias, dictionary_src, buffers, batch):
    """Forward of sae_ensemble.py:340-369 / :409-438 for one member: codes of the unused features are
    zeroed, and the loss has no bias-decay term."""
    c, _ = _encode_native(batch, encoder, bias)
    c = c.masked_fill(buffers["coef_mask"], 0.0)
    _, sq = _decode_native(c, dictionary_src, normalise=True, target=batch, want_x_hat=False)
    B, d = batch.shape
    l_rec = sq / (B * d)
    l_l1 = buffers["l1_alpha"] * c.abs().sum(-1).mean()
    return c, l_rec, l_l1


def _masked_init(tied, activation_size, n_dict_components, n_components_stack, l1_alpha, bias_decay, device, dtype):
    params, buffers = {}, {}
    params["encoder"] = torch.empty((n_components_stack, activation_size), device=device, dtype=dtype)
    nn.init.xavier_uniform_(params["encoder"])
    params["encoder_bias"] = torch.empty((n_components_stack,), device=device, dtype=dtype)
    nn.init.zeros_(params["encoder_bias"])
    if not tied:
        params["decoder"] = torch.empty((n_components_stack, activation_size), device=device, dtype=dtype)
        nn.init.xavier_uniform_(params["decoder"])
    buffers["l1_alpha"] = torch.tensor(l1_alpha, device=device, dtype=dtype)
    buffers["bias_decay"] = torch.tensor(bias_decay, device=device, dtype=dtype)
    buffers["dict_size"] = torch.tensor(n_dict_components, device=device, dtype=torch.long)
    buffers["coef_mask"] = torch.ones(n_components_stack, device=device, dtype=torch.bool)
    buffers["coef_mask"][:n_dict_components] = False
    return params, buffers


class FunctionalMaskedTiedSAE(DictSignature):
    """sae_ensemble.py:307-371: tied SAEs of different dictionary sizes stacked to a common width; member m
    only uses its first dict_size[m] features (the fused step masks them in the encode epilogue)."""
    _sc_kind = "masked_tied"

    @staticmethod
    def init(activation_size, n_dict_components, n_components_stack, l1_alpha, bias_decay=0.0, device=None, dtype=None):
        return _masked_init(True, activation_size, n_dict_components, n_components_stack, l1_alpha, bias_decay, device, dtype)

    @staticmethod
    def to_learned_dict(params, buffers):
        n = buffers["dict_size"].item()
        return TiedSAE(params["encoder"][:n], params["encoder_bias"][:n], norm_encoder=True)

    @staticmethod
    def loss(params, buffers, batch):
        from .learned_dict import _row_normalise_native
        D = _row_normalise_native(params["encoder"])
        c, l_rec, l_l1 = _masked_loss(D, params["encoder_bias"], D, buffers, batch)
        return l_rec + l_l1, ({"loss": l_rec + l_l1, "l_reconstruction": l_rec, "l_l1": l_l1}, {"c": c})


class FunctionalMaskedSAE(DictSignature):
    """sae_ensemble.py:375-442, the untied twin."""
    _sc_kind = "masked_untied"

    @staticmethod
    def init(activation_size, n_dict_components, n_components_stack, l1_alpha, bias_decay=0.0, device=None, dtype=None):
        return _masked_init(False, activation_size, n_dict_components, n_components_stack, l1_alpha, bias_decay, device, dtype)

    @staticmethod
    def to_learned_dict(params, buffers):
        n = buffers["dict_size"].item()
        return UntiedSAE(params["encoder"][:n], params["decoder"][:n], params["encoder_bias"][:n])

    @staticmethod
    def loss(params, buffers, batch):
        c, l_rec, l_l1 = _masked_loss(params["encoder"], params["encoder_bias"], params["decoder"], buffers, batch)
        return l_rec + l_l1, ({"loss": l_rec + l_l1, "l_reconstruction": l_rec, "l_l1": l_l1}, {"c": c})
