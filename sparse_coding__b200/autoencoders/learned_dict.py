"""Inference-side dictionaries - mirror of the reference's autoencoders/learned_dict.py:13-50,
111-196 (LearnedDict, UntiedSAE, TiedSAE).  encode / decode / get_learned_dict run through
libsc_b200 (sc_encode, sc_decode, sc_row_normalise); tensors must live on a CUDA sm_100a device.

Objects pickle under the reference's module path (`autoencoders.learned_dict`), so files written
by this engine load in the reference and the reference's own files (output_basic_test/*.pt) load
here after `sparse_coding__b200.install_reference_aliases()`.
"""
from __future__ import annotations

import ctypes as C
from abc import ABC, abstractmethod

import torch

from .. import _native as nat

_ws = nat.Workspace()



def center_native(batch, rot, trans, scale):
    """((batch - trans) Rot^T) * scale (learned_dict.py:164-165, sae_ensemble.py:125-127) with the d x d product on the
    tcgen05 core at fp32 accuracy (sc_gemm_f32).  rot may be [d,d] or, for an ensemble, [M,d,d] with batch [M,B,d]."""
    if rot.dim() == 3:
        return nat.gemm_f32(batch - trans[:, None, :], rot) * scale[:, None, :]
    return nat.gemm_f32(batch - trans[None, :], rot) * scale[None, :]


def uncenter_native(batch, rot, trans, scale):
    """(batch / scale) Rot + trans (learned_dict.py:167-168, sae_ensemble.py:129-131)."""
    if rot.dim() == 3:
        return nat.gemm_f32(batch / scale[:, None, :], rot, b_mn=True) + trans[:, None, :]
    return nat.gemm_f32(batch / scale[None, :], rot, b_mn=True) + trans[None, :]

def _prep(t: torch.Tensor, name: str) -> torch.Tensor:
    nat.require_cuda(t, name)
    t = t.detach()
    if not t.is_contiguous() or t.data_ptr() % 16:
        t = t.contiguous().clone() if t.data_ptr() % 16 else t.contiguous()
    return t


def _operand_format():
    """GEMM operand format of the inference kernels: fp16 (same 11-bit significand as the TF32 mode, twice the
    tensor-core rate) whenever the shape allows it, unless SC_B200_OPERANDS=tf32."""
    import os
    return os.environ.get("SC_B200_OPERANDS", "auto")


def _encode_native(batch, encoder, bias, want_counts=False, want_moments=False, want_c=True, want_half=False):
    """c = relu(batch @ encoder^T + bias) (learned_dict.py:127-131).  encoder may be [N,d] or a
    stack [G,N,d]; optional per-feature statistics come from the same GEMM's epilogue.
    want_half: return the codes as fp16 [G,B,N] (None when the shape cannot run with fp16 operands) - the form
    _decode_native(code_half=...) takes, so predict / FVU never write fp32 codes."""
    L = nat.lib()
    x = _prep(batch, "batch")
    W = _prep(encoder, "encoder")
    b = None if bias is None else _prep(bias, "encoder_bias")
    stacked = W.dim() == 3
    G = W.shape[0] if stacked else 1
    N, d = W.shape[-2], W.shape[-1]
    per_group = x.dim() == 3
    if x.shape[-1] != d:
        raise ValueError(f"batch width {x.shape[-1]} != dictionary width {d}")
    B = x.shape[-2]
    dev = x.device
    c = torch.empty((G, B, N), dtype=torch.float32, device=dev) if want_c else None
    nz = torch.zeros((G, N), dtype=torch.float32, device=dev) if (want_counts or want_moments) else None
    mom = torch.zeros((G, 4, N), dtype=torch.float32, device=dev) if want_moments else None
    need = L.sc_encode_workspace_bytes(G, B, d, N, int(per_group))
    ws_ptr, ws_bytes = _ws.aligned(need, dev)
    f16 = _operand_format() != "tf32" and bool(L.sc_infer_f16_supported(B, d, N, int(nz is not None)))
    c_half = None
    if want_half:
        if not f16 or nz is not None:
            return None, {}
        c_half = torch.empty((G, B, N), dtype=torch.float16, device=dev)
    args = nat.ScEncodeArgs(G=G, B=B, d=d, N=N, x_per_group=int(per_group), x=nat.ptr(x), W=nat.ptr(W), bias=nat.ptr(b),
                            out_c=nat.ptr(c), out_nonzero=nat.ptr(nz), out_moments=nat.ptr(mom),
                            workspace=ws_ptr, workspace_bytes=ws_bytes, stream=nat.stream_ptr(dev),
                            operand_format=int(f16), out_c_half=nat.ptr(c_half))
    with torch.cuda.device(dev):
        nat.check(L.sc_encode(C.byref(args)), "sc_encode")
    if want_half:
        return c_half, {}
    stats = {}
    if nz is not None:
        stats["nonzero"] = nz if stacked else nz[0]
    if mom is not None:
        m = mom if stacked else mom[0]
        stats["sum"], stats["sum2"], stats["sum3"], stats["sum4"] = m[..., 0, :], m[..., 1, :], m[..., 2, :], m[..., 3, :]
    if c is not None and not stacked:
        c = c[0]
    return c, stats


def _decode_native(code, decoder, normalise=True, target=None, want_x_hat=True, code_half=None):
    """x_hat = code @ rownorm(decoder) (learned_dict.py:29-32); with `target` also
    sum((x_hat - target)^2).  Returns (x_hat or None, sq_err tensor or None).  code_half: fp16 codes [G,B,N] from
    _encode_native(want_half=True), used instead of `code`."""
    L = nat.lib()
    c = code_half if code_half is not None else _prep(code, "code")
    W = _prep(decoder, "decoder")
    stacked = W.dim() == 3
    G = W.shape[0] if stacked else 1
    N, d = W.shape[-2], W.shape[-1]
    if c.shape[-1] != N:
        raise ValueError(f"code width {c.shape[-1]} != dictionary size {N}")
    B = c.shape[-2]
    dev = c.device
    tgt = None if target is None else _prep(target, "target")
    x_hat = torch.empty((G, B, d), dtype=torch.float32, device=dev) if want_x_hat else None
    sq = torch.empty((G,), dtype=torch.float32, device=dev) if tgt is not None else None
    need = L.sc_decode_workspace_bytes(G, B, d, N)
    ws_ptr, ws_bytes = _ws.aligned(need, dev)
    f16 = code_half is not None or (_operand_format() != "tf32" and bool(L.sc_infer_f16_supported(B, d, N, 0)))
    args = nat.ScDecodeArgs(G=G, B=B, d=d, N=N, normalise=int(normalise),
                            target_per_group=int(tgt is not None and tgt.dim() == 3),
                            c=None if code_half is not None else nat.ptr(c), Wd=nat.ptr(W), target=nat.ptr(tgt),
                            out_x_hat=nat.ptr(x_hat), out_sq_err=nat.ptr(sq), workspace=ws_ptr, workspace_bytes=ws_bytes,
                            stream=nat.stream_ptr(dev), operand_format=int(f16),
                            c_half=nat.ptr(code_half) if code_half is not None else None)
    with torch.cuda.device(dev):
        nat.check(L.sc_decode(C.byref(args)), "sc_decode")
    if not stacked:
        x_hat = None if x_hat is None else x_hat[0]
        sq = None if sq is None else sq[0]
    return x_hat, sq


def _row_normalise_native(W):
    """W / clamp(||W||_row, 1e-8) (learned_dict.py:118-120)."""
    L = nat.lib()
    Wc = _prep(W, "dictionary")
    G = Wc.shape[0] if Wc.dim() == 3 else 1
    N, d = Wc.shape[-2], Wc.shape[-1]
    out = torch.empty_like(Wc)
    with torch.cuda.device(Wc.device):
        nat.check(L.sc_row_normalise(nat.ptr(Wc), nat.ptr(out), G, N, d, nat.stream_ptr(Wc.device)), "sc_row_normalise")
    return out


class LearnedDict(ABC):
    """learned_dict.py:13-50."""
    n_feats: int
    activation_size: int

    @abstractmethod
    def get_learned_dict(self):
        pass

    @abstractmethod
    def encode(self, batch):
        pass

    @abstractmethod
    def to_device(self, device):
        pass

    def decode(self, code):
        """learned_dict.py:29-32: get_learned_dict() renormalises, then code @ dict; fused here."""
        return _decode_native(code, self._dictionary_source(), normalise=True)[0]

    def _dictionary_source(self):
        raise NotImplementedError

    def center(self, batch):
        return batch

    def uncenter(self, batch):
        return batch

    def _encode_operands(self):
        """(encoder matrix as encode() uses it, bias) for dictionaries whose encode is relu(x W^T + b), else None."""
        return None

    def predict(self, batch):
        """learned_dict.py:42-47.  When encode is the plain affine + ReLU form the codes go from the encode kernel to
        the decode kernel as fp16 (the decode GEMM's operand format anyway) and are never written in fp32."""
        batch_centered = self.center(batch)
        eb = self._encode_operands()
        if eb is not None:
            c_half, _ = _encode_native(batch_centered, eb[0], eb[1], want_c=False, want_half=True)
            if c_half is not None:
                x_hat = _decode_native(None, self._dictionary_source(), normalise=True, code_half=c_half)[0]
                return self.uncenter(x_hat)
        c = self.encode(batch_centered)
        x_hat_centered = self.decode(c)
        return self.uncenter(x_hat_centered)

    def n_dict_components(self):
        return self.get_learned_dict().shape[0]


class UntiedSAE(LearnedDict):
    """learned_dict.py:111-131."""

    def __init__(self, encoder, decoder, encoder_bias):
        self.encoder = encoder
        self.decoder = decoder
        self.encoder_bias = encoder_bias
        self.n_feats, self.activation_size = self.encoder.shape

    def get_learned_dict(self):
        return _row_normalise_native(self.decoder)

    def _dictionary_source(self):
        return self.decoder

    def _encode_operands(self):
        return self.encoder, self.encoder_bias

    def to_device(self, device):
        self.encoder = self.encoder.to(device)
        self.decoder = self.decoder.to(device)
        self.encoder_bias = self.encoder_bias.to(device)

    def encode(self, batch):
        return _encode_native(batch, self.encoder, self.encoder_bias)[0]


class TiedSAE(LearnedDict):
    """learned_dict.py:134-196."""

    def __init__(self, encoder, encoder_bias, centering=(None, None, None), norm_encoder=False):
        self.encoder = encoder
        self.encoder_bias = encoder_bias
        self.norm_encoder = norm_encoder
        self.n_feats, self.activation_size = self.encoder.shape
        center_trans, center_rot, center_scale = centering
        if center_trans is None:
            center_trans = torch.zeros(self.activation_size)
        if center_rot is None:
            center_rot = torch.eye(self.activation_size)
        if center_scale is None:
            center_scale = torch.ones(self.activation_size)
        self.center_trans = center_trans
        self.center_rot = center_rot
        self.center_scale = center_scale

    def initialize_missing(self):
        if not hasattr(self, "center_trans"):
            self.center_trans = torch.zeros(self.activation_size, device=self.encoder.device)
        if not hasattr(self, "center_rot"):
            self.center_rot = torch.eye(self.activation_size, device=self.encoder.device)
        if not hasattr(self, "center_scale"):
            self.center_scale = torch.ones(self.activation_size, device=self.encoder.device)

    def _identity_centering(self) -> bool:
        self.initialize_missing()
        rot, tr, sc = self.center_rot, self.center_trans, self.center_scale
        d = rot.shape[0]
        return bool((rot == torch.eye(d, device=rot.device, dtype=rot.dtype)).all() and (tr == 0).all()
                    and (sc == 1).all())

    def center(self, batch):
        """learned_dict.py:164-165; identity centring (the default) is skipped."""
        if self._identity_centering():
            return batch
        return center_native(batch, self.center_rot.to(batch.device), self.center_trans.to(batch.device),
                             self.center_scale.to(batch.device))

    def uncenter(self, batch):
        """learned_dict.py:167-168."""
        if self._identity_centering():
            return batch
        return uncenter_native(batch, self.center_rot.to(batch.device), self.center_trans.to(batch.device),
                               self.center_scale.to(batch.device))

    def get_learned_dict(self):
        return _row_normalise_native(self.encoder)

    def _dictionary_source(self):
        return self.encoder

    def to_device(self, device):
        self.initialize_missing()
        self.encoder = self.encoder.to(device)
        self.encoder_bias = self.encoder_bias.to(device)
        self.center_trans = self.center_trans.to(device)
        self.center_rot = self.center_rot.to(device)
        self.center_scale = self.center_scale.to(device)

    def _encode_operands(self):
        enc = _row_normalise_native(self.encoder) if self.norm_encoder else self.encoder
        return enc, self.encoder_bias

    def encode(self, batch):
        """learned_dict.py:186-196."""
        enc, bias = self._encode_operands()
        return _encode_native(batch, enc, bias)[0]


# pickles must resolve under the reference's module path (big_sweep.py:424 torch.save of these objects)
for _cls in (LearnedDict, UntiedSAE, TiedSAE):
    _cls.__module__ = "autoencoders.learned_dict"
