"""FVU / L0 / dead-feature / moment reductions - mirror of the reference's standard_metrics.py:
303-340, 439-452, 480-509 (same names and signatures).  The encode GEMM's epilogue produces the
per-feature counts and raw moments directly, and the decode GEMM's epilogue produces the squared
reconstruction error, so no [n, N] code matrix is materialised for the statistics.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as nat
from .autoencoders.learned_dict import LearnedDict, TiedSAE, UntiedSAE, _decode_native, _encode_native, _prep, \
    _row_normalise_native

_ws = nat.Workspace()


def _encoder_and_bias(model):
    """(encoder as used by model.encode, bias) when the model is one of ours, else None."""
    if isinstance(model, TiedSAE):
        enc = _row_normalise_native(model.encoder) if model.norm_encoder else model.encoder
        return enc, model.encoder_bias
    if isinstance(model, UntiedSAE):
        return model.encoder, model.encoder_bias
    return None


def _feature_stats(model, batch, moments=False):
    eb = _encoder_and_bias(model)
    if eb is None:
        c = model.encode(batch)
        st = {"nonzero": (c != 0).float().sum(0)}
        if moments:
            st.update(sum=c.sum(0), sum2=(c ** 2).sum(0), sum3=(c ** 3).sum(0), sum4=(c ** 4).sum(0))
        return st
    _, st = _encode_native(batch, eb[0], eb[1], want_counts=True, want_moments=moments, want_c=False)
    return st


def mean_nonzero_activations(model: LearnedDict, batch):
    """standard_metrics.py:303-306: (encode(center(batch)) != 0).float().mean(0)."""
    batch_centered = model.center(batch)
    return _feature_stats(model, batch_centered)["nonzero"] / batch_centered.shape[0]


def total_variance_sum(batch) -> torch.Tensor:
    """sum((batch - batch.mean(0))^2) through sc_column_variance."""
    L = nat.lib()
    x = _prep(batch, "batch")
    B, d = x.shape
    out = torch.empty((1,), dtype=torch.float32, device=x.device)
    ws_ptr, _ = _ws.aligned((d + 1) * 4, x.device)
    with torch.cuda.device(x.device):
        nat.check(L.sc_column_variance(nat.ptr(x), nat.ptr(out), B, d, ws_ptr, nat.stream_ptr(x.device)),
                  "sc_column_variance")
    return out[0]


def residual_sum_of_squares(model: LearnedDict, batch) -> torch.Tensor:
    """sum((batch - predict(batch))^2): the numerator of fraction_variance_unexplained before the two means cancel
    (row-sharded evaluation adds these per rank, dist.fraction_variance_unexplained_sharded)."""
    eb = _encoder_and_bias(model)
    identity = not isinstance(model, TiedSAE) or model._identity_centering()
    if eb is not None and identity:
        c_half, _ = _encode_native(batch, eb[0], eb[1], want_c=False, want_half=True)
        if c_half is not None:
            _, sq = _decode_native(None, model._dictionary_source(), normalise=True, target=batch, want_x_hat=False,
                                   code_half=c_half)
        else:
            c, _ = _encode_native(batch, eb[0], eb[1])
            _, sq = _decode_native(c, model._dictionary_source(), normalise=True, target=batch, want_x_hat=False)
        return sq.reshape(())
    return (batch - model.predict(batch)).pow(2).sum()


def feature_nonzero_counts(model: LearnedDict, batch) -> torch.Tensor:
    """[N] number of rows of `batch` (already centred) on which each feature is non-zero, codes never materialised."""
    return _feature_stats(model, batch)["nonzero"]


def fraction_variance_unexplained(model: LearnedDict, batch):
    """standard_metrics.py:308-312: mean((batch - predict(batch))^2) / mean((batch - batch.mean(0))^2)."""
    eb = _encoder_and_bias(model)
    identity = not isinstance(model, TiedSAE) or model._identity_centering()
    if eb is not None and identity:
        c_half, _ = _encode_native(batch, eb[0], eb[1], want_c=False, want_half=True)  # fp16 codes, never in fp32
        if c_half is not None:
            _, sq = _decode_native(None, model._dictionary_source(), normalise=True, target=batch, want_x_hat=False,
                                   code_half=c_half)
        else:
            c, _ = _encode_native(batch, eb[0], eb[1])
            _, sq = _decode_native(c, model._dictionary_source(), normalise=True, target=batch, want_x_hat=False)
        return sq / total_variance_sum(batch)
    x_hat = model.predict(batch)
    residuals = (batch - x_hat).pow(2).mean()
    total = (batch - batch.mean(dim=0)).pow(2).mean()
    return residuals / total


def r_squared(model: LearnedDict, batch):
    """standard_metrics.py:342-343."""
    return 1.0 - fraction_variance_unexplained(model, batch)


def fraction_variance_unexplained_top_activating(model: LearnedDict, batch, n_top=2):
    """standard_metrics.py:314-340."""
    c = model.encode(model.center(batch))
    mean_activation = c.mean(dim=0)
    idxs = torch.argsort(mean_activation, descending=True)
    top_n_idxs = idxs[:n_top]
    rest_idxs = idxs[n_top:]
    c_top = torch.zeros_like(c)
    c_top[:, top_n_idxs] = c[:, top_n_idxs]
    c_rest = torch.zeros_like(c)
    c_rest[:, rest_idxs] = c[:, rest_idxs]
    x_hat_top = model.center(model.decode(c_top))
    x_hat_rest = model.center(model.decode(c_rest))
    residuals_top = (batch - x_hat_top).pow(2).mean()
    residuals_rest = (batch - x_hat_rest).pow(2).mean()
    variance = (batch - batch.mean(dim=0)).pow(2).mean()
    return residuals_top / variance, residuals_rest / variance


def calc_feature_n_active(batch):
    """standard_metrics.py:439-442."""
    return torch.sum(batch != 0, dim=0)


def batched_calc_feature_n_ever_active(learned_dict: LearnedDict, activations: torch.Tensor, batch_size: int = 1000,
                                       threshold: int = 10) -> int:
    """standard_metrics.py:444-452.  The reference loops 1000-row slices to bound memory; the fused
    epilogue never materialises the codes, so slices are only as small as the GEMM likes."""
    n_active_count = torch.zeros(learned_dict.n_feats, device=activations.device)
    step = max(batch_size, 65536)
    for i in range(0, len(activations), step):
        n_active_count += _feature_stats(learned_dict, activations[i:i + step])["nonzero"]
    return int((n_active_count > threshold).sum().item())


def calc_feature_mean(batch):
    return torch.mean(batch, dim=0)


def calc_feature_variance(batch):
    return torch.var(batch, dim=0)


def calc_feature_skew(batch):
    variance = torch.var(batch, dim=0)
    return torch.mean(batch ** 3, dim=0) / torch.clamp(variance ** 1.5, min=1e-8)


def calc_feature_kurtosis(batch):
    variance = torch.var(batch, dim=0)
    return torch.mean(batch ** 4, dim=0) / torch.clamp(variance ** 2, min=1e-8)


def calc_moments_streaming(learned_dict, activations, batch_size=1000):
    """standard_metrics.py:480-509, including its weighting of every slice by `batch_size` even when
    the last one is shorter, and `times_active` counting slices whose mean activation is non-zero."""
    n_feats = learned_dict.n_feats
    dev = activations.device
    times_active = torch.zeros(n_feats, device=dev)
    mean = torch.zeros(n_feats, device=dev)
    m2 = torch.zeros(n_feats, device=dev)
    m3 = torch.zeros(n_feats, device=dev)
    m4 = torch.zeros(n_feats, device=dev)
    n = 0
    for i in range(0, len(activations), batch_size):
        batch = activations[i:i + batch_size]
        st = _feature_stats(learned_dict, batch, moments=True)
        rows = batch.shape[0]
        batch_mean, batch_m2, batch_m3, batch_m4 = st["sum"] / rows, st["sum2"] / rows, st["sum3"] / rows, st["sum4"] / rows
        times_active += (batch_mean != 0).float()
        mean = (n * mean + batch_size * batch_mean) / (n + batch_size)
        m2 = (n * m2 + batch_size * batch_m2) / (n + batch_size)
        m3 = (n * m3 + batch_size * batch_m3) / (n + batch_size)
        m4 = (n * m4 + batch_size * batch_m4) / (n + batch_size)
        n += batch_size
    var = m2 - mean ** 2
    skew = m3 / torch.clamp(var ** 1.5, min=1e-8)
    kurtosis = m4 / torch.clamp(var ** 2, min=1e-8)
    return times_active, mean, var, skew, kurtosis, m4


# ---------------------------------------------------------------------------------------------
# Max-cosine-similarity family (standard_metrics.py:268-300).  The [m, g] similarity matrix is one GEMM of
# two row-normalised dictionaries; it runs through the encode kernel (operands rounded to nearest, fp32
# accumulate) with a constant bias of 2 so that its ReLU never clips (|cos| <= 1), removed after the row max.
# ---------------------------------------------------------------------------------------------
def _max_cosine_sim(rows: torch.Tensor, against: torch.Tensor) -> torch.Tensor:
    """max_g <rows[m], against[g]> for every m; both [*, d], used as given (callers pass unit-norm rows)."""
    bias = torch.full((against.shape[0],), 2.0, dtype=torch.float32, device=rows.device)
    sim, _ = _encode_native(rows, against, bias)
    return sim.max(dim=-1).values - 2.0


def mcs_duplicates(ground: LearnedDict, model: LearnedDict):
    """standard_metrics.py:268-272: max cosine similarity of each model atom with all ground atoms."""
    return _max_cosine_sim(model.get_learned_dict(), ground.get_learned_dict())


def mmcs(model: LearnedDict, model2: LearnedDict):
    """standard_metrics.py:274-275."""
    return mcs_duplicates(model, model2).mean()


def mcs_to_fixed(model: LearnedDict, truth):
    """standard_metrics.py:277-280."""
    return _max_cosine_sim(model.get_learned_dict(), truth)


def mmcs_to_fixed(model: LearnedDict, truth):
    """standard_metrics.py:282-283."""
    return mcs_to_fixed(model, truth).mean()


def mmcs_from_list(ld_list):
    """standard_metrics.py:285-295: symmetric matrix of mmcs between all pairs (lower triangle computed)."""
    n_dicts = len(ld_list)
    mmcs_t = torch.eye(n_dicts)
    for i in range(n_dicts):
        for j in range(i):
            mmcs_t[i, j] = mmcs(ld_list[i], ld_list[j])
            mmcs_t[j, i] = mmcs_t[i, j]
    return mmcs_t


def representedness(features, model: LearnedDict):
    """standard_metrics.py:297-301: mmcs the other way around."""
    return _max_cosine_sim(features, model.get_learned_dict())
