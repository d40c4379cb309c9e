"""CPU restatement of the reference hot path.  TEST INFRASTRUCTURE - NOT PRODUCT.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this file, and only as the checker or as the
timed CPU baseline.  The product path (sparse_coding__b200) never imports it.

Parity status: PINNED against the reference's own code run in the build container
(`oracle/ref_loader.py` imports /root/reference unmodified; `tests/test_oracle_*`
compare every function here with it, and `oracle/make_golden.py` froze reference
outputs into tests/golden/).  The reference ships no golden vectors or
known-answer tests for this path.  The one third-party piece,
torchopt 0.7.1's Adam, is restated from its published algorithm in
`oracle/shims/torchopt` and is UNPINNED against a real wheel (none available).

Everything is written with explicit forward AND backward formulas (no autograd)
so that it states exactly what the CUDA kernels must compute.  Plain torch ops on
CPU tensors, fp32 by default (fp64 if the inputs are fp64).

Notation: x [B,d] batch, We/Wd [N,d], b [N], lam = l1_alpha, beta = bias_decay.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch

Tensor = torch.Tensor


# --------------------------------------------------------------------------- #
# FunctionalSAE  (autoencoders/sae_ensemble.py:13-77)                          #
# --------------------------------------------------------------------------- #
def row_normalise(W: Tensor) -> Tuple[Tensor, Tensor]:
    """sae_ensemble.py:58-59 / learned_dict.py:118-120: D = W / clamp(||W||_row, 1e-8)."""
    norms = torch.sqrt((W * W).sum(-1))
    cl = torch.clamp(norms, min=1e-8)
    return W / cl[:, None], cl


def sae_forward(We: Tensor, b: Tensor, Wd: Tensor, lam, beta, x: Tensor, mask: Optional[Tensor] = None):
    """FunctionalSAE.loss forward, sae_ensemble.py:53-77.  With `mask` ([N] bool, True = unused feature) it is
    FunctionalMaskedSAE.loss, :409-438 (codes of unused features zeroed; that loss has no bias-decay term, so
    callers pass beta = 0)."""
    z = x @ We.T + b                       # :54-55
    c = torch.clamp(z, min=0.0)            # :56
    if mask is not None:
        c = c.masked_fill(mask, 0.0)       # :425
    D, cl = row_normalise(Wd)              # :58-59
    x_hat = c @ D                          # :61
    r = x_hat - x
    l_rec = (r * r).mean()                 # :62
    l_l1 = lam * c.abs().sum(-1).mean()    # :63
    bn = torch.sqrt((b * b).sum())
    l_bd = beta * bn                       # :64
    losses = {"loss": l_rec + l_l1 + l_bd, "l_reconstruction": l_rec, "l_l1": l_l1, "l_bias_decay": l_bd}
    return losses, dict(z=z, c=c, D=D, cl=cl, r=r, bn=bn, mask=mask)


def sae_backward(We, b, Wd, lam, beta, x, fw):
    """Gradient of sae_forward's total loss w.r.t. (We, b, Wd); what
    torch.func.grad(FunctionalSAE.loss) returns (ensemble.py:119-122).
    Edge semantics of the reference's autograd: clamp passes gradient at z == 0,
    d|c|/dc = sign(c) = 0 at c == 0, d||b||/db = 0 at b == 0."""
    B, d = x.shape
    z, c, D, cl, r = fw["z"], fw["c"], fw["D"], fw["cl"], fw["r"]
    R = r * (2.0 / (B * d))
    q = R @ D.T                                               # dL/dc from the MSE term
    dz = (q + (lam / B) * (c > 0).to(x.dtype)) * (z >= 0).to(x.dtype)
    if fw.get("mask") is not None:
        dz = dz.masked_fill(fw["mask"], 0.0)                  # masked_fill_ blocks the gradient
    gWe = dz.T @ x
    gb = dz.sum(0)
    if float(fw["bn"]) > 0.0:
        gb = gb + beta * b / fw["bn"]
    gD = c.T @ R
    s = (gD * D).sum(-1)                                      # == sum_b c*q
    gWd = (gD - s[:, None] * D) / cl[:, None]
    # rows whose norm is below the clamp are divided by the constant 1e-8: no projection term
    small = torch.sqrt((Wd * Wd).sum(-1)) < 1e-8
    if small.any():
        gWd[small] = gD[small] / 1e-8
    return gWe, gb, gWd


# --------------------------------------------------------------------------- #
# FunctionalTiedSAE  (autoencoders/sae_ensemble.py:80-160)                     #
# --------------------------------------------------------------------------- #
def tied_center(rot, trans, scale, x):
    """sae_ensemble.py:126-127."""
    return ((x - trans[None, :]) @ rot.T) * scale[None, :]


def tied_uncenter(rot, trans, scale, xc):
    """sae_ensemble.py:130-131."""
    return (xc / scale[None, :]) @ rot + trans[None, :]


def tied_forward(W, b, lam, beta, x, rot=None, trans=None, scale=None, mask=None):
    """FunctionalTiedSAE.loss forward, sae_ensemble.py:134-160.  `loss_data` has
    no l_bias_decay key (:150-154) although the total includes it (:160)."""
    D, cl = row_normalise(W)                                   # :135-136
    xc = x if rot is None else tied_center(rot, trans, scale, x)  # :138
    z = xc @ D.T + b                                           # :140-141
    c = torch.clamp(z, min=0.0)                                # :142
    if mask is not None:
        c = c.masked_fill(mask, 0.0)                           # FunctionalMaskedTiedSAE, :356
    x_hat_c = c @ D                                            # :144
    r = x_hat_c - xc
    l_rec = (r * r).mean()                                     # :146
    l_l1 = lam * c.abs().sum(-1).mean()                        # :147
    bn = torch.sqrt((b * b).sum())
    l_bd = beta * bn                                           # :148
    losses = {"loss": l_rec + l_l1 + l_bd, "l_reconstruction": l_rec, "l_l1": l_l1}
    return losses, dict(z=z, c=c, D=D, cl=cl, r=r, bn=bn, xc=xc, mask=mask)


def tied_backward(W, b, lam, beta, x, fw):
    B, d = fw["xc"].shape
    z, c, D, cl, r, xc = fw["z"], fw["c"], fw["D"], fw["cl"], fw["r"], fw["xc"]
    R = r * (2.0 / (B * d))
    q = R @ D.T
    dz = (q + (lam / B) * (c > 0).to(x.dtype)) * (z >= 0).to(x.dtype)
    if fw.get("mask") is not None:
        dz = dz.masked_fill(fw["mask"], 0.0)
    gD = c.T @ R + dz.T @ xc                                   # decode + encode use the same D
    s = (gD * D).sum(-1)
    gW = (gD - s[:, None] * D) / cl[:, None]
    small = torch.sqrt((W * W).sum(-1)) < 1e-8
    if small.any():
        gW[small] = gD[small] / 1e-8
    gb = dz.sum(0)
    if float(fw["bn"]) > 0.0:
        gb = gb + beta * b / fw["bn"]
    return gW, gb


# --------------------------------------------------------------------------- #
# Adam as the reference actually executes it                                   #
# (ensemble.py:175-193 + torchopt 0.7.1, see oracle/shims/torchopt)            #
# --------------------------------------------------------------------------- #
def adam_step_(p: Tensor, g: Tensor, mu: Tensor, nu: Tensor, lr, b1=0.9, b2=0.999, eps=1e-8,
               eps_root=0.0, t: int = 1):
    """In place.  t is the bias-correction step: the reference never advances its
    stored count (ensemble.py:185-189 writes into a clone), so t == 1 always;
    t > 1 gives textbook Adam for the engine's `count_mode=1`."""
    mu.mul_(b1).add_(g, alpha=1 - b1)
    nu.mul_(b2).addcmul_(g, g, value=1 - b2)
    mu_hat = mu / (1 - b1 ** t)
    nu_hat = nu / (1 - b2 ** t)
    upd = mu_hat / (torch.sqrt(nu_hat + eps_root) + eps)
    p.add_(upd, alpha=-lr)


class EnsembleOracle:
    """FunctionalEnsemble.step_batch (ensemble.py:175-193) for sig in
    {FunctionalSAE, FunctionalFista (same loss, fista.py:60-84), FunctionalTiedSAE},
    written member by member with the explicit formulas above."""

    def __init__(self, params: Dict[str, Tensor], buffers: Dict[str, Tensor], tied: bool, lr=1e-3,
                 b1=0.9, b2=0.999, eps=1e-8, frozen_count=True):
        self.params = {k: v.clone() for k, v in params.items()}
        self.buffers = {k: v.clone() for k, v in buffers.items()}
        self.tied = tied
        self.M = self.params["encoder"].shape[0]
        self.lr, self.b1, self.b2, self.eps = lr, b1, b2, eps
        self.frozen = frozen_count
        self.t = 0
        self.mu = {k: torch.zeros_like(v) for k, v in self.params.items()}
        self.nu = {k: torch.zeros_like(v) for k, v in self.params.items()}

    def step_batch(self, x: Tensor, expand_dims=True):
        P, Bf = self.params, self.buffers
        self.t += 1
        t = 1 if self.frozen else self.t
        out = {}
        cs = []
        for m in range(self.M):
            xm = x if expand_dims else x[m]
            lam, beta = Bf["l1_alpha"][m], Bf["bias_decay"][m]
            mask = Bf["coef_mask"][m] if "coef_mask" in Bf else None
            if mask is not None:
                beta = 0.0  # the masked losses have no bias-decay term (sae_ensemble.py:360, :429)
            if self.tied and mask is not None:
                losses, fw = tied_forward(P["encoder"][m], P["encoder_bias"][m], lam, beta, xm, mask=mask)
                gW, gb = tied_backward(P["encoder"][m], P["encoder_bias"][m], lam, beta, xm, fw)
                grads = {"encoder": gW, "encoder_bias": gb}
            elif self.tied:
                ident = _is_identity_centering(Bf, m)
                args = (None, None, None) if ident else (Bf["center_rot"][m], Bf["center_trans"][m], Bf["center_scale"][m])
                losses, fw = tied_forward(P["encoder"][m], P["encoder_bias"][m], lam, beta, xm, *args)
                gW, gb = tied_backward(P["encoder"][m], P["encoder_bias"][m], lam, beta, xm, fw)
                grads = {"encoder": gW, "encoder_bias": gb}
            else:
                losses, fw = sae_forward(P["encoder"][m], P["encoder_bias"][m], P["decoder"][m], lam, beta, xm, mask)
                gWe, gb, gWd = sae_backward(P["encoder"][m], P["encoder_bias"][m], P["decoder"][m], lam, beta, xm, fw)
                grads = {"encoder": gWe, "encoder_bias": gb, "decoder": gWd}
            for k, g in grads.items():
                adam_step_(P[k][m], g, self.mu[k][m], self.nu[k][m], self.lr, self.b1, self.b2, self.eps, 0.0, t)
            if mask is not None:
                losses = {k: losses[k] for k in ("loss", "l_reconstruction", "l_l1")}
            for k, v in losses.items():
                out.setdefault(k, []).append(v)
            cs.append(fw["c"])
        return {k: torch.stack(v) for k, v in out.items()}, {"c": torch.stack(cs)}


def _is_identity_centering(buffers, m) -> bool:
    rot, tr, sc = buffers["center_rot"][m], buffers["center_trans"][m], buffers["center_scale"][m]
    d = rot.shape[0]
    return bool(torch.equal(rot, torch.eye(d, dtype=rot.dtype, device=rot.device)) and not tr.any() and torch.equal(sc, torch.ones_like(sc)))


# --------------------------------------------------------------------------- #
# LearnedDict inference side  (autoencoders/learned_dict.py:13-50,111-196)     #
# --------------------------------------------------------------------------- #
def untied_encode(We, b, x):
    """UntiedSAE.encode, learned_dict.py:127-131."""
    return torch.clamp(x @ We.T + b, min=0.0)


def tied_encode(W, b, x, norm_encoder=True):
    """TiedSAE.encode, learned_dict.py:186-196."""
    enc = row_normalise(W)[0] if norm_encoder else W
    return torch.clamp(x @ enc.T + b, min=0.0)


def decode(Wd, c):
    """LearnedDict.decode, learned_dict.py:29-32 (get_learned_dict renormalises, :118-120 / :172-174)."""
    return c @ row_normalise(Wd)[0]


def untied_predict(We, b, Wd, x):
    """LearnedDict.predict for UntiedSAE, learned_dict.py:42-47 (center/uncenter are the identity)."""
    return decode(Wd, untied_encode(We, b, x))


def tied_predict(W, b, x, rot, trans, scale, norm_encoder=True):
    xc = tied_center(rot, trans, scale, x)
    return tied_uncenter(rot, trans, scale, decode(W, tied_encode(W, b, xc, norm_encoder)))


# --------------------------------------------------------------------------- #
# FISTA  (autoencoders/fista.py:99-128; method twin :274-301)                  #
# --------------------------------------------------------------------------- #
def fista_eta(D: Tensor) -> Tensor:
    """fista.py:105-106: 1 / lambda_max(D D^T).  Same largest eigenvalue as the
    d x d Gram D^T D (non-zero spectra coincide); the reference uses the N x N one."""
    n, d = D.shape
    G = D @ D.T if n <= d else D.T @ D
    return 1.0 / torch.linalg.eigvalsh(G).max().to(D.dtype)


def fista_t_sequence(num_iter: int):
    """fista.py:108-118: t_{k+1} = (1 + sqrt(1 + 4 t_k^2)) / 2 in Python/numpy float64;
    returns the per-iteration momentum factors (t_k - 1) / t_{k+1}."""
    tk_n = 1.0
    out = []
    for _ in range(num_iter):
        tk = tk_n
        tk_n = (1 + np.sqrt(1 + 4 * tk ** 2)) / 2
        out.append((tk - 1) / tk_n)
    return out


def fista(x: Tensor, D: Tensor, l1_coef, a0: Tensor, num_iter: int, eta=None):
    """FunctionalFista.fista, fista.py:99-128.  Returns (ahat, Res)."""
    if eta is None:
        eta = fista_eta(D)
    a = a0
    y = a0
    for mom in fista_t_sequence(num_iter):
        a_pre = a
        res = x - y @ D                                       # :120
        y = y + eta * (res @ D.T)                             # :121
        a = torch.clamp(y - eta * l1_coef, min=0.0)           # :122
        y = a + (a - a_pre) * mom                             # :123
    res = x - a @ D                                           # :126
    return a, res


def quadratic_basis_update(D, res, ahat, lowest_activation, hessian_diag, step_size=0.001):
    """FunctionalFista.quadraticBasisUpdate, fista.py:131-138.  Only shape-valid when
    N == d (dBasis is [d, N], divided by an [N] vector, transposed and added to [N, d]);
    the final normalisation is over dim 0, i.e. COLUMN norms."""
    dB = step_size * (res.T @ ahat) / ahat.shape[0]
    dB = dB / (hessian_diag + lowest_activation)
    D2 = D + dB.T
    return D2 / torch.sqrt((D2 * D2).sum(0))


def dictionary_update(D, x, c, l1_coef, hessian_diag):
    """FunctionalFista.dictionary_update, fista.py:88-96.  The Hessian running mean is
    assigned into a throw-away dict by the caller (big_sweep.py:187-192 passes
    `separate_tensors` copies), so it never persists; it is still used within the call."""
    a, res = fista(x, D, l1_coef, c, 500)
    hd = hessian_diag * ((300 - 1.0) / 300) + (a * a).mean(0) / 300
    return quadratic_basis_update(D, res, a, 0.001, hd), res, hd


# --------------------------------------------------------------------------- #
# Metrics  (standard_metrics.py:303-340,439-452,480-509)                       #
# --------------------------------------------------------------------------- #
def mean_nonzero_activations(c: Tensor) -> Tensor:
    """standard_metrics.py:303-306 given c = encode(center(x))."""
    return (c != 0).to(torch.float32).mean(0)


def fraction_variance_unexplained(x: Tensor, x_hat: Tensor) -> Tensor:
    """standard_metrics.py:308-312 given x_hat = predict(x)."""
    resid = ((x - x_hat) ** 2).mean()
    total = ((x - x.mean(0)) ** 2).mean()
    return resid / total


def n_ever_active(c: Tensor, threshold=10, batch_size=1000) -> int:
    """standard_metrics.py:444-452: counts accumulate in fp32 over 1000-row slices."""
    cnt = torch.zeros(c.shape[1])
    for i in range(0, c.shape[0], batch_size):
        cnt += (c[i:i + batch_size] != 0).sum(0)
    return int((cnt > threshold).sum().item())


def moments_streaming(c: Tensor, batch_size=1000):
    """standard_metrics.py:480-509 (including its quirk of weighting every slice by
    `batch_size` even when the last one is shorter)."""
    N = c.shape[1] if c.dim() > 1 else 1
    times_active = torch.zeros(N)
    mean = torch.zeros(N)
    m2 = torch.zeros(N)
    m3 = torch.zeros(N)
    m4 = torch.zeros(N)
    n = 0
    for i in range(0, len(c), batch_size):
        f = c[i:i + batch_size]
        bm, b2, b3, b4 = f.mean(0), (f ** 2).mean(0), (f ** 3).mean(0), (f ** 4).mean(0)
        times_active += (bm != 0).float()
        mean = (n * mean + batch_size * bm) / (n + batch_size)
        m2 = (n * m2 + batch_size * b2) / (n + batch_size)
        m3 = (n * m3 + batch_size * b3) / (n + batch_size)
        m4 = (n * m4 + batch_size * b4) / (n + batch_size)
        n += batch_size
    var = m2 - mean ** 2
    skew = m3 / torch.clamp(var ** 1.5, min=1e-8)
    kurt = m4 / torch.clamp(var ** 2, min=1e-8)
    return times_active, mean, var, skew, kurt, m4


# --------------------------------------------------------------------------- #
# Init (same RNG consumption order as sae_ensemble.py:15-38 / :82-124)         #
# --------------------------------------------------------------------------- #
def xavier_uniform(n, d, generator=None, dtype=torch.float32):
    bound = math.sqrt(6.0 / (n + d))
    return (torch.rand(n, d, generator=generator, dtype=dtype) * 2 - 1) * bound


# --------------------------------------------------------------------------- #
# TF32-operand emulation of the untied step (what the GPU engine is DESIGNED to  #
# compute: every GEMM operand rounded to TF32 round-to-nearest, fp32 accumulate). #
# A sharper target than the fp32 reference: differences to it are accumulation    #
# order only.  Used next to - never instead of - the fp32 parity checks.          #
# --------------------------------------------------------------------------- #
def rn_tf32(t: Tensor) -> Tensor:
    i = t.contiguous().view(torch.int32)
    return ((i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF).view(torch.float32)


def sae_grads_tf32(We, b, Wd, lam, beta, x):
    q = rn_tf32
    B, d = x.shape
    z = q(x) @ q(We).T + b
    c = torch.clamp(z, min=0.0)
    D, cl = row_normalise(Wd)
    cr, Dr = q(c), q(D)
    r = cr @ Dr - x
    Rm = q(r * (2.0 / (B * d)))
    qq = Rm @ Dr.T
    dz = (qq + (lam / B) * (c > 0).to(x.dtype)) * (z >= 0).to(x.dtype)
    gWe = q(dz).T @ q(x)
    bn = torch.sqrt((b * b).sum())
    gb = dz.sum(0) + (beta * b / bn if float(bn) > 0 else 0.0)
    gD = cr.T @ Rm
    s = (cr * qq).sum(0)
    gWd = (gD - s[:, None] * D) / cl[:, None]
    return gWe, gb, gWd, cr
