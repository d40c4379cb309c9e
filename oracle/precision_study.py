"""Which tensor-core operand format keeps the SAE step inside the parity gates?

TEST INFRASTRUCTURE (a one-off study, results quoted in DESIGN.md).  Emulates the
GPU engine on CPU: every GEMM operand is rounded to the candidate format, products
are accumulated in fp32 (what tcgen05 does), everything else stays fp32.  Compares
100-step trajectories with the fp32 restatement on identical inputs.

    python oracle/precision_study.py
"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import restated as R


def rn_tf32(t):
    i = t.contiguous().view(torch.int32)
    i = (i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF
    return i.view(torch.float32)


def trunc_tf32(t):
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def rn_bf16(t):
    return t.to(torch.bfloat16).to(torch.float32)


def rn_fp16(t):
    return t.to(torch.float16).to(torch.float32)


FMT = {"fp32": lambda t: t, "tf32_rn": rn_tf32, "tf32_trunc": trunc_tf32, "bf16_rn": rn_bf16, "fp16_rn": rn_fp16}
# fp16 has a narrow exponent: the engine's fp16 mode stores the residual r (not R = 2r/(B d)) and the
# correspondingly scaled dz, and applies 2/(B d) in fp32 afterwards - emulated the same way below.
UNSCALED_RESIDUAL = {"fp16_rn"}


def step(P, S, lam, x, lr, q, unscaled=False):
    """one untied member step with operand rounding q (explicit formulas of restated.sae_*)"""
    We, b, Wd = P
    B, d = x.shape
    if unscaled:
        return step_unscaled(P, S, lam, x, lr, q)
    z = q(x) @ q(We).T + b
    c = torch.clamp(z, min=0)
    D, cl = R.row_normalise(Wd)
    xh = q(c) @ q(D)
    r = xh - x
    l_rec = (r * r).mean(); l_l1 = lam * c.sum(-1).mean()
    Rm = r * (2.0 / (B * d))
    qq = q(Rm) @ q(D).T
    dz = (qq + lam / B * (c > 0).float()) * (z >= 0).float()
    gWe = q(dz).T @ q(x)
    gb = dz.sum(0)
    gD = q(c).T @ q(Rm)
    s = (c * qq).sum(0)
    gWd = (gD - s[:, None] * D) / cl[:, None]
    for p, g, (mu, nu) in zip(P, (gWe, gb, gWd), S):
        R.adam_step_(p, g, mu, nu, lr)
    l0 = (c != 0).float().sum(-1).mean()
    fvu = (r * r).mean() / ((x - x.mean(0)) ** 2).mean()
    return float(l_rec + l_l1), float(l_rec), float(l_l1), float(l0), float(fvu)


def step_unscaled(P, S, lam, x, lr, q):
    We, b, Wd = P
    B, d = x.shape
    z = q(x) @ q(We).T + b
    c = torch.clamp(z, min=0)
    D, cl = R.row_normalise(Wd)
    r = q(c) @ q(D) - x
    l_rec = (r * r).mean(); l_l1 = lam * c.sum(-1).mean()
    sc = 2.0 / (B * d)
    qq = q(r) @ q(D).T                                   # q / sc
    dz = (qq + lam / B / sc * (c > 0).float()) * (z >= 0).float()
    gWe = (q(dz).T @ q(x)) * sc
    gb = dz.sum(0) * sc
    gD = (q(c).T @ q(r)) * sc
    s = (c * qq).sum(0) * sc
    gWd = (gD - s[:, None] * D) / cl[:, None]
    for p, g, (mu, nu) in zip(P, (gWe, gb, gWd), S):
        R.adam_step_(p, g, mu, nu, lr)
    l0 = (c != 0).float().sum(-1).mean()
    fvu = (r * r).mean() / ((x - x.mean(0)) ** 2).mean()
    return float(l_rec + l_l1), float(l_rec), float(l_l1), float(l0), float(fvu)


def make_data(n, d, n_feats, k_active, gen):
    feats = torch.randn(n_feats, d, generator=gen)
    feats /= feats.norm(dim=1, keepdim=True)
    codes = torch.zeros(n, n_feats)
    idx = torch.rand(n, n_feats, generator=gen).argsort(1)[:, :k_active]
    codes.scatter_(1, idx, torch.rand(n, k_active, generator=gen))
    return codes @ feats


def main(d=256, N=1024, B=1024, steps=100, lam=1e-3):
    gen = torch.Generator().manual_seed(0)
    data = make_data(B * 8, d, 2 * d, 8, gen)
    We0 = R.xavier_uniform(N, d, gen); Wd0 = R.xavier_uniform(N, d, gen)
    traj = {}
    for name, q in FMT.items():
        P = [We0.clone(), torch.zeros(N), Wd0.clone()]
        S = [(torch.zeros_like(p), torch.zeros_like(p)) for p in P]
        out = []
        g2 = torch.Generator().manual_seed(1)
        for i in range(steps):
            idx = torch.randint(0, data.shape[0], (B,), generator=g2)
            out.append(step(P, S, lam, data[idx], 1e-3, q, name in UNSCALED_RESIDUAL))
        traj[name] = np.array(out)
    ref = traj["fp32"]
    print(f"d={d} N={N} B={B} steps={steps} lam={lam}; final fp32: loss={ref[-1,0]:.5f} L0={ref[-1,3]:.1f} FVU={ref[-1,4]:.4f}")
    for name in FMT:
        if name == "fp32":
            continue
        t = traj[name]
        rel = np.abs(t - ref) / np.abs(ref)
        print(f"{name:11s} max-over-steps rel err: loss {rel[:,0].max():.2e}  l_rec {rel[:,1].max():.2e}  l_l1 {rel[:,2].max():.2e}  L0 {rel[:,3].max():.2e}  FVU {rel[:,4].max():.2e}")


if __name__ == "__main__":
    main()
    main(lam=1e-2)
    main(d=128, N=512, B=256, steps=300, lam=3e-3)
