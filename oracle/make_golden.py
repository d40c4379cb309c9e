"""Freeze outputs of the UNMODIFIED reference (imported from /root/reference through
oracle/ref_loader.py) into small fixtures under tests/golden/.  TEST INFRASTRUCTURE.

    python oracle/make_golden.py        # only where /root/reference exists

The fixtures travel to the GPU box (which has no /root/reference): tests compare both the CPU
restatement (oracle/restated.py) and the CUDA engine with them.
"""
import os
import sys
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")


def sparse_data(n, d, n_feats, k, seed):
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(n_feats, d, generator=g)
    feats /= feats.norm(dim=1, keepdim=True)
    codes = torch.zeros(n, n_feats)
    idx = torch.rand(n, n_feats, generator=g).argsort(1)[:, :k]
    codes.scatter_(1, idx, torch.rand(n, k, generator=g))
    return codes @ feats


def ensemble_case(ns, kind, seed):
    torch.manual_seed(seed)
    d, N, B, steps = 32, 64, 48, 6
    l1s, bds = [1e-3, 1e-2, 3e-2], [0.0, 0.01, 0.0]
    if kind in ("masked_untied", "masked_tied"):  # mixed dictionary sizes stacked to N (sae_ensemble.py:307-442)
        sig = ns.sae_ensemble.FunctionalMaskedTiedSAE if kind == "masked_tied" else ns.sae_ensemble.FunctionalMaskedSAE
        models = [sig.init(d, n, N, l1) for n, l1 in zip((24, 64, 40), l1s)]
    elif kind == "tied":
        def mk(l1, bd, i):
            if i == 0:
                return ns.FunctionalTiedSAE.init(d, N, l1, bias_decay=bd)
            rot = torch.linalg.qr(torch.randn(d, d))[0]
            return ns.FunctionalTiedSAE.init(d, N, l1, bias_decay=bd, translation=0.1 * torch.randn(d), rotation=rot,
                                             scaling=torch.rand(d) + 0.5)
        models = [mk(l1, bd, i) for i, (l1, bd) in enumerate(zip(l1s, bds))]
        sig = ns.FunctionalTiedSAE
    else:
        models = [ns.FunctionalSAE.init(d, N, l1, bias_decay=bd) for l1, bd in zip(l1s, bds)]
        sig = ns.FunctionalSAE
    for p, _ in models:
        p["encoder_bias"] += 0.01 * torch.randn(N)
    ens = ns.FunctionalEnsemble(models, sig, ns.torchopt.adam, {"lr": 1e-3}, device="cpu")
    init_params = {k: v.clone() for k, v in ens.params.items()}
    buffers = {k: v.clone() for k, v in ens.buffers.items()}
    data = sparse_data(B * steps, d, 2 * d, 4, seed + 1)
    losses = []
    for s in range(steps):
        l, aux = ens.step_batch(data[s * B:(s + 1) * B])
        losses.append({k: v.clone() for k, v in l.items()})
    st = ens.optim_states[0]
    return {
        "kind": kind, "lr": 1e-3, "init_params": init_params, "buffers": buffers, "data": data, "batch": B,
        "losses": losses, "final_params": {k: v.clone() for k, v in ens.params.items()},
        "final_mu": {k: v.clone() for k, v in st.mu.items()}, "final_nu": {k: v.clone() for k, v in st.nu.items()},
        "final_count": {k: v.clone() for k, v in st.count.items()}, "last_c": aux["c"].clone(),
    }


def inference_case(ns, seed):
    torch.manual_seed(seed)
    d, N, n = 32, 96, 1100
    We = torch.randn(N, d) / d ** 0.5
    Wd = torch.randn(N, d)
    b = 0.05 * torch.randn(N) - 0.1
    x = sparse_data(n, d, 2 * d, 4, seed + 1)
    ld = ns.UntiedSAE(We, Wd, b)
    rot = torch.linalg.qr(torch.randn(d, d))[0]
    tied = ns.TiedSAE(We.clone(), b.clone(), centering=(0.1 * torch.randn(d), rot, torch.rand(d) + 0.5), norm_encoder=True)
    out = {"We": We, "Wd": Wd, "b": b, "x": x,
           "tied_centering": (tied.center_trans, tied.center_rot, tied.center_scale)}
    for name, m in (("untied", ld), ("tied", tied)):
        c = m.encode(m.center(x))
        out[name] = {
            "c_head": c[:256].clone(), "predict_head": m.predict(x)[:256].clone(), "learned_dict": m.get_learned_dict(),
            "fvu": ns.fraction_variance_unexplained(m, x), "mean_nonzero": ns.mean_nonzero_activations(m, x),
            "n_ever_active_10": ns.batched_calc_feature_n_ever_active(m, m.center(x), threshold=10),
            "n_ever_active_1": ns.batched_calc_feature_n_ever_active(m, m.center(x)[:1000], threshold=1),
            "moments": ns.calc_moments_streaming(m, m.center(x), batch_size=1000),
        }
    return out


def fista_case(ns, seed):
    torch.manual_seed(seed)
    d, N, B = 32, 96, 40
    D = torch.randn(N, d)
    D = D / D.norm(dim=1, keepdim=True)
    x = sparse_data(B, d, 2 * d, 4, seed + 1)
    a0 = torch.relu(x @ D.T)
    out = {"D": D, "x": x, "a0": a0, "l1": 1e-2}
    for iters in (1, 10, 60):
        a, res = ns.FunctionalFista.fista(x, D, torch.tensor(1e-2), a0, iters, "cpu")
        out[f"a_{iters}"], out[f"res_{iters}"] = a, res
    out["eta"] = 1.0 / torch.linalg.eigvalsh(D @ D.T).max()
    # dictionary_update only runs for N == d (fista.py:131-138 shapes)
    Dq = torch.randn(d, d)
    Dq = Dq / Dq.norm(dim=1, keepdim=True)
    cq = torch.relu(x @ Dq.T)
    buffers = {"l1_alpha": torch.tensor(1e-2), "hessian_diag": torch.zeros(d)}
    new_dict, res = ns.FunctionalFista.dictionary_update({"encoder": Dq}, buffers, x, cq, Dq)
    out["du"] = {"D": Dq, "c": cq, "new_dict": new_dict, "res": res, "hessian_diag": buffers["hessian_diag"]}
    return out


def metrics_extra_case(ns):
    """fraction_variance_unexplained_top_activating (standard_metrics.py:314-340) and r_squared (:342-343) on the models
    and data of inference.pt (read back from that fixture, so the two files describe the same inputs)."""
    g = torch.load(os.path.join(OUT, "inference.pt"), weights_only=False)
    ld = ns.UntiedSAE(g["We"], g["Wd"], g["b"])
    tr, rot, scale = g["tied_centering"]
    tied = ns.TiedSAE(g["We"].clone(), g["b"].clone(), centering=(tr, rot, scale), norm_encoder=True)
    out = {}
    for name, m in (("untied", ld), ("tied", tied)):
        out[name] = {"r_squared": ns.r_squared(m, g["x"]),
                     "fvu_top_2": ns.fraction_variance_unexplained_top_activating(m, g["x"], n_top=2),
                     "fvu_top_7": ns.fraction_variance_unexplained_top_activating(m, g["x"], n_top=7)}
    return out


def main():
    ns = ref_loader.load()
    os.makedirs(OUT, exist_ok=True)
    if "--metrics-extra-only" in sys.argv:  # added later: leaves the older fixtures byte-identical
        torch.save(metrics_extra_case(ns), os.path.join(OUT, "metrics_extra.pt"))
        return
    torch.save(ensemble_case(ns, "untied", 100), os.path.join(OUT, "ensemble_untied.pt"))
    torch.save(ensemble_case(ns, "tied", 200), os.path.join(OUT, "ensemble_tied.pt"))
    torch.save(ensemble_case(ns, "masked_untied", 500), os.path.join(OUT, "ensemble_masked_untied.pt"))
    torch.save(ensemble_case(ns, "masked_tied", 600), os.path.join(OUT, "ensemble_masked_tied.pt"))
    torch.save(inference_case(ns, 300), os.path.join(OUT, "inference.pt"))
    torch.save(fista_case(ns, 400), os.path.join(OUT, "fista.pt"))
    torch.save(metrics_extra_case(ns), os.path.join(OUT, "metrics_extra.pt"))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
