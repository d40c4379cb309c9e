"""GPU parity cases added in round 2: the two metrics that had no test, the fp16 range guard, the no_stacking /
2-D batch combination, and BASELINE.json's configs 3, 4 and 5 at their real dictionary shapes."""
import os
import warnings

import numpy as np
import pytest
import torch

from conftest import make_sparse_data, rel_err
from oracle import restated as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_fvu_top_activating_and_r_squared_against_reference_fixture():
    """standard_metrics.py:314-343, frozen from the reference by oracle/make_golden.py (metrics_extra.pt)."""
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import standard_metrics as sm
    g = torch.load(os.path.join(GOLD, "inference.pt"), weights_only=False)
    w = torch.load(os.path.join(GOLD, "metrics_extra.pt"), weights_only=False)
    We, Wd, b, x = (g[k].cuda() for k in ("We", "Wd", "b", "x"))
    tr, rot, scale = (t.cuda() for t in g["tied_centering"])
    models = {"untied": sc.UntiedSAE(We, Wd, b),
              "tied": sc.TiedSAE(We.clone(), b.clone(), centering=(tr, rot, scale), norm_encoder=True)}
    for name, m in models.items():
        r2 = float(sm.r_squared(m, x))
        assert abs(r2 - float(w[name]["r_squared"])) <= 1e-3 * (1.0 + abs(float(w[name]["r_squared"]))), (name, r2)
        for n_top in (2, 7):
            top, rest = sm.fraction_variance_unexplained_top_activating(m, x, n_top=n_top)
            wt, wr = w[name][f"fvu_top_{n_top}"]
            assert abs(float(top) - float(wt)) <= 2e-3 * float(wt), (name, n_top, float(top), float(wt))
            assert abs(float(rest) - float(wr)) <= 2e-3 * float(wr), (name, n_top, float(rest), float(wr))


def test_fp16_range_guard_saturates_flags_and_falls_back():
    """A batch with |x| > 65504 in fp16-operand mode: the step stays finite (every fp16 conversion saturates), the
    device flag is raised, and "auto" continues with TF32 operands, whose results match the fp32 oracle again."""
    import sparse_coding__b200 as sc
    M, B, d, N = 2, 256, 128, 256
    torch.manual_seed(0)
    models = [sc.FunctionalSAE.init(d, N, l1) for l1 in (1e-3, 1e-2)]
    ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    if ens.operand_format != "auto":
        pytest.skip("operand format forced by the environment")
    x = make_sparse_data(B, d, 2 * d, 6, seed=1).cuda()
    ens.step_batch(x)
    assert ens.last_operand_format == "f16" and ens.check_operand_range()
    bad = x.clone()
    bad[3, 5] = 1.0e5
    bad[7, 9] = -3.0e5
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        losses, _ = ens.step_batch(bad)
        assert all(torch.isfinite(v).all() for v in losses.values())
        assert all(torch.isfinite(v).all() for v in ens.params.values())
        assert not ens.check_operand_range()
    assert any("fp16 range" in str(c.message) for c in caught)
    assert ens.operand_format == "tf32"
    # from here on TF32 operands: the same outlier batch now matches the fp32 oracle
    orc = R.EnsembleOracle({k: v.cpu().clone() for k, v in ens.params.items()},
                           {k: v.cpu().clone() for k, v in ens.buffers.items()}, tied=False, lr=1e-3)
    lg, _ = ens.step_batch(bad)
    lc, _ = orc.step_batch(bad.cpu())
    assert ens.last_operand_format == "tf32"
    for k in ("loss", "l_reconstruction"):
        assert torch.allclose(lg[k].cpu(), lc[k], rtol=1e-3), (k, lg[k], lc[k])


def test_no_stacking_accepts_a_shared_2d_batch():
    """ensemble.py:100-116,175-178: with no_stacking the reference expands a [B,d] batch to [M,B,d] and indexes it per
    member, so a shared 2-D batch must work and give what the stacked ensemble gives."""
    import sparse_coding__b200 as sc
    M, B, d, N = 3, 130, 64, 128
    l1s = (1e-3, 3e-3, 1e-2)
    torch.manual_seed(4)
    models = [sc.FunctionalSAE.init(d, N, l1) for l1 in l1s]
    clone = [({k: v.clone() for k, v in p.items()}, {k: v.clone() for k, v in b.items()}) for p, b in models]
    a = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    b = sc.FunctionalEnsemble(clone, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda", no_stacking=True)
    x = make_sparse_data(B, d, 2 * d, 5, seed=6).cuda()
    la, _ = a.step_batch(x)
    lb, _ = b.step_batch(x)                                   # 2-D batch, no_stacking
    lc, _ = b.step_batch(x.unsqueeze(0).expand(M, B, d))      # and the explicit per-member form still works
    for k in la:
        assert torch.allclose(la[k], lb[k], rtol=1e-5, atol=1e-8), k
    assert all(torch.isfinite(v).all() for v in lc.values())


@pytest.mark.parametrize("kind", ["tied", "untied"])
def test_config3_share_of_one_gpu(kind):
    """BASELINE.json configs[2]: GPT-2 small residual, d = 768, 6144 features, 32 members over 8 GPUs = 4 per GPU.
    One step of one GPU's share against the fp32 oracle (loss terms 1e-3 relative, L0 0.5 %)."""
    import sparse_coding__b200 as sc
    M, B, d, N = 4, 256, 768, 6144
    l1s = np.logspace(-4, -2, M)
    torch.manual_seed(12)
    sig = sc.FunctionalTiedSAE if kind == "tied" else sc.FunctionalSAE
    models = [sig.init(d, N, float(l1)) for l1 in l1s]
    ens = sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    orc = R.EnsembleOracle({k: v.cpu().clone() for k, v in ens.params.items()},
                           {k: v.cpu().clone() for k, v in ens.buffers.items()}, tied=(kind == "tied"), lr=1e-3)
    data = make_sparse_data(2 * B, d, 2 * d, 12, seed=13)
    for s in range(2):
        x = data[s * B:(s + 1) * B]
        lg, ag = ens.step_batch(x.cuda())
        lc, ac = orc.step_batch(x)
        for k in ("loss", "l_reconstruction", "l_l1"):
            assert torch.allclose(lg[k].cpu(), lc[k], rtol=1e-3, atol=1e-9), (s, k, lg[k].cpu(), lc[k])
    l0g = (ag["c"] != 0).float().sum(-1).mean(-1).cpu()
    l0c = (ac["c"] != 0).float().sum(-1).mean(-1)
    assert float(((l0g - l0c).abs() / l0c.clamp_min(1.0)).max()) < 5e-3


def test_config5_dictionary_shape_gradients():
    """BASELINE.json configs[4]: one dictionary of d = 2048 x 65536 features.  Gradients of one (small) batch through
    the fused step in gradient-only mode - what every data-parallel rank computes - against the fp32 oracle."""
    import sparse_coding__b200 as sc
    B, d, N = 128, 2048, 65536
    torch.manual_seed(14)
    params, buffers = sc.FunctionalSAE.init(d, N, 1e-3)
    ens = sc.FunctionalEnsemble([(params, buffers)], sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    x = make_sparse_data(B, d, 2 * d, 16, seed=15)
    grads, losses, _ = ens.compute_grads(x.cuda())
    P = {k: v[0].cpu() for k, v in ens.params.items()}
    ls, fw = R.sae_forward(P["encoder"], P["encoder_bias"], P["decoder"], buffers["l1_alpha"], buffers["bias_decay"], x)
    gWe, gb, gWd = R.sae_backward(P["encoder"], P["encoder_bias"], P["decoder"], buffers["l1_alpha"], buffers["bias_decay"], x, fw)
    for k in ("loss", "l_reconstruction", "l_l1"):
        assert abs(float(losses[k][0]) - float(ls[k])) <= 1e-3 * abs(float(ls[k])), k
    assert rel_err(grads["decoder"][0], gWd) < 3e-3
    assert rel_err(grads["encoder_bias"][0], gb) < 2e-2
    assert rel_err(grads["encoder"][0], gWe) < 5e-2   # ReLU-mask flips at B = 128 dominate (see test_gradients_match_oracle)
    # and one full step leaves finite parameters and the same loss
    l2, _ = ens.step_batch(x.cuda())
    assert abs(float(l2["loss"][0]) - float(ls["loss"])) <= 1e-3 * abs(float(ls["loss"]))
    assert all(torch.isfinite(v).all() for v in ens.params.values())


def test_fista_config4_full_batch_sampled_rows():
    """BASELINE.json configs[3] at its full size: B = 8192 rows, d = 512, N = 4096, 300 iterations on the GPU; rows
    are independent given the dictionary and the step size, so 96 rows spread over the batch (every tile schedule
    position class) are recomputed by the float64 oracle with the same eta and must agree to 1e-4 absolute."""
    import sparse_coding__b200 as sc
    B, d, N, iters = 8192, 512, 4096, 300
    torch.manual_seed(20)
    D = R.row_normalise(torch.randn(N, d))[0]
    x = make_sparse_data(B, d, 2 * d, 10, seed=21)
    a0 = torch.relu(x @ D.T - 0.05)
    l1 = 1e-3
    eta = float(R.fista_eta(D.double()))
    a, res = sc.FunctionalFista.fista(x.cuda(), D.cuda(), l1, a0.cuda(), iters, "cuda", eta=eta)
    rows = torch.cat([torch.arange(0, 24), torch.arange(4000, 4024), torch.arange(8168, 8192),
                      torch.randint(0, B, (24,), generator=torch.Generator().manual_seed(22))])
    a_ref, res_ref = R.fista(x[rows].double(), D.double(), l1, a0[rows].double(), iters, eta=eta)
    assert float((a[rows.cuda()].cpu().double() - a_ref).abs().max()) < 1e-4
    assert float((res[rows.cuda()].cpu().double() - res_ref).abs().max()) < 1e-4
    # whole batch: the returned residual is consistent with the returned codes, and the codes are non-negative
    assert float((x.cuda() - a @ D.cuda() - res).abs().max()) < 1e-4
    assert float(a.min()) >= 0.0


def test_row_sharded_entry_points_on_one_rank_match_the_plain_functions():
    """dist.fista_rows / *_sharded with the CUDA engine as the local computation (world 1, no process group): the
    same numbers as the unsharded functions, and the two halves of a batch processed separately combine to them."""
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import dist as scd, standard_metrics as sm
    torch.manual_seed(30)
    n, d, N = 1000, 128, 256
    x = make_sparse_data(n, d, 2 * d, 6, seed=31).cuda()
    We, Wd, b = torch.randn(N, d).cuda() / d ** 0.5, torch.randn(N, d).cuda(), (0.05 * torch.randn(N) - 0.1).cuda()
    m = sc.UntiedSAE(We, Wd, b)
    assert torch.allclose(scd.mean_nonzero_activations_sharded(m, x), sm.mean_nonzero_activations(m, x), atol=1e-6)
    assert scd.n_ever_active_sharded(m, x, threshold=10) == sm.batched_calc_feature_n_ever_active(m, x, threshold=10)
    fvu = float(sm.fraction_variance_unexplained(m, x))
    assert abs(float(scd.fraction_variance_unexplained_sharded(m, x)) - fvu) <= 1e-4 * fvu
    # partial results of two row shards add up (what the all-reduce does across ranks)
    c0, c1 = sm.feature_nonzero_counts(m, x[:400]), sm.feature_nonzero_counts(m, x[400:])
    assert torch.equal(c0 + c1, sm.feature_nonzero_counts(m, x))
    r0, r1 = sm.residual_sum_of_squares(m, x[:400]), sm.residual_sum_of_squares(m, x[400:])
    assert abs(float(r0 + r1) - float(sm.residual_sum_of_squares(m, x))) <= 1e-4 * float(r0 + r1)
    D = R.row_normalise(torch.randn(N, d))[0].cuda()
    a0 = torch.relu(x @ D.T)
    a, res = scd.fista_rows(x, D, 1e-2, a0, 20, device="cuda")
    a_full, res_full = sc.FunctionalFista.fista(x, D, 1e-2, a0, 20, "cuda")
    assert float((a - a_full).abs().max()) < 1e-5 and float((res - res_full).abs().max()) < 1e-5
    a_half, _ = scd.fista_rows(x[:500], D, 1e-2, a0[:500], 20, device="cuda")
    assert float((a_half - a_full[:500]).abs().max()) < 1e-5


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, True), (False, True), (True, False)])
def test_gemm_f32_is_fp32_accurate(a_mn, b_mn):
    """sc_gemm_f32 (3xTF32): the products the reference does with torch.einsum / torch.mm around the hot path."""
    from sparse_coding__b200 import _native as nat
    torch.manual_seed(40)
    G, M, N, K = 3, 203, 64, 100
    A, B = torch.randn(G, M, K), torch.randn(G, N, K)
    ref = torch.einsum("gmk,gnk->gmn", A.double(), B.double())
    Ad = (A.transpose(1, 2).contiguous() if a_mn else A).cuda()
    Bd = (B.transpose(1, 2).contiguous() if b_mn else B).cuda()
    if a_mn and M % 4:
        with pytest.raises(RuntimeError):
            nat.gemm_f32(Ad, Bd, a_mn=a_mn, b_mn=b_mn)
        A, Ad = A[:, :200], Ad[:, :, :200].contiguous()
        ref = ref[:, :200]
    out = nat.gemm_f32(Ad, Bd, a_mn=a_mn, b_mn=b_mn)
    # fp32-level (|ref| up to ~40, K = 100; plain TF32 would be ~1e-2): the lo*lo terms and the accumulator truncation remain
    assert float((out.cpu().double() - ref).abs().max()) < 1e-4


def test_fista_large_magnitudes_take_the_tf32_pairs():
    """fp16 operand pairs hold |v| < 65504: activations with row norms beyond 8192 make the host mirror ask for TF32
    pairs (ScFistaArgs.pair_format = 1), and the codes still match the float64 oracle at the scaled tolerance."""
    import sparse_coding__b200 as sc
    B, d, N = 128, 128, 256
    torch.manual_seed(50)
    D = R.row_normalise(torch.randn(N, d))[0]
    x = 3.0e4 * make_sparse_data(B, d, 2 * d, 6, seed=51)
    a0 = torch.relu(x @ D.T)
    a_ref, _ = R.fista(x.double(), D.double(), 1e-2, a0.double(), 20)
    a, _ = sc.FunctionalFista.fista(x.cuda(), D.cuda(), 1e-2, a0.cuda(), 20, "cuda")
    assert torch.isfinite(a).all()
    assert float((a.cpu().double() - a_ref).abs().max()) < 1e-6 * float(a_ref.abs().max()) + 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["untied", "tied"])
def test_cached_step_record_is_invisible(kind):
    """step_batch reuses the previous step's argument record while only the batch changes (ensemble._run_cached).  Two
    identical ensembles see the same sequence of steps and interventions - a replaced parameter tensor, an in-place
    edit of a buffer, another batch size, a gradient-only call, another operand format - one of them with the record
    thrown away before every step: every loss and the final state must be bit-identical."""
    import sparse_coding__b200 as sc
    M, d, N = 3, 128, 256
    sig = sc.FunctionalTiedSAE if kind == "tied" else sc.FunctionalSAE

    def build():
        torch.manual_seed(7)
        models = [sig.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-3, -2, M)]
        return sc.FunctionalEnsemble(models, sig, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")

    a, b = build(), build()
    x = make_sparse_data(1024, d, 2 * d, 6, seed=11).cuda()
    log = {id(a): [], id(b): []}

    def step(ens, batch, cached):
        if not cached:
            ens._step_ctx = None
        losses, aux = ens.step_batch(batch)
        log[id(ens)].append(torch.stack([losses[k] for k in sorted(losses)]).clone())
        return aux

    def both(fn):
        fn(a, True)
        fn(b, False)

    for i in range(3):
        both(lambda e, c: step(e, x[i * 256:(i + 1) * 256], c))
    assert a._step_ctx is not None                                   # the record exists and was used
    both(lambda e, c: e.params.__setitem__("encoder", e.params["encoder"].clone() * 1.01))   # replaced tensor
    both(lambda e, c: step(e, x[0:256], c))
    both(lambda e, c: e.buffers["l1_alpha"].mul_(2.0))               # buffer edited in place
    both(lambda e, c: step(e, x[256:512], c))
    both(lambda e, c: step(e, x[0:384], c))                          # another batch size
    both(lambda e, c: e.compute_grads(x[0:256]))                     # a gradient-only call in between
    both(lambda e, c: step(e, x[256:512], c))
    both(lambda e, c: setattr(e, "operand_format", "tf32"))          # another operand format
    both(lambda e, c: step(e, x[512:768], c))
    both(lambda e, c: step(e, x[768:1024], c))
    aux_a = step(a, x[0:256], True)
    aux_b = step(b, x[0:256], False)
    assert torch.equal(aux_a["c"], aux_b["c"])
    for la, lb in zip(log[id(a)], log[id(b)]):
        assert torch.equal(la, lb)
    for k in a.params:
        assert torch.equal(a.params[k], b.params[k]), k
        assert torch.equal(a.optim_states[0].mu[k], b.optim_states[0].mu[k]), k
        assert torch.equal(a.optim_states[0].nu[k], b.optim_states[0].nu[k]), k
