"""Parity of the inference-side kernels (encode / decode / metrics / FISTA / raw GEMM) with the
CPU oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from conftest import make_sparse_data, rel_err
from oracle import restated as R

pytestmark = pytest.mark.gpu


def _tf32(t):
    i = t.contiguous().view(torch.int32)
    return ((i + 0x0FFF + ((i >> 13) & 1)) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("shape", [(2, 200, 136, 100), (1, 128, 256, 64), (3, 33, 520, 36)])
def test_gemm_core_exact_on_tf32_inputs(a_mn, b_mn, shape):
    import ctypes as C
    from sparse_coding__b200 import _native as nat
    G, M, N, K = shape
    Mp, Np = (M + 3) // 4 * 4, (N + 3) // 4 * 4
    torch.manual_seed(0)
    A = _tf32(torch.randn(G, Mp, K))
    Bm = _tf32(torch.randn(G, Np, K))
    ref = torch.einsum("gmk,gnk->gmn", A.double(), Bm.double())
    Ad = (A.transpose(1, 2).contiguous() if a_mn else A).cuda()
    Bd = (Bm.transpose(1, 2).contiguous() if b_mn else Bm).cuda()
    out = torch.full((G, Mp, Np), float("nan"), device="cuda")
    L = nat.lib()
    nat.check(L.sc_gemm_tf32(nat.ptr(Ad), nat.ptr(Bd), nat.ptr(out), G, Mp, Np, K, a_mn, b_mn,
                             nat.stream_ptr(out.device)), "sc_gemm_tf32")
    assert float((out.cpu().double() - ref).abs().max()) < 1e-4 * np.sqrt(K)


@pytest.mark.parametrize("shape", [(300, 64, 256), (50, 500, 2000), (1000, 512, 512)])
def test_untied_encode_decode_predict(shape):
    import sparse_coding__b200 as sc
    B, d, N = shape
    torch.manual_seed(1)
    We, Wd, b = R.xavier_uniform(N, d), R.xavier_uniform(N, d), 0.01 * torch.randn(N)
    x = make_sparse_data(B, d, 2 * d, 6, seed=2)
    ld = sc.UntiedSAE(We.cuda(), Wd.cuda(), b.cuda())
    c_ref = R.untied_encode(We, b, x)
    c = ld.encode(x.cuda())
    assert rel_err(c, c_ref) < 2e-3
    assert rel_err(ld.get_learned_dict(), R.row_normalise(Wd)[0]) < 1e-6
    assert rel_err(ld.decode(c_ref.cuda()), R.decode(Wd, c_ref)) < 2e-3
    assert rel_err(ld.predict(x.cuda()), R.untied_predict(We, b, Wd, x)) < 3e-3
    assert ld.n_dict_components() == N and ld.n_feats == N and ld.activation_size == d


def test_tied_with_centering_and_metrics():
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import standard_metrics as sm
    B, d, N = 2000, 64, 256
    torch.manual_seed(3)
    W, b = R.xavier_uniform(N, d), 0.02 * torch.randn(N)
    rot = torch.linalg.qr(torch.randn(d, d))[0]
    trans, scale = 0.1 * torch.randn(d), torch.rand(d) + 0.5
    x = make_sparse_data(B, d, 2 * d, 5, seed=4)
    for centering in [(None, None, None), (trans, rot, scale)]:
        cent_gpu = tuple(None if t is None else t.cuda() for t in centering)
        ld = sc.TiedSAE(W.cuda(), b.cuda(), centering=cent_gpu, norm_encoder=True)
        ld.to_device("cuda")
        if centering[0] is None:
            r_, t_, s_ = torch.eye(d), torch.zeros(d), torch.ones(d)
        else:
            t_, r_, s_ = centering
        xc = R.tied_center(r_, t_, s_, x)
        c_ref = R.tied_encode(W, b, xc)
        pred_ref = R.tied_predict(W, b, x, r_, t_, s_)
        assert rel_err(ld.encode(ld.center(x.cuda())), c_ref) < 2e-3
        assert rel_err(ld.predict(x.cuda()), pred_ref) < 3e-3
        fvu_ref = float(R.fraction_variance_unexplained(x, pred_ref))
        fvu = float(sm.fraction_variance_unexplained(ld, x.cuda()))
        assert abs(fvu - fvu_ref) <= 1e-3 * fvu_ref
        l0_ref = float(R.mean_nonzero_activations(c_ref).sum())
        l0 = float(sm.mean_nonzero_activations(ld, x.cuda()).sum())
        assert abs(l0 - l0_ref) <= 5e-3 * l0_ref
        n_ref = R.n_ever_active(c_ref, threshold=10)
        n_gpu = sm.batched_calc_feature_n_ever_active(ld, ld.center(x.cuda()), threshold=10)
        assert abs(n_gpu - n_ref) <= 1


def test_moments_streaming_matches_reference_test():
    """test/test_stats_batched.py:11-27 tolerances (places 5/3/3/2) on an encoder-shaped problem."""
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import standard_metrics as sm
    n, d, N = 10000, 32, 64
    torch.manual_seed(5)
    We, b = torch.randn(N, d) / np.sqrt(d), torch.zeros(N)
    x = torch.randn(n, d)
    ld = sc.UntiedSAE(We.cuda(), We.cuda(), b.cuda())
    c = R.untied_encode(We, b, x)
    ref = R.moments_streaming(c, batch_size=1000)
    got = sm.calc_moments_streaming(ld, x.cuda(), batch_size=1000)
    names = ["times_active", "mean", "var", "skew", "kurtosis", "m4"]
    tol = [0.5, 2e-3, 4e-3, 2e-2, 5e-2, 4e-3]
    for nm, g, r, t in zip(names, got, ref, tol):
        assert float((g.cpu() - r).abs().max()) <= t * max(1.0, float(r.abs().max())), nm


@pytest.mark.parametrize("shape,iters", [((256, 64, 256), 50), ((50, 500, 2000), 30), ((512, 128, 128), 300)])
def test_fista_codes_match_oracle(shape, iters):
    """T3: FISTA codes within 1e-4 absolute of the fp32 reference algorithm."""
    import sparse_coding__b200 as sc
    B, d, N = shape
    torch.manual_seed(6)
    D = R.row_normalise(torch.randn(N, d))[0]
    x = make_sparse_data(B, d, 2 * d, 6, seed=7)
    a0 = torch.relu(x @ D.T)
    l1 = 1e-2
    a_ref, res_ref = R.fista(x.double(), D.double(), l1, a0.double(), iters)
    eta_ref = float(R.fista_eta(D.double()))
    a, res = sc.FunctionalFista.fista(x.cuda(), D.cuda(), l1, a0.cuda(), iters, "cuda")
    from sparse_coding__b200.autoencoders.fista import fista_native
    _, _, eta = fista_native(x.cuda(), D.cuda(), l1, a0.cuda(), 0)
    # tensor-core accumulation truncates: ~3e-8 relative per instruction on same-sign sums (K/8 of them)
    assert abs(float(eta) - eta_ref) <= 5e-5 * eta_ref
    assert float((a.cpu().double() - a_ref).abs().max()) < 1e-4
    assert float((res.cpu().double() - res_ref).abs().max()) < 1e-4
    # explicit eta path
    a2, _ = sc.FunctionalFista.fista(x.cuda(), D.cuda(), torch.tensor(l1), a0.cuda(), iters, "cuda", eta=eta_ref)
    assert float((a2.cpu().double() - a_ref).abs().max()) < 1e-4


def test_single_member_loss_matches_oracle():
    import sparse_coding__b200 as sc
    d, N, B = 64, 256, 300
    torch.manual_seed(8)
    params, buffers = sc.FunctionalSAE.init(d, N, 3e-3, bias_decay=0.01)
    params["encoder_bias"] += 0.01
    x = make_sparse_data(B, d, 2 * d, 6, seed=9)
    ls, _ = R.sae_forward(params["encoder"], params["encoder_bias"], params["decoder"], buffers["l1_alpha"],
                          buffers["bias_decay"], x)
    pg = {k: v.cuda() for k, v in params.items()}
    bg = {k: v.cuda() for k, v in buffers.items()}
    total, (ld, aux) = sc.FunctionalSAE.loss(pg, bg, x.cuda())
    for k in ("loss", "l_reconstruction", "l_l1", "l_bias_decay"):
        assert abs(float(ld[k]) - float(ls[k])) <= 1e-3 * abs(float(ls[k])) + 1e-9, k
    tp, tb = sc.FunctionalTiedSAE.init(d, N, 3e-3)
    ls, _ = R.tied_forward(tp["encoder"], tp["encoder_bias"], tb["l1_alpha"], tb["bias_decay"], x)
    total, (ld, aux) = sc.FunctionalTiedSAE.loss({k: v.cuda() for k, v in tp.items()}, {k: v.cuda() for k, v in tb.items()}, x.cuda())
    assert "l_bias_decay" not in ld
    assert abs(float(ld["loss"]) - float(ls["loss"])) <= 1e-3 * abs(float(ls["loss"]))


def test_fista_300_iterations_at_pythia_width():
    """T3 at the BASELINE.json configs[3] dictionary shape (d=512, N=4096, 300 iterations), on as many
    rows as the float64 CPU oracle finishes in seconds; rows are independent, so this is the full-size
    check per row."""
    import sparse_coding__b200 as sc
    B, d, N, iters = 320, 512, 4096, 300
    torch.manual_seed(10)
    D = R.row_normalise(torch.randn(N, d))[0]
    x = make_sparse_data(B, d, 2 * d, 10, seed=11)
    a0 = torch.relu(x @ D.T - 0.05)
    l1 = 1e-3
    a_ref, res_ref = R.fista(x.double(), D.double(), l1, a0.double(), iters)
    a, res = sc.FunctionalFista.fista(x.cuda(), D.cuda(), l1, a0.cuda(), iters, "cuda")
    assert float((a.cpu().double() - a_ref).abs().max()) < 1e-4
    assert float((res.cpu().double() - res_ref).abs().max()) < 1e-4
    # size-independent property: residual is consistent with the codes
    assert float((x.cuda() - a @ D.cuda() - res).abs().max()) < 1e-4


def test_max_cosine_similarity_family():
    """standard_metrics.py:268-300 against plain fp32 torch, including a row whose similarities are all negative
    (the ReLU of the encode kernel must not clip it)."""
    import sparse_coding__b200 as sc
    from sparse_coding__b200 import standard_metrics as sm
    torch.manual_seed(3)
    d, n1, n2 = 128, 384, 512
    W1, W2 = torch.randn(n1, d, device="cuda"), torch.randn(n2, d, device="cuda")
    W2[:, 0] = W2[:, 0].abs() + 3.0          # every atom of the second dictionary leans towards +e0 ...
    W1[5] = 0.0
    W1[5, 0] = -1.0                           # ... and this atom of the first one is -e0: all its cosines are negative
    ld1 = sc.UntiedSAE(torch.randn(n1, d, device="cuda"), W1, torch.zeros(n1, device="cuda"))
    ld2 = sc.TiedSAE(W2, torch.zeros(n2, device="cuda"), norm_encoder=True)
    D1 = W1 / W1.norm(dim=1, keepdim=True)
    D2 = W2 / W2.norm(dim=1, keepdim=True)
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        want = (D1 @ D2.T).max(dim=-1).values
        want_r = (D2 @ D1.T).max(dim=-1).values
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    got = sm.mcs_duplicates(ld2, ld1)
    assert got.shape == (n1,) and float(want[5]) < 0
    assert float((got - want).abs().max()) < 5e-4
    assert abs(float(sm.mmcs(ld2, ld1)) - float(want.mean())) < 1e-4      # mmcs(a, b) = mcs_duplicates(ground=a, model=b).mean()
    assert abs(float(sm.mmcs(ld1, ld2)) - float(want_r.mean())) < 1e-4
    assert float((sm.mcs_to_fixed(ld1, D2) - want).abs().max()) < 5e-4
    assert abs(float(sm.mmcs_to_fixed(ld1, D2)) - float(want.mean())) < 1e-4
    assert float((sm.representedness(D2, ld1) - want_r).abs().max()) < 5e-4
    m = sm.mmcs_from_list([ld1, ld2])
    assert m.shape == (2, 2) and float(m[0, 0]) == 1.0 and abs(float(m[1, 0]) - float(want.mean())) < 1e-4 and m[0, 1] == m[1, 0]
