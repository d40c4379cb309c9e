"""Resident-chunk input pipeline on the GPU, through the C ABI: the gathered batches are
bit-exact against the reference's host-side `chunk.to(float32)[idxs]` (minus means), and a train loop fed by a
ResidentChunk follows the one fed the reference way step for step."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("n,d,B", [(1000, 512, 256), (37, 8, 5), (5000, 776, 2048), (64, 1024, 64)])
def test_gather_is_bit_exact(dtype, n, d, B):
    import sparse_coding__b200 as sc
    g = torch.Generator().manual_seed(n + d)
    chunk = (3 * torch.randn(n, d, generator=g)).to(dtype)
    means = chunk.to(torch.float32).mean(dim=0)
    idx = torch.randint(-n, n, (B,), generator=g)
    for m in (None, means):
        rc = sc.ResidentChunk(chunk, device="cuda", means=m)
        want = chunk.to(torch.float32)[idx]          # big_sweep.py:401,170
        if m is not None:
            want = want - means                      # big_sweep.py:407
        for form in (idx.tolist(), idx.numpy(), idx, idx.cuda(), idx.cuda().to(torch.int32)):
            got = rc[form]
            assert got.dtype == torch.float32 and got.is_cuda and got.shape == (B, d)
            assert torch.equal(got.cpu(), want)
        rc.check()
    assert rc.shape == (n, d) and len(rc) == n and rc.to("cuda") is rc and rc.to(dtype=torch.float32) is rc
    assert rc[[]].shape == (0, d)
    assert torch.equal(rc[slice(0, 4)].cpu(), chunk.to(torch.float32)[:4] - means)


def test_out_of_range_rows_raise_like_the_reference():
    import sparse_coding__b200 as sc
    chunk = torch.randn(100, 16).half()
    rc = sc.ResidentChunk(chunk)
    with pytest.raises(IndexError):
        chunk[[3, 100]]
    with pytest.raises(IndexError):
        rc[[3, 100]]
    with pytest.raises(IndexError):
        rc[torch.tensor([-101])]
    out = rc[torch.tensor([3, 100, 5], device="cuda")]      # device indices: zero row + flag, raised at check()
    assert torch.equal(out[0].cpu(), chunk[3].float()) and torch.equal(out[2].cpu(), chunk[5].float())
    assert float(out[1].abs().max()) == 0.0
    with pytest.raises(IndexError):
        rc.check()
    rc.check()  # the flag is cleared once raised


def test_column_means_match_host():
    import sparse_coding__b200 as sc
    g = torch.Generator().manual_seed(3)
    for dtype, n, d in ((torch.float16, 20000, 512), (torch.float32, 3333, 64), (torch.float16, 7, 8)):
        chunk = (torch.randn(n, d, generator=g) + 0.5).to(dtype)
        rc = sc.ResidentChunk(chunk, means=True)
        want = chunk.to(torch.float64).mean(dim=0)
        assert float((rc.means.cpu().double() - want).abs().max()) < 2e-6
        # deterministic
        assert torch.equal(rc.column_means(), rc.means)


def test_train_loop_on_resident_chunk_equals_reference_style_loop(tmp_path):
    """Same index batches through (a) the reference-style host path `dataset[idxs].to(device)` with an fp32 CPU
    chunk and (b) the fp16 chunk resident in HBM: identical parameters afterwards (the inputs are bit-equal)."""
    import sparse_coding__b200 as sc

    class Cfg:
        use_wandb = False

    class Counter:
        value = 0

    d, N, n, B = 64, 128, 3000, 256
    g = torch.Generator().manual_seed(1)
    chunk_h = torch.randn(n, d, generator=g).half()
    means = chunk_h.float().mean(dim=0)
    batches = [torch.randperm(n, generator=g)[:B].tolist() for _ in range(12)] + [list(range(40))]
    finals = []
    for resident in (False, True):
        torch.manual_seed(0)
        models = [sc.FunctionalSAE.init(d, N, l1) for l1 in (1e-3, 3e-3)]
        ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
        ds = sc.ResidentChunk(chunk_h, means=means) if resident else (chunk_h.to(torch.float32) - means)
        cnt = Counter()
        sc.ensemble_train_loop(ens, Cfg(), {"device": "cuda", "batch_size": B}, "e", batches, ds, cnt, torch.float32,
                               fista_dictionary_update=False)
        assert cnt.value == len(batches) - 1
        finals.append({k: v.clone() for k, v in ens.params.items()})
    for k in finals[0]:
        assert torch.equal(finals[0][k], finals[1][k]), k


def test_train_on_chunks_writes_reference_layout(tmp_path):
    import sparse_coding__b200 as sc

    class Cfg(dict):
        __getattr__ = dict.get
        __setattr__ = dict.__setitem__

    d, N = 64, 128
    data_dir, out_dir = os.path.join(tmp_path, "acts"), os.path.join(tmp_path, "out")
    os.makedirs(data_dir)
    g = torch.Generator().manual_seed(2)
    for i in range(2):
        torch.save((torch.randn(1500, d, generator=g) + 1.0).half(), os.path.join(data_dir, f"{i}.pt"))
    cfg = Cfg(dataset_folder=data_dir, output_folder=out_dir, use_wandb=False, center_activations=True, n_repetitions=None,
              batch_size=500)
    torch.manual_seed(0)
    models = [sc.FunctionalTiedSAE.init(d, N, l1) for l1 in (1e-3, 1e-2)]
    ens = sc.FunctionalEnsemble(models, sc.FunctionalTiedSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
    p0 = ens.params["encoder"].clone()
    lds = sc.train_on_chunks([(ens, {"device": "cuda", "batch_size": 500, "dict_size": N}, "tied")], cfg,
                             ["dict_size"], ["l1_alpha"])
    assert not torch.equal(p0, ens.params["encoder"])
    assert len(lds) == 2 and lds[0][1]["dict_size"] == N and lds[1][1]["l1_alpha"] == pytest.approx(1e-2)
    assert os.path.isfile(os.path.join(out_dir, "means.pt")) and os.path.isfile(os.path.join(out_dir, "_1", "config.yaml"))
    back = sc.load_learned_dicts(os.path.join(out_dir, "_1", "learned_dicts.pt"))
    assert torch.equal(back[0][0].encoder, ens.params["encoder"][0].cpu())
    np.random.seed(0)
    first = int(np.random.permutation(2)[0])
    want = torch.load(os.path.join(data_dir, f"{first}.pt")).double().mean(dim=0)
    assert float((torch.load(os.path.join(out_dir, "means.pt")).double() - want).abs().max()) < 2e-6
