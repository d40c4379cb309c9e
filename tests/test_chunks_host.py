"""CPU tests of the caller-side pieces: checkpoint writer compatibility with the reference's
readers, sampler semantics, and the loud refusal of a CPU device (no fallback)."""
import os
import warnings

import pytest
import torch

from oracle import ref_loader

warnings.filterwarnings("ignore")
needs_ref = pytest.mark.skipif(not (ref_loader.available() and ref_loader.full_tree()), reason="/root/reference not present on this machine")


def test_device_sampler_covers_every_row_once():
    from sparse_coding__b200.chunks import DeviceBatchSampler
    s = DeviceBatchSampler(1003, 100, drop_last=False, device="cpu", seed=5)
    batches = list(s)
    assert len(batches) == len(s) == 11 and batches[-1].numel() == 3
    assert all(b.dtype == torch.int64 for b in batches)
    assert torch.equal(torch.cat(batches).sort().values, torch.arange(1003))
    s2 = DeviceBatchSampler(1003, 100, drop_last=True, device="cpu", seed=5)
    assert len(list(s2)) == len(s2) == 10
    # a second pass draws a new permutation, the same seed reproduces the first
    assert not torch.equal(torch.cat(list(s)), torch.cat(batches))
    assert torch.equal(torch.cat(list(DeviceBatchSampler(1003, 100, device="cpu", seed=5))), torch.cat(batches))
    with pytest.raises(ValueError):
        DeviceBatchSampler(10, 0, device="cpu")


def test_resident_chunk_refuses_cpu_and_bad_input():
    from sparse_coding__b200.chunks import ResidentChunk
    with pytest.raises(NotImplementedError):
        ResidentChunk(torch.zeros(4, 8, dtype=torch.float16), device="cpu")
    with pytest.raises(TypeError):
        ResidentChunk(torch.zeros(4, 8, dtype=torch.float64), device="cuda")
    with pytest.raises(TypeError):
        ResidentChunk(torch.zeros(8, dtype=torch.float16), device="cuda")
    with pytest.raises(ValueError):
        ResidentChunk(torch.zeros(4, 12, dtype=torch.float16), device="cuda")


def test_checkpoint_writer_round_trip(tmp_path):
    import sparse_coding__b200 as sc
    lds = [(sc.UntiedSAE(torch.randn(8, 4), torch.randn(8, 4), torch.zeros(8)), {"dict_size": 8, "l1_alpha": 1e-3}),
           (sc.TiedSAE(torch.randn(8, 4), torch.zeros(8), norm_encoder=True), {"dict_size": 8, "l1_alpha": 1e-2})]
    path = os.path.join(tmp_path, "_7", "learned_dicts.pt")
    sc.save_learned_dicts(lds, path)
    back = sc.load_learned_dicts(path)
    assert [h for _, h in back] == [h for _, h in lds]
    assert isinstance(back[0][0], sc.UntiedSAE) and torch.equal(back[0][0].decoder, lds[0][0].decoder)
    assert isinstance(back[1][0], sc.TiedSAE) and back[1][0].norm_encoder
    with pytest.raises(TypeError):
        sc.save_learned_dicts([(lds[0][0], "not a dict")], path)


@needs_ref
def test_reference_reads_our_checkpoints_and_we_read_theirs(tmp_path):
    """A `learned_dicts.pt` written here unpickles into the REFERENCE's classes (its plotting / interpretation
    scripts do exactly this), with identical tensors and methods that run; and the reference's shipped
    output_basic_test pickles load into this engine's classes."""
    import sparse_coding__b200 as sc
    torch.manual_seed(0)
    We, Wd, b = torch.randn(16, 8), torch.randn(16, 8), 0.1 * torch.randn(16)
    lds = [(sc.UntiedSAE(We, Wd, b), {"dict_size": 16, "l1_alpha": 1e-3}),
           (sc.TiedSAE(We, b, norm_encoder=True), {"dict_size": 16, "l1_alpha": 1e-2})]
    path = os.path.join(tmp_path, "learned_dicts.pt")
    sc.save_learned_dicts(lds, path)
    ns = ref_loader.load()
    with ref_loader.reference_modules():
        theirs = torch.load(path, weights_only=False, map_location="cpu")
    assert type(theirs[0][0]).__name__ == "UntiedSAE" and type(theirs[0][0]) is not sc.UntiedSAE
    assert type(theirs[0][0]) is ns.learned_dict.UntiedSAE and type(theirs[1][0]) is ns.learned_dict.TiedSAE
    x = torch.randn(5, 8)
    c = theirs[0][0].encode(x)
    assert torch.allclose(c, torch.relu(x @ We.T + b), atol=1e-6)
    assert torch.allclose(theirs[1][0].get_learned_dict(), We / We.norm(dim=1, keepdim=True), atol=1e-6)
    assert theirs[0][1] == {"dict_size": 16, "l1_alpha": 1e-3}
    sc.install_reference_aliases(force=True)
    root = os.path.join(ref_loader.REFERENCE_ROOT, "output_basic_test", "normal", "learned_dicts_epoch_0.pt")
    ours = sc.load_learned_dicts(root)
    assert isinstance(ours[0][0], sc.LearnedDict) and ours[0][1]["dict_size"] == 512
