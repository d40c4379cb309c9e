import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA sm_100a device (run with -m gpu on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def make_sparse_data(n, d, n_feats, k_active, seed):
    """Synthetic sparse-feature activations (in the spirit of sc_datasets/random_dataset.py)."""
    g = torch.Generator().manual_seed(seed)
    feats = torch.randn(n_feats, d, generator=g)
    feats /= feats.norm(dim=1, keepdim=True)
    codes = torch.zeros(n, n_feats)
    idx = torch.rand(n, n_feats, generator=g).argsort(1)[:, :k_active]
    codes.scatter_(1, idx, torch.rand(n, k_active, generator=g))
    return codes @ feats
