"""Optimizer factories with torchopt's call shape, for callers that do not have torchopt.

The reference builds its ensembles as `FunctionalEnsemble(models, sig, torchopt.adam, {"lr": ...})`
(ensemble.py:68-97; every recipe in big_sweep_experiments.py, basic_l1_sweep.py:71-76).  The engine
only needs to know WHICH optimizer and its hyper-parameters - Adam itself runs fused in the
weight-gradient GEMM epilogues - so `optimizer_func` is treated as a tag: anything whose
`__name__` is "adam" (torchopt.adam included) selects the fused Adam.
"""
from typing import Any, NamedTuple


class ScaleByAdamState(NamedTuple):
    """Same field names as torchopt's state so `state_dict()["optim_states"]` reads alike."""
    mu: Any
    nu: Any
    count: Any


class EmptyState(NamedTuple):
    pass


class AdamSpec(NamedTuple):
    lr: float
    b1: float
    b2: float
    eps: float
    eps_root: float


def adam(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, *, eps_root=0.0,
         moment_requires_grad=False, maximize=False, use_accelerated_op=False):
    if weight_decay != 0.0 or maximize:
        raise NotImplementedError("fused Adam supports weight_decay=0, maximize=False (what the reference uses)")
    return AdamSpec(float(lr), float(betas[0]), float(betas[1]), float(eps), float(eps_root))


def spec_from(optimizer_func, optimizer_kwargs) -> AdamSpec:
    name = getattr(optimizer_func, "__name__", str(optimizer_func))
    if name != "adam":
        raise NotImplementedError(
            f"optimizer {name!r}: the B200 engine fuses Adam only (the reference's recipes all pass torchopt.adam)")
    return adam(**optimizer_kwargs)
