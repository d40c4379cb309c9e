"""Load the UNMODIFIED reference hot-path modules on CPU.  TEST / BASELINE INFRASTRUCTURE.

Two sources: /root/reference (the build container only: `-m "not gpu"` tests validate `oracle/restated.py`
against the real reference code, `oracle/make_golden.py` generates the fixtures under tests/golden/), and
oracle/_ref (byte-for-byte copies of the four hot-path modules made by oracle/build_ref.py; they travel to the GPU
box, where `bench.py --impl reference` and its `cpu_baseline` leg time the reference's own
FunctionalEnsemble.step_batch on the host cores).  Nothing under sparse_coding__b200/ imports this module.

The reference imports three packages that are not installed (no network):
optree, torchopt, torchtyping -> `oracle/shims/`.  `big_sweep.py` and
`standard_metrics.py` import matplotlib / transformer_lens / wandb / baukit at
module level, so single `def`s are AST-extracted from them and executed in a
namespace that only holds what those functions touch.
"""
import ast
import importlib
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_SHIMS = os.path.join(_HERE, "shims")
_FULL = os.environ.get("SC_REFERENCE_ROOT", "/root/reference")
_VENDORED = os.path.join(_HERE, "_ref")
# the full tree when it is there (build container), else the four vendored modules
REFERENCE_ROOT = _FULL if os.path.isfile(os.path.join(_FULL, "autoencoders", "ensemble.py")) else _VENDORED


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "autoencoders", "ensemble.py"))


def full_tree() -> bool:
    """big_sweep.py / standard_metrics.py are there as well (only in the build container)."""
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "standard_metrics.py"))


_loaded = {}


def load():
    """Returns a namespace object with the reference modules / functions."""
    if "ns" in _loaded:
        return _loaded["ns"]
    if not available():
        raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
    # The engine package registers itself under the reference's import names (drop-in aliases).
    # Import the real reference under those names, keep the module objects, then put back whatever
    # was there so the rest of the process is unaffected.
    saved = {k: sys.modules.pop(k) for k in list(sys.modules)
             if k == "autoencoders" or k.startswith("autoencoders.") or k in ("torchopt", "optree", "torchtyping")}
    saved_path = list(sys.path)
    for p in (_SHIMS, REFERENCE_ROOT):
        sys.path.insert(0, p)

    class NS:
        pass

    ns = NS()
    ns.torchopt = importlib.import_module("torchopt")
    ns.optree = importlib.import_module("optree")
    ns.ensemble = importlib.import_module("autoencoders.ensemble")
    ns.sae_ensemble = importlib.import_module("autoencoders.sae_ensemble")
    ns.learned_dict = importlib.import_module("autoencoders.learned_dict")
    ns.fista = importlib.import_module("autoencoders.fista")
    assert ns.ensemble.__file__.startswith(REFERENCE_ROOT), ns.ensemble.__file__

    ns.FunctionalEnsemble = ns.ensemble.FunctionalEnsemble
    ns.FunctionalSAE = ns.sae_ensemble.FunctionalSAE
    ns.FunctionalTiedSAE = ns.sae_ensemble.FunctionalTiedSAE
    ns.FunctionalFista = ns.fista.FunctionalFista
    ns.UntiedSAE = ns.learned_dict.UntiedSAE
    ns.TiedSAE = ns.learned_dict.TiedSAE

    import numpy as np
    import torch

    def extract(path, names, extra_globals):
        with open(os.path.join(REFERENCE_ROOT, path)) as f:
            tree = ast.parse(f.read())
        body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
        found = {n.name for n in body}
        missing = set(names) - found
        if missing:
            raise RuntimeError(f"{path}: functions not found: {sorted(missing)}")
        mod = ast.Module(body=body, type_ignores=[])
        g = {"torch": torch, "np": np, "__builtins__": __builtins__}
        g.update(extra_globals)
        exec(compile(mod, os.path.join(REFERENCE_ROOT, path), "exec"), g)
        return {n: g[n] for n in names}

    if not full_tree():  # oracle/_ref: the four modules of the train step only
        ns._modules = {k: v for k, v in sys.modules.items()
                       if k == "autoencoders" or k.startswith("autoencoders.") or k in ("torchopt", "optree", "torchtyping")}
        for k in ns._modules:
            sys.modules.pop(k, None)
        sys.modules.update(saved)
        sys.path[:] = saved_path
        _loaded["ns"] = ns
        return ns

    class _Any:
        def __class_getitem__(cls, item):
            return cls

    def make_hyperparam_name(hv):
        return "_".join(f"{k}_{v}" for k, v in hv.items())

    bs = extract(
        "big_sweep.py",
        ["ensemble_train_loop", "separate_tensors", "unstacked_to_learned_dicts"],
        {"FunctionalFista": ns.FunctionalFista, "make_hyperparam_name": make_hyperparam_name},
    )
    # the extracted functions call each other through their module globals
    for fn in bs.values():
        fn.__globals__.update(bs)
    ns.ensemble_train_loop = bs["ensemble_train_loop"]
    ns.separate_tensors = bs["separate_tensors"]
    ns.unstacked_to_learned_dicts = bs["unstacked_to_learned_dicts"]

    sm_names = [
        "mean_nonzero_activations", "fraction_variance_unexplained",
        "fraction_variance_unexplained_top_activating", "r_squared",
        "calc_feature_n_active", "batched_calc_feature_n_ever_active",
        "calc_feature_mean", "calc_feature_variance", "calc_feature_skew",
        "calc_feature_kurtosis", "calc_moments_streaming",
    ]
    sm = extract(
        "standard_metrics.py", sm_names,
        {"LearnedDict": ns.learned_dict.LearnedDict, "TensorType": _Any},
    )
    for fn in sm.values():
        fn.__globals__.update(sm)
    for k, v in sm.items():
        setattr(ns, k, v)

    ns._modules = {k: v for k, v in sys.modules.items()
                   if k == "autoencoders" or k.startswith("autoencoders.") or k in ("torchopt", "optree", "torchtyping")}
    for k in ns._modules:
        sys.modules.pop(k, None)
    sys.modules.update(saved)
    sys.path[:] = saved_path
    _loaded["ns"] = ns
    return ns


class reference_modules:
    """Context manager: the reference's modules visible under their own import names (needed to
    unpickle the reference's own output_basic_test/*.pt files)."""

    def __enter__(self):
        ns = load()
        self.saved = {k: sys.modules.get(k) for k in ns._modules}
        sys.modules.update(ns._modules)
        return ns

    def __exit__(self, *exc):
        for k, v in self.saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
