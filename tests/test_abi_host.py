"""CPU-side checks: the C-ABI library loads without a GPU and exports every symbol the header
declares; the host mirror keeps the reference's API shape; the product path refuses CPU tensors."""
import ctypes
import os
import pickle
import re
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "sc_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sc_[a-z0-9_]+)\s*\(", text)))


def test_library_loads_and_exports_every_declared_symbol():
    from sparse_coding__b200 import _native as nat
    lib = nat.lib()  # loads without a GPU: the driver is only touched by the first CUDA call
    declared = _declared_symbols()
    assert len(declared) >= 17
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in include/sc_b200.h but not exported"
    assert set(nat.EXPORTS) <= set(declared)
    assert lib.sc_version() >= 100
    assert lib.sc_sae_step_workspace_bytes(8, 2048, 512, 4096, 0, 0) > 8 * 2048 * 4096 * 4
    assert lib.sc_fista_workspace_bytes(0, 1, 1) == 0


def test_argument_errors_are_reported_without_a_gpu():
    from sparse_coding__b200 import _native as nat
    lib = nat.lib()
    assert lib.sc_sae_step(None) == -1
    assert b"NULL" in lib.sc_last_error()
    args = nat.ScStepArgs(M=1, B=8, d=6, N=8)
    assert lib.sc_sae_step(ctypes.byref(args)) == -1
    assert b"multiples of 4" in lib.sc_last_error()


def test_struct_layouts_match_the_header_field_order():
    from sparse_coding__b200 import _native as nat
    text = open(os.path.join(ROOT, "include", "sc_b200.h")).read()
    for cname, cls in (("ScStepArgs", nat.ScStepArgs), ("ScEncodeArgs", nat.ScEncodeArgs),
                       ("ScDecodeArgs", nat.ScDecodeArgs), ("ScFistaArgs", nat.ScFistaArgs), ("ScAdam", nat.ScAdam)):
        head = text[:text.index("} %s;" % cname)]
        body = head[head.rindex("typedef struct {") + len("typedef struct {"):]
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            decl = re.sub(r"^(const\s+)?(float|int32_t|size_t|void|ScAdam|char)\s*\**", "", decl)
            names += [re.sub(r"\[\d+\]$", "", n.strip().lstrip("*")) for n in decl.split(",")]
        assert names == [f[0] for f in cls._fields_], cname


def test_no_cpu_path():
    import sparse_coding__b200 as sc
    models = [sc.FunctionalSAE.init(16, 32, 1e-3)]
    with pytest.raises(NotImplementedError):
        sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cpu")
    ld = sc.UntiedSAE(torch.zeros(8, 4), torch.zeros(8, 4), torch.zeros(8))
    with pytest.raises(NotImplementedError):
        ld.encode(torch.zeros(2, 4))
    with pytest.raises(NotImplementedError):
        sc.FunctionalFista.fista(torch.zeros(2, 4), torch.zeros(8, 4), 0.1, torch.zeros(2, 8), 3, "cpu")
    with pytest.raises(NotImplementedError):
        sc.torchopt_compat.spec_from(type("sgd", (), {"__name__": "sgd"}), {})


def test_init_consumes_rng_like_the_reference_restatement():
    import sparse_coding__b200 as sc
    from oracle import restated as R
    torch.manual_seed(3)
    p, b = sc.FunctionalSAE.init(24, 40, 1e-2, bias_decay=0.1)
    torch.manual_seed(3)
    bound = float(np.sqrt(6.0 / (24 + 40)))
    enc = torch.empty(40, 24).uniform_(-bound, bound)
    dec = torch.empty(40, 24).uniform_(-bound, bound)
    assert torch.equal(p["encoder"], enc) and torch.equal(p["decoder"], dec) and not p["encoder_bias"].any()
    assert float(b["l1_alpha"]) == pytest.approx(1e-2) and float(b["bias_decay"]) == pytest.approx(0.1)
    tp, tb = sc.FunctionalTiedSAE.init(24, 40, 1e-2)
    assert set(tp) == {"encoder", "encoder_bias"}
    assert set(tb) == {"center_rot", "center_trans", "center_scale", "l1_alpha", "bias_decay"}
    fp, fb = sc.FunctionalFista.init(24, 40, 1e-2)
    assert "hessian_diag" in fb and fb["hessian_diag"].shape == (40,)


def test_reference_import_names_and_pickles():
    import sparse_coding__b200 as sc
    assert sc.install_reference_aliases(force=True)
    from autoencoders.ensemble import FunctionalEnsemble
    from autoencoders.learned_dict import TiedSAE, UntiedSAE
    from autoencoders.fista import FunctionalFista, Fista
    assert FunctionalEnsemble is sc.FunctionalEnsemble and UntiedSAE is sc.UntiedSAE
    ld = sc.TiedSAE(torch.randn(8, 4), torch.zeros(8), norm_encoder=True)
    blob = pickle.dumps([(ld, {"dict_size": 8, "l1_alpha": 1e-3})])
    assert b"autoencoders.learned_dict" in blob and b"sparse_coding__b200" not in blob
    back = pickle.loads(blob)[0][0]
    assert isinstance(back, sc.TiedSAE) and torch.equal(back.encoder, ld.encoder)
    f = Fista(torch.randn(8, 4), torch.zeros(8))
    assert b"autoencoders.fista" in pickle.dumps(f)


def test_separate_tensors_and_unstack_errors_match_reference():
    from sparse_coding__b200.big_sweep import separate_tensors
    with pytest.raises(ValueError):
        separate_tensors({"a": torch.zeros(3, 2)}, 2)
    out = separate_tensors({"a": torch.arange(6.).reshape(3, 2)}, 3)
    assert len(out) == 3 and torch.equal(out[1]["a"], torch.tensor([2., 3.]))


def test_integration_stub_struct_matches_the_binding():
    """The ctypes stub printed in INTEGRATION.md (what a reference maintainer would paste) must describe the same
    ScStepArgs as the package's own binding: same field names in the same order, same size."""
    from sparse_coding__b200 import _native as nat
    src = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    code = src[src.index("class ScAdam(C.Structure):"):src.index("L.sc_sae_step_workspace_bytes.restype")]
    ns = {"C": ctypes}
    exec(code, ns)
    stub = ns["ScStepArgs"]
    assert [f[0] for f in stub._fields_] == [f[0] for f in nat.ScStepArgs._fields_]
    assert ctypes.sizeof(stub) == ctypes.sizeof(nat.ScStepArgs)
