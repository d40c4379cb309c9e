"""sc_encode_noalloc: the encode entry point for callers that cannot allocate device memory.

Some hosts of this library - an online feature monitor running beside a serving model, a bare C
harness with no torch allocator - have given every byte of device memory to their own pool, or
cannot accept the device-wide stall of a cudaMalloc inside their loop.  For them the library carries
a fixed 16 MiB scratch array in its own image and `sc_encode_noalloc` uses that instead of a caller
workspace.  This file checks it against the regular `sc_encode`, twice in a row with different
shapes (the array is shared and keeps the previous call's bytes), and checks the size limit.

Collected last on purpose (file name): with `pytest -x` every other test has run before this one.
"""
import ctypes as C

import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu

SCRATCH_FLOATS = 4 * 1024 * 1024  # SC_NOALLOC_SCRATCH_FLOATS in include/sc_b200.h


def _bind():
    from sparse_coding__b200 import _native as nat
    L = nat.lib()
    fn = L.sc_encode_noalloc
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    fn.restype = C.c_int
    return nat, L, fn


def _run_noalloc(x, W, b):
    nat, L, fn = _bind()
    B, d = x.shape
    N = W.shape[0]
    out = torch.full((B, N), float("nan"), device=x.device)
    rc = fn(nat.ptr(x), nat.ptr(W), nat.ptr(b), nat.ptr(out), B, d, N, nat.stream_ptr(x.device))
    assert rc == 0, L.sc_last_error().decode()
    return out


def test_noalloc_encode_matches_regular_encode_twice(monkeypatch):
    from sparse_coding__b200.autoencoders.learned_dict import _encode_native
    # sc_encode_noalloc runs the TF32-operand kernels; compare with sc_encode in the same operand format
    monkeypatch.setenv("SC_B200_OPERANDS", "tf32")
    torch.manual_seed(0)
    # two calls in a row, second one smaller: stale bytes of the first call stay behind it in the array
    for (B, d, N) in [(700, 256, 1024), (130, 64, 200)]:
        x = torch.randn(B, d, device="cuda")
        W = torch.randn(N, d, device="cuda") / d ** 0.5
        b = 0.1 * torch.randn(N, device="cuda")
        want, _ = _encode_native(x, W, b)            # existing entry point (sc_encode, caller workspace)
        got = _run_noalloc(x, W, b)                  # new entry point (built-in scratch)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert rel_err(got, want) < 2e-3             # tolerance of the other encode tests
        assert torch.equal(got, want)                # same kernels, same operands: bit identical
        ref = torch.relu(x.double() @ W.double().T + b.double())
        assert rel_err(got, ref) < 2e-3


def test_noalloc_encode_rejects_oversize_and_bad_shapes():
    nat, L, fn = _bind()
    d = 1024
    B = N = SCRATCH_FLOATS // d // 2 + 256           # B*d + N*d > scratch
    x = torch.empty(16, device="cuda")               # never dereferenced: the call must fail before any launch
    rc = fn(nat.ptr(x), nat.ptr(x), None, nat.ptr(x), B, d, N, nat.stream_ptr(x.device))
    assert rc == -4
    msg = L.sc_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "sc_encode" in msg
    rc = fn(nat.ptr(x), nat.ptr(x), None, nat.ptr(x), 8, 6, 8, nat.stream_ptr(x.device))
    assert rc == -1 and "multiples of 4" in L.sc_last_error().decode()
    torch.cuda.synchronize()
