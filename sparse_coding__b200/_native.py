"""ctypes binding of libsc_b200.so (the C ABI declared in include/sc_b200.h).

The product path has no fallback: if the library is missing or the device is not a
B200-class GPU, calls raise.  `build()` compiles it in-tree with nvcc for sm_100a.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsc_b200.so")
LIB_PATH = os.environ.get("SC_B200_LIB", LIB_PATH)  # A/B runs against another build of the same ABI
CSRC = os.path.join(_HERE, "csrc")

_lib = None
_lock = threading.Lock()

EXPORTS = [
    "sc_version", "sc_last_error", "sc_sae_step_workspace_bytes", "sc_sae_step",
    "sc_encode_workspace_bytes", "sc_encode", "sc_decode_workspace_bytes", "sc_decode",
    "sc_fista_workspace_bytes", "sc_fista", "sc_gemm_tf32", "sc_adam_apply", "sc_row_normalise", "sc_column_variance",
    "sc_profile_begin", "sc_profile_end", "sc_encode_noalloc", "sc_sae_step_f16_supported", "sc_codes_to_float",
    "sc_gemm_f32_workspace_bytes", "sc_gemm_f32",
    "sc_gather_rows", "sc_column_mean_workspace_bytes", "sc_column_mean", "sc_infer_f16_supported", "sc_dp_adam_broadcast",
]


class ScAdam(C.Structure):
    _fields_ = [("lr", C.c_float), ("b1", C.c_float), ("b2", C.c_float), ("eps", C.c_float),
                ("eps_root", C.c_float), ("bc1", C.c_float), ("bc2", C.c_float)]


class ScStepArgs(C.Structure):
    _fields_ = [
        ("M", C.c_int32), ("B", C.c_int32), ("d", C.c_int32), ("N", C.c_int32),
        ("tied", C.c_int32), ("x_per_member", C.c_int32),
        ("x", C.c_void_p),
        ("enc", C.c_void_p), ("enc_mu", C.c_void_p), ("enc_nu", C.c_void_p),
        ("bias", C.c_void_p), ("bias_mu", C.c_void_p), ("bias_nu", C.c_void_p),
        ("dec", C.c_void_p), ("dec_mu", C.c_void_p), ("dec_nu", C.c_void_p),
        ("l1_alpha", C.c_void_p), ("bias_decay", C.c_void_p),
        ("adam", ScAdam),
        ("out_losses", C.c_void_p), ("out_c", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t),
        ("stream", C.c_void_p),
        ("grad_enc", C.c_void_p), ("grad_bias", C.c_void_p), ("grad_dec", C.c_void_p),
        ("B_total", C.c_int32),
        ("operand_format", C.c_int32), ("out_c_half", C.c_void_p), ("dict_size", C.c_void_p),
        ("rs_world", C.c_int32), ("rs_rows", C.c_int32),
        ("rs_grad_enc", C.c_void_p * 8), ("rs_grad_dec", C.c_void_p * 8),
        ("rs_grad_bias", C.c_void_p * 8), ("rs_losses", C.c_void_p * 8),
        ("operands_valid", C.c_int32), ("out_code_scale", C.c_void_p),
        ("range_flag", C.c_void_p), ("step_tag", C.c_int32),
    ]


class ScEncodeArgs(C.Structure):
    _fields_ = [
        ("G", C.c_int32), ("B", C.c_int32), ("d", C.c_int32), ("N", C.c_int32),
        ("x_per_group", C.c_int32),
        ("x", C.c_void_p), ("W", C.c_void_p), ("bias", C.c_void_p),
        ("out_c", C.c_void_p), ("out_nonzero", C.c_void_p), ("out_moments", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("stream", C.c_void_p),
        ("operand_format", C.c_int32), ("out_c_half", C.c_void_p),
    ]


class ScDecodeArgs(C.Structure):
    _fields_ = [
        ("G", C.c_int32), ("B", C.c_int32), ("d", C.c_int32), ("N", C.c_int32),
        ("normalise", C.c_int32), ("target_per_group", C.c_int32),
        ("c", C.c_void_p), ("Wd", C.c_void_p), ("target", C.c_void_p),
        ("out_x_hat", C.c_void_p), ("out_sq_err", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("stream", C.c_void_p),
        ("operand_format", C.c_int32), ("c_half", C.c_void_p),
    ]


class ScFistaArgs(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("d", C.c_int32), ("N", C.c_int32), ("num_iter", C.c_int32),
        ("x", C.c_void_p), ("D", C.c_void_p), ("a0", C.c_void_p),
        ("l1_coef", C.c_float), ("eta", C.c_float),
        ("momentum", C.POINTER(C.c_float)),
        ("out_a", C.c_void_p), ("out_res", C.c_void_p), ("out_eta", C.c_void_p),
        ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t), ("stream", C.c_void_p),
        ("l1_coef_dev", C.c_void_p), ("pair_format", C.c_int32),
    ]


class ScGatherArgs(C.Structure):
    _fields_ = [
        ("chunk", C.c_void_p), ("elem_bytes", C.c_int32), ("n_rows", C.c_int64), ("d", C.c_int32),
        ("idx", C.c_void_p), ("B", C.c_int32), ("mean", C.c_void_p), ("out", C.c_void_p),
        ("err_flag", C.c_void_p), ("stream", C.c_void_p),
    ]


def build(verbose: bool = False) -> None:
    """Compile libsc_b200.so (and the GEMM self-test) in-tree with nvcc for sm_100a."""
    r = subprocess.run(["make", "-C", CSRC, "-j4"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libsc_b200.so failed")


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C sparse_coding__b200/csrc`).  There is no fallback path.")
        L = C.CDLL(LIB_PATH)
        L.sc_version.restype = C.c_int
        L.sc_last_error.restype = C.c_char_p
        for name in ("sc_sae_step_workspace_bytes",):
            getattr(L, name).restype = C.c_size_t
            getattr(L, name).argtypes = [C.c_int32] * 6
        L.sc_encode_workspace_bytes.restype = C.c_size_t
        L.sc_encode_workspace_bytes.argtypes = [C.c_int32] * 5
        L.sc_decode_workspace_bytes.restype = C.c_size_t
        L.sc_decode_workspace_bytes.argtypes = [C.c_int32] * 4
        L.sc_fista_workspace_bytes.restype = C.c_size_t
        L.sc_fista_workspace_bytes.argtypes = [C.c_int32] * 3
        L.sc_sae_step.argtypes = [C.POINTER(ScStepArgs)]
        L.sc_encode.argtypes = [C.POINTER(ScEncodeArgs)]
        L.sc_decode.argtypes = [C.POINTER(ScDecodeArgs)]
        L.sc_fista.argtypes = [C.POINTER(ScFistaArgs)]
        L.sc_gemm_tf32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int32] * 6 + [C.c_void_p]
        L.sc_gemm_f32_workspace_bytes.restype = C.c_size_t
        L.sc_gemm_f32_workspace_bytes.argtypes = [C.c_int32] * 4
        L.sc_gemm_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int32] * 6 + [C.c_void_p, C.c_size_t, C.c_void_p]
        L.sc_gemm_f32.restype = C.c_int
        L.sc_row_normalise.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        L.sc_column_variance.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        L.sc_sae_step_f16_supported.argtypes = [C.c_int32] * 3
        L.sc_sae_step_f16_supported.restype = C.c_int
        L.sc_codes_to_float.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        L.sc_codes_to_float.restype = C.c_int
        L.sc_adam_apply.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(ScAdam), C.c_void_p]
        L.sc_adam_apply.restype = C.c_int
        L.sc_gather_rows.argtypes = [C.POINTER(ScGatherArgs)]
        L.sc_dp_adam_broadcast.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_int32,
                                           C.c_int64, C.POINTER(ScAdam), C.c_void_p]
        L.sc_infer_f16_supported.argtypes = [C.c_int32] * 4
        L.sc_infer_f16_supported.restype = C.c_int
        L.sc_column_mean_workspace_bytes.restype = C.c_size_t
        L.sc_column_mean_workspace_bytes.argtypes = [C.c_int32]
        L.sc_column_mean.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        for name in ("sc_sae_step", "sc_encode", "sc_decode", "sc_fista", "sc_gemm_tf32", "sc_row_normalise",
                     "sc_column_variance"):
            getattr(L, name).restype = C.c_int
        _lib = L
    return _lib


def profile_begin() -> None:
    lib().sc_profile_begin()


def profile_end(cap: int = 4096):
    """[(kernel name, ms), ...] in launch order since profile_begin()."""
    names = (C.c_char_p * cap)()
    ms = (C.c_float * cap)()
    L = lib()
    L.sc_profile_end.argtypes = [C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.c_int]
    L.sc_profile_end.restype = C.c_int
    n = L.sc_profile_end(names, ms, cap)
    return [(names[i].decode(), float(ms[i])) for i in range(n)]


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().sc_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise NotImplementedError(
            f"{name} is on {t.device}: this engine only runs on a CUDA sm_100a device (no CPU path)")
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Workspace:
    """Grow-only device scratch owned by torch (so it lives in the caching allocator)."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self.buf is None or self.buf.numel() < nbytes + 1024 or self.buf.device != torch.device(device):
            self.buf = torch.empty(nbytes + 1024, dtype=torch.uint8, device=device)
        return self.buf

    def aligned(self, nbytes: int, device):
        b = self.get(nbytes, device)
        base = b.data_ptr()
        off = (-base) % 1024
        return C.c_void_p(base + off), b.numel() - off


_gemm_ws = Workspace()


def gemm_f32(A: torch.Tensor, B: torch.Tensor, a_mn: bool = False, b_mn: bool = False) -> torch.Tensor:
    """C[g] = A[g] B[g]^T at fp32 accuracy through sc_gemm_f32 (3xTF32 on the tcgen05 core): the engine's stand-in for
    the small torch.einsum / torch.mm products around the hot path.  A is [G,M,K] (or [G,K,M] with a_mn), B is [G,N,K]
    (or [G,K,N] with b_mn); 2-D operands are taken as G = 1 and the result is 2-D then."""
    squeeze = A.dim() == 2
    A3, B3 = (A.unsqueeze(0) if A.dim() == 2 else A), (B.unsqueeze(0) if B.dim() == 2 else B)
    require_cuda(A3, "A")
    require_cuda(B3, "B")
    A3, B3 = A3.contiguous(), B3.contiguous()
    G = A3.shape[0]
    M, K = (A3.shape[2], A3.shape[1]) if a_mn else (A3.shape[1], A3.shape[2])
    N, Kb = (B3.shape[2], B3.shape[1]) if b_mn else (B3.shape[1], B3.shape[2])
    if Kb != K or B3.shape[0] != G:
        raise ValueError(f"gemm_f32: operand shapes {tuple(A.shape)} / {tuple(B.shape)} do not match")
    L = lib()
    out = torch.empty((G, M, N), dtype=torch.float32, device=A3.device)
    need = L.sc_gemm_f32_workspace_bytes(G, M, N, K)
    ws_ptr, ws_bytes = _gemm_ws.aligned(need, A3.device)
    with torch.cuda.device(A3.device):
        check(L.sc_gemm_f32(ptr(A3), ptr(B3), ptr(out), G, M, N, K, int(a_mn), int(b_mn), ws_ptr, ws_bytes,
                            stream_ptr(A3.device)), "sc_gemm_f32")
    return out[0] if squeeze else out
