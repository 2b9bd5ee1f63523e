"""ctypes binding of ``_native/libdcad_b200.so`` (C ABI in ``include/dcad_b200.h``).

Design note: the kernels are plain ``extern "C"`` launchers taking raw device pointers and a stream, so binding
them needs no pybind11/torch headers (compile time stays in seconds) and every call is capturable in a CUDA graph
— the engine captures a whole training step, so the ctypes call overhead is paid once at capture.

The library is loaded lazily.  On a machine with a GPU a missing library is a hard error (no silent fallback to
eager PyTorch); on a CPU-only machine ``available()`` is simply False and the pure-torch reference paths are used
by the CPU plumbing tests.
"""
from __future__ import annotations

import ctypes
from ctypes import c_int, c_void_p, c_float, c_uint64, c_int64, c_uint32
from pathlib import Path
from typing import Optional

import torch

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "_native" / "libdcad_b200.so"

ERRORS = {
    -1: "bad shape", -2: "bad alignment", -3: "bad argument", -4: "unsupported tile", -5: "no driver entry point",
    -6: "tensor map encode failed", -7: "size beyond supported maximum", -8: "timeout",
}

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_DGELU = 0, 1, 2, 3
MAJOR_K, MAJOR_MN = 0, 1


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    """True when the native library can be used (built + a CUDA device is present)."""
    return torch.cuda.is_available() and _LIB_PATH.exists()


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        if not _LIB_PATH.exists():
            raise NativeError(
                f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(the CUDA path has no PyTorch fallback on a GPU machine)")
        _LIB = ctypes.CDLL(str(_LIB_PATH))
        _declare(_LIB)
    return _LIB


def check(rc: int, what: str) -> None:
    if rc == 0:
        return
    if rc < 0:
        raise NativeError(f"{what}: {ERRORS.get(rc, 'error')} ({rc})")
    raise NativeError(f"{what}: CUDA error {rc}")


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def _declare(L: ctypes.CDLL) -> None:
    def sig(name, argtypes, restype=c_int):
        if hasattr(L, name):
            f = getattr(L, name)
            f.argtypes = argtypes
            f.restype = restype

    P, I, F = c_void_p, c_int, c_float
    sig("dcad_version", [])
    sig("dcad_gemm_bf16", [P, P, P, I, I, I, I, I, I, I, I, P, P, P, I, I, I, I, I, I, P])
    sig("dcad_add_ln_fwd", [P, P, P, P, P, P, I, I, F, F, P, c_uint32, P])
    sig("dcad_add_ln_bwd", [P, P, P, P, P, P, P, P, P, P, P, I, I, F, P, c_uint32, P])
    sig("dcad_embed_ln_fwd", [P, P, P, P, P, P, P, P, I, I, I, F, F, P, c_uint32, P])
    sig("dcad_embed_ln_bwd", [P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, F, P, c_uint32, P])
    sig("dcad_colsum_bf16", [P, P, I, I, I, P])
    sig("dcad_cls_head_train", [P, I, P, P, P, I, F, P, c_uint32, P, P, P, P, P, I, P])
    sig("dcad_cls_head_eval", [P, I, P, P, P, I, P, P, P, P])
    sig("dcad_adam_step", [P, P, P, P, P, c_int64, c_int64, c_int64, F, F, F, F, P, P])
    sig("dcad_step_advance", [P, P])
    sig("dcad_cast_f32_bf16", [P, P, c_int64, P])
    sig("dcad_attn_fwd", [P, P, P, P, I, I, I, F, P, c_uint32, P])
    sig("dcad_attn_bwd", [P, P, P, P, P, P, I, I, I, F, P, c_uint32, P])
    sig("dcad_fedavg_local", [P, P, I, c_int64, I, F, P])
    sig("dcad_fedavg_p2p", [P, P, P, I, I, c_int64, I, I, c_uint32, P])
    sig("dcad_gather_rows_bf16", [P, P, I, I, I, I, P])
    sig("dcad_scatter_rows_bf16", [P, P, I, I, I, I, P])


# --------------------------------------------------------------------------------------------------------------
# GEMM
# --------------------------------------------------------------------------------------------------------------
def gemm_bf16(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_major: int = MAJOR_K, b_major: int = MAJOR_K,
              bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
              aux_out: Optional[torch.Tensor] = None, epi: int = EPI_NONE, accumulate: bool = False,
              block_n: int = 0, max_ctas: int = 0, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
    """``out[M,N] = epilogue(A·Bᵀ)`` on the tcgen05 kernel.  ``a``/``b`` are 2-D bf16 with unit inner stride.

    ``a`` is ``[M,K]`` (MAJOR_K) or ``[K,M]`` (MAJOR_MN); ``b`` is ``[N,K]`` (MAJOR_K) or ``[K,N]`` (MAJOR_MN).
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.dim() == 2 and b.dim() == 2 and out.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1 and out.stride(1) == 1
    M, N = out.shape
    if a_major == MAJOR_K:
        assert a.shape[0] == M
        K = a.shape[1]
    else:
        assert a.shape[1] == M
        K = a.shape[0]
    if b_major == MAJOR_K:
        assert tuple(b.shape) == (N, K), (b.shape, N, K)
    else:
        assert tuple(b.shape) == (K, N), (b.shape, K, N)
    out_fp32 = 1 if out.dtype == torch.float32 else 0
    if not out_fp32:
        assert out.dtype == torch.bfloat16
    ldaux = 0
    if aux is not None:
        assert aux.dtype == torch.bfloat16 and aux.stride(1) == 1
        ldaux = aux.stride(0)
    if aux_out is not None:
        assert aux_out.dtype == torch.bfloat16 and aux_out.stride(1) == 1
        ldaux = aux_out.stride(0)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N
    rc = lib().dcad_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a_major, b_major, a.stride(0),
                              b.stride(0), out.stride(0), ptr(bias), ptr(aux), ptr(aux_out), ldaux, epi, out_fp32,
                              1 if accumulate else 0, block_n, max_ctas, stream_ptr(stream))
    check(rc, "dcad_gemm_bf16")
    return out
