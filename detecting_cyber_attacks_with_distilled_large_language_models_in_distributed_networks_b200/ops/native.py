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
import os
from ctypes import c_int, c_void_p, c_float, c_uint64, c_int64, c_uint32
from pathlib import Path
from typing import Optional

import torch

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "_native" / "libdcad_b200.so"
if os.environ.get("DCAD_LIB"):  # A/B runs of two builds of the library inside one benchmark session
    _LIB_PATH = Path(os.environ["DCAD_LIB"]).resolve()

ERRORS = {
    -1: "bad shape", -2: "bad alignment", -3: "bad argument", -4: "unsupported tile", -5: "no driver entry point",
    -6: "tensor map encode failed", -7: "size beyond supported maximum", -8: "timeout",
}

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_DGELU = 0, 1, 2, 3
MAJOR_K, MAJOR_MN = 0, 1
TILE_PAIR = 0x1000  # OR into block_n: 256 x block_n tile on a CTA pair (tcgen05.mma.cta_group::2)


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
        if os.environ.get("DCAD_PDL", "1") == "0":
            _LIB.dcad_set_pdl(0)
        if os.environ.get("DCAD_ADAM_BLOCKS"):
            _LIB.dcad_set_adam_blocks_per_sm(int(os.environ["DCAD_ADAM_BLOCKS"]))
    return _LIB


def set_pdl(enabled: bool) -> bool:
    """Switch programmatic dependent launch of the step kernels on/off (default on; ``DCAD_PDL=0`` in the
    environment turns it off at load).  Returns the previous setting; affects launches and captures made afterwards."""
    return bool(lib().dcad_set_pdl(1 if enabled else 0))


def set_gemm_epilogue(mode: int) -> int:
    """0 = block-wide slab epilogue, 1 = per-warp 32-row boxes (see ``dcad_gemm_set_epilogue``).  Returns the previous
    mode.  ``block_n=384`` always runs the per-warp epilogue."""
    return int(lib().dcad_gemm_set_epilogue(int(mode)))


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
        f = getattr(L, name)
        f.argtypes = argtypes
        f.restype = restype

    P, I, F, L64, U32 = c_void_p, c_int, c_float, c_int64, c_uint32
    sig("dcad_version", [])
    sig("dcad_set_pdl", [I])
    sig("dcad_gemm_set_epilogue", [I])
    sig("dcad_set_adam_blocks_per_sm", [I])
    sig("dcad_gemm_bf16", [P, P, P, I, I, I, I, I, I, I, I, P, P, P, I, I, I, I, I, I, P])
    sig("dcad_gemm_bf16_colsum", [P, P, P, I, I, I, I, I, I, I, I, P, P, P, I, I, I, I, P, P, P])
    sig("dcad_gemm_bf16_group", [I, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, P])
    sig("dcad_gemm_splitk_workspace_bytes", [I, I, I], ctypes.c_size_t)
    sig("dcad_gemm_bf16_splitk", [P, P, P, I, I, I, I, I, I, I, I, P, P, P, I, I, I, I, I, P, ctypes.c_size_t, P])
    sig("dcad_add_ln_fwd", [P, P, P, P, P, P, I, I, F, F, P, U32, P])
    sig("dcad_add_ln_fwd_x", [P, P, P, P, P, P, P, I, I, F, F, P, U32, P])
    sig("dcad_add_ln_fwd_partials_x", [P, I, L64, P, P, P, P, P, P, P, I, I, F, F, P, U32, P])
    sig("dcad_add_ln_fwd_partials", [P, I, L64, P, P, P, P, P, P, I, I, F, F, P, U32, P])
    sig("dcad_add_ln_bwd_partials", [P, I, L64, P, P, P, P, P, P, P, P, P, P, I, I, F, P, U32, P])
    sig("dcad_gemm_bf16_splitk_partials", [P, P, I, I, I, I, I, I, I, I, I, I, P, ctypes.c_size_t, P, P, P])
    sig("dcad_add_ln_bwd", [P, P, P, P, P, P, P, P, P, P, P, I, I, F, P, U32, P])
    sig("dcad_embed_ln_fwd", [P, P, P, P, P, P, P, P, I, I, I, F, F, P, U32, P])
    sig("dcad_embed_ln_bwd", [P, P, P, P, P, P, P, P, P, P, P, P, P, I, I, I, F, P, U32, P, P])
    sig("dcad_colsum_bf16", [P, P, I, I, I, P])
    sig("dcad_cls_head_train", [P, I, P, P, P, I, F, P, U32, P, P, P, P, P, I, P])
    sig("dcad_tail_head", [P, I, L64, P, P, P, P, P, F, F, U32, P, P, P, F, U32, P, I, I, P, P, P, P, P, P, P, P, P, P])
    sig("dcad_cls_head_eval", [P, I, P, P, P, I, I, P, P, P, P, P])
    sig("dcad_adam_step", [P, P, P, P, P, L64, L64, L64, L64, L64, F, F, F, F, P, P])
    sig("dcad_adam_step_rows", [P, P, P, P, P, L64, L64, L64, L64, L64, F, F, F, F, P, P, L64, L64, I, P, P])
    sig("dcad_step_advance", [P, P])
    sig("dcad_adam_step_dp", [P, P, P, P, L64, L64, L64, L64, L64, L64, F, F, F, F, F, P, P, P, I, I, P])
    sig("dcad_adam_step_dp_rows", [P, P, P, P, P, P, I, I, I, I, F, F, F, F, F, P, P])
    sig("dcad_xgpu_barrier", [P, I, I, P, U32, P])
    sig("dcad_attn_fwd", [P, P, P, P, I, I, I, F, P, U32, P])
    sig("dcad_attn_bwd", [P, P, P, P, P, P, I, I, I, F, P, U32, P, P])
    sig("dcad_fedavg_local", [P, P, P, I, L64, I, P, L64, L64, I, P])
    sig("dcad_fedavg_p2p", [P, P, P, P, P, I, I, L64, I, I, L64, L64, U32, I, P])
    sig("dcad_fedavg_p2p_grouped", [P, I, P, P, P, I, P, I, L64, I, L64, L64, U32, I, P])
    sig("dcad_fedavg_nvls", [P, P, P, I, I, L64, L64, L64, U32, I, P])
    sig("dcad_p2p_alloc", [P, ctypes.c_size_t])
    sig("dcad_p2p_free", [P])
    sig("dcad_p2p_ipc_get", [P, P])
    sig("dcad_p2p_ipc_open", [P, P])
    sig("dcad_p2p_ipc_close", [P])
    sig("dcad_p2p_enable_peer_access", [I])
    sig("dcad_p2p_can_access", [I, I, P])
    sig("dcad_cast_f32_bf16", [P, P, L64, P])
    sig("dcad_gather_rows_bf16", [P, P, I, I, I, I, P])
    sig("dcad_scatter_rows_bf16", [P, P, I, I, I, I, P])


# --------------------------------------------------------------------------------------------------------------
# GEMM
# --------------------------------------------------------------------------------------------------------------
def gemm_bf16(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_major: int = MAJOR_K, b_major: int = MAJOR_K,
              bias: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
              aux_out: Optional[torch.Tensor] = None, epi: int = EPI_NONE, split_k: int = 1,
              reduce_add: bool = False, block_n: int = 0, max_ctas: int = 0, colsum: Optional[torch.Tensor] = None,
              row_lens: Optional[torch.Tensor] = None, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
    """``out[M,N] = epilogue(A·Bᵀ)`` on the tcgen05 kernel.  ``a``/``b`` are 2-D bf16 with unit inner stride.

    ``a`` is ``[M,K]`` (MAJOR_K) or ``[K,M]`` (MAJOR_MN); ``b`` is ``[N,K]`` (MAJOR_K) or ``[K,N]`` (MAJOR_MN).
    ``split_k > 1`` (fp32 ``out`` only, no epilogue) splits the reduction over several CTAs that accumulate into
    ``out`` with TMA reduce-add: ``out`` must be zero beforehand.  ``colsum`` (fp32 ``[N]``, bf16 ``out`` only): the
    epilogue also accumulates the column sums of ``out`` into it (the bias gradient when ``out`` is an output gradient).
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
    elif reduce_add:
        out_fp32 = 2  # accumulate into `out` with TMA reduce-add (it may live on a peer GPU)
    ldaux = 0
    if aux is not None:
        assert aux.dtype == torch.bfloat16 and aux.stride(1) == 1
        ldaux = aux.stride(0)
    if aux_out is not None:
        assert aux_out.dtype == torch.bfloat16 and aux_out.stride(1) == 1
        ldaux = aux_out.stride(0)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() == N
    if colsum is not None or row_lens is not None:
        assert not out_fp32 and split_k == 1
        assert colsum is None or (colsum.dtype == torch.float32 and colsum.numel() == N)
        assert row_lens is None or (row_lens.dtype == torch.int32 and row_lens.numel() * 128 >= M)
        rc = lib().dcad_gemm_bf16_colsum(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a_major, b_major,
                                         a.stride(0), b.stride(0), out.stride(0), ptr(bias), ptr(aux), ptr(aux_out),
                                         ldaux, epi, block_n, max_ctas, ptr(colsum), ptr(row_lens),
                                         stream_ptr(stream))
        check(rc, "dcad_gemm_bf16_colsum")
        return out
    rc = lib().dcad_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a_major, b_major, a.stride(0),
                              b.stride(0), out.stride(0), ptr(bias), ptr(aux), ptr(aux_out), ldaux, epi, out_fp32,
                              split_k, block_n, max_ctas, stream_ptr(stream))
    check(rc, "dcad_gemm_bf16")
    return out


def gemm_splitk_workspace_bytes(M: int, N: int, split_k: int) -> int:
    return int(lib().dcad_gemm_splitk_workspace_bytes(M, N, split_k))


GEMM_GROUP_MAX = 4


def gemm_bf16_group(problems, *, a_major: int = MAJOR_MN, b_major: int = MAJOR_MN, block_n: int = 256,
                    reduce_add: bool = False, max_ctas: int = 0, stream: Optional[torch.cuda.Stream] = None) -> None:
    """Up to four independent fp32-output GEMMs ``out_i = A_i . B_i^T`` (same operand layouts and tile width) in ONE
    persistent launch.  ``problems`` is a list of ``(a, b, out, split_k)``; ``split_k > 1`` (or ``reduce_add``)
    accumulates into ``out`` with TMA reduce-add, so ``out`` must hold zeros (or the running sum) beforehand."""
    n = len(problems)
    assert 1 <= n <= GEMM_GROUP_MAX
    PA, IA = c_void_p * n, c_int * n
    A, B, D = PA(), PA(), PA()
    M, N, K, lda, ldb, ldd, sk = IA(), IA(), IA(), IA(), IA(), IA(), IA()
    for i, (a, b, out, split_k) in enumerate(problems):
        assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and out.dtype == torch.float32
        assert a.stride(1) == 1 and b.stride(1) == 1 and out.stride(1) == 1
        A[i], B[i], D[i] = a.data_ptr(), b.data_ptr(), out.data_ptr()
        M[i], N[i] = out.shape
        K[i] = a.shape[1] if a_major == MAJOR_K else a.shape[0]
        lda[i], ldb[i], ldd[i], sk[i] = a.stride(0), b.stride(0), out.stride(0), split_k
    rc = lib().dcad_gemm_bf16_group(n, A, B, D, M, N, K, lda, ldb, ldd, sk, a_major, b_major, 2 if reduce_add else 1,
                                    block_n, max_ctas, stream_ptr(stream))
    check(rc, "dcad_gemm_bf16_group")


def gemm_bf16_splitk(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, workspace: torch.Tensor, *, split_k: int,
                     a_major: int = MAJOR_K, b_major: int = MAJOR_K, bias: Optional[torch.Tensor] = None,
                     aux: Optional[torch.Tensor] = None, aux_out: Optional[torch.Tensor] = None, epi: int = EPI_NONE,
                     block_n: int = 0, max_ctas: int = 0, stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
    """bf16-output GEMM with the reduction split over ``split_k`` CTAs per tile; ``workspace`` is caller-owned scratch
    (any dtype, >= ``gemm_splitk_workspace_bytes(M, N, split_k)`` bytes, 256-byte aligned)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and out.dtype == torch.bfloat16
    assert a.stride(1) == 1 and b.stride(1) == 1 and out.stride(1) == 1
    M, N = out.shape
    K = a.shape[1] if a_major == MAJOR_K else a.shape[0]
    ldaux = 0
    if aux is not None:
        ldaux = aux.stride(0)
    if aux_out is not None:
        ldaux = aux_out.stride(0)
    rc = lib().dcad_gemm_bf16_splitk(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a_major, b_major, a.stride(0),
                                     b.stride(0), out.stride(0), ptr(bias), ptr(aux), ptr(aux_out), ldaux, epi, split_k,
                                     block_n, max_ctas, workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                                     stream_ptr(stream))
    check(rc, "dcad_gemm_bf16_splitk")
    return out


# --------------------------------------------------------------------------------------------------------------
# LayerNorm family
# --------------------------------------------------------------------------------------------------------------
class Partials:
    """fp32 partial tiles a split-K GEMM left in its workspace: ``splits`` slices of ``slice_elems`` floats, row
    stride = number of columns.  Accepted by add_ln_fwd / add_ln_bwd in place of their first (bf16) operand."""

    def __init__(self, workspace: torch.Tensor, splits: int, slice_elems: int, rows: int, cols: int,
                 bias: Optional[torch.Tensor] = None):
        self.workspace, self.splits, self.slice_elems = workspace, splits, slice_elems
        self.shape = (rows, cols)
        self.bias = bias


def gemm_bf16_splitk_partials(a: torch.Tensor, b: torch.Tensor, M: int, N: int, workspace: torch.Tensor, *,
                              split_k: int, a_major: int = MAJOR_K, b_major: int = MAJOR_K, block_n: int = 0,
                              max_ctas: int = 0, bias: Optional[torch.Tensor] = None,
                              stream: Optional[torch.cuda.Stream] = None) -> Partials:
    """Pass 1 of the split-K GEMM only; ``bias`` is just carried along for the consumer of the partial sums."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.stride(1) == 1 and b.stride(1) == 1
    K = a.shape[1] if a_major == MAJOR_K else a.shape[0]
    splits, slice_elems = c_int(0), c_int64(0)
    rc = lib().dcad_gemm_bf16_splitk_partials(a.data_ptr(), b.data_ptr(), M, N, K, a_major, b_major, a.stride(0),
                                              b.stride(0), split_k, block_n, max_ctas, workspace.data_ptr(),
                                              workspace.numel() * workspace.element_size(), ctypes.byref(splits),
                                              ctypes.byref(slice_elems), stream_ptr(stream))
    check(rc, "dcad_gemm_bf16_splitk_partials")
    return Partials(workspace, splits.value, slice_elems.value, M, N, bias)


def add_ln_fwd(a, res, gamma, beta, out, rstd, *, eps=1e-12, p_drop=0.0, rng=None, stream_id=0, stream=None,
               xhat_out=None):
    """``xhat_out`` (bf16, like ``out``): also keep x_hat for ``add_ln_bwd(..., y=xhat, beta=None)``."""
    if xhat_out is not None:
        if isinstance(a, Partials):
            rows, cols = a.shape
            check(lib().dcad_add_ln_fwd_partials_x(a.workspace.data_ptr(), a.splits, a.slice_elems, ptr(a.bias),
                                                   ptr(res), ptr(gamma), ptr(beta), ptr(out), ptr(rstd), ptr(xhat_out),
                                                   rows, cols, eps, p_drop, ptr(rng), stream_id, stream_ptr(stream)),
                  "dcad_add_ln_fwd_partials_x")
            return out
        rows, cols = a.shape
        check(lib().dcad_add_ln_fwd_x(ptr(a), ptr(res), ptr(gamma), ptr(beta), ptr(out), ptr(rstd), ptr(xhat_out), rows,
                                      cols, eps, p_drop, ptr(rng), stream_id, stream_ptr(stream)), "dcad_add_ln_fwd_x")
        return out
    if isinstance(a, Partials):
        rows, cols = a.shape
        check(lib().dcad_add_ln_fwd_partials(a.workspace.data_ptr(), a.splits, a.slice_elems, ptr(a.bias), ptr(res),
                                             ptr(gamma), ptr(beta), ptr(out), ptr(rstd), rows, cols, eps, p_drop,
                                             ptr(rng), stream_id, stream_ptr(stream)), "dcad_add_ln_fwd_partials")
        return out
    rows, cols = a.shape
    check(lib().dcad_add_ln_fwd(ptr(a), ptr(res), ptr(gamma), ptr(beta), ptr(out), ptr(rstd), rows, cols, eps,
                                p_drop, ptr(rng), stream_id, stream_ptr(stream)), "dcad_add_ln_fwd")
    return out


def add_ln_bwd(dy1, dy2, y, gamma, beta, rstd, dz, da, dgamma, dbeta, dbias, *, p_drop=0.0, rng=None, stream_id=0,
               stream=None):
    """``beta=None``: ``y`` is the x_hat saved by ``add_ln_fwd(..., xhat_out=...)``; else x_hat is rebuilt from the
    LayerNorm output ``y``."""
    rows, cols = y.shape
    if isinstance(dy1, Partials):
        check(lib().dcad_add_ln_bwd_partials(dy1.workspace.data_ptr(), dy1.splits, dy1.slice_elems, ptr(dy2), ptr(y),
                                             ptr(gamma), ptr(beta), ptr(rstd), ptr(dz), ptr(da), ptr(dgamma),
                                             ptr(dbeta), ptr(dbias), rows, cols, p_drop, ptr(rng), stream_id,
                                             stream_ptr(stream)), "dcad_add_ln_bwd_partials")
        return
    check(lib().dcad_add_ln_bwd(ptr(dy1), ptr(dy2), ptr(y), ptr(gamma), ptr(beta), ptr(rstd), ptr(dz), ptr(da),
                                ptr(dgamma), ptr(dbeta), ptr(dbias), rows, cols, p_drop, ptr(rng), stream_id,
                                stream_ptr(stream)), "dcad_add_ln_bwd")


def embed_ln_fwd(ids, word, pos, gamma, beta, out, mean, rstd, seq, *, eps=1e-12, p_drop=0.0, rng=None, stream_id=0,
                 stream=None):
    rows, cols = out.shape
    assert ids.dtype == torch.int32
    check(lib().dcad_embed_ln_fwd(ptr(ids), ptr(word), ptr(pos), ptr(gamma), ptr(beta), ptr(out), ptr(mean),
                                  ptr(rstd), rows, seq, cols, eps, p_drop, ptr(rng), stream_id, stream_ptr(stream)),
          "dcad_embed_ln_fwd")
    return out


def embed_ln_bwd(dy1, dy2, ids, word, pos, gamma, mean, rstd, dword, dpos, dgamma, dbeta, seq, *, lens=None,
                 p_drop=0.0, rng=None, stream_id=0, touched=None, stream=None):
    """``touched`` (uint8 [vocab], optional): set to 1 for every vocabulary row that receives a gradient."""
    rows, cols = dy1.shape
    assert touched is None or touched.dtype == torch.uint8
    check(lib().dcad_embed_ln_bwd(ptr(dy1), ptr(dy2), ptr(ids), ptr(word), ptr(pos), ptr(gamma), ptr(mean),
                                  ptr(rstd), ptr(dword), ptr(dpos), ptr(dgamma), ptr(dbeta), ptr(lens), rows, seq,
                                  cols, p_drop, ptr(rng), stream_id, ptr(touched), stream_ptr(stream)),
          "dcad_embed_ln_bwd")


def colsum_bf16(x, out, stream=None):
    rows, cols = x.shape
    check(lib().dcad_colsum_bf16(ptr(x), ptr(out), rows, cols, x.stride(0), stream_ptr(stream)), "dcad_colsum_bf16")
    return out


# --------------------------------------------------------------------------------------------------------------
# classifier head
# --------------------------------------------------------------------------------------------------------------
def cls_head_train(h, sample_stride, W, bias, labels, batch, loss_accum, logits, dW, db, dcls, *, cols=768,
                   p_drop=0.0, rng=None, stream_id=0, stream=None):
    assert labels.dtype == torch.int32
    check(lib().dcad_cls_head_train(ptr(h), sample_stride, ptr(W), ptr(bias), ptr(labels), batch, p_drop, ptr(rng),
                                    stream_id, ptr(loss_accum), ptr(logits), ptr(dW), ptr(db), ptr(dcls), cols,
                                    stream_ptr(stream)), "dcad_cls_head_train")


def tail_head(a, res, gamma, beta, W, cls_bias, labels, batch, loss_accum, logits, dW, db, dz, da, dgamma, dbeta,
              dbias, *, cols=768, eps=1e-12, p_ln=0.0, ln_stream=0, p_head=0.0, head_stream=0, rng=None, stream=None):
    """Fused LayerNorm + classifier + cross-entropy forward and backward on ``batch`` CLS rows; ``a`` is a bf16
    ``[batch, cols]`` tensor or a ``Partials`` object (split-K pass-1 output + bias)."""
    assert labels.dtype == torch.int32
    if isinstance(a, Partials):
        part, splits, slice_elems, bias, a_ptr = a.workspace.data_ptr(), a.splits, a.slice_elems, ptr(a.bias), None
    else:
        part, splits, slice_elems, bias, a_ptr = None, 0, 0, None, ptr(a)
    check(lib().dcad_tail_head(part, splits, slice_elems, bias, a_ptr, ptr(res), ptr(gamma), ptr(beta), eps, p_ln,
                               ln_stream, ptr(W), ptr(cls_bias), ptr(labels), p_head, head_stream, ptr(rng), batch,
                               cols, ptr(loss_accum), ptr(logits), ptr(dW), ptr(db), ptr(dz), ptr(da), ptr(dgamma),
                               ptr(dbeta), ptr(dbias), stream_ptr(stream)), "dcad_tail_head")


def cls_head_eval(h, sample_stride, W, bias, labels, batch, loss_accum, confusion, probs=None, preds=None, *,
                  cols=768, stream=None):
    assert labels.dtype == torch.int32
    check(lib().dcad_cls_head_eval(ptr(h), sample_stride, ptr(W), ptr(bias), ptr(labels), batch, cols,
                                   ptr(loss_accum), ptr(confusion), ptr(probs), ptr(preds), stream_ptr(stream)),
          "dcad_cls_head_eval")


# --------------------------------------------------------------------------------------------------------------
# optimizer
# --------------------------------------------------------------------------------------------------------------
def adam_step(p, g, m, v, p_bf16, n, bf16_begin, bf16_end, zero_begin, zero_end, lr, beta1, beta2, eps, state,
              stream=None):
    check(lib().dcad_adam_step(ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_bf16), n, bf16_begin, bf16_end, zero_begin,
                               zero_end, lr, beta1, beta2, eps, ptr(state), stream_ptr(stream)), "dcad_adam_step")


def adam_step_rows(p, g, m, v, p_bf16, n, bf16_begin, bf16_end, zero_begin, zero_end, lr, beta1, beta2, eps, state,
                   row_flags, rows_begin, rows_end, row_len, advance_ticket=None, stream=None):
    """Adam over ``n`` elements where [rows_begin, rows_end) is a table of ``row_len``-element rows and rows whose
    byte in ``row_flags`` is 0 (never received a gradient: g = m = v = 0) are skipped -- an exact no-op for them.
    ``advance_ticket`` (zeroed int32 device word): the last block to finish also advances the step counter."""
    assert row_flags is None or row_flags.dtype == torch.uint8
    check(lib().dcad_adam_step_rows(ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_bf16), n, bf16_begin, bf16_end, zero_begin,
                                    zero_end, lr, beta1, beta2, eps, ptr(state), ptr(row_flags), rows_begin, rows_end,
                                    row_len, ptr(advance_ticket), stream_ptr(stream)), "dcad_adam_step_rows")


def adam_step_dp(p, g, m, v, begin, end, bf16_begin, bf16_end, push32_begin, push32_end, lr, beta1, beta2, eps,
                 grad_scale, state, peer_p_ptrs, peer_bf16_ptrs, self_index, stream=None):
    n = len(peer_p_ptrs) if peer_p_ptrs is not None else len(peer_bf16_ptrs)
    check(lib().dcad_adam_step_dp(ptr(p), ptr(g), ptr(m), ptr(v), begin, end, bf16_begin, bf16_end, push32_begin,
                                  push32_end, lr, beta1, beta2, eps, grad_scale, ptr(state), _ptr_array(peer_p_ptrs),
                                  _ptr_array(peer_bf16_ptrs), n, self_index, stream_ptr(stream)), "dcad_adam_step_dp")


def adam_step_dp_rows(m, v, g_own, p_ptrs, g_ptrs, flag_ptrs, row_lo, row_hi, row_len, lr, beta1, beta2, eps, grad_scale,
                      state, stream=None):
    """Row-sparse sharded Adam of the word-embedding table (``dcad_adam_step_dp_rows``); ``p_ptrs[0]`` is this rank's."""
    check(lib().dcad_adam_step_dp_rows(ptr(m), ptr(v), ptr(g_own), _ptr_array(p_ptrs), _ptr_array(g_ptrs),
                                       _ptr_array(flag_ptrs), len(p_ptrs), row_lo, row_hi, row_len, lr, beta1, beta2,
                                       eps, grad_scale, ptr(state), stream_ptr(stream)), "dcad_adam_step_dp_rows")


def xgpu_barrier(pad_ptrs, rank, state, phase, stream=None):
    check(lib().dcad_xgpu_barrier(_ptr_array(pad_ptrs), rank, len(pad_ptrs), ptr(state), phase, stream_ptr(stream)),
          "dcad_xgpu_barrier")


def step_advance(state, stream=None):
    check(lib().dcad_step_advance(ptr(state), stream_ptr(stream)), "dcad_step_advance")


# --------------------------------------------------------------------------------------------------------------
# attention
# --------------------------------------------------------------------------------------------------------------
def attn_fwd(qkv, ctx, lse2, lens, B, S, H, *, p_drop=0.0, rng=None, stream_id=0, stream=None):
    check(lib().dcad_attn_fwd(ptr(qkv), ptr(ctx), ptr(lse2), ptr(lens), B, S, H, p_drop, ptr(rng), stream_id,
                              stream_ptr(stream)), "dcad_attn_fwd")
    return ctx


def attn_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, *, p_drop=0.0, rng=None, stream_id=0, dbias=None, stream=None):
    """``dbias`` (fp32 [3*H*64], optional): accumulates the column sums of ``dqkv`` (fused QKV bias gradient)."""
    check(lib().dcad_attn_bwd(ptr(qkv), ptr(ctx), ptr(dctx), ptr(lse2), ptr(dqkv), ptr(lens), B, S, H, p_drop,
                              ptr(rng), stream_id, ptr(dbias), stream_ptr(stream)), "dcad_attn_bwd")
    return dqkv


# --------------------------------------------------------------------------------------------------------------
# FedAvg + peer runtime
# --------------------------------------------------------------------------------------------------------------
def _ptr_array(ptrs):
    if ptrs is None:
        return None
    arr = (c_void_p * len(ptrs))()
    for i, p_ in enumerate(ptrs):
        arr[i] = p_
    return arr


def _float_array(vals):
    if vals is None:
        return None
    arr = (c_float * len(vals))()
    for i, v_ in enumerate(vals):
        arr[i] = float(v_)
    return arr


def fedavg_local(src_ptrs, dst_ptr, count, *, dst_bf16_ptr=0, in_bf16=False, weights=None, bf16_begin=0, bf16_end=0,
                 max_ctas=0, stream=None):
    """dst = mean_k src_k over raw device pointers (all reachable from the current device)."""
    check(lib().dcad_fedavg_local(_ptr_array(src_ptrs), dst_ptr, dst_bf16_ptr, len(src_ptrs), count,
                                  1 if in_bf16 else 0, _float_array(weights), bf16_begin, bf16_end, max_ctas,
                                  stream_ptr(stream)), "dcad_fedavg_local")


def fedavg_p2p(arena_ptrs, pad_ptrs, rank, count, epoch, *, out_ptrs=None, mirror_ptrs=None, weights=None,
               in_bf16=False, mode=0, bf16_begin=0, bf16_end=0, max_ctas=0, stream=None):
    check(lib().dcad_fedavg_p2p(_ptr_array(arena_ptrs), _ptr_array(out_ptrs), _ptr_array(mirror_ptrs),
                                _ptr_array(pad_ptrs), _float_array(weights), rank, len(arena_ptrs), count,
                                1 if in_bf16 else 0, mode, bf16_begin, bf16_end, epoch, max_ctas,
                                stream_ptr(stream)), "dcad_fedavg_p2p")


def fedavg_p2p_grouped(src_ptrs, dst_ptrs, mirror_ptrs, pad_ptrs, rank, count, epoch, *, weights=None, in_bf16=False,
                       bf16_begin=0, bf16_end=0, max_ctas=0, stream=None):
    check(lib().dcad_fedavg_p2p_grouped(_ptr_array(src_ptrs), len(src_ptrs), _ptr_array(dst_ptrs),
                                        _ptr_array(mirror_ptrs), _ptr_array(pad_ptrs), len(dst_ptrs),
                                        _float_array(weights), rank, count, 1 if in_bf16 else 0, bf16_begin, bf16_end,
                                        epoch, max_ctas, stream_ptr(stream)), "dcad_fedavg_p2p_grouped")


def fedavg_nvls(mc_arena_ptr, mc_mirror_ptr, pad_ptrs, rank, count, epoch, *, bf16_begin=0, bf16_end=0, max_ctas=0,
                stream=None):
    check(lib().dcad_fedavg_nvls(mc_arena_ptr, mc_mirror_ptr, _ptr_array(pad_ptrs), rank, len(pad_ptrs), count,
                                 bf16_begin, bf16_end, epoch, max_ctas, stream_ptr(stream)), "dcad_fedavg_nvls")


def p2p_alloc(nbytes: int) -> int:
    out = c_void_p()
    check(lib().dcad_p2p_alloc(ctypes.byref(out), nbytes), "dcad_p2p_alloc")
    return out.value


def p2p_free(p: int) -> None:
    check(lib().dcad_p2p_free(p), "dcad_p2p_free")


def p2p_ipc_get(p: int) -> bytes:
    buf = ctypes.create_string_buffer(64)
    check(lib().dcad_p2p_ipc_get(p, buf), "dcad_p2p_ipc_get")
    return buf.raw


def p2p_ipc_open(handle: bytes) -> int:
    out = c_void_p()
    buf = ctypes.create_string_buffer(handle, 64)
    check(lib().dcad_p2p_ipc_open(buf, ctypes.byref(out)), "dcad_p2p_ipc_open")
    return out.value


def p2p_ipc_close(p: int) -> None:
    check(lib().dcad_p2p_ipc_close(p), "dcad_p2p_ipc_close")


def p2p_enable_peer_access(peer: int) -> None:
    check(lib().dcad_p2p_enable_peer_access(peer), "dcad_p2p_enable_peer_access")


def p2p_can_access(dev: int, peer: int) -> bool:
    out = c_int(0)
    check(lib().dcad_p2p_can_access(dev, peer, ctypes.byref(out)), "dcad_p2p_can_access")
    return bool(out.value)


class RawDeviceBuffer:
    """A device allocation made by the native runtime, viewable as torch tensors (no copy) through
    ``__cuda_array_interface__``.  Used for peer-mappable arenas and signal pads."""

    def __init__(self, ptr_: int, nbytes: int, owner: bool = True, ipc: bool = False):
        self.ptr, self.nbytes, self.owner, self.ipc = ptr_, nbytes, owner, ipc

    def view(self, dtype: torch.dtype, offset_bytes: int = 0, numel: Optional[int] = None) -> torch.Tensor:
        itemsize = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - offset_bytes) // itemsize
        typestr = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.int32: "<i4", torch.uint8: "|u1",
                   torch.int64: "<i8", torch.uint32: "<u4"}[dtype]

        class _Iface:
            pass

        o = _Iface()
        o.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (self.ptr + offset_bytes, False),
                                      "version": 3, "strides": None}
        o._keepalive = self
        t = torch.as_tensor(o, device=torch.device("cuda", torch.cuda.current_device()))
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        t._dcad_owner = self  # keep the allocation alive as long as any view lives
        return t

    def free(self) -> None:
        if self.ptr:
            if self.ipc:
                p2p_ipc_close(self.ptr)
            elif self.owner:
                p2p_free(self.ptr)
            self.ptr = 0


def cast_f32_bf16(src, dst, stream=None):
    check(lib().dcad_cast_f32_bf16(ptr(src), ptr(dst), src.numel(), stream_ptr(stream)), "dcad_cast_f32_bf16")
    return dst


def gather_rows_bf16(src, dst, rows, cols, every, stream=None):
    check(lib().dcad_gather_rows_bf16(ptr(src), ptr(dst), rows, cols, every, src.stride(0), stream_ptr(stream)),
          "dcad_gather_rows_bf16")
    return dst


def scatter_rows_bf16(src, dst, rows, cols, every, stream=None):
    check(lib().dcad_scatter_rows_bf16(ptr(src), ptr(dst), rows, cols, every, dst.stride(0), stream_ptr(stream)),
          "dcad_scatter_rows_bf16")
    return dst
