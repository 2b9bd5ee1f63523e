"""GPU test of dcad_gemm_bf16_splitk_static: the split-K GEMM for hosts that may not allocate device memory once they
are running (a stream capture in progress, a memory pool frozen at start-up, an embedding application that bans
run-time cudaMalloc).  Its fp32 scratch is a fixed 24 MiB region inside the library instead of a caller buffer.

The library is driven through its C ABI with ctypes, the way such a host would call it.  Two different problems are
run back to back -- so the second call finds the first one's partial sums in the shared scratch -- and each result
must equal, bit for bit, what the established dcad_gemm_bf16_splitk entry point produces with a caller-allocated
workspace (same kernels, same summation order, hence no tolerance is needed).  A problem too large for the built-in
scratch must be refused.  The file name makes pytest collect it last.
"""
from __future__ import annotations

import ctypes
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = (Path(__file__).resolve().parent.parent
       / "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200" / "_native"
       / "libdcad_b200.so")
P, I, SZ = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
EPI_BIAS, EPI_DGELU = 1, 3
ERR_TOO_LARGE = -7


def _lib():
    lib = ctypes.CDLL(str(LIB))
    common = [P, P, P, I, I, I, I, I, I, I, I, P, P, P, I, I, I, I, I]
    lib.dcad_gemm_bf16_splitk.argtypes = common + [P, SZ, P]
    lib.dcad_gemm_bf16_splitk.restype = I
    lib.dcad_gemm_bf16_splitk_static.argtypes = common + [P]
    lib.dcad_gemm_bf16_splitk_static.restype = I
    lib.dcad_gemm_splitk_workspace_bytes.argtypes = [I, I, I]
    lib.dcad_gemm_splitk_workspace_bytes.restype = SZ
    lib.dcad_gemm_splitk_static_capacity.argtypes = []
    lib.dcad_gemm_splitk_static_capacity.restype = SZ
    return lib


def _run_both(lib, M, N, K, split, epi, b_major, seed, check_mem=False):
    g = torch.Generator(torch.device("cuda", torch.cuda.current_device())).manual_seed(seed)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn((N, K) if b_major == 0 else (K, N), device="cuda", generator=g) * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    aux = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    stream = torch.cuda.current_stream().cuda_stream
    args = [a.data_ptr(), b.data_ptr(), None, M, N, K, 0, b_major, a.stride(0), b.stride(0), N,
            bias.data_ptr() if epi == EPI_BIAS else None, aux.data_ptr() if epi == EPI_DGELU else None, None, N, epi,
            split, 256, 0]
    need = lib.dcad_gemm_splitk_workspace_bytes(M, N, split)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    args[2] = want.data_ptr()
    assert lib.dcad_gemm_bf16_splitk(*args, ws.data_ptr(), need, stream) == 0
    got = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    args[2] = got.data_ptr()
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    assert lib.dcad_gemm_bf16_splitk_static(*args, stream) == 0
    torch.cuda.synchronize()
    if check_mem:
        # the claim itself: the call took no device memory from the driver (nothing allocated, nothing freed)
        assert torch.cuda.mem_get_info()[0] == free_before
    assert not torch.isnan(got.float()).any()
    assert torch.equal(got, want)
    # sanity against plain fp32 math so that "equal" cannot mean "equally wrong"
    ref = a.float() @ (b.float().t() if b_major == 0 else b.float())
    if epi == EPI_BIAS:
        ref = ref + bias
    else:
        x = aux.float().requires_grad_(True)
        torch.nn.functional.gelu(x).backward(torch.ones_like(x))
        ref = ref * x.grad
    assert ((got.float() - ref).abs().max() / ref.abs().max()).item() < 1e-2


def test_built_in_scratch_matches_caller_workspace_on_consecutive_calls():
    lib = _lib()
    cap = lib.dcad_gemm_splitk_static_capacity()
    assert cap == 24 << 20
    # call 1: the FFN lin2 forward shape of the training step (18.9 MB of scratch, bias epilogue)
    _run_both(lib, 2048, 768, 3072, 3, EPI_BIAS, 0, seed=1)
    # call 2, immediately after, a smaller ragged problem with another epilogue and operand layout: it reuses the
    # scratch the first call filled
    _run_both(lib, 520, 768, 1536, 4, EPI_DGELU, 1, seed=2, check_mem=True)
    # a problem whose partial sums cannot fit in the built-in scratch is refused, not truncated
    assert lib.dcad_gemm_splitk_workspace_bytes(8192, 3072, 4) > cap
    dummy = torch.empty(8, device="cuda", dtype=torch.bfloat16)
    rc = lib.dcad_gemm_bf16_splitk_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 8192, 3072, 768, 0, 0,
                                          768, 768, 3072, None, None, None, 0, 0, 4, 256, 0,
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == ERR_TOO_LARGE


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_every_gpu_has_its_own_built_in_scratch():
    """One process driving two GPUs: the call made while GPU 1 is current must use GPU 1's copy of the scratch (and
    GPU 1's kernel attributes), after GPU 0 has already been served."""
    lib = _lib()
    with torch.cuda.device(0):
        _run_both(lib, 520, 768, 1536, 4, EPI_BIAS, 0, seed=3)
    with torch.cuda.device(1):
        _run_both(lib, 520, 768, 1536, 4, EPI_BIAS, 0, seed=4)
        _run_both(lib, 2048, 768, 3072, 3, EPI_DGELU, 1, seed=5, check_mem=True)
