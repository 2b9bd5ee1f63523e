"""GPU tests of the GEMM's per-warp epilogue (``dcad_gemm_set_epilogue(1)``) and of the 128 x 384 tile, against fp32
PyTorch matmul element by element — the oracle SURVEY.md W5 names for the cuBLAS GEMMs the reference reaches through
``nn.Linear`` (client1.py:61 -> transformers modeling_distilbert.py:187-226).  Collected after the other kernel tests
(file name) so that a failure here cannot hide them."""
from __future__ import annotations

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nv():
    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native

    native.lib()
    return native


@pytest.fixture(params=[1, 2])
def warp_epilogue(nv, request):
    """1 = per-warp boxes beside the operand ring, 2 = per-warp boxes inside the ring (one more operand stage)"""
    prev = nv.set_gemm_epilogue(request.param)
    yield
    nv.set_gemm_epilogue(prev)


def _close(out, ref, tol):
    """|err| <= tol * |ref| + tol * rms(ref), element by element"""
    ref = ref.float()
    return torch.allclose(out.float(), ref, rtol=tol, atol=tol * ref.pow(2).mean().sqrt().item())


def _operands(M, N, K, a_major, b_major, seed):
    g = torch.Generator("cuda").manual_seed(seed)
    a = torch.randn((M, K) if a_major == 0 else (K, M), device="cuda", generator=g).bfloat16()
    b = (torch.randn((N, K) if b_major == 0 else (K, N), device="cuda", generator=g) * 0.05).bfloat16()
    A = a.float() if a_major == 0 else a.float().t()
    B = b.float() if b_major == 0 else b.float().t()
    return a, b, A @ B.t()


@pytest.mark.parametrize("a_major,b_major", [(0, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("bn", [64, 128, 192, 256])
def test_warp_epilogue_matches_matmul(nv, warp_epilogue, a_major, b_major, bn):
    """ragged M / N, several tiles per CTA, fp32 and bf16 outputs"""
    for (M, N, K) in [(392, 776, 320), (16, 768, 768)]:
        a, b, ref = _operands(M, N, K, a_major, b_major, seed=bn + a_major)
        for dt, tol in ((torch.float32, 1e-3), (torch.bfloat16, 1e-2)):
            out = torch.full((M, N), float("nan"), device="cuda", dtype=dt)
            nv.gemm_bf16(a, b, out, a_major=a_major, b_major=b_major, block_n=bn, max_ctas=5)
            assert _close(out, ref, tol), (M, N, K, dt)


def test_warp_epilogue_equals_block_epilogue_bitwise(nv):
    """same accumulators, same arithmetic: the two epilogues must agree bit for bit (bias+GELU with the kept
    pre-activation, dGELU with column sums, split-K reduce-add, workspace partials, the grouped launch, CTA pairs)"""
    M, N, K = 520, 1536, 448
    a, w, _ = _operands(M, N, K, 0, 0, seed=1)
    bias = torch.randn(N, device="cuda")
    aux = torch.randn(M, N, device="cuda").bfloat16()
    wt = w.t().contiguous()
    dy = torch.randn(640, 768, device="cuda").bfloat16()
    x = torch.randn(640, 512, device="cuda").bfloat16()
    res = {}
    for mode in (0, 1, 2):
        prev = nv.set_gemm_epilogue(mode)
        try:
            out, pre = (torch.empty(M, N, device="cuda", dtype=torch.bfloat16) for _ in range(2))
            nv.gemm_bf16(a, w, out, bias=bias, aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=256)
            dg = torch.empty_like(out)
            cs = torch.zeros(N, device="cuda")
            nv.gemm_bf16(a, wt, dg, b_major=nv.MAJOR_MN, aux=aux, epi=nv.EPI_DGELU, block_n=192, colsum=cs)
            pair = torch.empty_like(out)
            nv.gemm_bf16(a, w, pair, bias=bias, epi=nv.EPI_BIAS, block_n=256 | nv.TILE_PAIR)
            dw = torch.zeros(768, 512, device="cuda")
            nv.gemm_bf16(dy, x, dw, a_major=nv.MAJOR_MN, b_major=nv.MAJOR_MN, block_n=256, split_k=1)
            ws = torch.empty(nv.gemm_splitk_workspace_bytes(M, N, 3) // 4, device="cuda")
            sk = torch.empty_like(out)
            nv.gemm_bf16_splitk(a, w, sk, ws, split_k=3, bias=bias, epi=nv.EPI_BIAS, block_n=256)
            grp = [torch.zeros(768, 512, device="cuda"), torch.zeros(512, 768, device="cuda")]
            nv.gemm_bf16_group([(dy, x, grp[0], 1), (x, dy, grp[1], 1)], block_n=256)
            torch.cuda.synchronize()
            res[mode] = (out, pre, dg, pair, dw, sk, grp[0], grp[1])
        finally:
            nv.set_gemm_epilogue(prev)
    names = ("gelu", "pre", "dgelu", "pair", "wgrad", "splitk", "group0", "group1")
    for mode in (1, 2):
        for n, x0, x1 in zip(names, res[0], res[mode]):
            assert torch.equal(x0, x1), (mode, n)
    assert _close(res[1][0], F.gelu(a.float() @ w.float().t() + bias), 1e-2)


@pytest.mark.parametrize("b_major", [0, 1])
def test_wide_tile_384(nv, b_major):
    """128 x 384 tiles: two N = 192 accumulators per k-step, staging boxes aliased onto the operand ring.  Ragged M and
    N; one tile per CTA and several per CTA (max_ctas = 3: the producer must wait for the epilogue's stores)."""
    for (M, N, K) in [(300, 1160, 320), (2048, 3072, 768)]:
        a, w, ref = _operands(M, N, K, 0, b_major, seed=21 + b_major)
        for max_ctas in (0, 3):
            for dt, tol in ((torch.float32, 1e-3), (torch.bfloat16, 1e-2)):
                out = torch.full((M, N), float("nan"), device="cuda", dtype=dt)
                nv.gemm_bf16(a, w, out, b_major=b_major, block_n=384, max_ctas=max_ctas)
                assert _close(out, ref, tol), (M, N, K, max_ctas, dt)
        bias = torch.randn(N, device="cuda")
        if b_major == 0:   # lin1 forward: bias + GELU, pre-activation kept
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            pre = torch.full_like(out, float("nan"))
            nv.gemm_bf16(a, w, out, bias=bias, aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=384)
            assert _close(pre, ref + bias, 1e-2) and _close(out, F.gelu(ref + bias), 1e-2)
            out2 = torch.empty_like(out)
            nv.gemm_bf16(a, w, out2, bias=bias, epi=nv.EPI_BIAS_GELU, block_n=384, max_ctas=5)
            assert torch.equal(out, out2)
        else:              # lin2 dgrad: dGELU epilogue + column sums
            aux = torch.randn(M, N, device="cuda").bfloat16()
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            cs = torch.full((N,), 2.0, device="cuda")
            nv.gemm_bf16(a, w, out, b_major=1, aux=aux, epi=nv.EPI_DGELU, block_n=384, colsum=cs)
            x = aux.float().requires_grad_(True)
            F.gelu(x).backward(torch.ones_like(x))
            assert _close(out, ref * x.grad, 1e-2)
            # the kernel sums the fp32 values before they are rounded to bf16: compare with the fp32 reference, within
            # the rounding noise of M bf16 inputs per column
            want = (ref * x.grad).sum(0)
            noise = 2.0 ** -8 * (ref * x.grad).pow(2).mean().sqrt().item() * M ** 0.5
            assert torch.allclose(cs - 2.0, want, rtol=1e-2, atol=4 * noise + 1e-3)
