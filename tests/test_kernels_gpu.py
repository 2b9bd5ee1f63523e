"""GPU numerics tests: every hand-written sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
from __future__ import annotations

import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"


@pytest.fixture(scope="module")
def nv():
    import importlib

    mod = importlib.import_module(f"{PKG}.ops.native")
    assert mod.lib().dcad_version() >= 100
    return mod


def _rng_state(seed=1234, step=0):
    return torch.tensor([seed, step], dtype=torch.int64, device="cuda")


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


# ---------------------------------------------------------------- GEMM
@pytest.mark.parametrize("a_major,b_major,bn", [(0, 0, 64), (0, 0, 256), (0, 1, 128), (0, 1, 192), (1, 1, 128), (1, 1, 64)])
def test_gemm_matches_matmul(nv, a_major, b_major, bn):
    torch.manual_seed(0)
    M, N, K = 384, 768, 320
    a = torch.randn((M, K) if a_major == 0 else (K, M), device="cuda").bfloat16()
    b = torch.randn((N, K) if b_major == 0 else (K, N), device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.float32)
    nv.gemm_bf16(a, b, out, a_major=a_major, b_major=b_major, block_n=bn)
    A = a.float() if a_major == 0 else a.float().t()
    B = b.float() if b_major == 0 else b.float().t()
    ref = A @ B.t()
    assert _rel(out, ref) < 1e-3


@pytest.mark.parametrize("a_major,b_major,bn", [(0, 0, 128), (0, 0, 192), (0, 0, 256), (0, 1, 128), (0, 1, 256),
                                                (1, 1, 128), (1, 1, 256)])
def test_gemm_cta_pair_matches_matmul(nv, a_major, b_major, bn):
    """cta_group::2 tiles (256 x bn on a CTA pair): odd number of 128-row blocks, ragged N, several tiles per pair."""
    torch.manual_seed(1)
    for (M, N, K) in [(384, 768, 320), (1152, 1304, 192), (16, 768, 768)]:
        a = torch.randn((M, K) if a_major == 0 else (K, M), device="cuda").bfloat16()
        b = torch.randn((N, K) if b_major == 0 else (K, N), device="cuda").bfloat16()
        A = a.float() if a_major == 0 else a.float().t()
        B = b.float() if b_major == 0 else b.float().t()
        ref = A @ B.t()
        for dt in (torch.float32, torch.bfloat16):
            out = torch.full((M, N), float("nan"), device="cuda", dtype=dt)
            nv.gemm_bf16(a, b, out, a_major=a_major, b_major=b_major, block_n=bn | nv.TILE_PAIR, max_ctas=8)
            assert _rel(out, ref) < (1e-3 if dt == torch.float32 else 1e-2), (M, N, K, dt)


def test_gemm_cta_pair_epilogues_and_splitk(nv):
    torch.manual_seed(2)
    M, N, K = 520, 768, 1536
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    pre = a.float() @ w.float().t() + bias
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    aux_out = torch.empty_like(out)
    nv.gemm_bf16(a, w, out, bias=bias, aux_out=aux_out, epi=nv.EPI_BIAS_GELU, block_n=256 | nv.TILE_PAIR)
    assert _rel(out, F.gelu(pre)) < 1e-2 and _rel(aux_out, pre) < 1e-2
    for split in (2, 3):
        ws = torch.empty(nv.gemm_splitk_workspace_bytes(M, N, split) // 4, device="cuda")
        out2 = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
        nv.gemm_bf16_splitk(a, w, out2, ws, split_k=split, epi=nv.EPI_BIAS, bias=bias, block_n=256 | nv.TILE_PAIR)
        assert _rel(out2, pre) < 1e-2, split
    # weight-gradient form with TMA reduce-add split-K
    T, Nw, Kw = 2048, 768, 512
    dy = torch.randn(T, Nw, device="cuda").bfloat16()
    x = torch.randn(T, Kw, device="cuda").bfloat16()
    ref = dy.float().t() @ x.float()
    for split in (1, 2, 4):
        g = torch.zeros(Nw, Kw, device="cuda")
        nv.gemm_bf16(dy, x, g, a_major=nv.MAJOR_MN, b_major=nv.MAJOR_MN, block_n=256 | nv.TILE_PAIR, split_k=split)
        assert _rel(g, ref) < 1e-3, split


def test_gemm_group_matches_separate_launches(nv):
    """Four weight-gradient shaped problems of different sizes and split-K factors in one persistent launch."""
    torch.manual_seed(3)
    T = 640
    shapes = [(768, 3072, 2), (3072, 768, 2), (768, 768, 4), (2304, 768, 1)]   # dW[M, N] = dY[T, M]^T X[T, N]
    probs, refs = [], []
    for M, N, sk in shapes:
        dy = torch.randn(T, M, device="cuda").bfloat16()
        x = torch.randn(T, N, device="cuda").bfloat16()
        out = torch.zeros(M, N, device="cuda")
        probs.append((dy, x, out, sk))
        refs.append(dy.float().t() @ x.float())
    for n in (4, 3, 1):
        for _, _, out, _ in probs:
            out.zero_()
        nv.gemm_bf16_group(probs[:n], block_n=256)
        for (_, _, out, _), ref in zip(probs[:n], refs[:n]):
            assert _rel(out, ref) < 1e-3, n
    # accumulate mode on top of existing contents, few CTAs (several units per CTA across problem boundaries)
    for _, _, out, _ in probs:
        out.fill_(1.0)
    nv.gemm_bf16_group(probs, block_n=128, reduce_add=True, max_ctas=5)
    for (_, _, out, _), ref in zip(probs, refs):
        assert _rel(out, ref + 1.0) < 1e-3


@pytest.mark.parametrize("bn", [64, 192, 256])
def test_gemm_epilogue_column_sums(nv, bn):
    """dGELU GEMM whose epilogue also accumulates the column sums of its output (bias gradient of the consumer)."""
    torch.manual_seed(4)
    M, N, K = 300, 776, 320          # ragged rows and columns
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(K, N, device="cuda") * 0.05).bfloat16()     # MN-major B, like a dgrad
    aux = torch.randn(M, N, device="cuda").bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    cs = torch.full((N,), 2.0, device="cuda")
    nv.gemm_bf16(a, w, out, b_major=nv.MAJOR_MN, aux=aux, epi=nv.EPI_DGELU, block_n=bn, colsum=cs)
    x = aux.float().requires_grad_(True)
    F.gelu(x).backward(torch.ones_like(x))
    ref = (a.float() @ w.float()) * x.grad
    assert _rel(out, ref) < 1e-2
    assert _rel(cs - 2.0, ref.sum(0)) < 1e-2
    plain = torch.empty_like(out)
    nv.gemm_bf16(a, w, plain, b_major=nv.MAJOR_MN, aux=aux, epi=nv.EPI_DGELU, block_n=bn)
    assert torch.equal(plain, out)


def test_gemm_activation_epilogues_skip_padded_rows(nv):
    """row_lens: rows past a sequence's length (tile = one 128-token sequence) get zeros instead of GELU / dGELU."""
    torch.manual_seed(6)
    B, N, K = 3, 512, 256
    M = B * 128
    lens = torch.tensor([128, 70, 5], device="cuda", dtype=torch.int32)
    live = (torch.arange(128, device="cuda")[None, :] < lens[:, None]).reshape(M, 1)
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    for bn in (192, 256):
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        pre = torch.empty_like(out)
        ref_o, ref_p = torch.empty_like(out), torch.empty_like(out)
        nv.gemm_bf16(a, w, out, bias=bias, aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=bn, row_lens=lens)
        nv.gemm_bf16(a, w, ref_o, bias=bias, aux_out=ref_p, epi=nv.EPI_BIAS_GELU, block_n=bn)
        assert torch.equal(out[live.expand_as(out)], ref_o[live.expand_as(out)])
        assert torch.all(out[~live.expand_as(out)] == 0) and torch.equal(pre, ref_p)
        d = torch.empty_like(out)
        cs, cs_ref = torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
        nv.gemm_bf16(a, w, d, aux=pre, epi=nv.EPI_DGELU, block_n=bn, row_lens=lens, colsum=cs)
        nv.gemm_bf16(a, w, ref_o, aux=pre, epi=nv.EPI_DGELU, block_n=bn, colsum=cs_ref)
        assert torch.equal(d[live.expand_as(d)], ref_o[live.expand_as(d)]) and torch.all(d[~live.expand_as(d)] == 0)
        assert _rel(cs, (ref_o.float() * live).sum(0)) < 1e-2


def test_gemm_split_k_reduce_add(nv):
    torch.manual_seed(11)
    T, Nw, Kw = 2048, 768, 512      # dW[Nw, Kw] = dY[T, Nw]^T X[T, Kw]
    dy = torch.randn(T, Nw, device="cuda").bfloat16()
    x = torch.randn(T, Kw, device="cuda").bfloat16()
    ref = dy.float().t() @ x.float()
    for split in (1, 2, 5, 8):
        out = torch.zeros(Nw, Kw, device="cuda")
        nv.gemm_bf16(dy, x, out, a_major=nv.MAJOR_MN, b_major=nv.MAJOR_MN, block_n=256, split_k=split)
        assert _rel(out, ref) < 1e-3, split


@pytest.mark.parametrize("epi", ["bias", "gelu", "dgelu", "none"])
def test_gemm_splitk_workspace(nv, epi):
    """Two-pass split-K (fp32 partial tiles in a scratch workspace, then an ordered reduce + epilogue)."""
    torch.manual_seed(13)
    M, N, K = 520, 768, 1536      # M not a multiple of the 128-row tile
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    aux = torch.randn(M, N, device="cuda").bfloat16()
    ref = a.float() @ w.float().t()
    kw = {}
    if epi in ("bias", "gelu"):
        ref = ref + bias
        kw["bias"] = bias
    pre = ref
    if epi == "gelu":
        ref = F.gelu(pre)
        kw["aux_out"] = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    if epi == "dgelu":
        x = aux.float().requires_grad_(True)
        F.gelu(x).backward(torch.ones_like(x))
        ref = ref * x.grad
        kw["aux"] = aux
    code = {"none": nv.EPI_NONE, "bias": nv.EPI_BIAS, "gelu": nv.EPI_BIAS_GELU, "dgelu": nv.EPI_DGELU}[epi]
    for split in (2, 3, 4):
        ws = torch.empty(nv.gemm_splitk_workspace_bytes(M, N, split) // 4, device="cuda")
        out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
        nv.gemm_bf16_splitk(a, w, out, ws, split_k=split, epi=code, block_n=256, **kw)
        assert _rel(out, ref) < 1e-2, (epi, split)
        if epi == "gelu":
            assert _rel(kw["aux_out"], pre) < 1e-2
    # same result as the single-pass kernel up to bf16 rounding, and a too-small workspace is refused
    one = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    nv.gemm_bf16(a, w, one, block_n=256, epi=code, **{k: v for k, v in kw.items()})
    assert _rel(out, one) < 1e-2
    with pytest.raises(nv.NativeError):
        nv.gemm_bf16_splitk(a, w, out, torch.empty(1024, device="cuda"), split_k=2, epi=code, **kw)


def test_gemm_ragged_shapes_bf16_and_fp32(nv):
    torch.manual_seed(12)
    M, N, K = 200, 776, 72
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = torch.randn(N, K, device="cuda").bfloat16()
    for dt in (torch.bfloat16, torch.float32):
        out = torch.full((M, N), float("nan"), device="cuda", dtype=dt)
        nv.gemm_bf16(a, b, out, block_n=128)
        assert _rel(out, a.float() @ b.float().t()) < (1e-2 if dt == torch.bfloat16 else 1e-3)
    # strided output view (leading dimension larger than N)
    big = torch.zeros(M, 1024, device="cuda", dtype=torch.bfloat16)
    nv.gemm_bf16(a, b, big[:, 8:8 + N], block_n=256)
    assert _rel(big[:, 8:8 + N], a.float() @ b.float().t()) < 1e-2 and torch.all(big[:, :8] == 0)


def test_gemm_epilogues(nv):
    torch.manual_seed(1)
    M, N, K = 256, 512, 256
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    pre_ref = a.float() @ w.float().t() + bias
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    pre = torch.empty_like(out)
    nv.gemm_bf16(a, w, out, bias=bias, aux_out=pre, epi=nv.EPI_BIAS_GELU)
    assert _rel(pre, pre_ref) < 1e-2
    assert _rel(out, F.gelu(pre_ref)) < 1e-2
    # dgelu: D = (dY @ W) * gelu'(pre)
    dy = torch.randn(M, N, device="cuda").bfloat16()
    w2 = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()  # [N,K] row-major == B[K_red=N][N_out=K] MN-major
    pre2 = torch.randn(M, K, device="cuda").bfloat16()
    d = torch.empty(M, K, device="cuda", dtype=torch.bfloat16)
    nv.gemm_bf16(dy, w2, d, b_major=nv.MAJOR_MN, aux=pre2, epi=nv.EPI_DGELU, block_n=128)
    x = pre2.float().requires_grad_(True)
    F.gelu(x).backward(torch.ones_like(x))
    ref = (dy.float() @ w2.float()) * x.grad
    assert _rel(d, ref) < 1e-2


# ---------------------------------------------------------------- LayerNorm
def test_add_ln_fwd_bwd(nv):
    torch.manual_seed(2)
    rows, cols = 300, 768
    a = torch.randn(rows, cols, device="cuda").bfloat16()
    res = torch.randn(rows, cols, device="cuda").bfloat16()
    gamma = torch.rand(cols, device="cuda") + 0.5
    beta = torch.randn(cols, device="cuda") * 0.1
    out = torch.empty_like(a)
    rstd = torch.empty(rows, device="cuda")
    nv.add_ln_fwd(a, res, gamma, beta, out, rstd, eps=1e-12)
    z = (a.float() + res.float()).requires_grad_(True)
    g_ = gamma.clone().requires_grad_(True)
    b_ = beta.clone().requires_grad_(True)
    ref = F.layer_norm(z, (cols,), g_, b_, 1e-12)
    assert _rel(out, ref) < 1e-2
    dy1 = torch.randn(rows, cols, device="cuda").bfloat16()
    dy2 = torch.randn(rows, cols, device="cuda").bfloat16()
    ref.backward(dy1.float() + dy2.float())
    dz = torch.empty_like(a)
    dgamma = torch.zeros(cols, device="cuda")
    dbeta = torch.zeros(cols, device="cuda")
    dbias = torch.zeros(cols, device="cuda")
    nv.add_ln_bwd(dy1, dy2, out, gamma, beta, rstd, dz, None, dgamma, dbeta, dbias)
    assert _rel(dz, z.grad) < 2e-2
    assert _rel(dgamma, g_.grad) < 2e-2
    assert _rel(dbeta, b_.grad) < 1e-2
    assert _rel(dbias, z.grad.sum(0)) < 3e-2


def test_add_ln_consumes_splitk_partials(nv):
    """Split-K pass 1 + LayerNorm kernels that sum the fp32 partial tiles themselves (forward and backward)."""
    torch.manual_seed(21)
    M, N, K = 300, 768, 1536
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").bfloat16()
    gamma = torch.rand(N, device="cuda") + 0.5
    beta = torch.randn(N, device="cuda") * 0.1
    ws = torch.empty(nv.gemm_splitk_workspace_bytes(M, N, 4) // 4, device="cuda")
    for split in (2, 3, 4):
        part = nv.gemm_bf16_splitk_partials(a, w, M, N, ws, split_k=split, block_n=256, bias=bias)
        assert part.splits == split
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        rstd = torch.empty(M, device="cuda")
        nv.add_ln_fwd(part, res, gamma, beta, out, rstd)
        x = a.float() @ w.float().t() + bias + res.float()
        ref = F.layer_norm(x, (N,), gamma, beta, 1e-12)
        assert _rel(out, ref) < 1e-2, split
    # backward: dy1 = partial sums of a GEMM, compared with the bf16 path fed the same (rounded) sum
    dy2 = torch.randn(M, N, device="cuda").bfloat16()
    part = nv.gemm_bf16_splitk_partials(a, w, M, N, ws, split_k=3, block_n=256)
    dsum = (a.float() @ w.float().t())
    outs = []
    for first in (part, dsum.bfloat16()):
        dz = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        dg, db, dbias = (torch.zeros(N, device="cuda") for _ in range(3))
        nv.add_ln_bwd(first, dy2, out, gamma, beta, rstd, dz, None, dg, db, dbias)
        outs.append((dz, dg, db, dbias))
    for x, y in zip(*outs):
        assert _rel(x, y) < 2e-2


def test_add_ln_dropout_consistency(nv):
    """Dropout masks are regenerated in backward from (seed, step, stream, index): da must be zero exactly where the
    forward dropped, and the keep rate must match 1-p."""
    torch.manual_seed(3)
    rows, cols, p = 256, 768, 0.1
    a = torch.ones(rows, cols, device="cuda").bfloat16()
    gamma = torch.ones(cols, device="cuda")
    beta = torch.zeros(cols, device="cuda")
    out = torch.empty_like(a)
    rstd = torch.empty(rows, device="cuda")
    rng = _rng_state()
    nv.add_ln_fwd(a, None, gamma, beta, out, rstd, p_drop=p, rng=rng, stream_id=7)
    # with a == 1 everywhere, dropped elements are the row minimum after LN
    dropped = out.float() < 0
    rate = dropped.float().mean().item()
    assert abs(rate - p) < 0.01
    dy = torch.ones_like(a)
    dz = torch.empty_like(a)
    da = torch.empty_like(a)
    dg = torch.zeros(cols, device="cuda")
    db = torch.zeros(cols, device="cuda")
    nv.add_ln_bwd(torch.randn_like(a.float()).bfloat16(), None, out, gamma, beta, rstd, dz, da, dg, db, None, p_drop=p,
                  rng=rng, stream_id=7)
    assert torch.all(da[dropped] == 0)
    kept = ~dropped
    assert torch.allclose(da[kept].float(), dz[kept].float() / (1 - p), rtol=2e-2, atol=1e-3)
    # a different step gives a different mask
    rng2 = _rng_state(step=1)
    out2 = torch.empty_like(a)
    nv.add_ln_fwd(a, None, gamma, beta, out2, rstd, p_drop=p, rng=rng2, stream_id=7)
    assert ((out2.float() < 0) != dropped).any()


def test_embed_ln_fwd_bwd(nv):
    torch.manual_seed(4)
    B, S, cols, vocab = 4, 128, 768, 1000
    ids = torch.randint(0, vocab, (B * S,), device="cuda", dtype=torch.int32)
    word = (torch.randn(vocab, cols, device="cuda") * 0.02).requires_grad_(True)
    pos = (torch.randn(512, cols, device="cuda") * 0.02).requires_grad_(True)
    gamma = (torch.rand(cols, device="cuda") + 0.5).requires_grad_(True)
    beta = (torch.randn(cols, device="cuda") * 0.1).requires_grad_(True)
    out = torch.empty(B * S, cols, device="cuda", dtype=torch.bfloat16)
    mean = torch.empty(B * S, device="cuda")
    rstd = torch.empty(B * S, device="cuda")
    nv.embed_ln_fwd(ids, word.detach(), pos.detach(), gamma.detach(), beta.detach(), out, mean, rstd, S)
    x = word[ids.long()] + pos[torch.arange(B * S, device="cuda") % S]
    ref = F.layer_norm(x, (cols,), gamma, beta, 1e-12)
    assert _rel(out, ref) < 1e-2
    dy = torch.randn(B * S, cols, device="cuda").bfloat16()
    ref.backward(dy.float())
    dword = torch.zeros(vocab, cols, device="cuda")
    dpos = torch.zeros(512, cols, device="cuda")
    dg = torch.zeros(cols, device="cuda")
    db = torch.zeros(cols, device="cuda")
    touched = torch.zeros(vocab, device="cuda", dtype=torch.uint8)
    nv.embed_ln_bwd(dy, None, ids, word.detach(), pos.detach(), gamma.detach(), mean, rstd, dword, dpos, dg, db, S,
                    touched=touched)
    exp_touched = torch.zeros(vocab, device="cuda", dtype=torch.uint8)
    exp_touched[ids.long()] = 1
    assert torch.equal(touched, exp_touched)
    assert _rel(dword, word.grad) < 1e-2
    assert _rel(dpos, pos.grad) < 1e-2
    assert _rel(dg, gamma.grad) < 1e-2
    assert _rel(db, beta.grad) < 1e-2


def test_colsum(nv):
    x = torch.randn(2048, 2304, device="cuda").bfloat16()
    out = torch.zeros(2304, device="cuda")
    nv.colsum_bf16(x, out)
    assert _rel(out, x.float().sum(0)) < 1e-3


# ---------------------------------------------------------------- classifier head
def test_cls_head_train_and_eval(nv):
    torch.manual_seed(5)
    B, S, cols = 16, 128, 768
    h = torch.randn(B * S, cols, device="cuda").bfloat16()
    W = (torch.randn(2, cols, device="cuda") * 0.05).requires_grad_(True)
    bias = torch.randn(2, device="cuda").requires_grad_(True)
    labels = torch.randint(0, 2, (B,), device="cuda", dtype=torch.int32)
    loss_accum = torch.zeros(2, device="cuda")
    logits = torch.empty(B, 2, device="cuda")
    dW = torch.zeros(2, cols, device="cuda")
    db = torch.zeros(2, device="cuda")
    dcls = torch.empty(B, cols, device="cuda", dtype=torch.bfloat16)
    nv.cls_head_train(h, S * cols, W.detach(), bias.detach(), labels, B, loss_accum, logits, dW, db, dcls)
    x = h.view(B, S, cols)[:, 0, :].float().requires_grad_(True)
    ref_logits = x @ W.t() + bias
    ref_loss = F.cross_entropy(ref_logits, labels.long())
    ref_loss.backward()
    assert torch.allclose(logits, ref_logits, rtol=1e-4, atol=1e-4)
    assert abs(loss_accum[0].item() - ref_loss.item()) < 1e-4 and loss_accum[1].item() == 1.0
    assert _rel(dW, W.grad) < 1e-4
    assert _rel(db, bias.grad) < 1e-4
    assert _rel(dcls, x.grad) < 1e-2
    # eval variant: confusion matrix + loss + probabilities
    la = torch.zeros(2, device="cuda")
    conf = torch.zeros(4, device="cuda", dtype=torch.int32)
    probs = torch.empty(B, device="cuda")
    preds = torch.empty(B, device="cuda", dtype=torch.int32)
    nv.cls_head_eval(h, S * cols, W.detach(), bias.detach(), labels, B, la, conf, probs, preds)
    rp = ref_logits.argmax(1)
    assert torch.equal(preds.long(), rp)
    assert torch.allclose(probs, torch.softmax(ref_logits, 1)[:, 1].detach(), atol=1e-5)
    cm = torch.zeros(2, 2, dtype=torch.int64)
    for t, p_ in zip(labels.tolist(), rp.tolist()):
        cm[t, p_] += 1
    assert conf.cpu().view(2, 2).long().equal(cm)
    assert abs(la[0].item() - ref_loss.item()) < 1e-4


# ---------------------------------------------------------------- Adam
def test_adam_matches_torch(nv):
    torch.manual_seed(6)
    n = 4096 * 3
    p0 = torch.randn(n, device="cuda")
    p = p0.clone()
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    pb = torch.zeros(n // 2, device="cuda", dtype=torch.bfloat16)
    state = _rng_state(step=0)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=2e-5)
    for step in range(5):
        g = torch.randn(n, device="cuda")
        gk = g.clone()
        ref_p.grad = g.clone()
        opt.step()
        nv.adam_step(p, gk, m, v, pb, n, n // 2, n, 0, n // 2, 2e-5, 0.9, 0.999, 1e-8, state)
        nv.step_advance(state)
        assert torch.all(gk[: n // 2] == 0) and torch.equal(gk[n // 2:], g[n // 2:])
    assert torch.allclose(p, ref_p.detach(), rtol=1e-5, atol=1e-7)
    assert torch.equal(pb, p[n // 2:].bfloat16())
    assert state[1].item() == 5


def test_adam_row_skipping_is_exact(nv):
    """Rows of an embedding-like table without a gradient so far are skipped; the result equals dense torch Adam."""
    torch.manual_seed(5)
    head, rows, cols, tailn = 1024, 300, 256, 512        # [head | table rows x cols | tail]
    n = head + rows * cols + tailn
    p0 = torch.randn(n, device="cuda")
    p, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    state = _rng_state(step=0)
    flags = torch.zeros(rows, device="cuda", dtype=torch.uint8)
    ref_p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref_p], lr=1e-3)
    for step in range(6):
        g = torch.randn(n, device="cuda")
        hit = torch.randperm(rows, device="cuda")[: 40 if step < 4 else 0]  # later steps: moments only
        mask = torch.zeros(rows, device="cuda", dtype=torch.bool)
        mask[hit] = True
        g[head: head + rows * cols].view(rows, cols)[~mask] = 0.0
        flags[hit] = 1
        ref_p.grad = g.clone()
        opt.step()
        gk = g.clone()
        ticket = torch.zeros(1, device="cuda", dtype=torch.int32)
        if step % 2:   # the kernel's last block advances the step counter itself ...
            nv.adam_step_rows(p, gk, m, v, None, n, 0, 0, 0, n, 1e-3, 0.9, 0.999, 1e-8, state, flags, head,
                              head + rows * cols, cols, ticket)
            assert int(ticket.item()) == 0
        else:          # ... or a separate launch does
            nv.adam_step_rows(p, gk, m, v, None, n, 0, 0, 0, n, 1e-3, 0.9, 0.999, 1e-8, state, flags, head,
                              head + rows * cols, cols)
            nv.step_advance(state)
        assert int(state[1].item()) == step + 1
        assert torch.all(gk == 0)
    assert torch.allclose(p, ref_p.detach(), rtol=1e-5, atol=1e-7)
    untouched = flags == 0
    assert untouched.any()
    assert torch.equal(p[head: head + rows * cols].view(rows, cols)[untouched],
                       p0[head: head + rows * cols].view(rows, cols)[untouched])


# ---------------------------------------------------------------- attention
def _attn_ref(qkv, lens, B, S, H):
    D = 64
    q, k, v = qkv.float().view(B, S, 3, H, D).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) * 0.125
    mask = torch.arange(S, device=qkv.device)[None, :] >= lens[:, None].long()
    s = s.masked_fill(mask[:, None, None, :], float("-inf"))
    p = torch.softmax(s, -1)
    o = p @ v
    return o.permute(0, 2, 1, 3).reshape(B * S, H * D)


@pytest.mark.parametrize("S", [128, 96])
def test_attention_fwd_bwd(nv, S):
    torch.manual_seed(7)
    B, H = 3, 12
    qkv = torch.randn(B * S, 3 * H * 64, device="cuda").bfloat16()
    lens = torch.tensor([S, 57, 5], device="cuda", dtype=torch.int32)
    ctx = torch.empty(B * S, H * 64, device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * 128, device="cuda")
    nv.attn_fwd(qkv, ctx, lse2, lens, B, S, H)
    q32 = qkv.float().requires_grad_(True)
    ref = _attn_ref(q32, lens, B, S, H)
    # query rows past a sample's length are padding: their context is unspecified (whole warps of them are skipped and
    # write zeros) and, as in training, they carry no gradient
    valid = (torch.arange(S, device="cuda")[None, :] < lens[:, None]).reshape(B * S, 1)
    assert _rel(ctx * valid, ref * valid) < 2e-2 and bool(torch.isfinite(ctx.float()).all())
    dctx = (torch.randn(B * S, H * 64, device="cuda") * valid).bfloat16()
    ref.backward(dctx.float())
    dqkv = torch.empty_like(qkv)
    dbias = torch.full((3 * H * 64,), 0.5, device="cuda")
    nv.attn_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, dbias=dbias)
    assert _rel(dqkv, q32.grad) < 3e-2
    # fused bias gradient = column sums of dqkv (accumulated on top of what was there)
    assert _rel(dbias - 0.5, q32.grad.sum(0)) < 2e-2
    # padded keys receive exactly zero dK/dV
    g = dqkv.view(B, S, 3, H * 64)
    assert torch.all(g[1, 57:, 1:] == 0) and torch.all(g[2, 5:, 1:] == 0)


def test_attention_dropout_statistics(nv):
    torch.manual_seed(8)
    B, S, H, p = 2, 128, 12, 0.1
    qkv = torch.zeros(B * S, 3 * H * 64, device="cuda").bfloat16()
    qkv.view(B * S, 3, H * 64)[:, 2, :] = 1.0  # V == 1, uniform attention -> ctx == sum of kept probs / (1-p)
    ctx = torch.empty(B * S, H * 64, device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * 128, device="cuda")
    rng = _rng_state()
    nv.attn_fwd(qkv, ctx, lse2, None, B, S, H, p_drop=p, rng=rng, stream_id=3)
    m = ctx.float().mean().item()
    assert abs(m - 1.0) < 0.01          # unbiased
    assert ctx.float().std().item() > 1e-3  # but actually random


# ---------------------------------------------------------------- FedAvg (single GPU local kernel)
@pytest.mark.parametrize("n_clients", [2, 3, 8])
def test_fedavg_local_matches_reference_loop(nv, n_clients):
    torch.manual_seed(9)
    count = 1 << 20
    srcs = [torch.randn(count, device="cuda") for _ in range(n_clients)]
    dst = torch.empty(count, device="cuda")
    mirror = torch.zeros(count // 2, device="cuda", dtype=torch.bfloat16)
    nv.fedavg_local([s.data_ptr() for s in srcs], dst.data_ptr(), count, dst_bf16_ptr=mirror.data_ptr(),
                    bf16_begin=count // 2, bf16_end=count)
    # the reference's arithmetic (server.py:72-76): base += others in order, then true division
    # (run on the CPU: torch's CUDA scalar division multiplies by the reciprocal, which differs by 1 ulp for N=3)
    base = srcs[0].cpu().clone()
    for s in srcs[1:]:
        base += s.cpu()
    base /= n_clients
    assert torch.equal(dst.cpu(), base)  # bit parity
    assert torch.equal(mirror.cpu(), base[count // 2:].bfloat16())


def test_fedavg_local_weighted_and_bf16(nv):
    count = 1 << 16
    srcs = [torch.randn(count, device="cuda").bfloat16() for _ in range(4)]
    dst = torch.empty(count, device="cuda", dtype=torch.bfloat16)
    w = [0.1, 0.2, 0.3, 0.4]
    nv.fedavg_local([s.data_ptr() for s in srcs], dst.data_ptr(), count, in_bf16=True, weights=w)
    ref = sum(wi * s.float() for wi, s in zip(w, srcs))
    assert _rel(dst, ref) < 1e-2


def test_raw_buffer_views_and_cast(nv):
    p = nv.p2p_alloc(4096)
    buf = nv.RawDeviceBuffer(p, 4096)
    t = buf.view(torch.float32)
    assert t.numel() == 1024 and t.data_ptr() == p and torch.all(t == 0)
    t.fill_(1.5)
    b = torch.empty(1024, device="cuda", dtype=torch.bfloat16)
    nv.cast_f32_bf16(t, b)
    assert torch.all(b.float() == 1.5)
    del t
    buf.free()


@pytest.mark.parametrize("p_ln,p_head", [(0.0, 0.0), (0.1, 0.3)])
def test_tail_head_matches_the_three_kernel_sequence(nv, p_ln, p_head):
    """Fused LayerNorm + classifier + CE (+ backward) vs add_ln_fwd -> cls_head_train -> add_ln_bwd, and vs autograd
    when dropout is off."""
    torch.manual_seed(31)
    B, C = 16, 768
    a = torch.randn(B, C, device="cuda").bfloat16()
    res = torch.randn(B, C, device="cuda").bfloat16()
    gamma = torch.rand(C, device="cuda") + 0.5
    beta = torch.randn(C, device="cuda") * 0.1
    W = torch.randn(2, C, device="cuda") * 0.05
    cb = torch.randn(2, device="cuda") * 0.1
    labels = torch.randint(0, 2, (B,), device="cuda", dtype=torch.int32)
    rng = _rng_state(step=3)

    def outs():
        return dict(loss=torch.zeros(2, device="cuda"), logits=torch.zeros(B, 2, device="cuda"),
                    dW=torch.zeros(2, C, device="cuda"), db=torch.zeros(2, device="cuda"),
                    dz=torch.zeros(B, C, device="cuda", dtype=torch.bfloat16),
                    da=torch.zeros(B, C, device="cuda", dtype=torch.bfloat16),
                    dg=torch.zeros(C, device="cuda"), dbt=torch.zeros(C, device="cuda"), dbias=torch.zeros(C, device="cuda"))

    f = outs()
    nv.tail_head(a, res, gamma, beta, W, cb, labels, B, f["loss"], f["logits"], f["dW"], f["db"], f["dz"],
                 f["da"] if p_ln > 0 else None, f["dg"], f["dbt"], f["dbias"], cols=C, p_ln=p_ln, ln_stream=7,
                 p_head=p_head, head_stream=9, rng=rng)
    u = outs()
    y = torch.empty(B, C, device="cuda", dtype=torch.bfloat16)
    rstd = torch.empty(B, device="cuda")
    dcls = torch.empty(B, C, device="cuda", dtype=torch.bfloat16)
    nv.add_ln_fwd(a, res, gamma, beta, y, rstd, p_drop=p_ln, rng=rng, stream_id=7)
    nv.cls_head_train(y, C, W, cb, labels, B, u["loss"], u["logits"], u["dW"], u["db"], dcls, cols=C, p_drop=p_head,
                      rng=rng, stream_id=9)
    nv.add_ln_bwd(dcls, None, y, gamma, beta, rstd, u["dz"], u["da"] if p_ln > 0 else None, u["dg"], u["dbt"],
                  u["dbias"], p_drop=p_ln, rng=rng, stream_id=7)
    torch.cuda.synchronize()
    assert abs(f["loss"][0].item() - u["loss"][0].item()) < 5e-3 and f["loss"][1].item() == 1.0
    # the unfused path rounds y and d(cls) to bf16 in between: compare with bf16-level tolerances
    for k, tol in (("logits", 2e-2), ("dW", 3e-2), ("db", 1e-2), ("dz", 5e-2), ("dg", 5e-2), ("dbt", 5e-2), ("dbias", 5e-2)):
        assert _rel(f[k], u[k]) < tol, (k, _rel(f[k], u[k]))
    if p_ln > 0:
        assert _rel(f["da"], u["da"]) < 5e-2
        assert ((f["da"].float() == 0) == (u["da"].float() == 0)).float().mean() > 0.999   # same dropout mask
    else:
        x = (a.float() + res.float()).requires_grad_(True)
        g_, b_, W_, cb_ = (t.clone().requires_grad_(True) for t in (gamma, beta, W, cb))
        loss = F.cross_entropy(F.layer_norm(x, (C,), g_, b_, 1e-12) @ W_.t() + cb_, labels.long())
        loss.backward()
        assert abs(f["loss"][0].item() - loss.item()) < 1e-4
        for got, want in ((f["dW"], W_.grad), (f["db"], cb_.grad), (f["dg"], g_.grad), (f["dbt"], b_.grad),
                          (f["dz"], x.grad), (f["dbias"], x.grad.sum(0))):
            assert _rel(got, want) < 1e-2
