"""GPU tests at the REAL size (DistilBERT-base: 768 / 6 layers / 12 heads, batch 16, seq 128) against Hugging Face's
``DistilBertModel`` in fp32 — the model the reference trains (client1.py:53-65) — plus a kernel-level check of the
attention backward with dropout on.  Collected after the other GPU tests (file name)."""
from __future__ import annotations

import importlib

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"


def _hf_classifier(sd, dims):
    """HF DistilBertModel (fp32, eager attention, all dropouts 0) + Linear head carrying the engine's weights."""
    from transformers import DistilBertConfig, DistilBertModel

    cfg = DistilBertConfig(vocab_size=dims.vocab_size, max_position_embeddings=dims.max_position, dim=dims.dim,
                           n_layers=dims.n_layers, n_heads=dims.n_heads, hidden_dim=dims.hidden_dim, dropout=0.0,
                           attention_dropout=0.0)
    try:
        enc = DistilBertModel(cfg, attn_implementation="eager")
    except TypeError:
        enc = DistilBertModel(cfg)
    enc.load_state_dict({k[len("distilbert."):]: v for k, v in sd.items() if k.startswith("distilbert.")})
    head = torch.nn.Linear(dims.dim, dims.n_classes)
    head.load_state_dict({"weight": sd["classifier.weight"], "bias": sd["classifier.bias"]})
    return enc.cuda().train(), head.cuda().train()


def fullsize_parity_report(seed: int = 0):
    """(logit errors, per-tensor gradient errors) of the engine against HF fp32 on one full-size batch."""
    eng_mod = importlib.import_module(f"{PKG}.models.engine")
    arena_mod = importlib.import_module(f"{PKG}.models.arena")
    nv = importlib.import_module(f"{PKG}.ops.native")
    synth = importlib.import_module(f"{PKG}.data.synth")
    import dataclasses

    dims = dataclasses.replace(arena_mod.ModelDims(), dropout=0.0, attention_dropout=0.0, head_dropout=0.0)
    B, S = 16, 128
    eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=seed)
    ids, lens, labels = synth.synth_token_batch(B, S, dims.vocab_size, seed=seed)
    ids, lens, labels = ids.cuda(), lens.cuda(), labels.cuda()
    eng.set_batch(ids, lens, labels)
    sd = eng.state_dict()
    enc, head = _hf_classifier(sd, dims)
    mask = (torch.arange(S, device="cuda")[None, :] < lens[:, None]).long()
    torch.backends.cuda.matmul.allow_tf32 = False
    logits_ref = head(enc(input_ids=ids.long(), attention_mask=mask)[0][:, 0])
    loss_ref = F.cross_entropy(logits_ref, labels.long())
    loss_ref.backward()
    # engine: forward + head + backward (no optimizer step), every kernel hand-written
    h, stride = eng._forward(B, train=True)
    nv.cls_head_train(h, stride, eng.cw, eng.cb, eng.labels, B, eng.loss_accum, eng.logits, eng.d_cw, eng.d_cb,
                      eng.dcls, cols=dims.dim)
    eng._backward()
    torch.cuda.synchronize()
    ref_grads = {"distilbert." + k: p.grad for k, p in enc.named_parameters()}
    ref_grads["classifier.weight"], ref_grads["classifier.bias"] = head.weight.grad, head.bias.grad
    rows = []
    for k, _ in eng.arena.layout.keys:
        g, e = eng.arena.view(k, eng.arena.grad), ref_grads[k]
        nz = e[e != 0]
        rms = nz.pow(2).mean().sqrt().item() if nz.numel() else 0.0     # (embedding rows of unused tokens are 0)
        rows.append((k, ((g - e).norm() / e.norm().clamp_min(1e-30)).item(), e.norm().item(),
                     ((g - e).abs().max().item() / max(rms, 1e-30)), (g - e).norm().item(),
                     (g - e).abs().max().item() / max(e.abs().max().item(), 1e-30)))
    return {"loss": (float(eng.loss_accum[0].item()), float(loss_ref.item())),
            "logits": (eng.logits.clone(), logits_ref.detach()), "grads": rows}


def test_full_size_logits_and_all_gradients_match_hf_fp32():
    """bf16 tensor-core compute with fp32 accumulation against HF in strict fp32: the loss and the logits agree to bf16
    precision and EVERY one of the 102 gradient tensors agrees in relative norm; tensors whose true gradient is
    (numerically) zero -- k_lin.bias: softmax is invariant to a per-query shift -- are held to the scale of their kind."""
    rep = fullsize_parity_report()
    loss, loss_ref = rep["loss"]
    assert abs(loss - loss_ref) < 5e-3 * max(1.0, abs(loss_ref)), (loss, loss_ref)
    out, exp = rep["logits"]
    assert torch.allclose(out, exp, atol=2e-2, rtol=2e-2), (out - exp).abs().max().item()
    import statistics

    kind = {b: statistics.median(r[2] for r in rep["grads"] if r[0].endswith(".bias") == b) for b in (True, False)}
    bad = []
    for k, rel, norm, max_over_rms, err_norm, max_over_max in rep["grads"]:
        # norm-wise: 3 % of the tensor's own gradient, plus 0.1 % of the typical gradient norm of its kind (the true
        # k_lin.bias gradient is ~1e-9: only the second term can hold it); measured: <= 1.9 % on every tensor
        ok_norm = err_norm <= 3e-2 * norm + 1e-3 * kind[k.endswith(".bias")]
        # element-wise: no single element is off by more than 5 % of the tensor's largest element
        ok_elem = norm < 1e-3 * kind[k.endswith(".bias")] or max_over_max <= 5e-2
        if not (ok_norm and ok_elem):
            bad.append((k, round(rel, 4), round(max_over_max, 4)))
    assert not bad, bad


def _rng_state(seed=1234, step=0):
    return torch.tensor([seed, step], dtype=torch.int64, device="cuda")


def test_attention_backward_with_dropout_matches_autograd():
    """attn_bwd regenerates the dropout mask of attn_fwd from (seed, step, stream, element).  The mask the kernels use
    is recovered from two forwards whose V selects keys (V = [I; 0] and [0; I]), then autograd differentiates
    softmax -> mask / (1 - p) -> @V with exactly that mask."""
    nv = importlib.import_module(f"{PKG}.ops.native")
    torch.manual_seed(17)
    B, S, H, D, p = 2, 128, 12, 64, 0.1
    lens = torch.tensor([S, 77], device="cuda", dtype=torch.int32)
    qkv = torch.randn(B * S, 3 * H * D, device="cuda").bfloat16()
    rng = _rng_state()
    lse2 = torch.empty(B * H * 128, device="cuda")

    def fwd(x, pd):
        ctx = torch.empty(B * S, H * D, device="cuda", dtype=torch.bfloat16)
        nv.attn_fwd(x, ctx, lse2, lens, B, S, H, p_drop=pd, rng=rng, stream_id=5)
        return ctx

    # recover P_dropped[b, h, q, key] column block by column block
    probe = qkv.clone().view(B, S, 3, H, D)
    eye = torch.eye(D, device="cuda", dtype=torch.bfloat16)
    pdrop = torch.zeros(B, H, S, S, device="cuda")
    pfull = torch.zeros(B, H, S, S, device="cuda")
    for half in range(2):
        probe[:, :, 2] = 0
        probe[:, half * D:(half + 1) * D, 2] = eye[None, :, None, :]          # V[key, h, :] = e_(key - 64*half)
        x = probe.view(B * S, 3 * H * D)
        pdrop[..., half * D:(half + 1) * D] = fwd(x, p).float().view(B, S, H, D).permute(0, 2, 1, 3)
        pfull[..., half * D:(half + 1) * D] = fwd(x, 0.0).float().view(B, S, H, D).permute(0, 2, 1, 3)
    keep = (pdrop != 0) | (pfull == 0)
    valid_k = (torch.arange(S, device="cuda")[None, :] < lens[:, None])
    frac = 1.0 - keep[:, :, :, :][valid_k[:, None, None, :].expand_as(keep)].float().mean().item()
    assert 0.08 < frac < 0.12, frac                                            # ~10 % of the real keys are dropped
    # the kept probabilities are scaled by exactly 1 / (1 - p)
    both = (pdrop != 0) & (pfull > 1e-3)
    assert torch.allclose(pdrop[both], pfull[both] / (1 - p), rtol=3e-2, atol=1e-3)

    ctx = fwd(qkv, p)
    valid_q = valid_k.reshape(B * S, 1)
    dctx = (torch.randn(B * S, H * D, device="cuda") * valid_q).bfloat16()
    dqkv = torch.empty_like(qkv)
    nv.attn_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, p_drop=p, rng=rng, stream_id=5)

    q32 = qkv.float().requires_grad_(True)
    q, k, v = q32.view(B, S, 3, H, D).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) * 0.125
    s = s.masked_fill(~valid_k[:, None, None, :], float("-inf"))
    pd = torch.softmax(s, -1) * keep.float() / (1 - p)
    ref = (pd @ v).permute(0, 2, 1, 3).reshape(B * S, H * D)
    assert ((ctx.float() - ref) * valid_q).norm() / (ref * valid_q).norm() < 2e-2
    ref.backward(dctx.float())
    err = (dqkv.float() - q32.grad).norm() / q32.grad.norm()
    assert err < 3e-2, err.item()
    assert torch.allclose(dqkv.float(), q32.grad, rtol=0.1, atol=0.05 * q32.grad.pow(2).mean().sqrt().item() * 4)


def test_layernorm_backward_with_small_gamma_uses_saved_xhat():
    """Pretrained encoders have LayerNorm channels with small gamma and |beta| ~ 1.  Rebuilding x_hat from the bf16
    output, (y - beta) / gamma, then amplifies y's rounding by |beta| / |gamma|; with the x_hat the forward saves
    (what the training engine does) the backward stays at bf16 accuracy for every channel."""
    nv = importlib.import_module(f"{PKG}.ops.native")
    torch.manual_seed(3)
    rows, cols = 512, 768
    a = torch.randn(rows, cols, device="cuda").bfloat16()
    res = torch.randn(rows, cols, device="cuda").bfloat16()
    gamma = torch.exp(torch.empty(cols, device="cuda").uniform_(-6.9, 0.4))      # 1e-3 .. 1.5, log-uniform
    beta = torch.randn(cols, device="cuda")
    dy = torch.randn(rows, cols, device="cuda").bfloat16()
    x = (a.float() + res.float()).requires_grad_(True)
    g32, b32 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    F.layer_norm(x, (cols,), g32, b32, 1e-12).backward(dy.float())

    def run(save):
        y = torch.empty_like(a)
        xh = torch.empty_like(a) if save else None
        rstd = torch.empty(rows, device="cuda")
        nv.add_ln_fwd(a, res, gamma, beta, y, rstd, xhat_out=xh)
        dz = torch.empty_like(a)
        dgamma, dbeta, dbias = (torch.zeros(cols, device="cuda") for _ in range(3))
        nv.add_ln_bwd(dy, None, xh if save else y, gamma, None if save else beta, rstd, dz, None, dgamma, dbeta, dbias)
        torch.cuda.synchronize()
        return dz.float(), dgamma, dbeta

    dz, dgamma, dbeta = run(save=True)
    assert (dz - x.grad).norm() / x.grad.norm() < 1e-2
    assert torch.allclose(dbeta, b32.grad, rtol=1e-3, atol=1e-2)
    # d gamma_c = sum_rows dy * x_hat: accurate for EVERY channel, the tiny-gamma ones included
    scale = g32.grad.abs().mean().item()
    assert torch.allclose(dgamma, g32.grad, rtol=2e-2, atol=2e-2 * scale), (dgamma - g32.grad).abs().max().item()
    # the rebuild is visibly worse on the small-gamma channels (documented limitation of the cheaper path)
    _, dgamma_rebuilt, _ = run(save=False)
    small = gamma < 0.02
    err_saved = (dgamma - g32.grad)[small].abs().max().item()
    err_rebuilt = (dgamma_rebuilt - g32.grad)[small].abs().max().item()
    assert err_rebuilt > 5 * err_saved, (err_saved, err_rebuilt)


def _attn_ref_generic(qkv, lens, B, S, H, keep=None, p=0.0):
    D = 64
    q, k, v = qkv.view(B, S, 3, H, D).permute(2, 0, 3, 1, 4)
    s = (q @ k.transpose(-1, -2)) * 0.125
    valid = torch.arange(S, device=qkv.device)[None, :] < lens[:, None].long()
    s = s.masked_fill(~valid[:, None, None, :], float("-inf"))
    pr = torch.softmax(s, -1)
    if keep is not None:
        pr = pr * keep / (1 - p)
    return (pr @ v).permute(0, 2, 1, 3).reshape(B * S, H * D), valid


@pytest.mark.parametrize("S", [136, 256, 512])
def test_long_sequence_attention_fallback(S):
    """128 < S <= 512 (the reference's max_len is a parameter, client1.py:27, and DistilBERT has 512 positions): the
    attention entry points fall back to the CUDA-core kernels; forward and backward against PyTorch autograd."""
    nv = importlib.import_module(f"{PKG}.ops.native")
    torch.manual_seed(S)
    B, H = 2, 4
    lens = torch.tensor([S, S // 2 + 3], device="cuda", dtype=torch.int32)
    qkv = torch.randn(B * S, 3 * H * 64, device="cuda").bfloat16()
    ctx = torch.full((B * S, H * 64), float("nan"), device="cuda", dtype=torch.bfloat16)
    lse2 = torch.empty(B * H * S, device="cuda")
    nv.attn_fwd(qkv, ctx, lse2, lens, B, S, H)
    q32 = qkv.float().requires_grad_(True)
    ref, valid = _attn_ref_generic(q32, lens, B, S, H)
    vq = valid.reshape(B * S, 1)
    assert bool(torch.isfinite(ctx.float()).all())
    assert torch.all(ctx.float()[~vq.expand_as(ctx)] == 0)                        # padded query rows are zeros
    assert ((ctx.float() - ref) * vq).norm() / (ref * vq).norm() < 1e-2
    dctx = (torch.randn(B * S, H * 64, device="cuda") * vq).bfloat16()
    ref.backward(dctx.float())
    dqkv = torch.full_like(qkv, float("nan"))
    dbias = torch.full((3 * H * 64,), 0.5, device="cuda")
    nv.attn_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, dbias=dbias)
    assert bool(torch.isfinite(dqkv.float()).all())
    assert (dqkv.float() - q32.grad).norm() / q32.grad.norm() < 2e-2
    assert torch.allclose(dbias - 0.5, dqkv.float().sum(0), rtol=1e-2, atol=0.05)
    g = dqkv.view(B, S, 3, H * 64)
    assert torch.all(g[1, S // 2 + 3:] == 0)                                      # nothing flows into padding


def test_long_sequence_attention_fallback_dropout_and_engine():
    """Dropout in the fallback: the backward kernels regenerate the forward's mask (recovered here with V = one-hot
    probes); and the engine trains at seq 256 end to end."""
    nv = importlib.import_module(f"{PKG}.ops.native")
    torch.manual_seed(5)
    B, S, H, D, p = 1, 192, 2, 64, 0.1
    lens = torch.tensor([S], device="cuda", dtype=torch.int32)
    qkv = torch.randn(B * S, 3 * H * D, device="cuda").bfloat16()
    rng = _rng_state()
    lse2 = torch.empty(B * H * S, device="cuda")

    def fwd(x, pd):
        ctx = torch.empty(B * S, H * D, device="cuda", dtype=torch.bfloat16)
        nv.attn_fwd(x, ctx, lse2, lens, B, S, H, p_drop=pd, rng=rng, stream_id=7)
        return ctx

    probe = qkv.clone().view(B, S, 3, H, D)
    eye = torch.eye(D, device="cuda", dtype=torch.bfloat16)
    pdrop = torch.zeros(B, H, S, S, device="cuda")
    for blk in range(S // D):
        probe[:, :, 2] = 0
        probe[:, blk * D:(blk + 1) * D, 2] = eye[None, :, None, :]
        pdrop[..., blk * D:(blk + 1) * D] = fwd(probe.view(B * S, 3 * H * D), p).float().view(B, S, H, D).permute(0, 2, 1, 3)
    keep = (pdrop != 0).float()
    assert 0.08 < 1.0 - keep.mean().item() < 0.12
    ctx = fwd(qkv, p)
    q32 = qkv.float().requires_grad_(True)
    ref, _ = _attn_ref_generic(q32, lens, B, S, H, keep=keep, p=p)
    assert (ctx.float() - ref).norm() / ref.norm() < 1e-2
    dctx = torch.randn(B * S, H * D, device="cuda").bfloat16()
    ref.backward(dctx.float())
    dqkv = torch.empty_like(qkv)
    nv.attn_bwd(qkv, ctx, dctx, lse2, dqkv, lens, B, S, H, p_drop=p, rng=rng, stream_id=7)
    assert (dqkv.float() - q32.grad).norm() / q32.grad.norm() < 2e-2

    # engine at seq 256: loss goes down on a learnable rule, eval logits match the PyTorch module
    eng_mod = importlib.import_module(f"{PKG}.models.engine")
    arena_mod = importlib.import_module(f"{PKG}.models.arena")
    db = importlib.import_module(f"{PKG}.models.distilbert")
    dims = arena_mod.ModelDims(vocab_size=1000, max_position=512, dim=256, n_layers=2, n_heads=4, hidden_dim=512,
                               dropout=0.0, attention_dropout=0.0, head_dropout=0.0)
    Bq, Sq = 4, 256
    eng = eng_mod.B200Engine(dims, batch=Bq, seq=Sq, init_seed=3, lr=1e-3)
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 1000, (Bq, Sq), generator=g).cuda()
    lens_e = torch.tensor([256, 200, 131, 77]).cuda()
    labels = (ids[:, 1] % 2)
    mask = (torch.arange(Sq, device="cuda")[None, :] < lens_e[:, None]).long()
    eng.set_batch(ids.int(), lens_e.int(), labels.int())
    ref_m = db.DDoSClassifier(dims).cuda().eval()
    ref_m.load_state_dict(eng.state_dict())
    with torch.no_grad():
        exp = ref_m(ids, mask)
    assert torch.allclose(eng.logits_eval(), exp, atol=3e-2, rtol=3e-2)
    losses = []
    for it in range(30):
        eng.train_step()
        if it % 10 == 9:
            losses.append(eng.read_loss()[0])
    assert losses[-1] < losses[0], losses


def test_engine_at_the_executed_sequence_length_matches_the_padded_run():
    """The engine run at seq 96 on data whose samples are at most 88 tokens long learns what the engine run at the
    padded seq 128 learns: same evaluation logits, same parameters after training steps (dropout off; differences are
    bf16 accumulation order only)."""
    eng_mod = importlib.import_module(f"{PKG}.models.engine")
    arena_mod = importlib.import_module(f"{PKG}.models.arena")
    dims = arena_mod.ModelDims(vocab_size=1000, max_position=128, dim=256, n_layers=2, n_heads=4, hidden_dim=512,
                               dropout=0.0, attention_dropout=0.0, head_dropout=0.0)
    B = 8
    g = torch.Generator().manual_seed(2)
    lens = torch.randint(70, 89, (B,), generator=g)
    ids = torch.randint(5, 1000, (B, 128), generator=g)
    ids[torch.arange(128)[None, :] >= lens[:, None]] = 0
    labels = (ids[:, 1] % 2)
    outs = []
    for S in (128, 96):
        eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=3, lr=1e-3)
        eng.set_batch(ids[:, :S].contiguous().int().cuda(), lens.int().cuda(), labels.int().cuda())
        logits = eng.logits_eval().clone()
        p0 = eng.arena.flat.clone()
        for _ in range(3):
            eng.train_step()
        torch.cuda.synchronize()
        outs.append((logits, eng.arena.flat - p0, eng.read_loss()[0]))
    (l128, u128, loss128), (l96, u96, loss96) = outs
    assert torch.allclose(l128, l96, atol=2e-2, rtol=2e-2)
    assert abs(loss128 - loss96) < 2e-2
    moved = u128.abs() > 0.5e-3
    cos = torch.nn.functional.cosine_similarity(u128[moved], u96[moved], dim=0).item()
    assert cos > 0.98, cos
