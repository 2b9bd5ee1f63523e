"""GPU tests of the B200 step engine against the pure-PyTorch DDoSClassifier (same state_dict)."""
from __future__ import annotations

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"


def _mods():
    import importlib

    return (importlib.import_module(f"{PKG}.models.engine"), importlib.import_module(f"{PKG}.models.arena"),
            importlib.import_module(f"{PKG}.models.distilbert"))


def _small_dims(arena_mod, dropout=0.0):
    return arena_mod.ModelDims(vocab_size=1000, max_position=128, dim=256, n_layers=2, n_heads=4, hidden_dim=512,
                               dropout=dropout, attention_dropout=dropout, head_dropout=dropout)


def _batch(B, S, vocab, device="cuda", seed=0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, vocab, (B, S), generator=g)
    lens = torch.randint(S // 4, S + 1, (B,), generator=g)
    lens[0] = S
    labels = torch.randint(0, 2, (B,), generator=g)
    mask = (torch.arange(S)[None, :] < lens[:, None]).long()
    return ids.to(device), lens.to(device), labels.to(device), mask.to(device)


def test_eval_logits_match_torch_module():
    eng_mod, arena_mod, db = _mods()
    dims = _small_dims(arena_mod)
    B, S = 8, 128
    eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=3)
    ref = db.DDoSClassifier(dims).cuda().eval()
    ref.load_state_dict(eng.state_dict())
    ids, lens, labels, mask = _batch(B, S, dims.vocab_size)
    eng.set_batch(ids.int(), lens.int(), labels.int())
    out = eng.logits_eval()
    with torch.no_grad():
        exp = ref(ids, mask)
    assert torch.allclose(out, exp, atol=3e-2, rtol=3e-2), (out - exp).abs().max()
    # partial batch
    out5 = eng.logits_eval(5)
    assert torch.allclose(out5, exp[:5], atol=3e-2, rtol=3e-2)


def test_gradients_match_autograd_without_dropout():
    eng_mod, arena_mod, db = _mods()
    from importlib import import_module
    nv = import_module(f"{PKG}.ops.native")
    dims = _small_dims(arena_mod)
    B, S = 8, 128
    eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=4)
    ref = db.DDoSClassifier(dims).cuda().train()
    ref.load_state_dict(eng.state_dict())
    ids, lens, labels, mask = _batch(B, S, dims.vocab_size, seed=1)
    eng.set_batch(ids.int(), lens.int(), labels.int())
    h, stride = eng._forward(B, train=True)
    nv.cls_head_train(h, stride, eng.cw, eng.cb, eng.labels, B, eng.loss_accum, eng.logits, eng.d_cw, eng.d_cb,
                      eng.dcls, cols=dims.dim)
    eng._backward()
    torch.cuda.synchronize()
    loss = F.cross_entropy(ref(ids, mask), labels)
    loss.backward()
    assert abs(eng.loss_accum[0].item() - loss.item()) < 2e-2
    sd_grads = {k: p.grad for k, p in ref.named_parameters()}
    bad = []
    for k, _ in eng.arena.layout.keys:
        g = eng.arena.view(k, eng.arena.grad)
        e = sd_grads[k]
        # k_lin.bias has an exactly-zero true gradient (softmax is invariant to a per-query shift of the scores),
        # so errors are measured against the gradient scale of tensors of the same kind, not against ~0
        kind_scale = max(e.norm().item(), 1e-3 * max(v.norm().item() for kk, v in sd_grads.items()
                                                       if kk.endswith(".bias") == k.endswith(".bias")))
        rel = ((g - e).norm() / kind_scale).item()
        if rel > 8e-2:
            bad.append((k, rel))
    assert not bad, bad


def test_training_reduces_loss_and_graph_matches_eager():
    eng_mod, arena_mod, db = _mods()
    dims = _small_dims(arena_mod, dropout=0.1)
    B, S = 16, 128
    eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=5, lr=1e-3)
    ids, lens, labels, _ = _batch(B, S, dims.vocab_size, seed=2)
    labels = (ids[:, 1] % 2).to(labels.dtype)  # learnable rule
    eng.set_batch(ids.int(), lens.int(), labels.int())
    eng.capture()
    assert eng.launches_per_train_step > 40
    losses = []
    for it in range(30):
        eng.train_step()
        if it % 10 == 9:
            losses.append(eng.read_loss()[0])
    assert losses[-1] < losses[0], losses
    assert int(eng.rng[1].item()) == 30
    # end-to-end host path returns a finite running loss and advances the step counter
    v = eng.train_step_from_host(ids.int().cpu(), lens.int().cpu(), labels.int().cpu())
    assert v == v and int(eng.rng[1].item()) == 31
    # evaluation accumulators
    eng.eval_reset()
    eng.eval_batch()
    eng.eval_batch(7)
    loss_sum, nb, cm = eng.eval_read()
    assert nb == 2 and int(cm.sum()) == B + 7


def test_full_size_model_step_runs():
    eng_mod, arena_mod, db = _mods()
    eng = eng_mod.B200Engine(arena_mod.ModelDims(), batch=16, seq=128, init_seed=0)
    ids, lens, labels, _ = _batch(16, 128, 30522, seed=3)
    eng.set_batch(ids.int(), lens.int(), labels.int())
    eng.capture()
    for _ in range(3):
        eng.train_step()
    loss, n = eng.read_loss()
    assert n == 3 and 0.1 < loss < 3.0
    sd = eng.state_dict()
    assert len(sd) == 102 and sum(v.numel() for v in sd.values()) == 66_364_418


def test_pdl_overlap_does_not_change_gradients():
    """Programmatic dependent launch only moves prologues ahead of the predecessor's tail: the gradients of the
    full-size model must agree with the fully serialised launch order (differences: atomic/reduce-add order only)."""
    eng_mod, arena_mod, db = _mods()
    from importlib import import_module
    nv = import_module(f"{PKG}.ops.native")
    dims = arena_mod.ModelDims()
    B, S = 16, 128
    eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=7)
    ids, lens, labels, _ = _batch(B, S, dims.vocab_size, seed=5)
    eng.set_batch(ids.int(), lens.int(), labels.int())

    def grads():
        eng.arena.grad.zero_()
        eng.loss_accum.zero_()
        h, stride = eng._forward(B, train=True)
        nv.cls_head_train(h, stride, eng.cw, eng.cb, eng.labels, B, eng.loss_accum, eng.logits, eng.d_cw, eng.d_cb,
                          eng.dcls, cols=dims.dim, p_drop=dims.head_dropout, rng=eng.rng)
        eng._backward()
        torch.cuda.synchronize()
        return eng.arena.grad.clone(), float(eng.loss_accum[0].item())

    prev = nv.set_pdl(False)
    try:
        g_off, l_off = grads()
        nv.set_pdl(True)
        outs = [grads() for _ in range(3)]
    finally:
        nv.set_pdl(prev)
    for g_on, l_on in outs:
        assert abs(l_on - l_off) < 1e-5
        err = (g_on - g_off).abs().max().item()
        assert err <= 1e-3 * g_off.abs().max().item() + 1e-6, err


def test_training_state_checkpoint_resumes_b200_backend(tmp_path):
    """Model + Adam moments + touched-row flags + dropout counter restored: the resumed run tracks the uninterrupted
    one (differences: fp32 atomic / reduce-add order only), while a model-only restore (the reference's .pth) does not."""
    from importlib import import_module
    tr = import_module(f"{PKG}.train.trainer")
    ck = import_module(f"{PKG}.utils.checkpoint")
    pack = import_module(f"{PKG}.data.pack")
    eng_mod, arena_mod, db = _mods()
    dims = _small_dims(arena_mod, dropout=0.1)
    B, S, n = 8, 128, 32
    g = torch.Generator().manual_seed(11)
    split = pack.PackedSplit(torch.randint(1, dims.vocab_size, (n, S), generator=g, dtype=torch.int32),
                             torch.randint(S // 4, S + 1, (n,), generator=g, dtype=torch.int32),
                             torch.randint(0, 2, (n,), generator=g, dtype=torch.int32)).to("cuda")

    def backend(seed):
        return tr.B200Backend(dims, batch=B, seq=S, lr=1e-3, seed=seed, init_seed=seed)

    a = backend(3)
    gen = torch.Generator().manual_seed(5)
    a.train_epoch(split, gen)
    path = ck.save_training_state(str(tmp_path / "s.pt"), a, round_idx=1, epoch=1, shuffle_generator=gen)
    a.train_epoch(split, gen)
    want = a.engine.arena.flat.clone()

    b = backend(9)
    gen_b = torch.Generator()
    prog = ck.load_training_state(path, b, shuffle_generator=gen_b)
    assert prog["round"] == 1 and int(b.engine.rng[1].item()) == n // B
    b.train_epoch(split, gen_b)
    got = b.engine.arena.flat
    # mean, not max: where a gradient is ~0 the early Adam update is +-lr by the sign of rounding noise
    err_resume = (got - want).abs().mean().item()

    c = backend(3)                                        # model only, fresh Adam state and dropout counter
    c.load_state_dict(torch.load(path, weights_only=False)["model"])
    gen_c = torch.Generator()
    gen_c.set_state(torch.load(path, weights_only=False)["progress"]["shuffle_generator"])
    c.train_epoch(split, gen_c)
    err_model_only = (c.engine.arena.flat - want).abs().mean().item()
    assert err_resume < 2e-5 and err_model_only > 10 * max(err_resume, 1e-7), (err_resume, err_model_only)
