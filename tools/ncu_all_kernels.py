"""Launches EVERY kernel of the library once or twice at its production shape on ONE GPU — the target of
`ncu --set full` (tools/_run_ncu.sh).  Two eager full-size training steps (batch 16, seq 128, DistilBERT-base: GEMMs
of every instantiation the plan uses, attention, LayerNorm, embedding, head, Adam), an evaluation step, the
workspace split-K GEMM with its reduce kernel, the long-sequence attention fallback, and the multi-GPU kernels in
their one-rank form (the cross-GPU handshake then runs against this GPU's own signal pad — kernels that wait for
ANOTHER rank must not run under ncu, B200_PROFILING.md): fedavg_local over two arenas, fedavg_p2p two-shot / one-shot
with one participant, xgpu_barrier, adam_dp, step_advance.  fedavg_nvls needs a multicast object spanning several GPUs
and is profiled from the multi-GPU bench timings instead."""
import importlib
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
synth = importlib.import_module(f"{PKG}.data.synth")
nv = importlib.import_module(f"{PKG}.ops.native")
p2p = importlib.import_module(f"{PKG}.parallel.p2p")

dims = arena_mod.ModelDims()
B, S = 16, 128
eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=0, overlap_wgrad=False)
ids, lens, labels = synth.synth_token_batch(B, S, dims.vocab_size, seed=1)
eng.set_batch(ids.cuda(), lens.cuda(), labels.cuda())
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    eng._train_step_kernels()
eng.eval_reset()
eng._eval_kernels(B, True)
torch.cuda.synchronize()

# split-K through the workspace + reduce kernel (the no-allocation entry shares this path)
T, D, F = B * S, dims.dim, dims.hidden_dim
a = torch.randn(T, F, device="cuda").bfloat16()
w = (torch.randn(D, F, device="cuda") * 0.05).bfloat16()
out = torch.empty(T, D, device="cuda", dtype=torch.bfloat16)
ws = torch.empty(nv.gemm_splitk_workspace_bytes(T, D, 3) // 4, device="cuda")
nv.gemm_bf16_splitk(a, w, out, ws, split_k=3, bias=torch.zeros(D, device="cuda"), epi=nv.EPI_BIAS, block_n=256)
nv.colsum_bf16(out, torch.zeros(D, device="cuda"))

# per-warp epilogue and the 128 x 384 tile (options; see profiles/gemm_round2)
x = torch.randn(T, D, device="cuda").bfloat16()
w1 = (torch.randn(F, D, device="cuda") * 0.05).bfloat16()
h = torch.empty(T, F, device="cuda", dtype=torch.bfloat16)
pre = torch.empty_like(h)
nv.gemm_bf16(x, w1, h, bias=torch.zeros(F, device="cuda"), aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=384)
prev = nv.set_gemm_epilogue(1)
nv.gemm_bf16(x, w1, h, bias=torch.zeros(F, device="cuda"), aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=256)
nv.set_gemm_epilogue(prev)

# long-sequence attention fallback (S = 256)
Sg, H = 256, dims.n_heads
qkv = torch.randn(4 * Sg, 3 * D, device="cuda").bfloat16()
ctx = torch.empty(4 * Sg, D, device="cuda", dtype=torch.bfloat16)
lse2 = torch.empty(4 * H * Sg, device="cuda")
lg = torch.tensor([256, 200, 150, 90], device="cuda", dtype=torch.int32)
nv.attn_fwd(qkv, ctx, lse2, lg, 4, Sg, H)
nv.attn_bwd(qkv, ctx, torch.randn_like(ctx), lse2, torch.empty_like(qkv), lg, 4, Sg, H)

# FedAvg kernels, one-rank forms, full 66M-element arena
lay = eng.arena.layout
other = eng.arena.flat.clone()
dst = torch.empty_like(other)
mirror = torch.empty(lay.gemm_end, device="cuda", dtype=torch.bfloat16)
nv.fedavg_local([eng.arena.flat.data_ptr(), other.data_ptr()], dst.data_ptr(), lay.total, dst_bf16_ptr=mirror.data_ptr(),
                bf16_begin=0, bf16_end=lay.gemm_end)
pad = torch.zeros(64, device="cuda", dtype=torch.int32)
nv.fedavg_p2p([other.data_ptr()], [pad.data_ptr()], 0, lay.total, 1, mirror_ptrs=[mirror.data_ptr()], mode=p2p.TWO_SHOT,
              bf16_begin=0, bf16_end=lay.gemm_end)
nv.fedavg_p2p([other.data_ptr()], [pad.data_ptr()], 0, lay.total, 2, out_ptrs=[dst.data_ptr()],
              mirror_ptrs=[mirror.data_ptr()], mode=p2p.ONE_SHOT, bf16_begin=0, bf16_end=lay.gemm_end)
xb = other.bfloat16()
nv.fedavg_p2p([xb.data_ptr()], [pad.data_ptr()], 0, lay.total, 3, in_bf16=True, mode=p2p.TWO_SHOT)

# data-parallel optimizer kernels with a group of one
pad2 = torch.zeros(64, device="cuda", dtype=torch.int32)
nv.xgpu_barrier([pad2.data_ptr()], 0, eng.rng, 0)
ar = eng.arena
nv.adam_step_dp(ar.flat, ar.grad, ar.exp_avg, ar.exp_avg_sq, 0, lay.gemm_end, 0, lay.gemm_end, 0, 0, 2e-5, 0.9, 0.999, 1e-8,
                1.0, eng.rng, None, [ar.bf16.data_ptr()], 0)
nv.adam_step_dp(ar.flat, ar.grad, ar.exp_avg, ar.exp_avg_sq, lay.gemm_end, lay.total, 0, 0, lay.gemm_end, lay.total, 2e-5,
                0.9, 0.999, 1e-8, 1.0, eng.rng, [ar.flat.data_ptr()], None, 0)
nv.step_advance(eng.rng)
nv.cast_f32_bf16(ar.flat[: lay.gemm_end], ar.bf16)
torch.cuda.synchronize()
assert int(pad[33].item()) == 0 and int(pad2[33].item()) == 0
print("ok", eng.launches_per_train_step)
