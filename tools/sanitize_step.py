"""Small-model training steps + one call of every GEMM variant: the target of compute-sanitizer runs
(memcheck / racecheck / synccheck), which are far too slow for the full-size step."""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
import importlib  # noqa: E402

eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
nv = importlib.import_module(f"{PKG}.ops.native")

dims = arena_mod.ModelDims(vocab_size=1000, max_position=128, dim=256, n_layers=2, n_heads=4, hidden_dim=512,
                           dropout=0.1, attention_dropout=0.1, head_dropout=0.3)
B, S = 4, 128
eng = eng_mod.B200Engine(dims, batch=B, seq=S, init_seed=0)
g = torch.Generator().manual_seed(0)
ids = torch.randint(1, 1000, (B, S), generator=g, dtype=torch.int32).cuda()
lens = torch.tensor([128, 90, 64, 17], dtype=torch.int32).cuda()
labels = torch.randint(0, 2, (B,), generator=g, dtype=torch.int32).cuda()
eng.set_batch(ids, lens, labels)
for _ in range(2):
    eng._train_step_kernels()          # eager: every kernel of the step incl. the side-stream branch
eng.eval_reset()
eng.eval_batch()
torch.cuda.synchronize()

# GEMM variants the step above does not reach at these sizes
a = torch.randn(300, 192, device="cuda").bfloat16()
w = torch.randn(520, 192, device="cuda").bfloat16()
out = torch.empty(300, 520, device="cuda", dtype=torch.bfloat16)
for bn in (64, 128, 192, 256, 128 | nv.TILE_PAIR, 192 | nv.TILE_PAIR, 256 | nv.TILE_PAIR):
    nv.gemm_bf16(a, w, out, block_n=bn)
ws = torch.empty(nv.gemm_splitk_workspace_bytes(300, 520, 3) // 4, device="cuda")
nv.gemm_bf16_splitk(a, w, out, ws, split_k=3, block_n=256)
dy = torch.randn(256, 384, device="cuda").bfloat16()
x = torch.randn(256, 264, device="cuda").bfloat16()
outs = [torch.zeros(384, 264, device="cuda") for _ in range(3)]
nv.gemm_bf16_group([(dy, x, o, sk) for o, sk in zip(outs, (1, 2, 4))], block_n=256)
torch.cuda.synchronize()
print("ok", eng.launches_per_train_step)
