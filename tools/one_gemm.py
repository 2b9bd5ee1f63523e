"""Launch one named GEMM of the training step a few times (for `ncu --set full` captures)."""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv  # noqa: E402
from tools.tune_gemm import GEMMS  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--name", default="lin2_fwd")
ap.add_argument("--bn", type=int, default=96)
ap.add_argument("--iters", type=int, default=4)
a = ap.parse_args()
name, M, N, K, am, bm, epi = [g for g in GEMMS if g[0] == a.name][0]
dev = "cuda"
A = torch.randn((M, K) if am == 0 else (K, M), device=dev).bfloat16()
B = (torch.randn((N, K) if bm == 0 else (K, N), device=dev) * 0.05).bfloat16()
out = torch.empty(M, N, device=dev, dtype=torch.float32 if am == 1 else torch.bfloat16)
bias = torch.randn(N, device=dev) if epi in (1, 2) else None
aux = torch.randn(M, N, device=dev).bfloat16() if epi == 3 else None
aux_out = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == 2 else None
for _ in range(a.iters):
    nv.gemm_bf16(A, B, out, a_major=am, b_major=bm, bias=bias, aux=aux, aux_out=aux_out, epi=epi, block_n=a.bn)
torch.cuda.synchronize()
print("ok", name, a.bn)
