"""The two wide GEMMs of a block (N = 3072: lin1 forward with GELU + pre-activation output, lin2 dgrad with dGELU) have
192-256 tiles for 148 SMs.  Sweep BLOCK_N x persistent-grid size: an exact two-tiles-per-CTA grid (96 or 128 CTAs)
has the same critical path as 148 CTAs with a ragged second wave, but less L2 contention, and leaves SMs to the
weight-gradient branch."""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv  # noqa: E402
from tools.tune_gemm import time_graph  # noqa: E402

T, D, F = 2048, 768, 3072
dev = "cuda"
x = torch.randn(T, D, device=dev).bfloat16()
w1 = (torch.randn(F, D, device=dev) * 0.05).bfloat16()
b1 = torch.randn(F, device=dev)
hmid = torch.empty(T, F, device=dev, dtype=torch.bfloat16)
pre = torch.empty(T, F, device=dev, dtype=torch.bfloat16)
da = torch.randn(T, D, device=dev).bfloat16()
w2 = (torch.randn(D, F, device=dev) * 0.05).bfloat16()   # lin2 weight [768, 3072]: dgrad B operand is MN-major
res = {}
for name, fn in (
    ("lin1_fwd", lambda bn, mc: nv.gemm_bf16(x, w1, hmid, bias=b1, aux_out=pre, epi=nv.EPI_BIAS_GELU, block_n=bn, max_ctas=mc)),
    ("lin2_dgrad", lambda bn, mc: nv.gemm_bf16(da, w2, hmid, b_major=nv.MAJOR_MN, aux=pre, epi=nv.EPI_DGELU, block_n=bn, max_ctas=mc)),
    ("qkv_fwd", lambda bn, mc: nv.gemm_bf16(x, w1[:2304], hmid[:, :2304], bias=b1[:2304], epi=nv.EPI_BIAS, block_n=bn, max_ctas=mc)),
):
    for bn in (128, 192, 256):
        for mc in (148, 144, 128, 96, 72, 64):
            us = time_graph(lambda: fn(bn, mc))
            res[f"{name}/bn{bn}/ctas{mc}"] = round(us, 2)
            print(name, bn, mc, round(us, 2), flush=True)
(ROOT / "gpurun_out").mkdir(exist_ok=True)
(ROOT / "gpurun_out" / "tune_wide.json").write_text(json.dumps(res, indent=1))
