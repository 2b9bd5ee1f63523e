"""Is the tcgen05 GEMM main loop bound by the tensor pipe or by L2 -> SM operand traffic?

Runs one large K-major GEMM (592 tiles of 128x256) with the persistent grid capped at 148 / 74 / 37 / 16 CTAs.  If
the per-CTA tile rate is independent of how many CTAs run, the loop is issue/tensor bound; if fewer CTAs each run
faster, the shared resource (L2 slices, ~6300 B/clk chip-wide per the microarchitecture notes) is the limit.
Prints one JSON line per configuration.
"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv  # noqa: E402


def time_gemm(M, N, K, bn, max_ctas, iters=10, out_dtype=torch.bfloat16):
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=out_dtype)
    for _ in range(3):
        nv.gemm_bf16(A, B, out, block_n=bn, max_ctas=max_ctas)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        nv.gemm_bf16(A, B, out, block_n=bn, max_ctas=max_ctas)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    nv.set_pdl(False)
    pair = "--pair" in sys.argv
    cg = 2 if pair else 1
    for bn in (256, 128):
        for K in (768, 3072):
            M, N = 128 * 37 * cg, bn * 16 // cg
            tiles = 37 * 16 // cg  # tiles of (128*cg) x bn
            for ctas in (148, 74, 36, 16):
                us = time_gemm(M, N, K, bn | (nv.TILE_PAIR if pair else 0), ctas)
                workers = ctas // cg
                tiles_per_cta = (tiles + workers - 1) // workers
                kb = K // 64
                per_kb_ns = us * 1e3 / (tiles_per_cta * kb)
                bytes_per_kb = (128 + bn // cg) * 64 * 2
                print(json.dumps({
                    "pair": pair, "bn": bn, "M": M, "N": N, "K": K, "ctas": ctas, "us": round(us, 2),
                    "tiles_per_cta": tiles_per_cta, "ns_per_kblock_per_cta": round(per_kb_ns, 1),
                    "tflops": round(2.0 * M * N * K / us / 1e6, 1),
                    "l2_to_sm_TBps": round(ctas * bytes_per_kb / per_kb_ns / 1e3, 2),
                }), flush=True)


if __name__ == "__main__":
    main()
