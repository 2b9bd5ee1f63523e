"""Times every GEMM of the DDoSClassifier training step (tokens = batch*seq) for each BLOCK_N, inside a CUDA graph
(kernel time, not Python launch overhead), next to cuBLAS (torch.matmul) on the same math.  Prints a table and the
best BLOCK_N per GEMM; writes gpurun_out/tune_gemm.json."""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv  # noqa: E402

T = int(os.environ.get("TOKENS", 2048))
D, F = 768, 3072
# name, M, N, K, a_major, b_major, epi
GEMMS = [
    ("qkv_fwd", T, 3 * D, D, 0, 0, 1), ("out_fwd", T, D, D, 0, 0, 1), ("lin1_fwd", T, F, D, 0, 0, 2),
    ("lin2_fwd", T, D, F, 0, 0, 1),
    ("lin2_dgrad", T, F, D, 0, 1, 3), ("lin1_dgrad", T, D, F, 0, 1, 0), ("out_dgrad", T, D, D, 0, 1, 0),
    ("qkv_dgrad", T, D, 3 * D, 0, 1, 0),
    ("lin2_wgrad", D, F, T, 1, 1, 0), ("lin1_wgrad", F, D, T, 1, 1, 0), ("out_wgrad", D, D, T, 1, 1, 0),
    ("qkv_wgrad", 3 * D, D, T, 1, 1, 0),
]


def time_graph(fn, iters=20, reps=5):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters):
                fn()
        best = 1e9
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            g.replay()
            e1.record(s)
            s.synchronize()
            best = min(best, e0.elapsed_time(e1) / iters)
    torch.cuda.current_stream().wait_stream(s)
    return best * 1e3  # us


def main():
    # GEMM_EPI=1: time everything with the per-warp epilogue (dcad_gemm_set_epilogue); results go to *_epi1.json
    epi_mode = int(os.environ.get("GEMM_EPI", "0"))
    nv.set_gemm_epilogue(epi_mode)
    torch.manual_seed(0)
    dev = "cuda"
    res = []
    for name, M, N, K, am, bm, epi in GEMMS:
        a = torch.randn((M, K) if am == 0 else (K, M), device=dev).bfloat16()
        b = (torch.randn((N, K) if bm == 0 else (K, N), device=dev) * 0.05).bfloat16()
        wgrad = am == 1
        out = torch.empty(M, N, device=dev, dtype=torch.float32 if wgrad else torch.bfloat16)
        bias = torch.randn(N, device=dev) if epi in (1, 2) else None
        aux = torch.randn(M, N, device=dev).bfloat16() if epi == 3 else None
        aux_out = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == 2 else None
        A = a if am == 0 else a.t()
        B = b if bm == 0 else b.t()
        cub = time_graph(lambda: torch.matmul(A, B.t()))
        # the same WORK through the library: cuBLAS GEMM + the elementwise kernels its epilogue-less call needs for
        # what our epilogue fuses (bias, GELU with the pre-activation kept, dGELU) -- the like-for-like bar
        if epi == 1:
            fair = time_graph(lambda: torch.addmm(bias.bfloat16(), A, B.t()))
        elif epi == 2:
            bb = bias.bfloat16()

            def lib_gelu():
                pre = torch.addmm(bb, A, B.t())
                return pre, torch.nn.functional.gelu(pre)
            fair = time_graph(lib_gelu)
        elif epi == 3:
            ax = aux.float().requires_grad_(True)
            torch.nn.functional.gelu(ax).backward(torch.ones_like(ax))
            dg = ax.grad.bfloat16()   # (the derivative itself is precomputed: only the multiply is charged)
            fair = time_graph(lambda: torch.matmul(A, B.t()) * dg)
        else:
            fair = cub
        row = {"name": name, "M": M, "N": N, "K": K, "cublas_us": cub, "cublas_plus_epilogue_kernels_us": fair, "bn": {}}
        P = nv.TILE_PAIR  # "p" prefix below = 256 x bn tile on a CTA pair (cta_group::2)
        lab = lambda bn: ("p" if bn & P else "") + str(bn & ~P)  # noqa: E731
        cands = [(bn, 1) for bn in (64, 128, 192, 256) if N % bn == 0]
        if N % 384 == 0 and am == 0 and N >= 1536:
            cands.append((384, 1))  # 128 x 384 tiles (two N = 192 accumulators), per-warp epilogue
        cands += [(bn | P, 1) for bn in (128, 192, 256) if N % bn == 0 and not (bn == 192 and bm == 1)]
        if wgrad:
            cands = [(bn, sk) for bn in (128, 256, 128 | P, 256 | P) for sk in (1, 2, 3, 4, 8)]
        for bn, sk in cands:
            key = f"{lab(bn)}/{sk}" if wgrad else lab(bn)
            try:
                if wgrad:
                    out.zero_()
                us = time_graph(lambda: nv.gemm_bf16(a, b, out, a_major=am, b_major=bm, bias=bias, aux=aux,
                                                     aux_out=aux_out, epi=epi, block_n=bn, split_k=sk))
                row["bn"][key] = us
            except Exception as e:  # noqa: BLE001
                row["bn"][key] = None
                row.setdefault("errors", {})[key] = repr(e)[:120]
        if name in ("lin2_fwd", "lin1_dgrad", "out_fwd", "out_dgrad", "qkv_dgrad") or (T <= 128 and not wgrad):
            ws_cands = ((256, 2), (256, 3), (256, 4), (128, 2), (128, 3), (256 | P, 2), (256 | P, 3), (256 | P, 4),
                        (256 | P, 6), (128 | P, 2), (128 | P, 3))
            if T <= 128:  # the CLS-row tail of the last block: few output tiles, latency-bound
                ws_cands = ((64, 2), (64, 4), (64, 8), (128, 4), (128, 8), (128, 12), (256, 4), (256, 8), (256, 12),
                            (256, 16))
            for bn, sk in ws_cands:
                ws = torch.empty(nv.gemm_splitk_workspace_bytes(M, N, sk) // 4, device=dev)
                try:
                    us = time_graph(lambda: nv.gemm_bf16_splitk(a, b, out, ws, split_k=sk, a_major=am, b_major=bm,
                                                                bias=bias, aux=aux, aux_out=aux_out, epi=epi, block_n=bn))
                    row["bn"][f"ws{lab(bn)}/{sk}"] = us
                except Exception as e:  # noqa: BLE001
                    row["bn"][f"ws{lab(bn)}/{sk}"] = None
        ok = {k: v for k, v in row["bn"].items() if v}
        row["best_bn"] = min(ok, key=ok.get) if ok else None
        row["best_us"] = ok[row["best_bn"]] if ok else None
        row["tflops_best"] = 2.0 * M * N * K / (row["best_us"] * 1e-6) / 1e12 if ok else None
        res.append(row)
        print(f"{name:11s} {M:5d}x{N:5d}x{K:5d} cublas {cub:7.1f}us (+epilogue kernels {fair:5.1f}) | " +
              " ".join(f"{bn}:{(v if v else -1):6.1f}" for bn, v in row["bn"].items()) +
              f" | best {row['best_bn']} {row['best_us']:.1f}us {row['tflops_best']:.0f}TF", flush=True)
    tot = sum(r["best_us"] for r in res) * 6
    totc = sum(r["cublas_us"] for r in res) * 6
    totf = sum(r["cublas_plus_epilogue_kernels_us"] for r in res) * 6
    print(f"sum over 6 layers: ours(best) {tot:.0f} us, cuBLAS (no epilogues) {totc:.0f} us, "
          f"cuBLAS + epilogue kernels {totf:.0f} us")
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    name = (f"tune_gemm_T{T}" if T != 2048 else "tune_gemm") + (f"_epi{epi_mode}" if epi_mode else "") + ".json"
    (ROOT / "gpurun_out" / name).write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
