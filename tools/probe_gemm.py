"""GPU probe for the tcgen05 GEMM: checks every operand-major combination, tile width and epilogue against
torch.matmul in fp32, one subprocess per group so a faulting configuration cannot poison the others.

    python tools/probe_gemm.py            # run all groups, write gpurun_out/probe_gemm.json
    python tools/probe_gemm.py --group kk # run one group in-process
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

GROUPS = ["kk", "kmn", "mnmn", "epi", "perf"]


def _ref_gemm(a, b, a_major, b_major):
    A = a.float() if a_major == 0 else a.float().t()
    B = b.float() if b_major == 0 else b.float().t()
    return A @ B.t()


def run_group(group: str) -> dict:
    import torch
    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.ops import native as nv

    torch.manual_seed(0)
    dev = "cuda"
    out = {"group": group, "cases": []}

    def case(M, N, K, a_major, b_major, bn, epi=0, out_fp32=False, tag=""):
        a = torch.randn((M, K) if a_major == 0 else (K, M), device=dev).bfloat16()
        b = torch.randn((N, K) if b_major == 0 else (K, N), device=dev).bfloat16()
        bias = torch.randn(N, device=dev) if epi in (1, 2) else None
        aux = torch.randn(M, N, device=dev).bfloat16() if epi == 3 else None
        aux_out = torch.empty(M, N, device=dev, dtype=torch.bfloat16) if epi == 2 else None
        d = torch.full((M, N), float("nan"), device=dev, dtype=torch.float32 if out_fp32 else torch.bfloat16)
        rec = {"M": M, "N": N, "K": K, "a_major": a_major, "b_major": b_major, "bn": bn, "epi": epi,
               "out_fp32": out_fp32, "tag": tag}
        try:
            nv.gemm_bf16(a, b, d, a_major=a_major, b_major=b_major, bias=bias, aux=aux, aux_out=aux_out, epi=epi,
                         block_n=bn)
            torch.cuda.synchronize()
            ref = _ref_gemm(a, b, a_major, b_major)
            if epi in (1, 2):
                ref = ref + bias
            if epi == 2:
                pre = ref
                ref = torch.nn.functional.gelu(pre)
                rec["aux_err"] = float((aux_out.float() - pre).abs().max() / (pre.abs().max() + 1e-6))
            if epi == 3:
                x = aux.float()
                cdf = 0.5 * (1 + torch.erf(x / 2 ** 0.5))
                pdf = torch.exp(-0.5 * x * x) / (2 * 3.141592653589793) ** 0.5
                ref = ref * (cdf + x * pdf)
            err = (d.float() - ref).abs().max().item()
            scale = ref.abs().max().item() + 1e-6
            rec["max_abs_err"] = err
            rec["rel_err"] = err / scale
            rec["nan"] = int(torch.isnan(d.float()).sum().item())
            rec["ok"] = bool(rec["rel_err"] < (2e-2 if not out_fp32 else 2e-3) and rec["nan"] == 0)
        except Exception as e:  # noqa: BLE001
            rec["ok"] = False
            rec["error"] = repr(e)[:300]
        out["cases"].append(rec)
        print(json.dumps(rec), flush=True)

    if group == "kk":
        for bn in (64, 128, 192, 256):
            case(256, 768, 128, 0, 0, bn, tag="small")
        case(2048, 768, 768, 0, 0, 128, tag="out_lin")
        case(2048, 2304, 768, 0, 0, 256, tag="qkv")
        case(2048, 768, 3072, 0, 0, 128, tag="lin2")
        case(200, 776, 72, 0, 0, 64, tag="ragged")
        case(16, 768, 768, 0, 0, 64, tag="tiny_m")
    elif group == "kmn":
        for bn in (64, 128, 192, 256):
            case(256, 768, 128, 0, 1, bn, tag="small")
        case(2048, 768, 2304, 0, 1, 128, tag="dgrad_qkv")
        case(2048, 3072, 768, 0, 1, 256, tag="dgrad_lin2")
    elif group == "mnmn":
        for bn in (64, 128, 192, 256):
            case(256, 768, 128, 1, 1, bn, out_fp32=True, tag="small")
        case(768, 768, 2048, 1, 1, 128, out_fp32=True, tag="wgrad_out")
        case(3072, 768, 2048, 1, 1, 128, out_fp32=True, tag="wgrad_lin1")
        case(768, 3072, 2048, 1, 1, 256, out_fp32=True, tag="wgrad_lin2")
    elif group == "epi":
        case(2048, 3072, 768, 0, 0, 128, epi=1, tag="bias")
        case(2048, 3072, 768, 0, 0, 128, epi=2, tag="bias_gelu")
        case(2048, 3072, 768, 0, 1, 128, epi=3, tag="dgelu")
        case(256, 768, 128, 0, 0, 128, epi=0, out_fp32=True, tag="f32_kk")
    elif group == "perf":
        import torch
        shapes = [(2048, 2304, 768, 0, 0), (2048, 768, 768, 0, 0), (2048, 3072, 768, 0, 0), (2048, 768, 3072, 0, 0),
                  (2048, 768, 2304, 0, 1), (2048, 3072, 768, 0, 1), (2048, 768, 3072, 0, 1),
                  (768, 768, 2048, 1, 1), (3072, 768, 2048, 1, 1), (768, 3072, 2048, 1, 1), (2304, 768, 2048, 1, 1),
                  (8192, 8192, 8192, 0, 0)]
        for (M, N, K, am, bm) in shapes:
            bns = [b for b in (64, 128, 192, 256) if N % b == 0]
            for bn in bns:
                a = torch.randn((M, K) if am == 0 else (K, M), device=dev).bfloat16()
                b = torch.randn((N, K) if bm == 0 else (K, N), device=dev).bfloat16()
                d = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
                rec = {"M": M, "N": N, "K": K, "a_major": am, "b_major": bm, "bn": bn}
                try:
                    for _ in range(3):
                        nv.gemm_bf16(a, b, d, a_major=am, b_major=bm, block_n=bn)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    iters = 20
                    e0.record()
                    for _ in range(iters):
                        nv.gemm_bf16(a, b, d, a_major=am, b_major=bm, block_n=bn)
                    e1.record()
                    torch.cuda.synchronize()
                    ms = e0.elapsed_time(e1) / iters
                    rec["us"] = ms * 1e3
                    rec["tflops"] = 2.0 * M * N * K / (ms * 1e-3) / 1e12
                    # cuBLAS (torch) for the same math, as the library baseline
                    A = a if am == 0 else a.t()
                    B = b if bm == 0 else b.t()
                    for _ in range(3):
                        torch.matmul(A, B.t())
                    torch.cuda.synchronize()
                    e0.record()
                    for _ in range(iters):
                        torch.matmul(A, B.t())
                    e1.record()
                    torch.cuda.synchronize()
                    rec["cublas_us"] = e0.elapsed_time(e1) / iters * 1e3
                    rec["ok"] = True
                except Exception as e:  # noqa: BLE001
                    rec["ok"] = False
                    rec["error"] = repr(e)[:300]
                out["cases"].append(rec)
                print(json.dumps(rec), flush=True)
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--group", default=None)
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "probe_gemm.json"))
    args = ap.parse_args()
    if args.group:
        res = run_group(args.group)
        print("##RESULT##" + json.dumps(res))
        return
    results = []
    for g in GROUPS:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, "--group", g], capture_output=True, text=True, timeout=300)
            res = None
            for line in r.stdout.splitlines():
                if line.startswith("##RESULT##"):
                    res = json.loads(line[len("##RESULT##"):])
            if res is None:
                res = {"group": g, "crashed": True, "rc": r.returncode, "stdout": r.stdout[-3000:],
                       "stderr": r.stderr[-3000:]}
        except subprocess.TimeoutExpired as e:
            res = {"group": g, "timeout": True, "stdout": (e.stdout or b"")[-2000:].decode("utf8", "ignore")
                   if isinstance(e.stdout, bytes) else str(e.stdout)[-2000:]}
        res["seconds"] = time.time() - t0
        results.append(res)
        nok = sum(1 for c in res.get("cases", []) if c.get("ok"))
        print(f"[probe] group {g}: {nok}/{len(res.get('cases', []))} ok, {res['seconds']:.1f}s "
              f"{'CRASHED' if res.get('crashed') else ''}{'TIMEOUT' if res.get('timeout') else ''}", flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    Path(args.out).write_text(json.dumps(results, indent=1))


if __name__ == "__main__":
    main()
