"""Prints the full-size engine-vs-HF-fp32 parity numbers (loss, logits, per-tensor gradient errors) that
tests/test_x_fullsize_gpu.py asserts on; writes gpurun_out/fullsize_parity.json."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_x_fullsize_gpu import fullsize_parity_report  # noqa: E402

rep = fullsize_parity_report()
out, exp = rep["logits"]
res = {"loss": rep["loss"], "logits_max_abs_err": (out - exp).abs().max().item(),
       "grads": [{"key": k, "rel_norm_err": r, "ref_norm": n, "max_err_over_rms": m, "max_err_over_max": _mm} for k, r, n, m, _e, _mm in rep["grads"]]}
worst = sorted(res["grads"], key=lambda g: -g["rel_norm_err"])[:12]
print("loss (engine, hf):", rep["loss"], " logits max |err|:", res["logits_max_abs_err"])
for g in worst:
    print(f"{g['key']:70s} rel {g['rel_norm_err']:.4f}  |ref| {g['ref_norm']:.3e}  max/rms {g['max_err_over_rms']:.3f}")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "fullsize_parity.json"), "w"), indent=1)
