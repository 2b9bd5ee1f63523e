"""Times the end-of-step Adam launch over the non-GEMM region (biases / LayerNorm / classifier / embeddings) of the
arena for different numbers of touched vocabulary rows."""
import importlib
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
nv = importlib.import_module(f"{PKG}.ops.native")
dims = arena_mod.ModelDims()
eng = eng_mod.B200Engine(dims, batch=16, seq=96, init_seed=0)
lay = eng.arena.layout
for touched in [int(t) for t in os.environ.get("PROBE_ROWS", "0,400,2232,30522").split(",")]:
    eng.touched_rows.zero_()
    if touched:
        eng.touched_rows[torch.randperm(dims.vocab_size, device="cuda")[:touched]] = 1
    # 20 launches inside one CUDA graph: eager launches through ctypes are host-bound (about 25 us each), which
    # would hide the kernel time for the small cases
    for _ in range(3):
        eng._adam_range(lay.gemm_end, lay.total, advance=True)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(20):
            eng._adam_range(lay.gemm_end, lay.total, advance=True)
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    print(f"touched rows {touched:6d}: {e0.elapsed_time(e1) / 20 * 1e3:7.1f} us per launch (graph of 20)", flush=True)
