"""In-context cost of every launch of the training step: capture the step cut off after k launches for k = 1..L,
replay each graph, and report T(k) - T(k-1).  Unlike serialized per-kernel timings (ncu) this sees the real cache
state, programmatic-dependent-launch overlap and the side-stream branch, i.e. what each launch adds to the step's
critical path.  Writes gpurun_out/prefix_profile.txt."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
import importlib  # noqa: E402

eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
synth = importlib.import_module(f"{PKG}.data.synth")


def main():
    stride = int(os.environ.get("STRIDE", "1"))
    reps = int(os.environ.get("REPS", "30"))
    dims = arena_mod.ModelDims()
    eng = eng_mod.B200Engine(dims, batch=16, seq=128, init_seed=0)
    ids, lens, labels = synth.synth_token_batch(16, 128, dims.vocab_size, seed=1)
    eng.set_batch(ids.cuda(), lens.cuda(), labels.cuda())
    eng._launch_log = []
    eng._train_step_kernels()
    torch.cuda.synchronize()
    names = list(eng._launch_log)
    eng._launch_log = None
    L = len(names)
    s = torch.cuda.Stream(priority=-1)
    times = {}
    with torch.cuda.stream(s):
        for k in list(range(stride, L + 1, stride)) + ([L] if L % stride else []):
            eng._launch_limit = k
            eng._train_step_kernels()  # warm (lazy attribute sets)
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                eng._train_step_kernels()
            for _ in range(3):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            for _ in range(reps):
                g.replay()
            e1.record(s)
            s.synchronize()
            times[k] = e0.elapsed_time(e1) * 1e3 / reps
    lines = [f"prefix profile: {L} launches, full step {times[L]:.1f} us (graph replay, mean of {reps})"]
    prev, prev_k = 0.0, 0
    for k in sorted(times):
        lines.append(f"{k:4d} {times[k]:9.1f} us  +{times[k] - prev:7.1f}   " + " | ".join(names[prev_k:k]))
        prev, prev_k = times[k], k
    out = "\n".join(lines)
    print(out)
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    (ROOT / "gpurun_out" / "prefix_profile.txt").write_text(out + "\n")


if __name__ == "__main__":
    main()
