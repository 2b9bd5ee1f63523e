"""Kernel timeline of one graph-replayed training step (CUPTI through torch.profiler): true start / end of every
kernel on both graph branches, so overlap and gaps are visible.  Diagnostic only — never a bench number.
Writes gpurun_out/timeline.txt."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
import importlib  # noqa: E402

eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
synth = importlib.import_module(f"{PKG}.data.synth")

dims = arena_mod.ModelDims()
SEQ = int(os.environ.get("SEQ", "96"))
eng = eng_mod.B200Engine(dims, batch=16, seq=SEQ, init_seed=0)
ids, lens, labels = synth.synth_token_batch(16, 128, dims.vocab_size, seed=1)
ids = ids[:, :SEQ].contiguous()
eng.set_batch(ids.cuda(), lens.cuda(), labels.cuda())
eng.capture()
for _ in range(5):
    eng.train_step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        eng.train_step()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
evs.sort(key=lambda e: e.time_range.start)
n = len(evs) // 3
last = evs[2 * n:]
t0 = last[0].time_range.start
lines = [f"{len(last)} kernels in the last of 3 replays; step span {(last[-1].time_range.end - t0):.1f} us"]
streams = {}
prev_end = {}
for e in last:
    st = streams.setdefault(e.device_index if not hasattr(e, 'stream') else getattr(e, 'stream', 0), len(streams))
    name = e.name.replace("dcad::", "").replace("void ", "")
    name = name[:60]
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    lines.append(f"{s:9.1f} +{d:6.1f}  {name}")
out = "\n".join(lines)
print(out)
os.makedirs(ROOT / "gpurun_out", exist_ok=True)
(ROOT / "gpurun_out" / "timeline.txt").write_text(out + "\n")
