"""torchrun worker: CUPTI kernel timeline (torch.profiler) of one graph-replayed step of a data-parallel client, rank 0's
view — which kernels of the step wait on the other GPUs and for how long.  Diagnostic only.
Writes gpurun_out/timeline_dp<G>.txt."""
import importlib
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
eng_mod = importlib.import_module(f"{PKG}.models.engine")
arena_mod = importlib.import_module(f"{PKG}.models.arena")
synth = importlib.import_module(f"{PKG}.data.synth")
dp_mod = importlib.import_module(f"{PKG}.parallel.dp")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dims = arena_mod.ModelDims()
grp = dp_mod.DPGroup(rank, world, arena_mod.ArenaLayout(dims))
eng = eng_mod.B200Engine(dims, batch=16, seq=128, device=local, init_seed=0, dp=grp)
ids, lens, labels = synth.synth_token_batch(16, 128, dims.vocab_size, seed=1 + rank)
eng.set_batch(ids.cuda(), lens.cuda(), labels.cuda())
eng.capture()
for _ in range(5):
    eng.train_step()
torch.cuda.synchronize()
dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        eng.train_step()
    torch.cuda.synchronize()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "memcpy" not in e.name.lower()]
    evs.sort(key=lambda e: e.time_range.start)
    n = len(evs) // 3
    last = evs[2 * n:]
    t0 = last[0].time_range.start
    lines = [f"{len(last)} kernels in the last of 3 replays; step span {(last[-1].time_range.end - t0):.1f} us; dp{world}"]
    for e in last:
        name = e.name.replace("dcad::", "").replace("void ", "")[:70]
        lines.append(f"{e.time_range.start - t0:9.1f} +{e.time_range.end - e.time_range.start:6.1f}  {name}")
    out = "\n".join(lines)
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    (ROOT / "gpurun_out" / f"timeline_dp{world}.txt").write_text(out + "\n")
    print(out[-3000:])
assert grp.error_flag() == 0
dist.barrier()
dist.destroy_process_group()
