"""FedAvg aggregation sweep (BASELINE.json config 5): the 66M-parameter DistilBERT arena, fp32 and bf16, at the
launched world size; fused peer-memory kernel (two-shot / one-shot) vs NCCL all-reduce baseline.  Run under torchrun.
Device-timed, max over ranks.  Appends one JSON line per configuration to gpurun_out/fedavg_sweep.jsonl."""
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.arena import ArenaLayout, ModelDims  # noqa: E402
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import nccl_baseline, p2p  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lay = ArenaLayout(ModelDims())
    iters = 20

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = []
    for dtype, name in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
        esz = 4 if dtype == torch.float32 else 2
        elems = lay.total
        mirror = lay.gemm_end if dtype == torch.float32 else 0
        grp = p2p.PeerGroup(rank, world, elems, mirror, with_out_buffer=True, arena_dtype=dtype)
        grp.arena_tensor().copy_(torch.randn(elems, device=dev).to(dtype))
        nbytes = elems * esz
        for mode, mname in ((p2p.TWO_SHOT, "two_shot"), (p2p.ONE_SHOT, "one_shot")):
            for ctas in (0, 64, 296):
                ms = timed(lambda: grp.fedavg(mode, max_ctas=ctas))
                # bytes arriving at each GPU over NVLink: two-shot = pulled shards + peers' pushed shards (+ bf16 mirror)
                link = ((world - 1) / world * (2 * nbytes + mirror * 2)) if mode == p2p.TWO_SHOT else (world - 1) * nbytes
                out.append({"world": world, "dtype": name, "kernel": f"p2p_{mname}", "max_ctas": ctas or 148, "ms": ms,
                            "arena_MB": nbytes / 1e6, "ingress_bytes_per_gpu": link,
                            "ingress_GBps": link / (ms * 1e-3) / 1e9, "roofline_ms_770GBps": link / 770e9 * 1e3,
                            "err": grp.error_flag()})
        if dtype == torch.float32:
            try:
                sg = p2p.SymmPeerGroup(rank, world, elems, mirror)
                sg.arena_tensor().copy_(grp.arena_tensor())
                if sg.has_nvls:
                    for ctas in (0, 64, 32):
                        ms = timed(lambda: sg.fedavg(p2p.NVLS, max_ctas=ctas))
                        link = nbytes / world + (world - 1) / world * (nbytes + mirror * 2)
                        out.append({"world": world, "dtype": name, "kernel": "nvls_multimem", "max_ctas": ctas or 148,
                                    "ms": ms, "arena_MB": nbytes / 1e6, "ingress_bytes_per_gpu": link,
                                    "ingress_GBps": link / (ms * 1e-3) / 1e9, "roofline_ms_770GBps": link / 770e9 * 1e3,
                                    "err": sg.error_flag()})
                sg.close()
            except Exception as e:  # noqa: BLE001
                out.append({"world": world, "dtype": name, "kernel": "nvls_multimem", "unavailable": repr(e)[:120]})
        scratch = grp.arena_tensor().clone()
        m = torch.empty(max(mirror, 8), device=dev, dtype=torch.bfloat16)
        ms = timed(lambda: nccl_baseline.nccl_fedavg_(scratch, m if mirror else None, mirror))
        out.append({"world": world, "dtype": name, "kernel": "nccl_allreduce+scale+cast", "ms": ms,
                    "arena_MB": nbytes / 1e6})
        dist.barrier()
        grp.close()
    if rank == 0:
        os.makedirs(ROOT / "gpurun_out", exist_ok=True)
        with open(ROOT / "gpurun_out" / "fedavg_sweep.jsonl", "a") as f:
            for r in out:
                f.write(json.dumps(r) + "\n")
                print(json.dumps(r), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
