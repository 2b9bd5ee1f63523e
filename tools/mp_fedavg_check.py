"""torchrun worker: multi-process (one rank per GPU) FedAvg over CUDA-IPC-shared arenas; checks bit parity with the
reference arithmetic and that weighted / quorum variants behave.  Prints FEDAVG_IPC_OK on rank 0."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import p2p  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    elems, mirror = 1 << 22, 1 << 20
    grp = p2p.PeerGroup(rank, world, elems, mirror)
    arena = grp.arena_tensor()
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    mine = torch.randn(elems, device=dev, generator=g)
    arena.copy_(mine)
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine)
    base = allv[0].cpu().clone()
    for a in allv[1:]:
        base += a.cpu()
    base /= world
    torch.cuda.synchronize()
    dist.barrier()
    grp.fedavg(p2p.TWO_SHOT)
    torch.cuda.synchronize()
    assert grp.error_flag() == 0
    assert torch.equal(arena.cpu(), base), f"rank {rank}: two-shot mismatch"
    assert torch.equal(grp.mirror_tensor().cpu(), base[:mirror].bfloat16())
    # weighted (n_k / n style) second round on fresh data
    arena.copy_(mine)
    torch.cuda.synchronize()
    dist.barrier()
    w = [(i + 1) / sum(range(1, world + 1)) for i in range(world)]
    grp.fedavg(p2p.TWO_SHOT, weights=w)
    torch.cuda.synchronize()
    exp = sum(wi * a.cpu() for wi, a in zip(w, allv))
    assert torch.allclose(arena.cpu(), exp, rtol=1e-5, atol=1e-6)
    dist.barrier()
    grp.close()
    # symmetric-memory group: two-shot over its unicast pointers, NVLS when the host exposes multicast
    try:
        sg = p2p.SymmPeerGroup(rank, world, elems, mirror)
    except Exception as e:  # noqa: BLE001
        sg = None
        if rank == 0:
            print("SYMM_UNAVAILABLE", repr(e)[:200], flush=True)
    if sg is not None:
        for mode, name in ((p2p.TWO_SHOT, "two_shot"), (p2p.NVLS, "nvls")):
            if mode == p2p.NVLS and not sg.has_nvls:
                if rank == 0:
                    print("NVLS_UNAVAILABLE (no multicast support on this host)", flush=True)
                continue
            sg.arena_tensor().copy_(mine)
            torch.cuda.synchronize()
            dist.barrier()
            sg.fedavg(mode)
            torch.cuda.synchronize()
            assert sg.error_flag() == 0
            got = sg.arena_tensor().cpu()
            if mode == p2p.TWO_SHOT:
                assert torch.equal(got, base), f"rank {rank}: symm two-shot mismatch"
            else:
                assert torch.allclose(got, base, rtol=1e-6, atol=1e-6), f"rank {rank}: nvls mismatch"
            assert torch.allclose(sg.mirror_tensor().float().cpu(), base[:mirror].bfloat16().float(), rtol=1e-2, atol=1e-2)
            if rank == 0:
                print(f"SYMM_{name.upper()}_OK", flush=True)
        dist.barrier()
        sg.close()
    if rank == 0:
        print("FEDAVG_IPC_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
