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
    extra_checks(rank, world, dev, mine, allv, base, elems, mirror)
    if rank == 0:
        print("FEDAVG_IPC_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


def extra_checks(rank, world, dev, mine, allv, base, elems, mirror):
    """Round-2 paths: NVLS with the locally cast mirror, a participant that never shows up (nothing may be written,
    the error word is reported once), rounds over a subset of the clients agreed through the store, the asynchronous
    transport handle, and the bf16 exchange."""
    import time
    from datetime import timedelta

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.fed import rounds, transport

    def say(msg):
        if rank == 0:
            print(msg, flush=True)

    # ---- 1. NVLS, every GPU casting its own bf16 copy ----
    try:
        sg = p2p.SymmPeerGroup(rank, world, elems, mirror)
    except Exception:  # noqa: BLE001
        sg = None
    if sg is not None and sg.has_nvls:
        sg.arena_tensor().copy_(mine)
        sg.mirror_tensor().zero_()
        torch.cuda.synchronize()
        dist.barrier()
        sg.fedavg(p2p.NVLS, local_cast=True)
        torch.cuda.synchronize()
        assert sg.error_flag() == 0
        got = sg.arena_tensor().cpu()
        assert torch.allclose(got, base, rtol=1e-6, atol=1e-6)
        assert torch.equal(sg.mirror_tensor().cpu(), got[:mirror].bfloat16()), "mirror is not the cast of the local arena"
        say("NVLS_LOCAL_CAST_OK")
    if sg is not None:
        dist.barrier()
        sg.close()

    grp = p2p.PeerGroup(rank, world, elems, mirror)
    arena = grp.arena_tensor()
    # ---- 2. a client that never launches: the others time out, write NOTHING, report the failure once ----
    arena.copy_(mine)
    torch.cuda.synchronize()
    dist.barrier()
    absent = world - 1
    t0 = time.time()
    if rank != absent:
        grp.fedavg(p2p.TWO_SHOT, epoch=1)          # expects every rank; rank `absent` never comes
        torch.cuda.synchronize()
        assert grp.error_flag() == 1, "a missing peer must raise the error word"
        assert grp.error_flag() == 0, "the error word is cleared once it has been reported"
    dist.barrier()
    torch.cuda.synchronize()
    assert torch.equal(arena, mine), f"rank {rank}: a timed-out round modified the parameters"
    say(f"ABSENT_PEER_LEAVES_ARENAS_UNTOUCHED_OK ({time.time() - t0:.1f} s)")
    # ... and the next round works again (explicit epochs keep the ranks in step although one skipped a launch)
    dist.barrier()
    grp.fedavg(p2p.TWO_SHOT, epoch=2)
    torch.cuda.synchronize()
    assert grp.error_flag() == 0 and torch.equal(arena.cpu(), base)
    say("ROUND_AFTER_FAILURE_OK")

    # ---- 3. quorum through the store + the asynchronous transport (its epochs count rounds from 1: fresh pads) ----
    from torch.distributed import distributed_c10d as c10d

    dist.barrier()
    grp.close()
    grp = p2p.PeerGroup(rank, world, elems, mirror)
    arena = grp.arena_tensor()

    store = dist.PrefixStore("mp_check", c10d._get_default_store())
    coord = rounds.RoundCoordinator(store, rank, world, deadline_s=20.0, min_clients=1)

    class Plan:  # the last client sits round 0 out (what utils.faults "drop=N@0" does)
        def before_upload(self, client_id, rnd):
            return not (client_id == world - 1 and rnd == 0 and world > 1)

    tr = transport.NvSwitchTransport(grp, transport.AUTO, coordinator=coord, faults=Plan(), client_id=rank)
    arena.copy_(mine)
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.time()
    pend = tr.aggregate_async(None)      # returns with the kernel in flight
    ok = pend.wait()
    torch.cuda.synchronize()
    members = list(range(world - 1)) if world > 1 else [0]
    sub = allv[members[0]].cpu().clone()
    for m in members[1:]:
        sub += allv[m].cpu()
    sub /= len(members)
    if rank in members:
        assert ok and torch.equal(arena.cpu(), sub), f"rank {rank}: subset round mismatch"
    else:
        assert not ok and torch.equal(arena, mine), "the client that sat out must keep its local model"
    assert time.time() - t0 < 15.0, "an announced absence must not cost the deadline"
    dist.barrier()
    arena.copy_(mine)
    torch.cuda.synchronize()
    dist.barrier()
    assert tr.aggregate(None) and torch.equal(arena.cpu(), base)      # round 1: everybody again
    say("QUORUM_SUBSET_AND_ASYNC_OK")
    dist.barrier()
    grp.close()

    # ---- 3b. the signal-pad protocol under skew: no host barrier between rounds, every rank delays its launch by a
    # random amount and refills its arena right after its own kernel; the start handshake must hold a fast rank back
    # until the slow ranks' arenas are final, the end handshake must keep a fast rank from refilling an arena a slow
    # rank is still reading ----
    import random

    grp = p2p.PeerGroup(rank, world, elems, mirror)
    arena = grp.arena_tensor()
    rnd_gen = random.Random(1000 + rank)
    gens = [torch.Generator(device=dev).manual_seed(7000 + r) for r in range(world)]
    for it in range(8):
        vals = [torch.randn(elems, device=dev, generator=gens[r]) for r in range(world)]  # every rank knows all inputs
        arena.copy_(vals[rank])
        time.sleep(rnd_gen.random() * 0.05)
        grp.fedavg(p2p.TWO_SHOT if it % 2 == 0 else p2p.TWO_SHOT, epoch=it + 1)
        exp = vals[0].cpu().clone()
        for v in vals[1:]:
            exp += v.cpu()
        exp /= world
        torch.cuda.synchronize()
        assert grp.error_flag() == 0, f"rank {rank} round {it}: handshake timed out"
        assert torch.equal(arena.cpu(), exp), f"rank {rank} round {it}: skewed launch broke the round"
    say("SKEWED_LAUNCHES_OK")
    dist.barrier()
    grp.close()

    # ---- 4. bf16 exchange arenas (half the NVLink bytes) ----
    xg = p2p.PeerGroup(rank, world, elems, 0, arena_dtype=torch.bfloat16)
    xg.arena_tensor().copy_(mine.bfloat16())
    torch.cuda.synchronize()
    dist.barrier()
    xg.fedavg(p2p.TWO_SHOT, with_mirror=False)
    torch.cuda.synchronize()
    assert xg.error_flag() == 0
    exp16 = allv[0].bfloat16().float().cpu()
    for a in allv[1:]:
        exp16 += a.bfloat16().float().cpu()
    exp16 = (exp16 / world).bfloat16()
    assert torch.equal(xg.arena_tensor().cpu(), exp16), "bf16 exchange: fp32 accumulation, one rounding of the mean"
    say("BF16_EXCHANGE_OK")
    dist.barrier()
    xg.close()


if __name__ == "__main__":
    main()
