"""torchrun worker: intra-client data parallelism (fused wgrad reduce-scatter, sharded Adam, fused all-gather) checked
against a single-GPU step on the concatenated batch.  Small model, dropout off.  Prints DP_OK on rank 0."""
import dataclasses
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.arena import ArenaLayout, ModelDims  # noqa: E402
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.engine import B200Engine  # noqa: E402
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel.dp import DPGroup  # noqa: E402


def batch(B, S, vocab, seed):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, vocab, (B, S), generator=g)
    lens = torch.randint(S // 4, S + 1, (B,), generator=g)
    labels = torch.randint(0, 2, (B,), generator=g)
    return ids.int(), lens.int(), labels.int()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    full = os.environ.get("DP_FULL", "0") == "1"
    dims = ModelDims() if full else ModelDims(vocab_size=1000, max_position=128, dim=256, n_layers=2, n_heads=4,
                                              hidden_dim=512, dropout=0.0, attention_dropout=0.0, head_dropout=0.0)
    if full:
        dims = dataclasses.replace(dims, dropout=0.0, attention_dropout=0.0, head_dropout=0.0)
    Bl, S = 8, 128
    lay = ArenaLayout(dims)
    G = int(os.environ.get("DP_GROUP", world))
    grp = DPGroup.build(rank, world, G, lay) if G != world else DPGroup(rank, world, lay)
    client = rank // G
    assert sum(hi - lo for r in range(grp.world) for lo, hi in grp.gemm_ranges(r)) == lay.gemm_end
    assert grp.r2_bounds[-1] == lay.total
    eng = B200Engine(dims, batch=Bl, seq=S, device=local, init_seed=11, lr=1e-3, dp=grp)
    ids, lens, labels = batch(Bl * G, S, dims.vocab_size, seed=5 + client)
    sl = slice(grp.rank * Bl, (grp.rank + 1) * Bl)
    eng.set_batch(ids[sl].to(dev), lens[sl].to(dev), labels[sl].to(dev))
    p0 = eng.arena.flat.clone()

    # single-GPU oracle on the concatenated batch (every rank computes it redundantly)
    ref = B200Engine(dims, batch=Bl * G, seq=S, device=local, init_seed=11, lr=1e-3)
    ref.set_batch(ids.to(dev), lens.to(dev), labels.to(dev))
    assert torch.equal(ref.arena.flat, p0)

    for step in range(3):
        eng.train_step()
        ref.train_step()
    torch.cuda.synchronize()
    dist.barrier()
    assert grp.error_flag() == 0
    grp.allgather_fp32_gemm_params()
    got, exp = eng.arena.flat, ref.arena.flat
    # Adam's first steps move every weight by ~lr regardless of the gradient's size, so compare the UPDATE directions:
    # the fraction of parameters whose 3-step update differs by more than half a step is what matters
    upd_g, upd_e = got - p0, exp - p0
    moved = upd_e.abs() > 0.5e-3
    bad = ((upd_g - upd_e).abs() > 1.0e-3) & moved
    frac_bad = bad.float().mean().item()
    cos = torch.nn.functional.cosine_similarity(upd_g[moved], upd_e[moved], dim=0).item()
    # bf16 mirrors must be identical on every rank and equal to the owner's fp32 values rounded
    mir = eng.arena.bf16.float()
    allm = [torch.empty_like(mir) for _ in range(world)]
    dist.all_gather(allm, mir)
    same_mirror = all(torch.equal(allm[client * G], m) for m in allm[client * G:(client + 1) * G])
    r2 = slice(lay.gemm_end, lay.total)
    allp = [torch.empty_like(got[r2]) for _ in range(world)]
    dist.all_gather(allp, got[r2].contiguous())
    same_r2 = all(torch.equal(allp[client * G], p) for p in allp[client * G:(client + 1) * G])
    loss_dp, _ = eng.read_loss()
    loss_ref, _ = ref.read_loss()
    ok = frac_bad < 0.02 and cos > 0.98 and same_mirror and same_r2
    print(f"rank {rank}: frac_bad={frac_bad:.4f} cos={cos:.4f} same_mirror={same_mirror} same_r2={same_r2} "
          f"loss_dp(local)={loss_dp:.4f} loss_ref={loss_ref:.4f} owned={[len(grp.gemm_ranges(r)) for r in range(grp.world)]}", flush=True)
    assert ok
    # merged optimizer state of the sharded Adam == the single-GPU oracle's (same layout; flags exactly: the union of
    # the rows the GPUs touched is the set of rows of the concatenated batch)
    m_dp, v_dp, fl_dp = eng.gathered_optimizer_state()
    assert torch.equal(fl_dp, ref.touched_rows), "touched-row flags differ from the single-GPU run"
    cos_m = torch.nn.functional.cosine_similarity(m_dp, ref.arena.exp_avg, dim=0).item()
    cos_v = torch.nn.functional.cosine_similarity(v_dp, ref.arena.exp_avg_sq, dim=0).item()
    assert cos_m > 0.98 and cos_v > 0.98, (cos_m, cos_v)
    assert int((m_dp != 0).sum()) > 0.5 * int((ref.arena.exp_avg != 0).sum())

    if world // G > 1:
        # inter-client FedAvg over client groups: every GPU must end with mean over clients of the clients' params
        full_params = [torch.empty_like(got) for _ in range(world)]
        dist.all_gather(full_params, got.contiguous())
        expect = full_params[0].cpu().clone()
        for c in range(1, world // G):
            expect += full_params[c * G].cpu()
        expect /= (world // G)
        grp.fedavg_clients()
        torch.cuda.synchronize()
        dist.barrier()
        assert grp.error_flag() == 0
        assert torch.equal(eng.arena.flat.cpu(), expect), "grouped FedAvg mismatch"
        assert torch.equal(eng.arena.bf16.cpu(), expect[: lay.gemm_end].bfloat16())
        if rank == 0:
            print("DP_FEDAVG_CLIENTS_OK", flush=True)
    # captured graph + timing
    eng.capture()
    for _ in range(5):
        eng.train_step()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        eng.train_step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 20], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert grp.error_flag() == 0
    # ---- training-state checkpoint of a data-parallel client: snapshot, two more steps, restore, the same two steps
    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.utils.checkpoint import (  # noqa: E501
        load_training_state, training_state)

    class _Backend:
        name = "b200"

        def __init__(self, engine, dp):
            self.engine, self.dp = engine, dp

        def state_dict(self):
            self.dp.allgather_fp32_gemm_params()
            return self.engine.state_dict()

        def load_state_dict(self, sd):
            self.engine.load_state_dict(sd)

    shim = _Backend(eng, grp)
    snap = training_state(shim, round_idx=3)                      # collective; complete on every rank
    assert snap["optimizer"]["exp_avg"].numel() == lay.total
    p_snap = eng.arena.flat.clone()
    step_snap = int(eng.rng[1].item())

    def two_steps():
        eng.train_step()
        eng.train_step()
        torch.cuda.synchronize()
        dist.barrier()
        grp.allgather_fp32_gemm_params()
        return eng.arena.flat.clone(), int(eng.rng[1].item())

    a_params, a_step = two_steps()
    prog = load_training_state(snap, shim)                        # collective: moments split by ownership again
    assert prog["round"] == 3 and int(eng.rng[1].item()) == step_snap
    assert torch.equal(eng.arena.flat, p_snap)
    own = torch.zeros(lay.total, dtype=torch.bool, device=dev)
    for lo, hi in eng.owned_optimizer_ranges():
        own[lo:hi] = True
    assert int((eng.arena.exp_avg[~own] != 0).sum()) == 0, "moments outside the owned ranges must stay zero"
    b_params, b_step = two_steps()
    assert a_step == b_step == step_snap + 2
    upd_a, upd_b = a_params - p_snap, b_params - p_snap
    rel = ((upd_a - upd_b).norm() / upd_a.norm().clamp_min(1e-12)).item()
    assert grp.error_flag() == 0
    assert rel < 0.02, f"resumed run diverges from the uninterrupted one: rel. update difference {rel:.4f}"
    if rank == 0:
        print(f"DP_CKPT_OK rel_update_diff={rel:.2e}", flush=True)
        print(f"DP graph step: {t.item():.3f} ms ({world} GPUs x batch {Bl})", flush=True)
        print("DP_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
