"""Single-box launcher: N federated clients = N GPUs (one process per GPU), aggregation over NVSwitch peer memory.

Run under torchrun (``python -m torch.distributed.run --nproc-per-node N -m <pkg>.cli launch ...``) or directly
(``... cli launch --num-clients N`` re-executes itself under torchrun).  Each rank is client ``rank + 1``; rank 0 also
writes the server artefact ``ddos_distilbert_model.pth`` (reference server.py:77) after every round.
``torch.distributed`` (NCCL or gloo) is plumbing only — IPC-handle exchange and barriers; the aggregation itself is
the peer-memory kernel unless ``--transport nccl`` (the measured baseline) is requested.
"""
from __future__ import annotations

import dataclasses
import os
import sys

import torch

from ..config import FedConfig
from ..models.arena import ArenaLayout
from ..ops import native as nv
from .client import run_client
from .rounds import RoundCoordinator
from .transport import (AUTO, LoopbackTransport, NcclTransport, NvSwitchBf16Transport, NvSwitchDPTransport,
                        NvSwitchTransport, Transport)


def _env_int(k: str, d: int) -> int:
    try:
        return int(os.environ.get(k, d))
    except ValueError:
        return d


def launch(cfg: FedConfig) -> None:
    import torch.distributed as dist

    G = max(1, cfg.gpus_per_client)
    if "RANK" not in os.environ and cfg.num_clients * G > 1:
        port = 29600 + os.getpid() % 1000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={cfg.num_clients * G}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), "-m", __package__.rsplit(".", 1)[0] + ".cli",
               "launch"] + sys.argv[2:]
        os.execv(sys.executable, cmd)
    rank, world, local = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1), _env_int("LOCAL_RANK", 0)
    use_cuda = nv.available() and cfg.backend != "torch"
    if world > 1:
        if use_cuda:
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group("gloo")
    if world % G:
        raise ValueError("world size must be a multiple of --gpus-per-client")
    ccfg = dataclasses.replace(cfg, client_id=rank // G + 1, num_clients=world // G)
    reporter = rank % G == 0
    transport: Transport
    backend_kwargs = {}
    group = None
    if use_cuda:
        backend_kwargs["device"] = local
    if G > 1:
        if not use_cuda:
            raise RuntimeError("--gpus-per-client > 1 needs the B200 backend")
        from ..parallel.dp import DPGroup

        dpg = DPGroup.build(rank, world, G, ArenaLayout(ccfg.model_dims()))
        backend_kwargs.update(dp=dpg, init_seed=ccfg.client_id)
        transport = NvSwitchDPTransport(dpg) if world // G > 1 else LoopbackTransport(1)
    elif world > 1 and use_cuda and cfg.transport in ("nvswitch", "nvswitch-oneshot", "nvls", "auto", "nvswitch-bf16"):
        from ..parallel import p2p
        from ..utils.faults import FaultPlan

        layout = ArenaLayout(ccfg.model_dims())
        plan = FaultPlan.parse(cfg.faults) if cfg.faults else FaultPlan.from_env()
        coord = _coordinator(cfg, rank, world)
        if cfg.transport == "nvswitch-bf16":
            xgroup = p2p.PeerGroup(rank, world, layout.total, 0, arena_dtype=torch.bfloat16)
            transport = NvSwitchBf16Transport(xgroup, coord, plan, ccfg.client_id)
        else:
            one = cfg.transport == "nvswitch-oneshot"
            group = None
            if cfg.transport in ("nvls", "auto"):
                # symmetric memory (allocator + handle exchange only) also carries the NVLink multicast mapping
                try:
                    group = p2p.SymmPeerGroup(rank, world, layout.total, layout.gemm_end)
                except Exception as e:  # noqa: BLE001
                    if rank == 0:
                        print(f"[launch] symmetric-memory group unavailable ({e!r}); using cudaMalloc + IPC (two-shot)",
                              file=sys.stderr)
                if group is not None and cfg.transport == "nvls" and not group.has_nvls and rank == 0:
                    print("[launch] this host exposes no NVLink multicast: --transport nvls runs the two-shot kernel",
                          file=sys.stderr)
            if group is None:
                group = p2p.PeerGroup(rank, world, layout.total, layout.gemm_end, with_out_buffer=one)
            backend_kwargs.update(flat_params=group.arena_tensor(), bf16_params=group.mirror_tensor())
            mode = {"nvswitch": p2p.TWO_SHOT, "nvswitch-oneshot": p2p.ONE_SHOT, "nvls": p2p.NVLS, "auto": AUTO}[cfg.transport]
            from ..report.log import EventLog

            transport = NvSwitchTransport(group, mode, coordinator=coord, faults=plan, client_id=ccfg.client_id,
                                          log=EventLog(f"CLIENT {ccfg.client_id}", cfg.event_log, quiet=True))
    elif world > 1 and use_cuda and cfg.transport == "nccl":
        transport = NcclTransport()
    elif world > 1:
        from ..utils.faults import FaultPlan

        plan = FaultPlan.parse(cfg.faults) if cfg.faults else FaultPlan.from_env()
        transport = _StoreTransport(_coordinator(cfg, rank, world), rank, plan, ccfg.client_id)
    else:
        transport = LoopbackTransport(1)
    try:
        run_client(ccfg, transport, backend_kwargs=backend_kwargs, reporter=reporter)
        if world > 1:
            dist.barrier()
        if rank == 0 and reporter:
            # the server artefact: the aggregated model as a reference-compatible state_dict
            pth = os.path.join(cfg.out_dir, f"client{ccfg.client_id}_model.pth")
            if os.path.exists(pth):
                torch.save(torch.load(pth, map_location="cpu"), os.path.join(cfg.out_dir, "ddos_distilbert_model.pth"))
    finally:
        transport.close()
        if world > 1:
            dist.destroy_process_group()


def _coordinator(cfg: FedConfig, rank: int, world: int) -> RoundCoordinator:
    """Round agreement over the job's rendezvous store (owned by the launcher process, not by any client)."""
    import torch.distributed as dist
    from torch.distributed import distributed_c10d as c10d

    store = dist.PrefixStore("dcad_rounds", c10d._get_default_store())
    deadline = cfg.round_timeout if cfg.round_timeout is not None else cfg.timeout
    return RoundCoordinator(store, rank, world, deadline_s=deadline, min_clients=cfg.min_clients)


class _StoreTransport(Transport):
    """CPU multi-process FedAvg (plumbing tier, no GPU): the participants of a round (fed/rounds.py: deadline +
    quorum, so a client that sits the round out does not block the others) publish their state_dicts in the
    rendezvous store and every participant averages them with the reference's arithmetic in client order."""
    name = "store"

    def __init__(self, coordinator: RoundCoordinator, rank: int, faults=None, client_id: int = 0):
        self.coord, self.rank, self.faults, self.client_id = coordinator, rank, faults, client_id
        self.round_idx = 0

    def aggregate(self, backend) -> bool:
        import io

        from ..parallel.fedavg import fedavg_state_dicts_cpu

        rnd, self.round_idx = self.round_idx, self.round_idx + 1
        take_part = not (self.faults and not self.faults.before_upload(self.client_id, rnd))
        store = self.coord.store
        chunk = 4 << 20  # the TCP store refuses values above 8 MiB
        if take_part:
            buf = io.BytesIO()
            torch.save({k: v.cpu() for k, v in backend.state_dict().items()}, buf)
            raw = buf.getvalue()
            n = (len(raw) + chunk - 1) // chunk
            for i in range(n):
                store.set(f"sd/r{rnd}/{self.rank}/{i}", raw[i * chunk: (i + 1) * chunk])
            store.set(f"sd/r{rnd}/{self.rank}/n", str(n).encode())
        members = self.coord.agree(rnd, take_part=take_part)
        if members is None:
            return False

        def fetch(m: int):
            n = int(store.get(f"sd/r{rnd}/{m}/n").decode())
            return torch.load(io.BytesIO(b"".join(store.get(f"sd/r{rnd}/{m}/{i}") for i in range(n))), weights_only=True)

        sds = [fetch(m) for m in members]
        backend.load_state_dict(fedavg_state_dicts_cpu(sds))
        return True
