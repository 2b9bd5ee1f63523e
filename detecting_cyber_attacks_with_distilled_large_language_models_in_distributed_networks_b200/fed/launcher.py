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
from .transport import LoopbackTransport, NcclTransport, NvSwitchDPTransport, NvSwitchTransport, Transport


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
    elif world > 1 and use_cuda and cfg.transport.startswith("nvswitch"):
        from ..parallel.p2p import ONE_SHOT, TWO_SHOT, PeerGroup

        layout = ArenaLayout(ccfg.model_dims())
        one = cfg.transport.endswith("oneshot")
        group = PeerGroup(rank, world, layout.total, layout.gemm_end, with_out_buffer=one)
        backend_kwargs.update(flat_params=group.arena_tensor(), bf16_params=group.mirror_tensor())
        transport = NvSwitchTransport(group, ONE_SHOT if one else TWO_SHOT)
    elif world > 1 and use_cuda and cfg.transport == "nccl":
        transport = NcclTransport()
    elif world > 1:
        transport = _GlooTransport()
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


class _GlooTransport(Transport):
    """CPU multi-process FedAvg (plumbing tier, no GPU): gather state_dicts on every rank, average with the
    reference's arithmetic in client order."""
    name = "gloo"

    def aggregate(self, backend) -> bool:
        import torch.distributed as dist

        from ..parallel.fedavg import fedavg_state_dicts_cpu

        sds = [None] * dist.get_world_size()
        dist.all_gather_object(sds, {k: v.cpu() for k, v in backend.state_dict().items()})
        backend.load_state_dict(fedavg_state_dicts_cpu(sds))
        return True
