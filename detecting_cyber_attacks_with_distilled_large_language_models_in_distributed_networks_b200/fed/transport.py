"""Aggregation transports behind one interface: ``aggregate(backend) -> bool`` makes the backend's parameters the
FedAvg of all clients' parameters (True on success).

    NvSwitchTransport   product path: peer-memory kernel over NVLink/NVSwitch (parallel/p2p.py), in place, no NCCL
    NcclTransport       measured baseline: ncclAllReduce + scale + cast
    SocketTransport     cross-host fallback: the reference's wire protocol (parallel/socket_transport.py)
    LoopbackTransport   in-process fake for tests: N backends in one process, CPU arithmetic of the reference
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..parallel import socket_transport as st
from ..report.log import EventLog


class Transport:
    name = "none"

    def aggregate(self, backend) -> bool:  # pragma: no cover - interface
        return False

    def close(self) -> None:
        pass


class SocketTransport(Transport):
    name = "socket"

    def __init__(self, host="localhost", port_recv=st.PORT_RECEIVE, port_send=st.PORT_SEND, timeout=st.TIMEOUT,
                 max_retries=5, gzip_level=0, log: Optional[EventLog] = None, faults=None, client_id: int = 0):
        self.host, self.port_recv, self.port_send = host, port_recv, port_send
        self.timeout, self.max_retries, self.gzip_level = timeout, max_retries, gzip_level
        self.log = log
        self.faults, self.client_id, self.round_idx = faults, client_id, 0  # utils.faults.FaultPlan (optional)

    def aggregate(self, backend) -> bool:
        rnd, self.round_idx = self.round_idx, self.round_idx + 1
        if not st.send_model(backend.state_dict(), self.host, self.port_recv, self.timeout, self.gzip_level, self.log,
                             faults=self.faults, client_id=self.client_id, round_idx=rnd):
            return False
        agg = st.receive_aggregated_model(self.host, self.port_send, self.max_retries, self.timeout, self.log)
        if agg is None:
            return False
        backend.load_state_dict(agg)
        return True


class NvSwitchTransport(Transport):
    """The backend's engine must have been built on ``group.arena_tensor()`` / ``group.mirror_tensor()``."""
    name = "nvswitch"

    def __init__(self, group, mode: int = 0, weights: Optional[Sequence[float]] = None):
        self.group, self.mode, self.weights = group, mode, weights

    def aggregate(self, backend) -> bool:
        self.group.fedavg(self.mode, weights=self.weights)
        torch.cuda.current_stream().synchronize()
        if self.group.error_flag():
            return False
        if self.mode == 1:  # one-shot wrote the separate output buffer: adopt it
            eng = backend.engine
            eng.arena.flat.copy_(self.group.out_tensor())
            eng.refresh_bf16()
        return True

    def close(self) -> None:
        self.group.close()


class NvSwitchDPTransport(Transport):
    """Clients that span several GPUs (``parallel/dp.py``): gather the sharded fp32 GEMM weights inside every client
    group, then one grouped FedAvg kernel per GPU averages one arena per client into all GPUs of all clients."""
    name = "nvswitch-dp"

    def __init__(self, dpg, weights: Optional[Sequence[float]] = None):
        self.dpg, self.weights = dpg, weights

    def aggregate(self, backend) -> bool:
        self.dpg.allgather_fp32_gemm_params()
        self.dpg.fedavg_clients(self.weights)
        torch.cuda.current_stream().synchronize()
        return self.dpg.error_flag() == 0

    def close(self) -> None:
        self.dpg.close()


class NcclTransport(Transport):
    name = "nccl"

    def aggregate(self, backend) -> bool:
        from ..parallel.nccl_baseline import nccl_fedavg_

        eng = backend.engine
        nccl_fedavg_(eng.arena.flat, eng.arena.bf16, eng.arena.layout.gemm_end)
        return True


class LoopbackTransport(Transport):
    """Test double: every client of one process registers here; ``aggregate`` of the last arriving client averages
    all registered state_dicts with the reference's CPU arithmetic and loads the result into each backend."""
    name = "loopback"

    def __init__(self, num_clients: int, faults=None, min_clients: Optional[int] = None):
        self.num_clients = num_clients
        self.min_clients = num_clients if min_clients is None else min_clients
        self.faults = faults            # utils.faults.FaultPlan: dropped clients never register
        self.pending: List = []
        self.round_idx = 0

    def aggregate(self, backend, client_id: int = 0) -> bool:
        if self.faults and not self.faults.before_upload(client_id, self.round_idx):
            return False
        self.pending.append(backend)
        if len(self.pending) < self.num_clients:
            return True
        return self.finish_round()

    def finish_round(self) -> bool:
        """Close the round with whoever registered (the deadline of a real server): aggregates when the quorum
        ``min_clients`` is met, otherwise leaves every model untouched and reports failure."""
        from ..parallel.fedavg import fedavg_state_dicts_cpu

        pending, self.pending = self.pending, []
        self.round_idx += 1
        if len(pending) < max(self.min_clients, 1):
            return False
        agg = fedavg_state_dicts_cpu([b.state_dict() for b in pending])
        for b in pending:
            b.load_state_dict(agg)
        return True
