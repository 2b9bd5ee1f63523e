"""Aggregation transports behind one interface: ``aggregate(backend) -> bool`` makes the backend's parameters the
FedAvg of all clients' parameters (True on success).  ``aggregate_async(backend)`` starts the same round and returns a
handle whose ``wait()`` gives the result: the NVSwitch transports launch their kernel on a side stream and return at
once, so the caller can stage the next round's batches and write the local round's files while the GPUs average
(BASELINE config 3; the reference is strictly serial, server.py:116-137 / client1.py:391-395).

    NvSwitchTransport   product path: peer-memory kernel over NVLink/NVSwitch (parallel/p2p.py), in place, no NCCL;
                        two-shot P2P, one-shot, or NVLS multimem (auto: NVLS from 4 clients up when the host has
                        NVLink multicast); deadline + quorum through fed/rounds.py
    NvSwitchBf16Transport  same, exchanging bf16 copies of the parameters (half the NVLink bytes)
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


class Pending:
    """Result of ``aggregate_async``: ``wait()`` blocks until the round is over and returns its success."""

    def __init__(self, result: Optional[bool] = None, finish=None):
        self._result, self._finish = result, finish

    def wait(self) -> bool:
        if self._result is None:
            self._result = bool(self._finish())
            self._finish = None
        return self._result


def _aggregate_async_default(self, backend) -> Pending:
    # transports without a device-side round have nothing to overlap: the round runs inside wait(), i.e. after the
    # caller has written the local model's files (the round replaces the parameters those files describe)
    return Pending(None, lambda: self.aggregate(backend))


Transport.aggregate_async = _aggregate_async_default

AUTO = -1  # NvSwitchTransport mode: NVLS when the group has a multicast mapping and >= 4 clients, else two-shot


class NvSwitchTransport(Transport):
    """The backend's engine must have been built on ``group.arena_tensor()`` / ``group.mirror_tensor()``.

    ``coordinator`` (fed/rounds.RoundCoordinator): every round first agrees on the participants (deadline + quorum), so
    a client that died or straggles costs the others one deadline instead of the round; without it every rank of the
    group is expected (bench, tests).  ``faults`` (utils/faults.FaultPlan) can make this client sit rounds out."""
    name = "nvswitch"

    def __init__(self, group, mode: int = 0, weights: Optional[Sequence[float]] = None, coordinator=None,
                 faults=None, client_id: int = 0, local_cast: bool = True, log: Optional[EventLog] = None):
        self.group, self.mode, self.weights = group, mode, weights
        self.coordinator, self.faults, self.client_id = coordinator, faults, client_id
        self.local_cast, self.log = local_cast, log
        self.round_idx = 0
        self._stream: Optional[torch.cuda.Stream] = None
        self.last_members: Optional[List[int]] = None

    def resolved_mode(self, n_members: int) -> int:
        from ..parallel.p2p import NVLS, TWO_SHOT

        if self.mode != AUTO:
            return self.mode
        full = n_members == self.group.world
        # measured (profiles/README.md): two-shot wins at 2 GPUs (0.52 vs 0.84 ms), NVLS from 4 up (0.73 vs 0.95 at 8)
        return NVLS if (getattr(self.group, "has_nvls", False) and full and n_members >= 4) else TWO_SHOT

    def aggregate_async(self, backend) -> Pending:
        rnd, self.round_idx = self.round_idx, self.round_idx + 1
        take_part = not (self.faults and not self.faults.before_upload(self.client_id, rnd))
        members = None
        if self.coordinator is not None:
            members = self.coordinator.agree(rnd, take_part=take_part)
            if self.log is not None:
                self.log.event("round_members", round=rnd + 1, members=members, waited_s=self.coordinator.last_wait_s)
            if members is None:
                return Pending(False)
        elif not take_part:
            return Pending(False)
        self.last_members = members
        n = len(members) if members is not None else self.group.world
        mode = self.resolved_mode(n)
        weights = self.weights
        if weights is not None and members is not None and len(weights) != n:
            tot = float(sum(weights[m] for m in members))
            weights = [weights[m] / tot for m in members]
        cur = torch.cuda.current_stream()
        if self._stream is None:
            self._stream = torch.cuda.Stream()
        side = self._stream
        side.wait_stream(cur)  # local training (and the state_dict snapshot taken for the local files) precede the round
        kw = dict(weights=weights, participants=members, epoch=rnd + 1, stream=side)
        if hasattr(self.group, "multicast_ptr"):
            kw["local_cast"] = self.local_cast
        t_launch = torch.cuda.Event(enable_timing=True)
        t_done = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(side):
            t_launch.record(side)
            self.group.fedavg(mode, **kw)
            t_done.record(side)

        def finish() -> bool:
            torch.cuda.current_stream().wait_event(t_done)  # whatever the caller enqueues next sees the averaged model
            t_done.synchronize()
            self.last_round_ms = t_launch.elapsed_time(t_done)
            if self.group.error_flag():
                return False
            if mode == 1:  # one-shot wrote the separate output buffer: adopt it
                eng = backend.engine
                eng.arena.flat.copy_(self.group.out_tensor())
                eng.refresh_bf16()
            return True

        return Pending(None, finish)

    def aggregate(self, backend) -> bool:
        return self.aggregate_async(backend).wait()

    def close(self) -> None:
        self.group.close()


class NvSwitchBf16Transport(Transport):
    """bf16 parameter exchange (BASELINE configs 3 and 5): every client casts its fp32 master arena into a bf16
    exchange arena in peer memory, the two-shot kernel averages those (fp32 accumulation, one rounding of the mean),
    and the mean is cast back into the fp32 master arena and the bf16 compute copy.  Half the NVLink bytes of the fp32
    round; the price is that the averaged master weights carry bf16 precision (8 mantissa bits) after every round --
    the reference averages in fp32 (server.py:73-76), so this is an opt-in (``--transport nvswitch-bf16``)."""
    name = "nvswitch-bf16"

    def __init__(self, xgroup, coordinator=None, faults=None, client_id: int = 0, log: Optional[EventLog] = None):
        self.xgroup, self.coordinator, self.faults, self.client_id, self.log = xgroup, coordinator, faults, client_id, log
        self.round_idx = 0

    def aggregate(self, backend) -> bool:
        from ..ops import native as nv
        from ..parallel.p2p import TWO_SHOT

        rnd, self.round_idx = self.round_idx, self.round_idx + 1
        take_part = not (self.faults and not self.faults.before_upload(self.client_id, rnd))
        members = None
        if self.coordinator is not None:
            members = self.coordinator.agree(rnd, take_part=take_part)
            if members is None:
                return False
        elif not take_part:
            return False
        eng = backend.engine
        x = self.xgroup.arena_tensor()                       # bf16 [total]
        nv.cast_f32_bf16(eng.arena.flat, x)
        self.xgroup.fedavg(TWO_SHOT, participants=members, with_mirror=False, epoch=rnd + 1)
        torch.cuda.current_stream().synchronize()
        if self.xgroup.error_flag():
            return False
        eng.arena.flat.copy_(x)                              # bf16 -> fp32 master weights
        eng.arena.bf16.copy_(x[: eng.arena.layout.gemm_end])
        return True

    def close(self) -> None:
        self.xgroup.close()


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
