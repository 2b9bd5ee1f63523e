"""Sample GPU clocks / throttle reasons DURING a timed region (B200_PROFILING.md recipe).

``ClockSampler`` polls NVML in a background thread every few milliseconds when ``pynvml`` is importable (a timed region
of a fraction of a second still gets dozens of samples) and falls back to an ``nvidia-smi -lms`` subprocess otherwise.
Both report the same dictionary: median SM clock under load, maximum SM clock, peak power, active slowdown reasons."""
from __future__ import annotations

import os
import statistics
import subprocess
import tempfile
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: Optional[int] = None, period_ms: int = 100, nvml_period_ms: int = 5):
        self.gpu_index, self.period_ms, self.nvml_period_ms = gpu_index, period_ms, nvml_period_ms
        self.proc: Optional[subprocess.Popen] = None
        self.path: Optional[str] = None
        self._thread = None
        self._stop = None
        self._nvml_rows: List[tuple] = []

    # ---- NVML path ----
    def _nvml_handle(self):
        import pynvml

        pynvml.nvmlInit()
        idx = self.gpu_index or 0
        try:  # honour CUDA_VISIBLE_DEVICES: map the CUDA ordinal to the physical device through its UUID
            import torch

            uuid = str(torch.cuda.get_device_properties(idx).uuid)
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:  # noqa: BLE001
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)

    def _nvml_loop(self, pynvml, h) -> None:
        reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(pynvml, "nvmlDeviceGetCurrentClocksThrottleReasons")
        bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
        while not self._stop.is_set():
            try:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
                pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                rs = int(reasons_fn(h))
                self._nvml_rows.append((float(sm), float(mx), pw, tuple(n for n, b in bits.items() if rs & b)))
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.nvml_period_ms / 1000.0)

    def start(self) -> "ClockSampler":
        if os.environ.get("DCAD_CLOCKS_NVML", "1") != "0":
            try:
                import threading

                pynvml, h = self._nvml_handle()
                self._stop = threading.Event()
                self._thread = threading.Thread(target=self._nvml_loop, args=(pynvml, h), daemon=True)
                self._thread.start()
                return self
            except Exception:  # noqa: BLE001
                self._thread = None
        fd, self.path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
        os.close(fd)
        cmd = ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms)]
        if self.gpu_index is not None:
            cmd += ["-i", str(self.gpu_index)]
        try:
            self.proc = subprocess.Popen(cmd, stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None
        return self

    def stop(self) -> Dict:
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
            rows = self._nvml_rows
            reasons = sorted({r for row in rows for r in row[3]})
            return {"sm_mhz": statistics.median(r[0] for r in rows) if rows else None,
                    "sm_max_mhz": max(r[1] for r in rows) if rows else None,
                    "power_w_max": round(max(r[2] for r in rows), 2) if rows else None,
                    "reasons": reasons, "samples": len(rows), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm: List[float] = []
        sm_max: List[float] = []
        power: List[float] = []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    sm_max.append(float(parts[2]))
                    power.append(float(parts[3]))
                except ValueError:
                    continue
                for n, v in zip(names, parts[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        finally:
            try:
                os.unlink(self.path)
            except OSError:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(sm_max) if sm_max else None,
                "power_w_max": max(power) if power else None, "reasons": sorted(reasons), "samples": len(sm)}
