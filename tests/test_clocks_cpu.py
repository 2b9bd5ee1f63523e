"""utils/clocks.py: the sampler that bench.py wraps around the timed region.  Without a GPU it must come back with a
well-formed record (no exception, no leaked temp file), and the nvidia-smi fallback must parse the query's CSV rows."""
import glob
import importlib
import os
import stat
import tempfile
import time

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
clocks = importlib.import_module(f"{PKG}.utils.clocks")


def test_sampler_degrades_without_a_gpu(monkeypatch):
    monkeypatch.setenv("DCAD_CLOCKS_NVML", "0")
    monkeypatch.setenv("PATH", "/nonexistent")                       # no nvidia-smi either
    before = set(glob.glob(os.path.join(tempfile.gettempdir(), "clocks_*.csv")))
    rec = clocks.ClockSampler(0).start().stop()
    assert rec["sm_mhz"] is None and rec["reasons"] == [] and rec["samples"] == 0
    leaked = set(glob.glob(os.path.join(tempfile.gettempdir(), "clocks_*.csv"))) - before
    for p in leaked:                                                  # tolerated, but do not let them pile up
        os.unlink(p)


def test_nvidia_smi_fallback_parses_clock_rows(tmp_path, monkeypatch):
    fake = tmp_path / "nvidia-smi"
    fake.write_text("#!/bin/sh\n"
                    "echo '0, 1965, 1965, 512.30, 0x0000000000000004, Not Active, Not Active, Not Active, Active'\n"
                    "echo '0, 1935, 1965, 640.10, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active'\n"
                    "echo '0, 1905, 1965, 600.00, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active'\n"
                    "sleep 30\n")
    fake.chmod(fake.stat().st_mode | stat.S_IEXEC)
    monkeypatch.setenv("DCAD_CLOCKS_NVML", "0")
    monkeypatch.setenv("PATH", f"{tmp_path}{os.pathsep}/usr/bin{os.pathsep}/bin")
    s = clocks.ClockSampler(0).start()
    time.sleep(0.5)
    rec = s.stop()
    assert rec["samples"] == 3
    assert rec["sm_mhz"] == 1935.0 and rec["sm_max_mhz"] == 1965.0 and rec["power_w_max"] == 640.10
    assert rec["reasons"] == ["sw_power_cap"]                         # kept and reported, not a rejection reason
