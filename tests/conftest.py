"""Test configuration: the ``gpu`` marker (tests that need a real B200) and shared fixtures.

The reference has no tests at all; the strategy here is: CPU tier = plumbing + parity against the
reference's own Python functions and HF DistilBERT; GPU tier = every CUDA kernel against a plain PyTorch fp32
reference of the same op, plus the end-to-end engine.
"""
from __future__ import annotations

import importlib.util
import sys
import types
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
REFERENCE_DIRS = [Path("/root/reference"), ROOT / "baseline" / "_ref"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def reference_dir():
    for d in REFERENCE_DIRS:
        if (d / "client1.py").exists():
            return d
    return None


def _stub_plotting_modules():
    """matplotlib / seaborn are not installed on this image; the reference imports them at module scope."""
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:  # noqa: BLE001
                m = types.ModuleType(name)
                m.use = lambda *a, **k: None
                sys.modules[name] = m
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def load_reference_module(name: str):
    """Import ``client1`` / ``client2`` / ``server`` from the read-only reference tree (oracle for parity tests)."""
    d = reference_dir()
    if d is None:
        pytest.skip("reference tree not available")
    _stub_plotting_modules()
    spec = importlib.util.spec_from_file_location(f"_reference_{name}", d / f"{name}.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def ref_client1():
    return load_reference_module("client1")


@pytest.fixture(scope="session")
def ref_server():
    return load_reference_module("server")
