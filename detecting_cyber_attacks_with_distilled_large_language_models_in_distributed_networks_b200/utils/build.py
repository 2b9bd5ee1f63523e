"""In-tree build of the native sm_100a library (``_native/libdcad_b200.so``).

The reference ships no native code at all; every device instruction it runs comes from stock
PyTorch.  This framework owns its hot kernels, so there is a build step: plain ``nvcc`` per ``.cu`` file (seconds
each, cross-compiles without a GPU), linked into one shared object with a C ABI (``include/dcad_b200.h``) that
``ops/native.py`` loads through ctypes.  The ``.so`` lives in-tree so it travels with a repository snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent.parent
REPO_ROOT = PKG_DIR.parent
CSRC = REPO_ROOT / "csrc"
INCLUDE = REPO_ROOT / "include"
NATIVE_DIR = PKG_DIR / "_native"
LIB_PATH = NATIVE_DIR / "libdcad_b200.so"
OBJ_DIR = NATIVE_DIR / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def find_nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC or put /usr/local/cuda/bin on PATH")
    return cand


def _sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _deps_digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))):
        h.update(p.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(nvcc: str, src: Path, digest: str, verbose: bool) -> tuple[Path, str]:
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".stamp")
    key = hashlib.sha256(src.read_bytes() + digest.encode()).hexdigest()
    if obj.exists() and stamp.exists() and stamp.read_text() == key:
        return obj, ""
    cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(key)
    (OBJ_DIR / (src.stem + ".ptxas.txt")).write_text(r.stderr)
    if verbose:
        print(f"[build] compiled {src.name}", file=sys.stderr)
    return obj, r.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libdcad_b200.so``.  Incremental (content-hashed)."""
    nvcc = find_nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for p in OBJ_DIR.glob("*.stamp"):
            p.unlink()
    digest = _deps_digest()
    srcs = _sources()
    if not srcs:
        raise RuntimeError(f"no CUDA sources under {CSRC}")
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile_one(nvcc, s, digest, verbose), srcs))
    objs = [o for o, _ in results]
    link_key = hashlib.sha256("".join(sorted((OBJ_DIR / (o.stem + ".stamp")).read_text() for o in objs)).encode())
    link_stamp = OBJ_DIR / "link.stamp"
    if LIB_PATH.exists() and link_stamp.exists() and link_stamp.read_text() == link_key.hexdigest():
        return LIB_PATH
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB_PATH), *map(str, objs),
           "-Xcompiler", "-fPIC", "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    link_stamp.write_text(link_key.hexdigest())
    if verbose:
        print(f"[build] linked {LIB_PATH}", file=sys.stderr)
    return LIB_PATH


HOST_SRC = CSRC / "host"
TOK_LIB_PATH = NATIVE_DIR / "libdcad_tok.so"
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall"]


def build_host(force: bool = False, verbose: bool = False) -> Path:
    """Compile the host-side native code (``csrc/host/*.cpp``: the WordPiece tokenizer) with the C++ compiler."""
    cxx = os.environ.get("CXX") or shutil.which("g++") or shutil.which("c++")
    if not cxx:
        raise RuntimeError("no C++ compiler found (set CXX)")
    srcs = sorted(HOST_SRC.glob("*.cpp"))
    NATIVE_DIR.mkdir(parents=True, exist_ok=True)
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    key = hashlib.sha256(b"".join(p.read_bytes() for p in srcs) + " ".join(CXX_FLAGS).encode()).hexdigest()
    stamp = OBJ_DIR / "host.stamp"
    if not force and TOK_LIB_PATH.exists() and stamp.exists() and stamp.read_text() == key:
        return TOK_LIB_PATH
    r = subprocess.run([cxx, *CXX_FLAGS, *map(str, srcs), "-o", str(TOK_LIB_PATH)], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"host build failed:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(key)
    if verbose:
        print(f"[build] built {TOK_LIB_PATH}", file=sys.stderr)
    return TOK_LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    print(build_host(force="--force" in sys.argv, verbose=True))
