"""Builds llamafile_b200/libtinyblas_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m llamafile_b200.build [--force] [--verbose]
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = PKG / "_obj"
LIB = PKG / "libtinyblas_b200.so"
HOST_SHIM = PKG / "libtb200_loader.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _flags():
    """FLAGS + optional -D overrides for experiments (TB200_NVCC_DEFS="-DTB200_PLAN_WARPS=16 -DTB200_PLAN_STAGES=2")."""
    return FLAGS + os.environ.get("TB200_NVCC_DEFS", "").split()


def _sig(paths) -> str:
    h = hashlib.sha1()
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(_flags()).encode())
    return h.hexdigest()


def sources():
    return sorted(CSRC.glob("*.cu"))


def build(force: bool = False, verbose: bool = False) -> Path:
    """Idempotent and safe to call from several processes at once (one rank per GPU under torchrun): an exclusive file lock
    serialises them, and whoever comes second finds the stamp up to date."""
    import fcntl
    OBJ.mkdir(exist_ok=True)
    with open(OBJ / ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            return _build_locked(force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(force: bool, verbose: bool) -> Path:
    srcs = sources()
    deps = srcs + sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.c")) + [PKG.parent / "include" / "tinyblas_b200.h"]
    sig = _sig(deps)
    stamp = OBJ / ".stamp"
    if not force and LIB.exists() and HOST_SHIM.exists() and stamp.exists() and stamp.read_text() == sig:
        return LIB
    OBJ.mkdir(exist_ok=True)

    def cc(src: Path):
        obj = OBJ / (src.stem + ".o")
        cmd = [NVCC, *_flags(), "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0 or verbose:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(cc, srcs))
    subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", str(LIB), *map(str, objs), "-Xcompiler", "-fPIC", "-Xlinker", "--no-undefined", "-lpthread", "-ldl"])
    # the plain-C dlopen() loader a llamafile maintainer would drop next to llamafile/cuda.c
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-Wall", "-I", str(PKG.parent / "include"),
                           "-o", str(HOST_SHIM), str(CSRC / "tb200_loader.c"), "-ldl"])
    stamp.write_text(sig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
