"""Builds csrc/ into cs744-distributed-data-parallel_b200/_native/libddpb200.so with nvcc.

sm_100a only (tcgen05 / TMEM / TMA need the arch-specific target).  Sources compile in parallel,
objects are cached by content hash so an unchanged file is not recompiled.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import pathlib
import shutil
import subprocess
import sys

ROOT = pathlib.Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "cs744-distributed-data-parallel_b200" / "_native"
OBJ_DIR = ROOT / "build" / "obj"
LIB = OUT_DIR / "libddpb200.so"

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def sources():
    return sorted(list((CSRC / "kernels").glob("*.cu")) + list((CSRC / "runtime").glob("*.cpp")) +
                  list((CSRC / "runtime").glob("*.cu")))


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(path: pathlib.Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(path.read_bytes())
    for hdr in sorted(list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list((ROOT / "include").glob("*.h"))):
        h.update(hdr.read_bytes())
    return h.hexdigest()[:16]


def _compile(src: pathlib.Path, verbose: bool) -> pathlib.Path:
    obj = OBJ_DIR / f"{src.stem}.{_digest(src)}.o"
    if obj.exists():
        return obj
    for old in OBJ_DIR.glob(f"{src.stem}.*.o"):
        old.unlink()
    cmd = [_nvcc(), *NVCC_FLAGS, "-x", "cu", "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = OBJ_DIR / f"{src.stem}.ptxas.log"
    log.write_text(res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stderr[-6000:]}")
    if verbose:
        print(f"[build] {src.name} ok")
    return obj


def build(verbose: bool = True, force: bool = False) -> pathlib.Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    if force:
        for o in OBJ_DIR.glob("*.o"):
            o.unlink()
    srcs = sources()
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    stamp = hashlib.sha256(" ".join(sorted(o.name for o in objs)).encode()).hexdigest()[:16]
    stamp_file = OUT_DIR / "build.stamp"
    if LIB.exists() and stamp_file.exists() and stamp_file.read_text() == stamp and not force:
        return LIB
    cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *map(str, objs), "-ldl"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stderr[-4000:]}")
    stamp_file.write_text(stamp)
    if verbose:
        print(f"[build] linked {LIB}")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
