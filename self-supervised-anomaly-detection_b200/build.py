"""Build libssad_gen.so in-tree with nvcc for sm_100a (no GPU needed: nvcc cross-compiles)."""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libssad_gen.so")
OBJ_DIR = os.path.join(HERE, "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-shared", "-Xptxas", "-v"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "ssad.h")]
    return any(os.path.getmtime(p) > t for p in deps)


def _compile_one(src: str, hdr_mtime: float, force: bool):
    """nvcc -c one .cu -> build/<name>.o (skipped when the object is newer than the source and every header)."""
    obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
    if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), hdr_mtime):
        return obj, 0, ""
    cmd = [NVCC] + [f for f in FLAGS if f != "-shared"] + ["-c", "-o", obj, src]
    res = subprocess.run(cmd, capture_output=True, text=True)
    return obj, res.returncode, res.stdout + res.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "ssad.h")]
    hdr_mtime = max(os.path.getmtime(p) for p in hdrs)
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(lambda s: _compile_one(s, hdr_mtime, force), sources()))
    log = "".join(r[2] for r in results)
    failed = any(r[1] != 0 for r in results)
    if verbose or failed:
        sys.stderr.write(log)
    if failed:
        raise RuntimeError("nvcc failed building libssad_gen.so")
    res = subprocess.run([NVCC, "-shared", "-o", LIB] + [r[0] for r in results], capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed linking libssad_gen.so")
    if log.strip():
        with open(os.path.join(LIB_DIR, "ptxas.log"), "a" if not force else "w") as f:
            f.write(log)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
