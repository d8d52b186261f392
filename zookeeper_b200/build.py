"""Compile the CUDA sources under ``zookeeper_b200/csrc`` into ``libzkb200.so`` (sm_100a only).

nvcc cross-compiles without a GPU, so this runs on the CPU-only build box as well as on the B200
box.  The shared library is placed in-tree next to this file so that it travels with a snapshot of
the repository.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_PATH = PKG_DIR / "libzkb200.so"
_STAMP = PKG_DIR / "csrc" / ".build_stamp"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "-DZKB_BUILD=1",
]
# ZKB_MEASURE=1: a measurement build whose GEMM launchers read tile-shape overrides and the
# "drop the MMA issue / drop the TMA loads" switches from the environment (scripts/bench_gemm.py).
# The product build (build(), the tests, bench.py) has none of them compiled in.
if os.environ.get("ZKB_MEASURE") == "1":
    NVCC_FLAGS.append("-DZKB_MEASURE=1")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libzkb200.so cannot be built")


def _sources():
    return sorted(CSRC.glob("*.cu"))


def _fingerprint() -> str:
    h = hashlib.sha256()
    for f in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_native(force: bool = False, verbose: bool = False) -> Path:
    """Build ``libzkb200.so`` if sources changed; return its path."""
    fp = _fingerprint()
    if not force and LIB_PATH.exists() and _STAMP.exists() and _STAMP.read_text() == fp:
        return LIB_PATH
    nvcc = _nvcc()
    objdir = PKG_DIR / "csrc" / "_obj"
    objdir.mkdir(exist_ok=True)
    objs = []
    procs = []
    for src in _sources():
        obj = objdir / (src.stem + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(str(obj))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}:\n{out.decode(errors='replace')}")
        if verbose and out:
            print(out.decode(errors="replace"))
    link = [nvcc, "-shared", "-o", str(LIB_PATH), *objs, "-cudart", "static",
            "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout.decode(errors='replace')}")
    _STAMP.write_text(fp)
    return LIB_PATH


if __name__ == "__main__":
    import sys

    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv))
