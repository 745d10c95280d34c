"""Build recipe for libpcseg_ops.so (sm_100a only) — explicit nvcc, in-tree output.

Replaces the reference's tf_ops/BuildTFOps.sh / BuildOpsV2.sh (nvcc -O2 with no -arch, linked
against TensorFlow).  Here: no TensorFlow, no torch headers, plain C ABI (include/pcseg_ops.h).
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
LIBDIR = PKG / "lib"
LIB = LIBDIR / "libpcseg_ops.so"
OBJDIR = LIBDIR / "obj"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# No --use_fast_math: ComputeVoxelIndex needs the IEEE divide the reference build had
# (tf_ops/BuildOpsV2.sh:6-13 passes no fast-math flag).
CFLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
          "-Wno-deprecated-gpu-targets"]


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "pcseg_ops.h"]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    h.update(" ".join(ARCH + CFLAGS).encode())
    return h.hexdigest()


def _src_digest(src: Path) -> str:
    h = hashlib.sha256(src.read_bytes())
    text = src.read_text()
    for p in sorted(CSRC.glob("*.cuh")):  # the headers this source includes (not the public header: prototypes only)
        if f'#include "{p.name}"' in text:
            h.update(p.read_bytes())
    h.update(" ".join(ARCH + CFLAGS).encode())
    return h.hexdigest()


def _compile(src: Path) -> Path:
    obj = OBJDIR / (src.stem + ".o")
    stamp = OBJDIR / (src.stem + ".sha256")
    dig = _src_digest(src)
    if obj.exists() and stamp.exists() and stamp.read_text().strip() == dig:
        return obj  # unchanged source + headers + flags: keep the object
    cmd = [NVCC, *ARCH, *CFLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    (OBJDIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stderr}")
    stamp.write_text(dig)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link libpcseg_ops.so. Idempotent (content hash)."""
    OBJDIR.mkdir(parents=True, exist_ok=True)
    stamp = LIBDIR / "build.sha256"
    dig = _digest()
    if not force and LIB.exists() and stamp.exists() and stamp.read_text().strip() == dig:
        return LIB
    if not Path(NVCC).exists():
        raise RuntimeError(f"nvcc not found at {NVCC}; libpcseg_ops.so cannot be built")
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(_compile, sources()))
    cmd = [NVCC, *ARCH, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stderr}")
    stamp.write_text(dig)
    if verbose:
        print(f"built {LIB}", file=sys.stderr)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
