"""Build recipe for libpcseg_ops.so (sm_100a only) — explicit nvcc, in-tree output.

Replaces the reference's tf_ops/BuildTFOps.sh / BuildOpsV2.sh (nvcc -O2 with no -arch, linked
against TensorFlow).  Here: no TensorFlow, no torch headers, plain C ABI (include/pcseg_ops.h).
"""
from __future__ import annotations

import contextlib
import fcntl
import hashlib
import os
import subprocess
import sys
import tempfile
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


def _atomic_write(path: Path, text: str) -> None:
    fd, tmp = tempfile.mkstemp(dir=str(path.parent), prefix=path.name + ".")
    with os.fdopen(fd, "w") as f:
        f.write(text)
    os.replace(tmp, path)


def _compile(src: Path, force: bool = False) -> Path:
    obj = OBJDIR / (src.stem + ".o")
    stamp = OBJDIR / (src.stem + ".sha256")
    dig = _src_digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text().strip() == dig:
        return obj  # unchanged source + headers + flags: keep the object
    tmp = OBJDIR / f".{src.stem}.{os.getpid()}.o"  # never leave a half-written object under the final name
    cmd = [NVCC, *ARCH, *CFLAGS, "-c", str(src), "-o", str(tmp)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    (OBJDIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
    if r.returncode != 0:
        tmp.unlink(missing_ok=True)
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stderr}")
    os.replace(tmp, obj)
    _atomic_write(stamp, dig)
    return obj


@contextlib.contextmanager
def _build_lock():
    """One builder at a time across processes (torchrun starts one process per GPU and each imports the
    package): the others block here, then find the stamp current and return without compiling."""
    LIBDIR.mkdir(parents=True, exist_ok=True)
    with open(LIBDIR / ".build.lock", "w") as f:
        fcntl.flock(f, fcntl.LOCK_EX)
        try:
            yield
        finally:
            fcntl.flock(f, fcntl.LOCK_UN)


def _current(dig: str) -> bool:
    stamp = LIBDIR / "build.sha256"
    return LIB.exists() and stamp.exists() and stamp.read_text().strip() == dig


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link libpcseg_ops.so.  Idempotent (content hash) unless
    `force`; serialised across processes by a file lock; objects and the library appear under their final
    names only when complete (os.replace), so a concurrent dlopen never sees a partial file."""
    dig = _digest()
    if not force and _current(dig):
        return LIB
    if not Path(NVCC).exists():
        raise RuntimeError(f"nvcc not found at {NVCC}; libpcseg_ops.so cannot be built")
    with _build_lock():
        if not force and _current(dig):  # another process built it while this one waited for the lock
            return LIB
        OBJDIR.mkdir(parents=True, exist_ok=True)
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
            objs = list(ex.map(lambda s: _compile(s, force), sources()))
        tmp = LIBDIR / f".libpcseg_ops.{os.getpid()}.so"
        cmd = [NVCC, *ARCH, "-shared", "-o", str(tmp), *map(str, objs), "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            tmp.unlink(missing_ok=True)
            raise RuntimeError(f"link failed:\n{r.stderr}")
        os.replace(tmp, LIB)
        _atomic_write(LIBDIR / "build.sha256", dig)
    if verbose:
        print(f"built {LIB}", file=sys.stderr)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
