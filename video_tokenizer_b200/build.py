"""Build libb200vq.so in-tree with nvcc for sm_100a (no torch headers involved).

    python -m video_tokenizer_b200.build [--force] [--verbose]

The library is a plain C-ABI shared object (include/b200vq.h); the Python side binds it
with ctypes.  nvcc cross-compiles without a GPU, so this also runs on the CPU-only
build container.

The translation units compile in parallel into per-process temporary objects and the
finished library is moved into place atomically.  Concurrent builders (one process per GPU
under torchrun on a fresh checkout) serialise on a file lock and re-check staleness after
acquiring it, so exactly one of them compiles and none can load a half-written file.
"""
from __future__ import annotations

import concurrent.futures
import contextlib
import fcntl
import os
import shutil
import subprocess
import sys
import tempfile

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libb200vq.so")
LOCK = os.path.join(PKG, ".libb200vq.lock")
SOURCES = ["b200vq.cu", "b200vq_tc.cu", "b200vq_linear.cu", "b200vq_entropy.cu"]
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libb200vq.so")


def _deps():
    out = [os.path.join(ROOT, "include", "b200vq.h")]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cu", ".cuh", ".h")):
            out.append(os.path.join(CSRC, f))
    return out


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in _deps())


@contextlib.contextmanager
def _build_lock():
    """Exclusive advisory lock shared by every process that may build into this tree."""
    fd = os.open(LOCK, os.O_CREAT | os.O_RDWR, 0o644)
    try:
        fcntl.flock(fd, fcntl.LOCK_EX)
        yield
    finally:
        try:
            fcntl.flock(fd, fcntl.LOCK_UN)
        finally:
            os.close(fd)


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    return res.stderr


def _compile_and_link(verbose: bool) -> None:
    nvcc = _nvcc()
    common = [nvcc, "-O3", "-std=c++17", "-lineinfo", *ARCH_FLAGS, "-Xcompiler", "-fPIC",
              "-I", os.path.join(ROOT, "include"), "-I", CSRC]
    if verbose:
        common += ["-Xptxas", "-v"]
    # objects and the linked library get names private to this process, in the package directory
    # (same filesystem as LIB, so the final os.replace is atomic)
    with tempfile.TemporaryDirectory(prefix=".b200vq_build_", dir=PKG) as tmp:
        objs = [os.path.join(tmp, os.path.splitext(s)[0] + ".o") for s in SOURCES]
        with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as pool:
            logs = list(pool.map(lambda so: _run(common + ["-c", os.path.join(CSRC, so[0]), "-o", so[1]]),
                                 zip(SOURCES, objs)))
        if verbose:
            print("\n".join(logs))
        out = os.path.join(tmp, "libb200vq.so")
        _run([nvcc, *ARCH_FLAGS, "-shared", "-Xcompiler", "-fPIC", *objs, "-o", out, "-lcudart"])
        os.replace(out, LIB)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    with _build_lock():
        if force or is_stale():   # another process may have finished the build while we waited
            _compile_and_link(verbose)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
