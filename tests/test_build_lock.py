"""CPU: concurrent builders (one process per GPU under torchrun on a fresh checkout) serialise on the build lock and
never leave or load a half-written library (video_tokenizer_b200/build.py)."""
import ctypes
import multiprocessing as mp
import os
import time

from video_tokenizer_b200 import build


def _builder(q):
    try:
        path = build.build()
        lib = ctypes.CDLL(path)
        q.put(int(lib.b200vq_version()))
    except Exception as e:   # pragma: no cover
        q.put(repr(e))


def test_concurrent_builders_share_one_build():
    build.build()
    src = os.path.join(build.CSRC, "b200vq_entropy.cu")
    st = os.stat(src)
    try:
        os.utime(src, (st.st_atime, time.time() + 1))       # the library is now stale for every process
        assert build.is_stale()
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        procs = [ctx.Process(target=_builder, args=(q,)) for _ in range(3)]
        for p in procs:
            p.start()
        results = [q.get(timeout=900) for _ in procs]
        for p in procs:
            p.join(60)
        assert all(isinstance(r, int) and r >= 122 for r in results), results
        assert not build.is_stale()
        leftovers = [f for f in os.listdir(build.PKG) if f.startswith(".b200vq_build_")]
        assert leftovers == [], leftovers
    finally:
        os.utime(src, (st.st_atime, st.st_mtime))
        if build.is_stale():                                 # mtime restored below the library's: never stale here
            build.build()
