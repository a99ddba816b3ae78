"""CPU: the C-ABI library builds, loads and exports every symbol include/b200vq.h declares."""
import ctypes
import os
import re

from video_tokenizer_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b200vq.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200vq_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    build.build()
    lib = ctypes.CDLL(build.LIB)
    declared = _declared_symbols()
    assert len(declared) >= 14
    for name in declared:
        assert hasattr(lib, name), f"libb200vq.so does not export {name}"
    assert sorted(_lib.SIGNATURES) == declared, "ctypes signatures out of sync with include/b200vq.h"


def test_host_only_entry_points():
    lib = _lib.load()
    assert lib.b200vq_version() == 122
    assert lib.b200vq_max_dim() == 1024
    assert lib.b200vq_error_string(0) == b"ok"
    assert lib.b200vq_error_string(-2) == b"unsupported code dimension"
    n = lib.b200vq_workspace_bytes(4096, 24, 8192)
    assert n >= 4096 * 8 + 16 and n % 256 == 0
    assert lib.b200vq_workspace_bytes(-1, 24, 8192) == 0
    assert lib.b200vq_tc_supported(24, 1, 2) == 1 and lib.b200vq_tc_supported(64, 1, 0) == 0
    assert lib.b200vq_tc_supported(16, 0, 0) == 0
    assert lib.b200vq_tc_image_bytes(8192, 24) == 32 * 256 * 6 * 16 + 16
    assert lib.b200vq_tc_image_bytes(300, 16) == 2 * 256 * 4 * 16 + 16
