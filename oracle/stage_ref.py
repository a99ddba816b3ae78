"""Stage the reference files the config-2 harness needs under baseline/_ref/ (git-ignored; travels to the GPU box).

TEST INFRASTRUCTURE.  `/root/reference` exists only in the build container; the GPU box sees `/root/repo` alone
(SURVEY.md 7.4).  `baseline/_ref/` is the reserved place for the reference install (.gitignore: "the reference
install is not product source"): the files are copied there byte for byte, never edited, never committed, and
are only ever imported by the test harness (`oracle/ref_tokenizer.py`) and the reference arm of `bench.py`.

    python -m oracle.stage_ref        # idempotent; no-op when /root/reference is absent
"""
from __future__ import annotations

import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, "baseline", "_ref")

# the path north_star names + what `models/larp_tokenizer.py` imports at module level
FILES = [
    "models/models.py",
    "models/bottleneck.py",
    "models/larp_tokenizer.py",
    "models/transformer.py",
    "models/embed.py",
    "models/model_new/quantizer/fsq.py",
    "models/model_new/base/blocks.py",
    "models/model_new/base/transformer.py",
    "models/model_new/base/utils.py",
    "models/model_new/base/rope.py",
    "models/model_new/base/simpletransformer.py",
    "utils/statistics.py",
    "cfgs/larp_tokenizer.yaml",
    "cfgs/larp_tokenizer_large.yaml",
    "cfgs/larp_tokenizerf256t1024.yaml",
]


def stage(src_root: str = "/root/reference") -> int:
    if not os.path.isdir(src_root):
        return 0
    n = 0
    for rel in FILES:
        src = os.path.join(src_root, rel)
        if not os.path.isfile(src):
            continue
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src) or os.path.getsize(dst) != os.path.getsize(src):
            shutil.copy2(src, dst)
        n += 1
    return n


if __name__ == "__main__":
    print(f"staged {stage()} reference files under {DEST}")
