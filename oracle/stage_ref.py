"""Stage the unmodified reference tree under ``oracle/_ref/reference``.

The reference is a bash script, so its "build" is a copy of the files it runs
from.  The copy lives under ``oracle/_ref/`` only: git-ignored (it never enters
history) but not gpurun-ignored, so it travels to the GPU box where
``/root/reference`` does not exist -- the same treatment a compiled
``oracle/_ref/*.so`` would get.  Screenshots (docs/img, 5 MB) are left out.
"""
from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference"
DST = os.path.join(HERE, "_ref", "reference")


def stage(src: str = SRC, dst: str = DST) -> bool:
    if not os.path.isfile(os.path.join(src, "setup.sh")):
        return False
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns(".git", "img"))
    return True


if __name__ == "__main__":
    ok = stage()
    print("staged %s -> %s" % (SRC, DST) if ok else "no reference at %s; nothing staged" % SRC)
    sys.exit(0)
