#!/usr/bin/env python
"""Materialise the reference's own sources under oracle/_ref/ so that they travel to the GPU box.

TEST INFRASTRUCTURE ONLY.  `/root/reference` exists in the build container but not on the GPU box; `oracle/_ref/` is
git-ignored (nothing of the reference enters the history) but NOT gpurun-ignored, so a copy made here is present
there.  The reference is pure Python 2.7 (no compile step): the "build" is a file copy of `src/*.py`,
`src/pyFastq/*.py` and LICENSE (GPL v2).  `oracle/ref_shim.py` then runs these files, unmodified on disk, under
Python 3 (textual Py2 -> 3 substitutions at load time, listed in its header).  `bench.py` times them
(`cpu_baseline.kind = "reference"`, `--impl reference`).

    python oracle/make_ref.py            # called by __graft_entry__.build()
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = os.environ.get("QUADE_REFERENCE_SRC", "/root/reference")


def make(verbose=False):
    """Returns the path of oracle/_ref if it holds the reference sources afterwards, else None."""
    if os.path.isfile(os.path.join(SRC, "src", "Quade.py")):
        for rel in ("src", os.path.join("src", "pyFastq")):
            os.makedirs(os.path.join(DST, rel), exist_ok=True)
            for f in sorted(os.listdir(os.path.join(SRC, rel))):
                if f.endswith(".py"):
                    shutil.copyfile(os.path.join(SRC, rel, f), os.path.join(DST, rel, f))
        if os.path.isfile(os.path.join(SRC, "LICENSE")):
            shutil.copyfile(os.path.join(SRC, "LICENSE"), os.path.join(DST, "LICENSE"))
        with open(os.path.join(DST, "README"), "w") as fh:
            fh.write("Unmodified copy of a-slide/Quade src/*.py (GPL v2, see LICENSE) made by oracle/make_ref.py.\n"
                     "Git-ignored; used only as the timed CPU baseline and as a cross-check of the oracle.\n")
        if verbose:
            print("oracle/_ref: reference sources copied from", SRC)
    return DST if os.path.isfile(os.path.join(DST, "src", "Quade.py")) else None


if __name__ == "__main__":
    p = make(verbose=True)
    print(p or "reference sources not available")
    sys.exit(0 if p else 1)
