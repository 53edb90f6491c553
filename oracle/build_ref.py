"""Stage the UNMODIFIED reference package under ``oracle/_ref/`` (TEST INFRASTRUCTURE).

    python -m oracle.build_ref        # build container only (/root/reference present)

The reference is pure Python (no C/C++/CUDA), so "building" it is a plain
copy of ``/root/reference/search/li`` to ``oracle/_ref/li`` -- byte for byte, checked by a
SHA-256 manifest.  ``oracle/_ref/`` is git-ignored (reference sources never enter this
repository's history) but not gpurun-ignored, so the copy travels to the GPU box with the
snapshot, where ``/root/reference`` does not exist.  It lets ``bench.py --impl reference`` time
the reference's own ``LearnedIndex.search`` (search/li/LearnedIndex.py:41-161) on the box's host
cores under the three shims of ``oracle/ref_harness.py`` (``cpu_baseline.kind = "reference"``).

``__graft_entry__.build()`` calls ``stage()`` whenever ``/root/reference`` is present.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.path.join(os.environ.get("LMI_REFERENCE_ROOT", "/root/reference"), "search", "li")
REF_DST = os.path.join(HERE, "_ref")
MANIFEST = os.path.join(REF_DST, "MANIFEST.json")


def _sha256(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 16), b""):
            h.update(blk)
    return h.hexdigest()


def _tree(root: str):
    out = {}
    for dirpath, dirnames, files in os.walk(root):
        dirnames[:] = [d for d in dirnames if d != "__pycache__"]
        for f in sorted(files):
            if f.endswith(".py"):
                p = os.path.join(dirpath, f)
                out[os.path.relpath(p, root)] = _sha256(p)
    return out


def staged() -> bool:
    return os.path.isfile(os.path.join(REF_DST, "li", "LearnedIndex.py"))


def stage(force: bool = False) -> str:
    """Copy the reference package; returns the directory to put on ``sys.path``."""
    if not os.path.isfile(os.path.join(REF_SRC, "LearnedIndex.py")):
        if staged():
            return REF_DST
        raise RuntimeError(f"reference package not found under {REF_SRC}")
    src = _tree(REF_SRC)
    if staged() and not force and os.path.isfile(MANIFEST):
        try:
            if json.load(open(MANIFEST)).get("files") == src:
                return REF_DST
        except Exception:
            pass
    dst = os.path.join(REF_DST, "li")
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    os.makedirs(REF_DST, exist_ok=True)
    shutil.copytree(REF_SRC, dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    got = _tree(dst)
    assert got == src, "staged copy differs from the reference"
    json.dump({"source": REF_SRC, "files": got}, open(MANIFEST, "w"), indent=1, sort_keys=True)
    return REF_DST


def verify() -> bool:
    """The staged files still hash to the manifest written at staging time (nothing edited them)."""
    if not (staged() and os.path.isfile(MANIFEST)):
        return False
    return json.load(open(MANIFEST)).get("files") == _tree(os.path.join(REF_DST, "li"))


if __name__ == "__main__":
    print(stage())
