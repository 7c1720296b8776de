"""TEST INFRASTRUCTURE ONLY -- recipe that makes the unmodified reference travel to the GPU box.

The reference (Optiland fork, pure Python) lives read-only at /root/reference in the build container and does not
exist on the GPU box. This script copies its `optiland/` package (sources + the vendored glass catalogue under
`optiland/database/`) VERBATIM into the git-ignored `oracle/_ref/` -- an output directory that ships with the working
tree (listed in .gitignore, not in .gpurunignore), exactly like the built libolrt.so. Nothing is edited; a manifest
of sha256 sums (`oracle/_ref/MANIFEST.sha256`) is written so that a test can check the copy is the reference.
Run by `__graft_entry__.build()`. Where /root/reference is absent (the GPU box) it does nothing.

Consumers: `oracle/ref_shim.py` (import shim) -> tests/test_*dropin*.py, bench.py's reference baselines.
The product package never imports from here.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

SRC = os.environ.get("OLRT_REFERENCE_PATH", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SKIP_DIRS = {"__pycache__", "visualization_assets"}


def _files(root):
    for d, dirs, files in os.walk(root):
        dirs[:] = sorted(x for x in dirs if x not in SKIP_DIRS)
        for f in sorted(files):
            if f.endswith((".pyc", ".pyo")):
                continue
            yield os.path.relpath(os.path.join(d, f), root)


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def main() -> int:
    src = os.path.join(SRC, "optiland")
    if not os.path.isdir(src):
        print(f"fetch_ref: {src} not present (GPU box?) -- keeping whatever is under {DST}")
        return 0
    dst = os.path.join(DST, "optiland")
    os.makedirs(dst, exist_ok=True)
    want = set(_files(src))
    copied = 0
    lines = []
    for rel in sorted(want):
        a, b = os.path.join(src, rel), os.path.join(dst, rel)
        sa = os.stat(a)
        if not (os.path.exists(b) and os.stat(b).st_size == sa.st_size and int(os.stat(b).st_mtime) == int(sa.st_mtime)):
            os.makedirs(os.path.dirname(b), exist_ok=True)
            shutil.copy2(a, b)
            copied += 1
        if not rel.startswith("database" + os.sep):
            lines.append(f"{_sha(b)}  optiland/{rel}")
    for rel in set(_files(dst)) - want:       # the copy holds exactly the reference's files
        os.remove(os.path.join(dst, rel))
    with open(os.path.join(DST, "MANIFEST.sha256"), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    print(f"fetch_ref: {len(want)} files under {dst} ({copied} copied), manifest of {len(lines)} sources")
    return 0


if __name__ == "__main__":
    sys.exit(main())
