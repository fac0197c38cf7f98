"""Recipe for oracle/_ref/: an UNMODIFIED copy of the reference's pure-Python package (TEST INFRASTRUCTURE).

The reference is stdlib Python with nothing to compile, so "building" it is copying `/root/reference/src`
(the `.py` files only) to `oracle/_ref/src`.  `oracle/_ref/` is git-ignored -- the reference's sources never enter
this repository's history -- but it is NOT gpurun-ignored, so the copy travels to the GPU box, where
`/root/reference` does not exist, and `bench.py --impl reference` can time the reference's own classes there
(`cpu_baseline.kind == "reference"`).  `__graft_entry__.build()` runs this whenever `/root/reference` is present.

    python oracle/fetch_ref.py [--src /root/reference]
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")


def fetch(src_root: str = "/root/reference") -> str:
    """Copies src_root/src/**/*.py -> oracle/_ref/src; returns the destination, or "" when src_root is absent."""
    src = os.path.join(src_root, "src")
    if not os.path.isdir(src):
        return ""
    dst = os.path.join(DST, "src")
    for d, _dirs, files in os.walk(src):
        rel = os.path.relpath(d, src)
        if "__pycache__" in rel:
            continue
        os.makedirs(os.path.join(dst, rel), exist_ok=True)
        for f in files:
            if not f.endswith(".py"):
                continue
            a, b = os.path.join(d, f), os.path.join(dst, rel, f)
            if not (os.path.exists(b) and filecmp.cmp(a, b, shallow=False)):
                shutil.copyfile(a, b)
    with open(os.path.join(DST, "PROVENANCE"), "w") as fh:
        fh.write(f"verbatim copy of {src} (*.py) made by oracle/fetch_ref.py; not tracked by git\n")
    return dst


if __name__ == "__main__":
    root = sys.argv[sys.argv.index("--src") + 1] if "--src" in sys.argv else "/root/reference"
    print(fetch(root) or f"{root}/src not found: nothing copied")
