"""TEST INFRASTRUCTURE ONLY -- stages the UNMODIFIED reference files the oracle executes into oracle/_ref/.

    python -m oracle.stage_ref          (also run by __graft_entry__.build() when /root/reference exists)

/root/reference does not exist on the GPU box.  oracle/_ref/ is git-ignored (the reference's sources never enter the
history) but NOT gpurun-ignored, so it travels with the snapshot like the built .so files; oracle/ref_loader.py falls
back to it.  With it, `bench.py --impl reference`, bench.py's `cpu_baseline` leg and the adata-level tests run the
reference's own code on the box's host cores instead of the NumPy restatement.

What is staged: the sub-packages the hot path and its adata wrappers import from (vectorfield, tools, prediction,
simulation, preprocessing) and the top-level modules of `dynamo/` -- *.py files only, byte for byte (MANIFEST.json keeps
a sha256 per file and is checked by ref_loader.verify_staged()).
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("DVF_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
PACKAGES = ("vectorfield", "tools", "prediction", "simulation", "preprocessing")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(verbose=True) -> bool:
    """Returns False (and stages nothing) when the reference tree is absent, e.g. on the GPU box."""
    root = os.path.join(SRC, "dynamo")
    if not os.path.isdir(os.path.join(root, "vectorfield")):
        return False
    files = [f for f in sorted(os.listdir(root)) if f.endswith(".py")]
    for pkg in PACKAGES:
        for dirpath, _, names in os.walk(os.path.join(root, pkg)):
            relp = os.path.relpath(dirpath, root)
            files += [os.path.join(relp, n) for n in sorted(names) if n.endswith(".py")]
    manifest = {}
    for rel in files:
        src, dst = os.path.join(root, rel), os.path.join(DST, "dynamo", rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        digest = _sha(src)
        if not (os.path.exists(dst) and _sha(dst) == digest):
            shutil.copyfile(src, dst)
        manifest[os.path.join("dynamo", rel)] = digest
    head = ""
    try:
        with open(os.path.join(SRC, ".git", "HEAD")) as f:
            head = f.read().strip()
    except OSError:
        pass
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump({"source": SRC, "head": head, "files": manifest}, f, indent=1, sort_keys=True)
    if verbose:
        print("staged %d reference files (%.1f MB) into %s" % (
            len(files), sum(os.path.getsize(os.path.join(DST, k)) for k in manifest) / 1e6, DST), file=sys.stderr)
    return True


if __name__ == "__main__":
    if not stage():
        print("reference tree not found at %s -- nothing staged" % SRC, file=sys.stderr)
