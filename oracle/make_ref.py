"""Recipe that materialises the UNMODIFIED reference under oracle/_ref/ so that it exists on the GPU box.

TEST INFRASTRUCTURE (see oracle/codenerf_oracle.py header).  /root/reference is read-only and does not exist on the
GPU box; oracle/_ref/ is git-ignored (never enters history) but travels with the gpurun snapshot, exactly like the
built libcnrf.so.  The copy is byte-for-byte: train.py, eval.py, config/*.yml and view_synthesis/**/*.py.  Consumers:

  * tests/test_gpu_reference_drivers.py runs oracle/_ref/train.py and oracle/_ref/eval.py, unchanged, in a subprocess
    against the drop-in package (row N1 of the coverage table: "train.py and eval.py ... run unchanged");
  * bench.py --impl reference and bench.py's cpu_baseline leg time the live reference functions on the host cores
    (kind = "reference");
  * bench.py's `eager_torch_b200` key runs the same reference functions through stock eager torch on cuda:0
    (BASELINE.md row R2).

    python oracle/make_ref.py            # idempotent; also called by __graft_entry__.build()

A MANIFEST (relative path, size, sha256 of every copied file) is written next to the copy so a test can prove the
tree on the GPU box is the reference and not an edited one.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
KEEP_EXT = (".py", ".yml")


def _files(src):
    for root, dirs, files in os.walk(src):
        dirs[:] = [d for d in dirs if d not in ("__pycache__", ".git", "assets")]
        for f in sorted(files):
            if f.endswith(KEEP_EXT):
                p = os.path.join(root, f)
                yield os.path.relpath(p, src)


def sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as fh:
        h.update(fh.read())
    return h.hexdigest()


def make(src=None, verbose=True):
    src = src or os.environ.get("CODENERF_REFERENCE", "/root/reference")
    if not os.path.isdir(src):
        if verbose:
            print(f"make_ref: {src} not present (GPU box): keeping the existing {DST}")
        return os.path.isdir(DST)
    manifest = {}
    for rel in _files(src):
        s, d = os.path.join(src, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        if not os.path.exists(d) or sha256(d) != sha256(s):
            shutil.copyfile(s, d)
        manifest[rel] = {"bytes": os.path.getsize(d), "sha256": sha256(d)}
    with open(os.path.join(DST, "MANIFEST.json"), "w") as fh:
        json.dump({"source": src, "files": manifest}, fh, indent=1, sort_keys=True)
    if verbose:
        print(f"make_ref: {len(manifest)} reference files under {DST}")
    return True


def verify():
    """True when every file listed in the manifest is present with the recorded hash (nothing edited after the copy)."""
    mf = os.path.join(DST, "MANIFEST.json")
    if not os.path.exists(mf):
        return False
    files = json.load(open(mf))["files"]
    return all(os.path.exists(os.path.join(DST, rel)) and sha256(os.path.join(DST, rel)) == meta["sha256"]
               for rel, meta in files.items())


if __name__ == "__main__":
    ok = make()
    sys.exit(0 if ok and verify() else 1)
