"""Stage the UNMODIFIED reference script tree where the GPU box can see it.

`/root/reference` exists only in the build container; the GPU box gets a snapshot of this repo.  The
reference is a plain script tree (no setup.py / pyproject.toml, so `pip install --target baseline/_ref
/root/reference` has nothing to install); its "install" is therefore a byte-for-byte copy of its Python
sources into `baseline/_ref/` — git-ignored (never in history), not gpurun-ignored (travels to the box like
the built .so files).  `baseline/REFERENCE_MANIFEST.json` (tracked) holds the sha256 of every staged file;
`dropin.reference_root()` refuses a tree whose hashes differ, so whatever the drop-in tests and
`bench.py`'s `reference_trainer` leg drive is provably the reference as shipped.

Only the drop-in harness (tests/test_gpu_dropin_reference.py, bench.py's reference_trainer leg) reads the
staged tree; the product package never imports it.

    python baseline/stage_reference.py [--src /root/reference] [--check]
"""
import argparse
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
MANIFEST = os.path.join(HERE, "REFERENCE_MANIFEST.json")
KEEP_DIRS = ("src",)
KEEP_FILES = ("requirements.txt", "README.md")


def _sha(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def _walk(root: str):
    for d in KEEP_DIRS:
        for base, _, files in os.walk(os.path.join(root, d)):
            for fn in sorted(files):
                if fn.endswith(".py"):
                    yield os.path.relpath(os.path.join(base, fn), root)
    for fn in KEEP_FILES:
        if os.path.exists(os.path.join(root, fn)):
            yield fn


def manifest_of(root: str) -> dict:
    return {rel: _sha(os.path.join(root, rel)) for rel in sorted(_walk(root))}


def stage(src: str) -> dict:
    man = manifest_of(src)
    if os.path.isdir(DEST):
        shutil.rmtree(DEST)
    for rel in man:
        dst = os.path.join(DEST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), dst)
    with open(MANIFEST, "w") as f:
        json.dump({"source": "ShayManor/3d-gaussian-splatting", "files": man}, f, indent=1, sort_keys=True)
        f.write("\n")
    return man


def verify(root: str = DEST) -> bool:
    """True iff `root` holds exactly the files of the tracked manifest with identical contents."""
    if not (os.path.exists(MANIFEST) and os.path.isdir(root)):
        return False
    with open(MANIFEST) as f:
        want = json.load(f)["files"]
    try:
        return all(_sha(os.path.join(root, rel)) == h for rel, h in want.items())
    except OSError:
        return False


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default="/root/reference")
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    if a.check:
        ok = verify()
        print("staged reference matches the manifest" if ok else "staged reference missing or modified")
        sys.exit(0 if ok else 1)
    if not os.path.isdir(a.src):
        print(f"{a.src} not present: nothing staged (the GPU box uses the tree staged in the build container)")
        sys.exit(0)
    m = stage(a.src)
    print(f"staged {len(m)} files into {DEST}")
