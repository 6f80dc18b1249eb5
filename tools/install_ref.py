"""Make the unmodified reference model travel to the GPU box.

    python tools/install_ref.py [--src /root/reference] [--check]

Copies the reference's ``core/`` package (raft.py, corr.py, update.py, extractor.py, utils/ — the
caller of the correlation path and the reference's own implementation of it) and the two
functions' worth of ``train.py`` the training benchmark needs, byte for byte, into
``baseline/_ref/``.  That directory is git-ignored (reference sources never enter this
repository's history) but not gpurun-ignored, so the files ride along with the snapshot that
``gpurun`` and the driver push to the B200 box, where ``/root/reference`` does not exist.

Consumers (all of them skip / fall back when ``baseline/_ref/core/raft.py`` is absent):
  * tests/test_gpu_reference_dropin.py — the reference's own ``RAFT(args)`` on cuda:0 with this
    library's ``corr.py`` ahead of ``core/`` on ``sys.path``, against the same model with the
    reference's own ``corr.py``;
  * tests/test_harness.py — pins harness/raft_stock.py to the reference model;
  * bench.py — ``--impl reference`` and ``cpu_baseline`` time the reference's own CPU code
    (``kind: "reference"``); ``dropin_unchanged`` times the unmodified ``raft.py`` on the GPU.

``__graft_entry__.build()`` runs this when ``/root/reference`` is present.  Nothing here is
imported by the product path.
"""
from __future__ import annotations

import argparse
import filecmp
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, "baseline", "_ref")
DEFAULT_SRC = "/root/reference"
# what travels: the model package, and train.py for sequence_loss / the optimiser recipe (cfg5)
TREES = ("core",)
FILES = ("train.py",)


def ref_core_dir() -> str | None:
    """Directory to put on sys.path to import the unmodified reference model, or None."""
    for base in (DEST, DEFAULT_SRC):
        d = os.path.join(base, "core")
        if os.path.isfile(os.path.join(d, "raft.py")):
            return d
    return None


def ref_root_dir() -> str | None:
    d = ref_core_dir()
    return os.path.dirname(d) if d else None


def install(src: str = DEFAULT_SRC, quiet: bool = False) -> bool:
    if not os.path.isdir(os.path.join(src, "core")):
        if not quiet:
            print(f"install_ref: {src}/core not found; nothing installed", file=sys.stderr)
        return False
    os.makedirs(DEST, exist_ok=True)
    for tree in TREES:
        dst = os.path.join(DEST, tree)
        if os.path.isdir(dst):
            shutil.rmtree(dst)
        shutil.copytree(os.path.join(src, tree), dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
        for dirpath, _, names in os.walk(dst):   # the source tree is read-only; the copy need not be
            os.chmod(dirpath, 0o755)
            for n in names:
                os.chmod(os.path.join(dirpath, n), 0o644)
    for f in FILES:
        shutil.copyfile(os.path.join(src, f), os.path.join(DEST, f))
    if not quiet:
        print(f"install_ref: copied {', '.join(TREES + FILES)} from {src} to {DEST}")
    return True


def check(src: str = DEFAULT_SRC) -> bool:
    """True when baseline/_ref is a byte-identical copy of the reference files it mirrors."""
    ok = True
    for tree in TREES:
        for dirpath, _, names in os.walk(os.path.join(src, tree)):
            rel = os.path.relpath(dirpath, src)
            for n in names:
                if n.endswith(".pyc"):
                    continue
                a, b = os.path.join(dirpath, n), os.path.join(DEST, rel, n)
                if not (os.path.isfile(b) and filecmp.cmp(a, b, shallow=False)):
                    print(f"install_ref: {os.path.join(rel, n)} differs or is missing", file=sys.stderr)
                    ok = False
    for f in FILES:
        if not (os.path.isfile(os.path.join(DEST, f)) and filecmp.cmp(os.path.join(src, f), os.path.join(DEST, f), shallow=False)):
            print(f"install_ref: {f} differs or is missing", file=sys.stderr)
            ok = False
    return ok


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default=DEFAULT_SRC)
    ap.add_argument("--check", action="store_true", help="verify the copy instead of making it")
    a = ap.parse_args()
    sys.exit(0 if (check(a.src) if a.check else install(a.src)) else 1)
