"""Loader for the UNMODIFIED reference model (core/raft.py) used by tests and bench.py.

The reference resolves its modules by bare name (``from corr import CorrBlock`` in raft.py:8,
``from utils.utils import ...`` in raft.py:9), so which correlation block it runs is decided by
``sys.path`` order alone.  ``load_raft_module(dropin=True)`` imports a private copy of ``raft`` with
this repository's ``pytorch-raft_b200/`` ahead of the reference's ``core/``;
``dropin=False`` imports it with the reference's own ``corr.py``.  Both copies can live in one
process: the bare-named modules are removed from ``sys.modules`` again after each import and the
returned module objects keep their own bindings.

The reference tree is found in ``baseline/_ref`` (tools/install_ref.py; travels to the GPU box)
or, in the build container, in ``/root/reference``.  Nothing here is on the product path.
"""
from __future__ import annotations

import argparse
import ast
import importlib
import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-raft_b200")
_BARE = ("raft", "corr", "update", "extractor", "utils", "utils.utils", "utils.augmentor", "utils.frame_utils",
         "utils.flow_viz", "alt_cuda_corr")


def ref_core_dir():
    for base in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        d = os.path.join(base, "core")
        if os.path.isfile(os.path.join(d, "raft.py")):
            return d
    return None


def available() -> bool:
    return ref_core_dir() is not None


def load_raft_module(dropin: bool):
    """Import the reference's raft.py; returns the module.  ``dropin`` selects whose corr.py it binds."""
    core = ref_core_dir()
    if core is None:
        raise FileNotFoundError("reference tree not found (run tools/install_ref.py in the build container)")
    saved_path = list(sys.path)
    saved_mods = {k: sys.modules.pop(k) for k in _BARE if k in sys.modules}
    try:
        # drop every entry that could shadow the choice, then put the chosen order in front
        sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") not in (PKG, core)]
        sys.path[:0] = [PKG, core] if dropin else [core]
        ours = saved_mods.get("corr")
        if dropin and ours is not None and os.path.abspath(getattr(ours, "__file__", "")) == os.path.join(PKG, "corr.py"):
            sys.modules["corr"] = ours          # already imported by the caller: bind the very same module object
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            importlib.invalidate_caches()
            mod = importlib.import_module("raft")
        corr_file = os.path.abspath(sys.modules["corr"].__file__)
        want = os.path.join(PKG if dropin else core, "corr.py")
        assert corr_file == os.path.abspath(want), (corr_file, want)
        if dropin:
            saved_mods["corr"] = sys.modules["corr"]   # the drop-in stays importable under its own name
        return mod
    finally:
        sys.path[:] = saved_path
        for k in _BARE:
            sys.modules.pop(k, None)
        sys.modules.update(saved_mods)


def make_args(small=False, mixed_precision=False, alternate_corr=False, dropout=0.0):
    # raft.py:41-45 tests `'dropout' not in self.args`: an argparse.Namespace supports that
    return argparse.Namespace(small=small, mixed_precision=mixed_precision, alternate_corr=alternate_corr,
                              dropout=dropout)


def load_train_functions():
    """``sequence_loss`` and ``fetch_optimizer`` of the reference's train.py (train.py:47-86), taken out of the
    file with ``ast`` because importing train.py as a module pulls in matplotlib, cv2, datasets and evaluate."""
    core = ref_core_dir()
    if core is None:
        raise FileNotFoundError("reference tree not found")
    path = os.path.join(os.path.dirname(core), "train.py")
    tree = ast.parse(open(path).read(), path)
    keep = [n for n in tree.body
            if (isinstance(n, ast.FunctionDef) and n.name in ("sequence_loss", "fetch_optimizer"))
            or (isinstance(n, ast.Assign) and any(getattr(t, "id", "") == "MAX_FLOW" for t in n.targets))]
    import torch
    import torch.optim as optim
    ns = {"torch": torch, "optim": optim}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["sequence_loss"], ns["fetch_optimizer"]
