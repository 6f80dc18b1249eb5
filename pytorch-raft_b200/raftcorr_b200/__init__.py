"""raftcorr_b200 — host side of the B200-native RAFT correlation path (see ../corr.py)."""
from . import _capi, ops  # noqa: F401
from ._capi import RaftCorrError  # noqa: F401
from .graph import GraphedForward, graphed  # noqa: F401

__all__ = ["_capi", "ops", "RaftCorrError", "graphed", "GraphedForward"]
