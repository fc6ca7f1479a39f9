"""iqlpref_b200: the IQL training step and reward-model relabelling pass of ml4ai/iqlpref,
re-built for B200 (sm_100a).  Python mirrors the reference surface; compute is libiqlb.so."""
from . import _lib  # noqa: F401

__all__ = ["iql", "relabel", "reward_models"]
