"""Distributed runtime: bootstrap, arenas, collectives, DP engine, eval pipeline."""
from .bootstrap import Comm, World, init_world, shutdown, spawn, world  # noqa: F401
