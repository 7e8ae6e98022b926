"""Topology, control-plane communicator, symmetric heap and stream/flag synchronisation."""
from .comm import Comm, World
from .topology import Env, init_env

__all__ = ["Comm", "World", "Env", "init_env"]
