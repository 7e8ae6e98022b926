"""Baseline-only stand-in package: see MPI.py.  Not part of ddl_b200."""
from . import MPI  # noqa: F401

__version__ = "0.0-standin"
