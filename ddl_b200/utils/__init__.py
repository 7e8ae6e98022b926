"""Cross-cutting helpers: callbacks, logging decorators, metrics, tracing, checkpoints."""
from .callbacks import execute_callbacks, for_all_methods, with_logging

__all__ = ["execute_callbacks", "for_all_methods", "with_logging"]
