"""The dataset wrapper users subclass (reference ddl/datasetwrapper.py:4-19).

Life cycle of a producer function, identical to the reference's:

1. it is constructed by the *consumer* and shipped to every producer (pickled in the reference,
   reference ddl/types.py:9; copied per producer here), so construction must stay cheap;
2. ``on_init(rank_global=...)`` runs on the producer: load or create the data and answer with a
   ``DataProducerOnInitReturn`` describing the window;
3. ``post_init(my_ary=window, rank_global=..., n_instance=...)`` runs once the window exists: fill it;
4. ``execute_function()`` runs every producer iteration: the in-place transform of the window,
   normally the local shuffle.  Optional hooks with the reference's names -- ``on_push_begin``,
   ``global_shuffle``, ``on_shuffle_end``, ``on_push_end`` -- are called when defined.

Additions: ``lengths`` (valid elements per row, for ``variable_length`` windows) and
``state_dict``/``load_state_dict`` so user state can ride along with a loader checkpoint.
"""
from __future__ import annotations

import os
from typing import Any, Optional


def _launcher_rank() -> int:
    for var in ("RANK", "SLURM_PROCID", "OMPI_COMM_WORLD_RANK"):
        val = os.environ.get(var)
        if val is not None and val.isdigit():
            return int(val)
    return 0


class ProducerFunctionSkeleton:
    #: set by ``post_init`` when the window holds variable-size samples
    lengths: Optional[Any] = None

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        self.rank_global: Optional[int] = None
        self.my_ary = None

    # -- required by the producer loop ------------------------------------------------------------
    def on_init(self, *args: Any, **kwargs: Any):
        """Remember which producer this is.  Subclasses call this first and then return a
        ``DataProducerOnInitReturn``."""
        self.rank_global = kwargs["rank_global"] if "rank_global" in kwargs else _launcher_rank()

    def post_init(self, *args: Any, **kwargs: Any) -> None:
        """Keep a handle on the window; subclasses fill it."""
        self.my_ary = kwargs["my_ary"]

    def execute_function(self, *args: Any, **kwargs: Any) -> None:
        raise NotImplementedError(f"{type(self).__name__} must define execute_function()")

    # -- optional user state for checkpoints ----------------------------------------------------------
    def state_dict(self) -> dict:
        return {}

    def load_state_dict(self, state: dict) -> None:
        del state
