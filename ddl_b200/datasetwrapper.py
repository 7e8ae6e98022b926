"""The dataset wrapper users subclass (reference ddl/datasetwrapper.py:4-19).

``on_init`` runs on the producer and must return a ``DataProducerOnInitReturn``; ``post_init``
receives the window ``my_ary`` and fills it; ``execute_function`` is the per-iteration in-place
transform, normally the local shuffle.  The object is pickled on its way to the producer
(reference ddl/types.py:9), so heavy loading belongs in ``on_init``.
"""
from __future__ import annotations

import os


class ProducerFunctionSkeleton:
    def __init__(self, *args, **kwargs):
        self.my_ary = None
        self.rank_global = None

    def on_init(self, *args, **kwargs):
        try:
            self.rank_global = kwargs["rank_global"]
        except KeyError:
            self.rank_global = int(os.environ.get("RANK", "0"))

    def post_init(self, *args, **kwargs):
        self.my_ary = kwargs["my_ary"]

    def execute_function(self, *args, **kwargs):
        raise NotImplementedError
