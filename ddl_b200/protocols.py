"""Names of the producer-loop hooks (reference ddl/protocols.py:4-18).

The reference protocol lists ``exec_function`` although its loop calls ``execute_function``
(reference ddl/datapusher.py:154); the working name is the one kept here.
"""
from typing import Any, Protocol, runtime_checkable

HOOKS = ("on_init", "post_init", "on_push_begin", "global_shuffle", "execute_function", "on_shuffle_end",
         "on_push_end")


@runtime_checkable
class CallbackProtocol(Protocol):
    def on_push_begin(self, **kwargs) -> Any: ...

    def global_shuffle(self, **kwargs) -> Any: ...

    def execute_function(self, **kwargs) -> Any: ...

    def on_push_end(self, **kwargs) -> Any: ...

    def on_shuffle_end(self, **kwargs) -> Any: ...
