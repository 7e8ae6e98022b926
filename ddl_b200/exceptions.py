"""Exceptions of the loader (reference ddl/exceptions.py:1-5).

The reference's ``DoesNotMatchError`` misspells ``__init__`` so its fields are never set; here
``value`` and ``message`` are real attributes.
"""


class DoesNotMatchError(Exception):
    def __init__(self, value=None, message: str = ""):
        self.value = value
        self.message = message
        super().__init__(value, message)


class ShutdownError(RuntimeError):
    """The loader was used after it shut down (all epochs consumed, or ``close()`` was called)."""


class LoaderTimeoutError(TimeoutError):
    """A bounded wait on a producer or a peer expired (dead peer detection)."""
