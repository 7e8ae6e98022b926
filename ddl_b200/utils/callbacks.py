"""Callback dispatch and logging decorators (reference ddl/utils.py:9-57).

``execute_callbacks`` of the reference returns after the *first* callback in the list because
its ``return`` sits inside the loop (reference ddl/utils.py:9-22); since the user's producer
function is always first, the global shuffler appended behind it never runs
(SURVEY.md 0.2-1).  Here every callback that defines the hook runs, in order, and the first
non-None result is returned.  ``compat=True`` restores the as-wired behaviour so that an
unmodified reference run can be reproduced.

``with_logging`` costs one ``isEnabledFor`` check when DEBUG logging is off; the reference
builds ``repr()`` of every argument unconditionally (reference ddl/utils.py:28-30).
"""
from __future__ import annotations

import functools
import logging
import os
from typing import Any, Callable, Iterable

log = logging.getLogger("ddl_b200")


def _rank() -> int:
    try:
        return int(os.environ.get("RANK", "0"))
    except ValueError:
        return 0


def execute_callbacks(position: str, callbacks: Iterable[Any], compat: bool = False, **kwargs) -> Any:
    result = None
    for cb in callbacks:
        method = getattr(cb, position, None)
        if compat:
            # first callback only, whether or not it defines the hook
            return method(**kwargs) if method is not None else None
        if method is None:
            continue
        if log.isEnabledFor(logging.DEBUG):
            log.debug("[%03d] --> '%s' callback '%s'", _rank(), position, cb.__class__.__name__)
        ret = method(**kwargs)
        if result is None:
            result = ret
    return result


def with_logging(func: Callable[..., Any]) -> Callable[..., Any]:
    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        if not log.isEnabledFor(logging.DEBUG):
            return func(*args, **kwargs)
        sig = ", ".join([repr(a) for a in args[1:]] + [f"{k}={v!r}" for k, v in kwargs.items()])
        try:
            log.debug("[%03d] --> %s: %s", _rank(), func.__qualname__, sig)
            result = func(*args, **kwargs)
            log.debug("[%03d] <-- %s: %r", _rank(), func.__qualname__, result)
            return result
        except Exception as e:  # logged and re-raised, as in the reference
            log.exception("Exception raised in %s. exception: %s", func.__qualname__, e)
            raise

    return wrapper


def for_all_methods(decorator: Callable[..., Any], exclude: str | list[str] | None = None):
    if exclude is None:
        exclude = []
    elif not isinstance(exclude, list):
        exclude = [exclude]

    def decorate(cls):
        for attr, val in list(cls.__dict__.items()):
            if callable(val) and attr not in exclude and not isinstance(val, (staticmethod, classmethod, type)):
                setattr(cls, attr, decorator(val))
        return cls

    return decorate
