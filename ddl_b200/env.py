"""``@distributed_dataloader``: runtime bring-up and role dispatch (reference ddl/ddl_env.py:100-128).

The decorated function runs on every consumer with two extra positional arguments, the
environment and the connection -- ``func(*args, env, conn)`` -- exactly as the reference calls
``func(*args, mpi_env, conn)``.  Every other rank becomes a ``DataPusher`` and runs
``push_data()`` until the consumer shuts it down.  Differences: keyword arguments are forwarded
(the reference drops them, ddl/ddl_env.py:116); a one-rank world gets a working connection; the
decorator also accepts options: ``@distributed_dataloader(n_instances="world", n_producers=2)``.
"""
from __future__ import annotations

import functools
import logging
from typing import Callable, Optional

from .config import LoaderConfig
from .connection import Connection
from .datapusher import DataPusher
from .parallel.topology import Env, init_env, instances_from_environment

log = logging.getLogger("ddl_b200")


def init_mpi(n_instances: int = 1, **kwargs) -> Env:
    """The reference's name for ``init_env`` (reference ddl/ddl_env.py:33-97)."""
    return init_env(n_instances, **kwargs)


def distributed_dataloader(func: Optional[Callable] = None, *, n_instances=None, n_producers: Optional[int] = None,
                           config: Optional[LoaderConfig] = None, world=None, device=None, n_storage: int = 0):
    def decorate(f: Callable) -> Callable:
        @functools.wraps(f)
        def wrapper(*args, **kwargs):
            n = n_instances
            if n is None:
                n = instances_from_environment(1)
            elif n == "world":
                import os

                n = world.size if world is not None else int(os.environ.get("WORLD_SIZE", "1"))
                n -= n_storage
            cfg = config or LoaderConfig.from_env()
            if cfg.exchange == "push":
                # the push form parks streams in cuStreamWaitValue32; with the default 8 hardware
                # queues an unrelated stream (NCCL, the table stream) can end up queued behind a
                # parked one and the ranks deadlock.  Must be set before the CUDA context exists.
                import os

                import torch

                if "CUDA_DEVICE_MAX_CONNECTIONS" not in os.environ:
                    if torch.cuda.is_initialized():
                        log.warning("exchange='push': CUDA is already initialised; start the process with "
                                    "CUDA_DEVICE_MAX_CONNECTIONS=32 to rule out false stream dependencies")
                    os.environ["CUDA_DEVICE_MAX_CONNECTIONS"] = "32"
            env = init_env(int(n), world=world, n_producers=n_producers, device=device, n_storage=n_storage)
            conn = Connection(env.comm_per_gpu_shm, env, cfg)
            result = None
            if env.is_storage:
                from .storage import StorageServer

                StorageServer(env, cfg).serve()
            elif env.is_consumer:
                result = f(*args, env, conn, **kwargs)
            else:
                pusher = DataPusher(conn, env.comm_per_gpu_shm, env.comm_nth_pusher, env.comm_global)
                pusher.push_data()
            log.info("[%03d]: finished", env.comm_global.rank)
            env.comm_global.barrier()
            return result

        return wrapper

    return decorate(func) if func is not None else decorate
