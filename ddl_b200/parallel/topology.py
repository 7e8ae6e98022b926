"""Runtime bring-up: who is a consumer, who produces, and which groups connect them.

Replaces ``init_mpi`` and the role dispatch of the reference (reference ddl/ddl_env.py:12-97).
The rank arithmetic is the reference's: the world is cut into ``n_instances`` contiguous
groups (``color = rank // group_size``); position 0 of a group is the consumer, the others are
producers; ``comm_nth_pusher`` links the k-th member of every group
(``color_nth_pusher = rank % group_size``) and is where the global shuffle runs.

Two transports share that arithmetic:

* ``process`` (group_size > 1) -- the reference's own layout: producers are OS processes that
  own host windows in shared memory.  The instance must live on one node, checked by hostname
  where the reference uses ``Split_type(COMM_TYPE_SHARED)`` (reference ddl/ddl_env.py:58-73).
* ``device`` (group_size == 1, one process per GPU as under torchrun) -- producers are roles
  inside the consumer process that drive kernels on side streams; shards and slot tables live in
  HBM.  The *virtual* layout (``n_producers + 1`` slots per instance) keeps ``rank_global`` and
  the partner-RNG seed of every producer identical to what the reference would assign, which is
  what makes permutations reproducible across the two frameworks.
"""
from __future__ import annotations

import dataclasses
import logging
import os
from typing import Any, List, Optional

import torch

from ..exceptions import DoesNotMatchError
from .comm import Comm, World

log = logging.getLogger("ddl_b200")


@dataclasses.dataclass
class Env:
    """The reference's ``MPI_Env`` (reference ddl/types.py:24-32) plus what a GPU run needs."""

    comm_global: Comm
    comm_per_gpu: Comm
    comm_per_gpu_shm: Comm
    comm_nth_pusher: Comm
    color: int
    color_nth_pusher: int
    n_instances: int
    # -- extensions ------------------------------------------------------------------------------
    world: Optional[World] = None
    transport: str = "device"
    n_producers: int = 1
    device: torch.device = torch.device("cpu")
    consumer_group: Any = None  # torch ProcessGroup of the consumers (the DDP world)
    loader_group: Any = None    # private ProcessGroup of the consumers for the loader's own stream barriers
    n_storage: int = 0          # dedicated storage ranks (GPUs that only hold shards), device transport
    is_storage: bool = False

    @property
    def is_consumer(self) -> bool:
        return self.comm_per_gpu_shm.rank == 0 and not self.is_storage

    def storage_rank_of(self, instance: int) -> int:
        """World rank whose GPU holds `instance`'s shards (the instance itself without storage ranks)."""
        return instance if self.n_storage == 0 else self.n_instances + instance % self.n_storage

    @property
    def group_size(self) -> int:
        return self.n_producers + 1

    @property
    def instance_idx(self) -> int:
        return self.color

    def virtual_rank_global(self, producer_idx: int) -> int:
        """``rank_global`` the reference would give producer ``producer_idx`` (1-based) here."""
        return self.color * self.group_size + producer_idx


def init_env_error_handling(comm_global: Comm, n_instances: int) -> None:
    """reference ddl/ddl_env.py:12-30 (raises instead of MPI_Abort so callers can clean up)."""
    if comm_global.size < 2 and comm_global.rank == 0:
        log.warning("Using only a single rank!")
    if n_instances < 1 or comm_global.size % n_instances != 0:
        if comm_global.rank == 0:
            log.error("Error! Number of processes must be a multiple of %d", n_instances)
        raise DoesNotMatchError((comm_global.size, n_instances),
                                f"world size {comm_global.size} is not a multiple of n_instances={n_instances}")


def _pick_device(local_index: int, want_cuda: Optional[bool]) -> torch.device:
    cuda = torch.cuda.is_available() if want_cuda is None else want_cuda
    if cuda and torch.cuda.is_available():
        return torch.device("cuda", local_index % torch.cuda.device_count())
    return torch.device("cpu")


def _init_env_with_storage(n_instances, n_storage, world, n_producers, device, make_consumer_group) -> Env:
    """Device transport with dedicated storage GPUs (BASELINE config 4, producers:consumers 1:3):
    ranks [0, n_instances) consume, ranks [n_instances, n_instances + n_storage) only hold shards.
    The reference cannot express fewer producers than consumers (ddl/connection.py:29-30)."""
    comm_global = world.comm
    rank, size = comm_global.rank, comm_global.size
    if size != n_instances + n_storage:
        raise DoesNotMatchError((size, n_instances, n_storage), "world size must be n_instances + n_storage")
    is_storage = rank >= n_instances
    role = comm_global.split(1 if is_storage else 0, key=rank, name="role")
    own = comm_global.split(rank, key=rank, name="per_gpu")
    comm_global.barrier()
    if device is None:
        device = _pick_device(int(os.environ.get("LOCAL_RANK", str(rank))), None)
    device = torch.device(device)
    if device.type == "cuda":
        torch.cuda.set_device(device)
    group = private = None
    try:
        import torch.distributed as dist

        if make_consumer_group and dist.is_initialized():
            group = dist.new_group(list(range(n_instances)))
            private = dist.new_group(list(range(n_instances)))
    except Exception as e:
        log.warning("could not create the consumer process group: %s", e)
    producers = int(n_producers if n_producers is not None else os.environ.get("DDL_B200_PRODUCERS", "1"))
    return Env(comm_global=comm_global, comm_per_gpu=own, comm_per_gpu_shm=own, comm_nth_pusher=role,
               color=rank if not is_storage else rank - n_instances, color_nth_pusher=0, n_instances=n_instances,
               world=world, transport="device", n_producers=producers, device=device, consumer_group=group,
               n_storage=n_storage, is_storage=is_storage, loader_group=private)


def init_env(n_instances: int = 1, world: Optional[World] = None, n_producers: Optional[int] = None,
             device: Optional[torch.device] = None, make_consumer_group: bool = True, n_storage: int = 0) -> Env:
    """Build the communicators and decide the transport.  Collective over the whole world."""
    world = world or World.from_environment()
    if n_storage > 0:
        return _init_env_with_storage(n_instances, n_storage, world, n_producers, device, make_consumer_group)
    comm_global = world.comm
    rank, size = comm_global.rank, comm_global.size
    init_env_error_handling(comm_global, n_instances)

    group_size = size // n_instances
    color = rank // group_size
    comm_per_gpu = comm_global.split(color, key=rank, name="per_gpu")
    hosts = comm_per_gpu.hostnames()
    if len(set(hosts)) != 1:
        raise DoesNotMatchError(hosts, f"instance {color} spans several nodes: {sorted(set(hosts))}")
    comm_per_gpu_shm = comm_per_gpu  # same members; the node check above is the SHARED split
    color_nth = rank % group_size
    comm_nth_pusher = comm_global.split(color_nth, key=rank, name="nth_pusher")
    comm_global.barrier()

    if group_size > 1:
        transport = "process"
        producers = group_size - 1
        if n_producers not in (None, producers):
            raise DoesNotMatchError((n_producers, producers), "n_producers is fixed by the world size here")
    else:
        transport = "device"
        producers = int(n_producers if n_producers is not None else os.environ.get("DDL_B200_PRODUCERS", "1"))
        if producers < 1:
            raise ValueError("at least one producer per instance is needed")

    if device is None:
        local = int(os.environ.get("LOCAL_RANK", str(rank)))
        if transport == "process":
            # one GPU per instance; producer processes stay on the host
            device = _pick_device(color, None) if comm_per_gpu.rank == 0 else torch.device("cpu")
        else:
            device = _pick_device(local, None)
    device = torch.device(device)
    if device.type == "cuda":
        torch.cuda.set_device(device)

    consumer_group = loader_group = None
    try:
        import torch.distributed as dist

        if make_consumer_group and dist.is_initialized() and n_instances > 1:
            consumers: List[int] = [i * group_size for i in range(n_instances)]
            consumer_group = dist.group.WORLD if transport == "device" else dist.new_group(consumers)
            if transport == "device":
                # the loader's stream barriers must never interleave with the user's collectives
                # (DDP buckets) on one communicator: give them a communicator of their own
                loader_group = dist.new_group(consumers)
    except Exception as e:  # a missing DDP group is not fatal for the loader itself
        log.warning("could not create the consumer process group: %s", e)

    return Env(comm_global=comm_global, comm_per_gpu=comm_per_gpu, comm_per_gpu_shm=comm_per_gpu_shm,
               comm_nth_pusher=comm_nth_pusher, color=color, color_nth_pusher=color_nth, n_instances=n_instances,
               world=world, transport=transport, n_producers=producers, device=device,
               consumer_group=consumer_group, loader_group=loader_group)


def instances_from_environment(default: int = 1) -> int:
    """reference ddl/ddl_env.py:103-107: SLURM decides, else one instance.  ``DDL_B200_INSTANCES``
    overrides; the value ``world`` means one instance per process (the torchrun model)."""
    override = os.environ.get("DDL_B200_INSTANCES")
    if override:
        if override == "world":
            return int(os.environ.get("WORLD_SIZE", "1"))
        return int(override)
    try:
        _ = os.environ["SLURM_JOBID"]
        return int(os.environ["SLURM_GPUS_PER_NODE"]) * int(os.environ["SLURM_NNODES"])
    except KeyError:
        return default
