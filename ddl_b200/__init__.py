"""ddl_b200 -- a Blackwell-native distributed dataloader for PyTorch.

Same public names as ``maximilian-tech/ddl`` (reference ddl/__init__.py:7-21):
``ProducerFunctionSkeleton``, ``DataProducerOnInitReturn``, ``distributed_dataloader``,
``DistributedDataLoader``, ``Marker`` -- built from scratch for B200: dataset shards resident in
HBM, per-epoch shuffles on 8-byte slot ids (bit-identical permutations to the reference's numpy
streams), and batches assembled by one hand-written sm_100a kernel that pulls samples over
NVLink, converts and collates them straight into a double-buffered batch tensor.
"""

__all__ = [
    "ProducerFunctionSkeleton",
    "DataProducerOnInitReturn",
    "distributed_dataloader",
    "DistributedDataLoader",
    "Marker",
    # extensions
    "LoaderConfig",
    "Connection",
    "DataPusher",
    "Convert",
    "init_mpi",
    "init_env",
    "DoesNotMatchError",
    "LoaderTimeoutError",
]

from .config import LoaderConfig
from .connection import Connection
from .dataloader import DistributedDataLoader, Marker, ProducerFunctionSkeleton
from .datapusher import DataProducerOnInitReturn, DataPusher
from .env import distributed_dataloader, init_mpi
from . import exceptions, ops
from .exceptions import DoesNotMatchError, LoaderTimeoutError
from .ops.gather import Convert
from .parallel.topology import init_env

__version__ = "0.1.0"
