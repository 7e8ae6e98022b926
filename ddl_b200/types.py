"""Plain data carried between the roles (reference ddl/types.py:7-37, ddl/datapusher.py:14-19)."""
from __future__ import annotations

import dataclasses
from enum import Enum, auto
from typing import Any, Optional, Tuple

import numpy as np


class Marker(Enum):
    """What the training loop reports to the loader (reference ddl/types.py:35-37)."""

    END_OF_BATCH = auto()
    END_OF_EPOCH = auto()


class WorkerInfo(Enum):
    """Outcome of a producer-side wait (reference ddl/connection.py:12-14)."""

    CONTINUE = auto()
    STOP = auto()


@dataclasses.dataclass
class DataProducerOnInitReturn:
    """What a producer function's ``on_init`` reports about its data
    (reference ddl/datapusher.py:14-19).  The four reference fields come first and keep their
    meaning; the rest lift the reference's float32-only restriction
    (reference ddl/connection.py:105).

    nData/nValues: rows and elements per row of the window; shape = (nData, nValues);
    splits: column-block widths of the fields of one sample, in *output* elements.
    dtype: element type of the window.  convert: a ``ddl_b200.ops.Convert`` applied while the
    batch is assembled (for example uint8 -> normalised bf16).  shuffle_seed: declare the
    per-iteration local shuffle as ``default_rng(shuffle_seed).shuffle(window)`` so that it runs
    on the GPU (perm_gen) instead of calling ``execute_function``.  variable_length: rows are
    padded to nValues in the window, and after ``post_init`` the producer function exposes
    ``self.lengths`` (valid elements per row); only the valid part of a sample crosses NVLink and
    the batch is padded with ``convert.pad_value`` while it is collated.  replicate: every
    consumer receives the batches of instance 0's window (a shared validation set, for example)
    instead of its own: with ``exchange="push"`` each sample is converted once by the GPU that owns
    it and written once through the NVLS multicast mapping of the batch rings (``multimem.st``;
    NVSwitch replicates it), or once per consumer where multicast is unavailable.
    """

    nData: int
    nValues: int
    shape: Tuple[int, ...]
    splits: Tuple[int, ...]
    dtype: Any = np.float32
    convert: Any = None
    shuffle_seed: Optional[int] = None
    variable_length: bool = False
    replicate: bool = False


@dataclasses.dataclass
class MetaData_Consumer_To_Producer:
    """reference ddl/types.py:7-12"""

    producer_function: Any
    global_shuffle_fraction_exchange: float
    global_shuffle_exchange_method: str
    batch_size: int


@dataclasses.dataclass
class MetaData_Producer_To_Consumer:
    """reference ddl/types.py:15-21 (+ dtype/convert/shuffle_seed, see DataProducerOnInitReturn)"""

    nData: int
    nValues: int
    shape: Tuple[int, ...]
    splits: Tuple[int, ...]
    batches_per_window: int
    dtype: Any = np.float32
    convert: Any = None
    shuffle_seed: Optional[int] = None
    variable_length: bool = False
    replicate: bool = False
