"""One frozen configuration object for the loader, with ``DDL_B200_*`` environment overrides.

The reference has no configuration system: constructor arguments, three SLURM variables and a
string switch with one legal value (reference ddl/mpi_dataloader.py:108-118,
ddl/ddl_env.py:103-105, ddl/datapusher.py:96-106).
"""
from __future__ import annotations

import dataclasses
import os
from typing import Any


@dataclasses.dataclass(frozen=True)
class LoaderConfig:
    # "designed": every callback runs, so the global shuffle is active;
    # "as_wired": the reference's first-callback-only dispatch, bit-for-bit its behaviour.
    compat: str = "designed"
    # How execute_function is realised on the device transport:
    #   "auto"   -> "device" when on_init declared shuffle_seed, else "host"
    #   "device" -> perm_gen kernel on the slot table (bit-exact default_rng(seed).shuffle)
    #   "trace"  -> run execute_function on a row-id proxy and upload the permutation
    #   "host"   -> run execute_function on the real pinned window and re-upload it
    shuffle_mode: str = "auto"
    ring_depth: int = 2            # batch buffers in flight (double buffering)
    # CUDA priority of the batch-assembly stream.  Equal priority (0) by default: a high-priority
    # all-SM gather was measured to hold back small kernels of the training stream until it ended.
    fetch_priority: int = 0
    prefetch: bool = True          # launch batch i+1 while batch i is being consumed
    prefetch_depth: int = 1        # batches assembled ahead of the consumer (needs ring_depth > prefetch_depth)
    fetch_streams: int = 1         # batch-assembly streams used round-robin (consecutive gathers overlap)
    barrier: str = "auto"          # cross-rank stream barrier: auto | nccl | flags | host
    gather_variant: int = 0        # BDDL_VARIANT_*
    gather_tune: int = 0
    gather_max_ctas: int = 0       # SM budget of the batch-assembly kernel (0 = all SMs)
    # pull: each consumer's kernel loads its samples from their owners (no cross-GPU hand-over per
    # step).  push: each owner's kernel stores the samples it owns into every consumer's ring slot
    # and signals `ready` from inside the kernel; all consumers must step in lock-step.
    exchange: str = "pull"
    # push form: allocate the batch rings with the VMM API and bind them to an NVLS multicast object
    # so that replicated windows are delivered with one multimem.st per 16 bytes.  "auto" = where the
    # platform supports it, "on" = required, "off" = cudaMalloc rings, replicated windows unicast.
    multicast: str = "auto"
    host_pull: bool = False        # process transport: kernel reads the pinned window directly
    residency: str = "hbm"         # device transport: shards in "hbm" or in pinned "host" memory
    # host residency: "gather" (threaded CPU gather into a pinned buffer + one cudaMemcpyAsync),
    # "memcpy" (one cudaMemcpyAsync per row run) or "pull" (kernel reads host memory, zero-copy)
    host_stage: str = "pull"
    host_threads: int = 4
    # sequence/context parallelism: deliver only columns [seq_rank, seq_rank+1) * nValues/seq_size of
    # every sample (single-field windows); the rest of the row never leaves its owner
    seq_rank: int = 0
    seq_size: int = 1
    timeout_s: float = 600.0       # bounded waits: a dead producer raises instead of hanging
    metrics: bool = True
    # integrity mode (the stand-in for a race detector): every batch is assembled a
    # second time by the independent LDG form of the kernel and the two fingerprints are compared
    verify: bool = False
    trace: bool = False            # CUDA-event timing of every loader stage
    # debug canaries: every ring slot is stamped (lend number, batch index) on the
    # fetch stream right behind its assembly; the stamp is checked when the batch is handed out and
    # again when it is marked, so a slot that is refilled while the training loop still holds it, or
    # handed out before its assembly finished, raises instead of corrupting a step silently.
    # Synchronises per batch: a debugging aid, like compute-sanitizer's racecheck would be.
    canary: bool = False

    def __post_init__(self) -> None:
        self.validate()  # also configs built directly (not through from_env) are checked

    @staticmethod
    def from_env(**overrides: Any) -> "LoaderConfig":
        vals: dict = {}
        for f in dataclasses.fields(LoaderConfig):
            raw = os.environ.get("DDL_B200_" + f.name.upper())
            if raw is None:
                continue
            if f.type in ("bool", bool):
                vals[f.name] = raw.lower() in ("1", "true", "yes", "on")
            elif f.type in ("int", int):
                vals[f.name] = int(raw)
            elif f.type in ("float", float):
                vals[f.name] = float(raw)
            else:
                vals[f.name] = raw
        vals.update({k: v for k, v in overrides.items() if v is not None})
        return LoaderConfig(**vals)

    def validate(self) -> None:
        if self.compat not in ("designed", "as_wired"):
            raise ValueError(f"compat must be 'designed' or 'as_wired', not {self.compat!r}")
        if self.shuffle_mode not in ("auto", "device", "trace", "host"):
            raise ValueError(f"unknown shuffle_mode {self.shuffle_mode!r}")
        if self.barrier not in ("auto", "nccl", "flags", "host"):
            raise ValueError(f"unknown barrier {self.barrier!r}")
        if self.multicast not in ("auto", "on", "off"):
            raise ValueError(f"multicast must be 'auto', 'on' or 'off', not {self.multicast!r}")
        if self.exchange not in ("pull", "push"):
            raise ValueError(f"unknown exchange {self.exchange!r}")
        if self.residency not in ("hbm", "host") or self.host_stage not in ("gather", "memcpy", "pull"):
            raise ValueError("residency is 'hbm' or 'host'; host_stage is 'gather', 'memcpy' or 'pull'")
        if self.prefetch_depth < 1 or self.fetch_streams < 1 or self.ring_depth <= self.prefetch_depth:
            raise ValueError("need prefetch_depth >= 1, fetch_streams >= 1 and ring_depth > prefetch_depth")
        if self.ring_depth < 2:
            raise ValueError("ring_depth must be at least 2")
