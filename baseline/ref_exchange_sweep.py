"""BASELINE config 5, reference side: the reference's own SendRecvReplaceGlobalShuffler.global_shuffle
(reference ddl/shuffle.py:92-108) timed per sample size on host rows.  Run under baseline/mpirun.py:

    python baseline/mpirun.py -np 8 baseline/ref_exchange_sweep.py

The numbers are "reference code over in-box MPI stand-in": the stand-in moves bulk payloads
through POSIX shared memory (two memcpys), which is what a shared-memory MPI BTL does too.
"""
import json
import os
import time

import numpy as np

from ddl.shuffle import SendRecvReplaceGlobalShuffler  # the reference
from mpi4py import MPI

world = MPI.COMM_WORLD
rank, size = world.Get_rank(), world.Get_size()
out = []
for sample_bytes in (4096, 65536, 1 << 20, 16 << 20, 64 << 20):
    rows = max(2, (64 << 20) // sample_bytes)
    ary = np.full((rows, sample_bytes // 4), rank, dtype=np.float32)
    sh = SendRecvReplaceGlobalShuffler(ary, rows, world, world)
    sh.global_shuffle()
    world.Barrier()
    iters = 4
    t0 = time.perf_counter()
    for _ in range(iters):
        sh.global_shuffle()
    world.Barrier()
    dt = (time.perf_counter() - t0) / iters
    rec = {"n_ranks": size, "sample_bytes": sample_bytes, "rows_exchanged": rows, "bytes_per_rank": rows * sample_bytes,
           "ms": round(dt * 1e3, 3), "gbs_per_rank": round(rows * sample_bytes / dt / 1e9, 2),
           "label": "reference code over in-box MPI stand-in (CPU, host rows)"}
    out.append(rec)
    if rank == 0:
        print(json.dumps(rec), flush=True)
if rank == 0 and os.environ.get("BDDL_SWEEP_OUT"):
    with open(os.environ["BDDL_SWEEP_OUT"], "a") as fh:
        for r in out:
            fh.write(json.dumps(r) + "\n")
