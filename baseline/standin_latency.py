"""How much of the reference's window hand-over is the MPI stand-in?  Zero-byte Ssend/Recv ping-pong and the
cancellable Issend + Waitany form the reference uses (reference ddl/connection.py:153-182), over the stand-in.

    python baseline/standin_latency.py
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
if os.environ.get("BDDL_MPI_RANK") is None:
    sys.path.insert(0, HERE)
    import mpirun

    sys.exit(mpirun.launch(2, [os.path.abspath(__file__)], timeout=120))

from mpi4py import MPI  # noqa: E402  (the stand-in)

comm = MPI.COMM_WORLD
rank = comm.Get_rank()
N = 300
comm.Barrier()
t0 = time.perf_counter()
for _ in range(N):
    if rank == 0:
        comm.Ssend([None, 0, MPI.INT], dest=1, tag=7)
        comm.Recv([None, 0, MPI.INT], source=1, tag=7)
    else:
        comm.Recv([None, 0, MPI.INT], source=0, tag=7)
        comm.Ssend([None, 0, MPI.INT], dest=0, tag=7)
blocking = (time.perf_counter() - t0) / N / 2
comm.Barrier()
never = comm.Ibarrier() if False else None
t0 = time.perf_counter()
for _ in range(N):
    if rank == 0:
        comm.Ssend([None, 0, MPI.INT], dest=1, tag=8)
    else:
        req = comm.Irecv([None, 0, MPI.INT], source=0, tag=8)
        MPI.Request.Waitany([req, MPI.Request()])
waitany = (time.perf_counter() - t0) / N
if rank == 0:
    print(json.dumps({"standin_zero_byte_ssend_recv_us": round(blocking * 1e6, 1),
                      "standin_ssend_to_waitany_us": round(waitany * 1e6, 1)}), flush=True)
