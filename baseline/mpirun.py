"""`mpirun -np N python script.py ...` for the baseline-only MPI stand-in: starts N ranks of
this interpreter with the stand-in `mpi4py` and the reference install on PYTHONPATH.

    python baseline/mpirun.py -np 4 tests/run_ddl.py
"""
import argparse
import os
import socket
import subprocess
import sys
import time
import uuid

HERE = os.path.dirname(os.path.abspath(__file__))


def launch(np_, argv, extra_env=None, timeout=None, cwd=None, quiet=False, log_dir=None):
    """Start `np_` ranks and wait for them.  With `log_dir` every rank's stderr goes to
    `<log_dir>/rank_<r>.err` (the caller reports the tail of a failing rank)."""
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    session = uuid.uuid4().hex[:8]
    procs = []
    for r in range(np_):
        env = dict(os.environ)
        env.update(extra_env or {})
        env.update(BDDL_MPI_RANK=str(r), BDDL_MPI_SIZE=str(np_), BDDL_MPI_PORT=str(port), BDDL_MPI_SESSION=session)
        env["PYTHONPATH"] = os.pathsep.join([os.path.join(HERE, "mpi_standin"), os.path.join(HERE, "_ref"),
                                             env.get("PYTHONPATH", "")])
        err = open(os.path.join(log_dir, f"rank_{r}.err"), "w") if log_dir else None
        procs.append(subprocess.Popen([sys.executable] + list(argv), env=env, cwd=cwd,
                                      stdout=subprocess.DEVNULL if quiet else None, stderr=err))
        if err is not None:
            err.close()
    rc = 0
    deadline = None if timeout is None else time.monotonic() + timeout
    try:
        for p in procs:  # one deadline for the whole world, not one per rank
            p.wait(timeout=None if deadline is None else max(0.1, deadline - time.monotonic()))
            rc = rc or p.returncode
    except subprocess.TimeoutExpired:
        rc = 124
    finally:
        for p in procs:  # only the processes started here
            if p.poll() is None:
                p.terminate()
        for p in procs:
            try:
                p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                p.kill()
    if rc != 0:
        _unlink_session(session)
    return rc


def _unlink_session(session):
    """Ranks that were killed cannot unlink their shared-memory windows: do it for them."""
    import glob

    for f in glob.glob(f"/dev/shm/bddlmpi_{session}_*"):
        try:
            os.unlink(f)
        except OSError:
            pass


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("-np", type=int, required=True)
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("argv", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    sys.exit(launch(a.np, a.argv, timeout=a.timeout))
