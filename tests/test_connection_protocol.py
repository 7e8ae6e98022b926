"""The lend/return token protocol and shutdown of the process transport under stress
(reference ddl/connection.py:153-187; SURVEY.md 3.4 state machine).  Ranks are threads; the flag
page and the windows are real POSIX shared memory, the waits are the native bounded waits."""
import random
import threading
import time

import numpy as np
import pytest

from ddl_b200.config import LoaderConfig
from ddl_b200.connection import Connection
from ddl_b200.parallel import init_env
from ddl_b200.types import WorkerInfo
from test_units_cpu import run_ranks


@pytest.mark.parametrize("n_producers,epochs", [(1, 40), (3, 60)])
def test_no_lost_or_duplicated_handover(n_producers, epochs):
    rows = 257

    def body(w):
        env = init_env(1, world=w, device="cpu", make_consumer_group=False)
        conn = Connection(env.comm_per_gpu_shm, env, LoaderConfig(timeout_s=30.0))
        rng = random.Random(w.rank)
        if env.is_consumer:
            arys = conn.init_windows([(rows, 3)] * n_producers, [np.float32] * n_producers)
            conn.Barrier()
            reads = [0] * n_producers
            target = 1
            for _ in range(epochs):
                conn.start_access_epoch(target)
                win = arys[target - 1]
                first = win[0, 0]
                assert (win == first).all(), "window changed while lent to the consumer"
                assert first == reads[target - 1] + 1, "an iteration was skipped or delivered twice"
                time.sleep(rng.random() * 0.002)
                assert (win == first).all()
                reads[target - 1] += 1
                conn.end_access_epoch(target)
                target = target % n_producers + 1
            conn.shutdown_operation()
            conn.Barrier()
            conn.finalize(arys)
            return reads
        arys = conn.init_windows((rows, 3), np.float32)
        mine = arys[conn.rank - 1]
        conn.Barrier()
        it = 0
        while True:
            it += 1
            for r in range(0, rows, 64):  # a slow, observable "shuffle"
                mine[r:r + 64] = it
                if rng.random() < 0.05:
                    time.sleep(0.0005)
            conn.heartbeat(conn.rank)
            if conn.Iend_access_epoch() is WorkerInfo.STOP:
                break
            if conn.Istart_access_epoch() is WorkerInfo.STOP:
                break
        conn.Barrier()
        conn.finalize(arys)
        return it

    out = run_ranks(n_producers + 1, body, timeout=120)
    reads = out[0]
    assert sum(reads) == epochs
    for k in range(n_producers):
        # a producer is at most one iteration ahead of what the consumer has read
        assert out[k + 1] in (reads[k], reads[k] + 1)


def test_shutdown_cancels_a_producer_that_is_never_borrowed():
    def body(w):
        env = init_env(1, world=w, device="cpu", make_consumer_group=False)
        conn = Connection(env.comm_per_gpu_shm, env, LoaderConfig(timeout_s=30.0))
        if env.is_consumer:
            arys = conn.init_windows([(8, 1)], [np.float32])
            conn.Barrier()
            time.sleep(0.05)
            conn.shutdown_operation()  # never borrowed anything
            conn.Barrier()
            conn.finalize(arys)
            return "done"
        arys = conn.init_windows((8, 1), np.float32)
        conn.Barrier()
        info = conn.Iend_access_epoch()  # offers, then waits: must be cancelled by the stop word
        conn.Barrier()
        conn.finalize(arys)
        return info

    out = run_ranks(2, body)
    assert out == ["done", WorkerInfo.STOP]
