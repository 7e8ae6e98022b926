"""Unit tests of the runtime layers on the CPU: callbacks (designed vs as-wired dispatch),
control-plane communicator, topology arithmetic, shared-memory heap, configuration.  The reference
has no unit tests at all; ranks here are threads sharing a HashStore."""
import logging
import threading

import numpy as np
import pytest
import torch

import ddl_b200 as ddl
from ddl_b200.config import LoaderConfig
from ddl_b200.exceptions import DoesNotMatchError
from ddl_b200.parallel import World, init_env
from ddl_b200.parallel.heap import SymmetricHeap
from ddl_b200.utils import execute_callbacks, for_all_methods, with_logging


def run_ranks(n, fn, timeout=60):
    """Run fn(world) on n threads (one rank each); re-raise the first failure."""
    worlds = World.local(n)
    errs, out = [], [None] * n

    def body(w):
        try:
            out[w.rank] = fn(w)
        except BaseException as e:  # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=body, args=(w,), daemon=True) for w in worlds]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout)
    assert not any(t.is_alive() for t in ts), "a rank hung"
    if errs:
        raise errs[0]
    return out


# -- callbacks -----------------------------------------------------------------------------------------
class A:
    def __init__(self):
        self.calls = []

    def on_init(self, **kw):
        self.calls.append("on_init")
        return "A"

    def execute_function(self, **kw):
        self.calls.append("execute_function")


class B:
    def __init__(self):
        self.calls = []

    def global_shuffle(self, **kw):
        self.calls.append("global_shuffle")

    def on_init(self, **kw):
        self.calls.append("on_init")
        return "B"


def test_designed_dispatch_runs_every_callback():
    a, b = A(), B()
    assert execute_callbacks("on_init", [a, b]) == "A"  # first non-None result
    execute_callbacks("global_shuffle", [a, b])
    assert b.calls == ["on_init", "global_shuffle"]


def test_as_wired_dispatch_stops_after_the_first_callback():
    """reference ddl/utils.py:9-22: the return sits inside the loop."""
    a, b = A(), B()
    assert execute_callbacks("global_shuffle", [a, b], compat=True) is None
    assert b.calls == []  # the shuffler behind the user's function is never reached
    assert execute_callbacks("on_init", [a, b], compat=True) == "A" and b.calls == []


def test_with_logging_is_free_when_debug_is_off(caplog):
    class Loud:
        def __repr__(self):
            raise AssertionError("repr must not be built when DEBUG is off")

    @for_all_methods(with_logging)
    class T:
        def f(self, x):
            return 7

    logging.getLogger("ddl_b200").setLevel(logging.WARNING)
    assert T().f(Loud()) == 7


# -- communicator --------------------------------------------------------------------------------------
def test_comm_point_to_point_fifo_and_collectives():
    def body(w):
        c = w.comm
        nxt, prv = (c.rank + 1) % c.size, (c.rank - 1) % c.size
        for i in range(5):
            c.send_obj((c.rank, i), nxt, tag=3)
        got = [c.recv_obj(prv, tag=3) for _ in range(5)]
        assert got == [(prv, i) for i in range(5)]
        assert c.allgather_obj(c.rank * 10) == [r * 10 for r in range(c.size)]
        assert c.bcast_obj("x" if c.rank == 2 else None, 2) == "x"
        c.barrier()
        sub = c.split(c.rank % 2, key=c.rank)
        assert sub.size == (c.size + 1 - c.rank % 2) // 2 and sub.members == [r for r in range(c.size) if r % 2 == c.rank % 2]
        a = np.full(1000, c.rank, dtype=np.float32)
        c.sendrecv_replace(a, dest=nxt, source=prv, sendtag=50, recvtag=50)
        assert (a == prv).all()
        if c.rank == 0:
            c.ssend_obj("sync", 1, tag=9)
        if c.rank == 1:
            assert c.recv_obj(0, tag=9) == "sync"
        return True

    assert all(run_ranks(4, body))


def test_sendrecv_replace_bulk_goes_through_shared_memory():
    def body(w):
        c = w.comm
        nxt, prv = (c.rank + 1) % c.size, (c.rank - 1) % c.size
        a = np.full((512, 1024), c.rank, dtype=np.float32)  # 2 MiB: the shared-memory path
        a[:, 0] = np.arange(512)
        c.sendrecv_replace(a, dest=nxt, source=prv, sendtag=51, recvtag=51)
        assert (a[:, 1:] == prv).all() and (a[:, 0] == np.arange(512)).all()
        return True

    assert all(run_ranks(3, body))


# -- topology ------------------------------------------------------------------------------------------
def test_topology_matches_the_reference_layout():
    """W=12, n_instances=3 (reference ddl/ddl_env.py:42-44,75-77)."""
    def body(w):
        env = init_env(3, world=w, device="cpu", make_consumer_group=False)
        return (env.color, env.color_nth_pusher, env.comm_per_gpu.members, env.comm_nth_pusher.members,
                env.is_consumer, env.transport, env.n_producers)

    res = run_ranks(12, body)
    assert [r[0] for r in res] == [0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2]
    assert [r[1] for r in res] == [0, 1, 2, 3] * 3
    assert res[5][2] == [4, 5, 6, 7] and res[5][3] == [1, 5, 9]
    assert [r[4] for r in res] == [True, False, False, False] * 3
    assert all(r[5] == "process" and r[6] == 3 for r in res)


def test_topology_rejects_a_world_that_does_not_divide():
    def body(w):
        with pytest.raises(DoesNotMatchError):
            init_env(3, world=w, device="cpu", make_consumer_group=False)
        return True

    assert all(run_ranks(4, body))


def test_topology_device_transport_and_storage_ranks():
    def body(w):
        env = init_env(3, world=w, device="cpu", n_producers=2, n_storage=1, make_consumer_group=False)
        return env.is_storage, env.is_consumer, env.storage_rank_of(1), env.virtual_rank_global(2)

    res = run_ranks(4, body)
    assert [r[0] for r in res] == [False, False, False, True]
    assert all(r[2] == 3 for r in res)
    assert [r[3] for r in res[:3]] == [2, 5, 8]  # instance * (P + 1) + k, the reference's rank_global


# -- heap ------------------------------------------------------------------------------------------------
def test_shared_memory_heap_is_visible_to_peers():
    def body(w):
        heap = SymmetricHeap(w.comm, torch.device("cpu"))
        mine = heap.alloc("x", 1024)
        mine.tensor(torch.int64, (128,)).fill_(w.rank + 1)
        heap.publish()
        w.comm.barrier()
        seen = [int(heap.peer_buffer(r, "x").tensor(torch.int64, (128,))[5]) for r in range(w.comm.size)]
        w.comm.barrier()
        heap.close()
        return seen

    assert run_ranks(3, body) == [[1, 2, 3]] * 3


# -- config ------------------------------------------------------------------------------------------------
def test_config_env_overrides_and_validation(monkeypatch):
    monkeypatch.setenv("DDL_B200_RING_DEPTH", "5")
    monkeypatch.setenv("DDL_B200_COMPAT", "as_wired")
    monkeypatch.setenv("DDL_B200_PREFETCH", "off")
    cfg = LoaderConfig.from_env(timeout_s=9.0)
    assert cfg.ring_depth == 5 and cfg.compat == "as_wired" and cfg.prefetch is False and cfg.timeout_s == 9.0
    with pytest.raises(ValueError):
        LoaderConfig.from_env(compat="sometimes")
    with pytest.raises(ValueError):
        LoaderConfig(ring_depth=2, prefetch_depth=2).validate()


def test_public_names_match_the_reference():
    """reference ddl/__init__.py:7-13"""
    for name in ("ProducerFunctionSkeleton", "DataProducerOnInitReturn", "distributed_dataloader",
                 "DistributedDataLoader", "Marker"):
        assert hasattr(ddl, name)
    assert {m.name for m in ddl.Marker} == {"END_OF_BATCH", "END_OF_EPOCH"}
    r = ddl.DataProducerOnInitReturn(4, 2, (4, 2), (1, 1))
    assert (r.nData, r.nValues, r.shape, r.splits) == (4, 2, (4, 2), (1, 1))
    with pytest.raises(NotImplementedError):
        ddl.ProducerFunctionSkeleton().execute_function()


def test_config_built_directly_is_validated():
    """A LoaderConfig that never went through from_env is checked too: ring_depth <= prefetch_depth
    would let a prefetch overwrite the ring slot the training loop is still reading."""
    import pytest

    with pytest.raises(ValueError):
        ddl.LoaderConfig(ring_depth=2, prefetch_depth=2)
    with pytest.raises(ValueError):
        ddl.LoaderConfig(exchange="teleport")
    assert ddl.LoaderConfig(ring_depth=3, prefetch_depth=2).ring_depth == 3
