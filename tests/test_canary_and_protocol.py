"""Race-detection stand-ins (the reference has none, its only checks are mypy and
ruff, reference .gitlab-ci.yml:80-95): the (lend, batch) canaries of the batch ring, and a
hypothesis *stateful* test that drives the consumer API through random interleavings while a
model of the window ownership machine (FILLING -> OFFERED -> LENT -> FILLING, reference
ddl/connection.py:153-182) checks every transition the producers make."""
import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, Phase, settings
from hypothesis import strategies as st
from hypothesis.stateful import RuleBasedStateMachine, initialize, invariant, precondition, rule

import ddl_b200 as ddl
from ddl_b200.datapusher import DevicePusher
from oracle import expected_epochs


class IdShard(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, v=4):
        super().__init__()
        self.n, self.v = n, v

    def on_init(self, **kw):
        super().on_init(**kw)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (1, self.v - 1),
                                            shuffle_seed=self.rank_global)

    def post_init(self, **kw):
        a = kw["my_ary"]
        gid = self.rank_global * 1_000_000 + np.arange(self.n)
        a[:, 0] = gid
        a[:, 1:] = (gid[:, None] % 97)


def _loader(n, B, P, epochs, cfg):
    from ddl_b200.connection import Connection
    from ddl_b200.parallel.topology import init_env

    env = init_env(1, n_producers=P, device=torch.device("cpu"))
    conn = Connection(env.comm_per_gpu_shm, env, cfg)
    return ddl.DistributedDataLoader(IdShard(n), B, conn, epochs, 0.5, "sendrecv_replace", 0, 1)


def test_canary_passes_on_a_correct_run_and_catches_a_refilled_slot():
    cfg = ddl.LoaderConfig(canary=True, ring_depth=3, prefetch_depth=2, metrics=False)
    loader = _loader(512, 64, 2, 3, cfg)
    for _ in range(2):
        for a, b in loader:
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)
    it = iter(loader)
    next(it)
    # what a too-shallow ring would do: refill the slot the training loop still holds
    loader._stamp(loader._cur_slot, loader._cur_batch + 1, loader.fetch)
    with pytest.raises(RuntimeError, match="canary"):
        loader.mark(ddl.Marker.END_OF_BATCH)
    loader.close()


class WindowModel:
    """Executable statement of the ownership machine of one producer's window (two table buffers)."""

    def __init__(self):
        self.state = "FILLING"   # of the newest iteration
        self.lent_buf = None
        self.ready_buf = None
        self.produced = 0

    def produce(self, buf):
        assert buf != self.lent_buf, "a producer rewrote the table buffer the consumer is reading"
        self.ready_buf, self.state = buf, "OFFERED"
        self.produced += 1

    def lend(self, buf):
        assert self.ready_buf == buf and self.state == "OFFERED", "borrowed a window that was not offered"
        self.lent_buf, self.state = buf, "LENT"

    def take_back(self):
        assert self.lent_buf is not None, "returned a window that was not lent"
        self.lent_buf = None
        if self.state == "LENT":
            self.state = "FILLING"


class LoaderMachine(RuleBasedStateMachine):
    """Random walks through the consumer protocol (batches, marks, checkpoints) on the CPU engine."""

    @initialize(n=st.sampled_from([96, 200, 333]), B=st.sampled_from([16, 32, 50]), P=st.integers(1, 3),
                depth=st.integers(1, 3))
    def start(self, n, B, P, depth):
        self.n, self.B, self.P, self.E = n, B, P, 7
        self.models = [WindowModel() for _ in range(P)]
        self._patch()
        cfg = ddl.LoaderConfig(ring_depth=depth + 1, prefetch_depth=depth, metrics=False, canary=True)
        self.loader = _loader(n, B, P, self.E, cfg)
        self.exp = expected_epochs(1, P, n, B, self.E, 0.5)[0]
        self.epoch, self.batch, self.held, self.done = 0, 0, False, False

    def _patch(self):
        """Wrap the producer role's three transitions so the model sees (and checks) each of them."""
        models = self.models
        self._orig = (DevicePusher.produce, DevicePusher.lend, DevicePusher.take_back)
        produce, lend, take_back = self._orig

        def produce_(p):
            produce(p)
            models[p.k].produce(p._ready[0])

        def lend_(p):
            models[p.k].lend(p._ready[0])  # run-ahead producers produce again inside lend()
            lend(p)

        def take_back_(p):
            take_back(p)
            models[p.k].take_back()

        DevicePusher.produce, DevicePusher.lend, DevicePusher.take_back = produce_, lend_, take_back_

    @precondition(lambda self: not self.done and not self.held and self.batch < self.n // self.B)
    @rule()
    def take_batch(self):
        a, b = self.loader[self.batch]
        want = self.exp[self.epoch][self.batch * self.B:(self.batch + 1) * self.B]
        assert np.array_equal(a[:, 0].numpy().astype(np.int64), want)
        self.held = True

    @precondition(lambda self: not self.done and self.held)
    @rule()
    def mark_batch(self):
        self.loader.mark(ddl.Marker.END_OF_BATCH)
        self.held, self.batch = False, self.batch + 1

    @precondition(lambda self: not self.done and not self.held and self.batch == self.n // self.B)
    @rule()
    def mark_epoch(self):
        self.loader.mark(ddl.Marker.END_OF_EPOCH)
        self.epoch, self.batch = self.epoch + 1, 0
        if self.epoch == self.E:
            self.done = True

    @precondition(lambda self: not self.done and not self.held and 0 < self.batch < self.n // self.B)
    @rule()
    def checkpoint_roundtrip(self):
        state = self.loader.state_dict()
        self.loader.load_state_dict(state)

    @rule()
    def idle(self):
        """Always enabled: a finished run (all epochs consumed) has nothing else left to do."""

    @invariant()
    def at_most_one_window_lent(self):
        if hasattr(self, "models"):
            assert sum(m.lent_buf is not None for m in self.models) <= 1

    def teardown(self):
        try:
            if hasattr(self, "loader") and not self.done:
                self.loader.close()
        finally:
            if hasattr(self, "_orig"):
                DevicePusher.produce, DevicePusher.lend, DevicePusher.take_back = self._orig


LoaderMachine.TestCase.settings = settings(max_examples=12, stateful_step_count=60, deadline=None,
                                            suppress_health_check=list(HealthCheck),
                                            phases=(Phase.explicit, Phase.reuse, Phase.generate))
TestLoaderMachine = LoaderMachine.TestCase


def test_reference_batch_matches_delivered_batch_on_the_cpu_engine():
    """`loader.reference_batch()` (independent row fetch + plain PyTorch convert) equals what the loader
    delivered, bit for bit -- the oracle the multi-GPU benchmarks use, exercised here on the CPU engine with a
    converting window (uint8 -> bf16, per-channel scale and bias)."""
    n, B, hw = 256, 32, 48
    cv = ddl.Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), hw)
    data = np.random.default_rng(4).integers(0, 256, size=(n, 3 * hw), dtype=np.uint8)

    class Img(ddl.ProducerFunctionSkeleton):
        def on_init(self, **kw):
            super().on_init(**kw)
            return ddl.DataProducerOnInitReturn(n, 3 * hw, (n, 3 * hw), (3 * hw,), dtype=np.uint8, convert=cv,
                                                shuffle_seed=9)

        def post_init(self, **kw):
            kw["my_ary"][...] = data

    from ddl_b200.connection import Connection
    from ddl_b200.parallel.topology import init_env

    env = init_env(1, n_producers=1, device=torch.device("cpu"))
    loader = ddl.DistributedDataLoader(Img(), B, Connection(env.comm_per_gpu_shm, env, ddl.LoaderConfig(metrics=False)), 1)
    for (x,) in loader:
        (ref,) = loader.reference_batch()
        assert x.dtype == torch.bfloat16 and torch.equal(x.view(torch.int16), ref.view(torch.int16))
        loader.mark(ddl.Marker.END_OF_BATCH)
    loader.mark(ddl.Marker.END_OF_EPOCH)
