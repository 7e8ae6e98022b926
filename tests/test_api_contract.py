"""Small behaviours of the public API that the reference also has (or should have had):
index protocol, marker validation, exchange-method check, geometry checks, single-rank worlds."""
import numpy as np
import pytest

import ddl_b200 as ddl


class Tiny(ddl.ProducerFunctionSkeleton):
    def __init__(self, n=64, v=4, splits=(1, 3), seed=True):
        super().__init__()
        self.n, self.v, self.splits, self.seed = n, v, splits, seed

    def on_init(self, **kw):
        super().on_init(**kw)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), self.splits,
                                            shuffle_seed=self.rank_global if self.seed else None)

    def post_init(self, **kw):
        kw["my_ary"][...] = np.arange(self.n * self.v, dtype=np.float32).reshape(self.n, self.v)

    def execute_function(self, **kw):
        np.random.default_rng(0).shuffle(self.my_ary)


def _with_loader(fn, pf=None, B=16, epochs=2, **kw):
    @ddl.distributed_dataloader(n_instances=1, n_producers=kw.pop("P", 1), device="cpu")
    def main(env, conn):
        with ddl.DistributedDataLoader(pf or Tiny(), B, conn, epochs, **kw) as loader:
            fn(loader, env)

    main()


def test_index_protocol_and_markers():
    def body(loader, env):
        assert len(loader) == 4
        with pytest.raises(IndexError):      # the legacy sequence protocol ends an epoch this way
            loader[4]
        with pytest.raises(ValueError):
            loader[-1]
        with pytest.raises(ValueError):
            loader.mark("END_OF_BATCH")
        a, b = loader[2]
        assert a.shape == (16, 1) and b.shape == (16, 3)
        a2, _ = loader[2]                    # random access to the same batch gives the same rows
        assert np.array_equal(a.numpy(), a2.numpy())

    _with_loader(body)


def test_single_rank_world_finishes_cleanly():
    """The reference raises in mark(END_OF_EPOCH) when the world has one rank
    (reference ddl/mpi_dataloader.py:173-174, 232-233)."""
    seen = []

    def body(loader, env):
        for _ in range(2):
            for a, b in loader:
                seen.append(a.shape[0])
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
        assert env.comm_global.size == 1 and env.transport == "device"

    _with_loader(body)
    assert seen == [16] * 8


def test_splits_must_cover_the_row():
    with pytest.raises(ddl.DoesNotMatchError):
        _with_loader(lambda l, e: None, pf=Tiny(splits=(1, 2)))


def test_batch_larger_than_window_is_rejected():
    with pytest.raises(AssertionError):
        _with_loader(lambda l, e: None, B=65)


def test_connection_none_builds_a_one_rank_world():
    loader = ddl.DistributedDataLoader(Tiny(), 16, None, 1, config=ddl.LoaderConfig())
    try:
        (a, b) = loader[0]
        assert a.shape == (16, 1)
    finally:
        loader.close()


def test_metrics_count_batches_and_bytes():
    def body(loader, env):
        for a, b in loader:
            loader.mark(ddl.Marker.END_OF_BATCH)
        m = loader.metrics.as_dict()
        assert m["batches"] >= 4 and m["samples"] >= 64 and m["bytes_in"] >= 64 * 16

    _with_loader(body)


def test_batches_after_shutdown_raise():
    """A closed loader says so instead of touching released buffers."""
    import numpy as np
    import pytest
    import torch

    import ddl_b200 as ddl
    from ddl_b200.connection import Connection
    from ddl_b200.exceptions import ShutdownError
    from ddl_b200.parallel.topology import init_env

    class P(ddl.ProducerFunctionSkeleton):
        def on_init(self, **kw):
            super().on_init(**kw)
            return ddl.DataProducerOnInitReturn(64, 4, (64, 4), (4,), shuffle_seed=1)

        def post_init(self, **kw):
            kw["my_ary"][...] = np.arange(256, dtype=np.float32).reshape(64, 4)

    env = init_env(1, n_producers=1, device=torch.device("cpu"))
    loader = ddl.DistributedDataLoader(P(), 16, Connection(env.comm_per_gpu_shm, env, ddl.LoaderConfig(metrics=False)), 1)
    _ = loader[0]
    loader.close()
    with pytest.raises(ShutdownError):
        loader[0]
