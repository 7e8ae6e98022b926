"""Variable-size samples through the public API: rows padded in the window, only the valid prefix
is moved, batches padded while collating (BASELINE north star: variable-size samples; the
reference has fixed float32 rows only, ddl/connection.py:105)."""
import numpy as np
import pytest
import torch

import ddl_b200 as ddl


class Ragged(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, max_len):
        super().__init__()
        self.n, self.max_len = n, max_len

    def on_init(self, **kw):
        super().on_init(**kw)
        cv = ddl.Convert(ddl.ops.gather.N.CVT_I32_I64, pad_value=-1)
        return ddl.DataProducerOnInitReturn(self.n, self.max_len, (self.n, self.max_len), (self.max_len,),
                                            dtype=np.int32, convert=cv, shuffle_seed=self.rank_global,
                                            variable_length=True)

    def post_init(self, **kw):
        a = kw["my_ary"]
        g = np.random.default_rng(3)
        self.lengths = g.integers(1, self.max_len + 1, size=self.n)
        a[...] = 77777  # garbage beyond the valid prefix must never be delivered
        for i, l in enumerate(self.lengths):
            a[i, :l] = i * 1000 + np.arange(l)


def run_case(device):
    n, L, B = 300, 40, 32

    @ddl.distributed_dataloader(n_instances=1, n_producers=1, device=device)
    def main(env, conn):
        loader = ddl.DistributedDataLoader(Ragged(n, L), B, conn, 2)
        lens_ref = np.random.default_rng(3).integers(1, L + 1, size=n)
        order = np.arange(n)
        rng = np.random.default_rng(1)  # rank_global of producer 1 on instance 0
        for _ in range(2):
            rng.shuffle(order)
            for i, (x,) in enumerate(loader):
                rows = order[i * B:(i + 1) * B]
                lens = loader.batch_lengths().cpu().numpy()
                assert np.array_equal(lens, lens_ref[rows])
                x = x.cpu().numpy()
                assert x.dtype == np.int64
                for j, r in enumerate(rows):
                    l = lens_ref[r]
                    assert np.array_equal(x[j, :l], r * 1000 + np.arange(l))
                    assert (x[j, l:] == -1).all()
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)

    main()


def test_variable_length_cpu():
    run_case("cpu")


@pytest.mark.gpu
def test_variable_length_gpu():
    run_case(None)


class Tokens(ddl.ProducerFunctionSkeleton):
    def on_init(self, **kw):
        super().on_init(**kw)
        return ddl.DataProducerOnInitReturn(256, 64, (256, 64), (64,), dtype=np.int32,
                                            convert=ddl.Convert.widen_tokens(), shuffle_seed=self.rank_global)

    def post_init(self, **kw):
        kw["my_ary"][...] = np.arange(256 * 64, dtype=np.int32).reshape(256, 64)


def run_sequence_shard(device):
    """Context/sequence parallelism: each consumer receives only its column slice."""
    for cp_rank in range(4):
        @ddl.distributed_dataloader(n_instances=1, n_producers=1, device=device,
                                    config=ddl.LoaderConfig(seq_rank=cp_rank, seq_size=4))
        def main(env, conn):
            loader = ddl.DistributedDataLoader(Tokens(), 32, conn, 1)
            order = np.arange(256)
            np.random.default_rng(1).shuffle(order)
            for i, (x,) in enumerate(loader):
                assert x.shape == (32, 16) and x.dtype == torch.int64
                want = order[i * 32:(i + 1) * 32, None] * 64 + cp_rank * 16 + np.arange(16)[None, :]
                assert np.array_equal(x.cpu().numpy(), want)
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)

        main()


def test_sequence_shard_cpu():
    run_sequence_shard("cpu")


@pytest.mark.gpu
def test_sequence_shard_gpu():
    run_sequence_shard(None)
