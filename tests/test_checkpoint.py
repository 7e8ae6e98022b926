"""Checkpoint / resume (absent in the reference): a loader restored from
``state_dict`` continues with exactly the samples the uninterrupted run would have delivered."""
import numpy as np
import torch

import ddl_b200 as ddl


class Shard(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, v):
        super().__init__()
        self.n, self.v = n, v

    def on_init(self, **kw):
        super().on_init(**kw)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (1, self.v - 1),
                                            shuffle_seed=self.rank_global)

    def post_init(self, **kw):
        a = kw["my_ary"]
        a[:, 0] = self.rank_global * 100000 + np.arange(self.n)
        a[:, 1:] = 1.0


def _consume(loader, epochs_left, stop_after=None):
    """ids per delivered batch; optionally stop after `stop_after` batches (between batches)."""
    out, count = [], 0
    for _ in range(epochs_left):
        for a, b in loader:
            out.append(a[:, 0].long().cpu().numpy().copy())
            loader.mark(ddl.Marker.END_OF_BATCH)
            count += 1
            if stop_after is not None and count == stop_after:
                return out, False
        loader.mark(ddl.Marker.END_OF_EPOCH)
    return out, True


def test_resume_continues_the_exact_sequence():
    run_resume_case("cpu")


def run_resume_case(device):
    n, v, B, E, P = 700, 4, 64, 7, 2
    bpw = n // B
    results = {}

    @ddl.distributed_dataloader(n_instances=1, n_producers=P, device=device)
    def straight(env, conn):
        loader = ddl.DistributedDataLoader(Shard(n, v), B, conn, E, 0.5)
        results["full"], _ = _consume(loader, E)

    straight()
    for stop in (3, bpw, 2 * bpw + 5):  # mid-window, exactly at a window change, later mid-window
        @ddl.distributed_dataloader(n_instances=1, n_producers=P, device=device)
        def first(env, conn):
            loader = ddl.DistributedDataLoader(Shard(n, v), B, conn, E, 0.5)
            results["head"], _ = _consume(loader, E, stop_after=stop)
            results["state"] = loader.state_dict()
            loader.close()

        @ddl.distributed_dataloader(n_instances=1, n_producers=P, device=device)
        def second(env, conn):
            loader = ddl.DistributedDataLoader(Shard(n, v), B, conn, E, 0.5)
            loader.load_state_dict(results["state"])
            results["tail"], done = _consume(loader, E - loader.epoch)
            assert done

        first()
        second()
        got = results["head"] + results["tail"]
        assert len(got) == len(results["full"])
        for x, y in zip(got, results["full"]):
            assert np.array_equal(x, y)


def test_pcg64_state_dict_is_numpy_compatible():
    from ddl_b200.ops import Pcg64

    r = Pcg64(77)
    r.next32()
    g = np.random.Generator(np.random.PCG64())
    g.bit_generator.state = r.numpy_state()
    a = np.arange(50)
    b = a.copy()
    g.shuffle(a)
    r.shuffle(b)
    assert np.array_equal(a, b)
