"""Randomised single-rank loaders (CPU execution path of the device transport) against the
reference semantics: whatever the number of producers, window size, batch size, ring depth,
prefetch depth and shuffle mode, the consumer must see exactly the oracle's order."""
import numpy as np
from hypothesis import given, settings, strategies as st

import ddl_b200 as ddl
from oracle import expected_epochs


class Ids(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, declare):
        super().__init__()
        self.n, self.declare, self.rng = n, declare, None

    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(self.rank_global)
        return ddl.DataProducerOnInitReturn(self.n, 2, (self.n, 2), (1, 1),
                                            shuffle_seed=self.rank_global if self.declare else None)

    def post_init(self, **kw):
        self.my_ary = kw["my_ary"]
        self.my_ary[:, 0] = self.rank_global * 1_000_000 + np.arange(self.n)
        self.my_ary[:, 1] = 0

    def execute_function(self, **kw):
        self.rng.shuffle(self.my_ary)


@given(P=st.integers(1, 3), n=st.integers(20, 300), B=st.integers(4, 40), epochs=st.integers(1, 7),
       ring=st.integers(2, 5), mode=st.sampled_from(["device", "trace", "host"]), data=st.data())
@settings(max_examples=25, deadline=None)
def test_random_configuration_matches_oracle(P, n, B, epochs, ring, mode, data):
    B = min(B, n)
    pd = data.draw(st.integers(1, ring - 1))
    cfg = ddl.LoaderConfig(ring_depth=ring, prefetch_depth=pd, shuffle_mode=mode, prefetch=data.draw(st.booleans()))
    seen = []

    @ddl.distributed_dataloader(n_instances=1, n_producers=P, device="cpu", config=cfg)
    def main(env, conn):
        with ddl.DistributedDataLoader(Ids(n, mode == "device"), B, conn, epochs, 0.5) as loader:
            for _ in range(epochs):
                ids = []
                assert len(loader) == n // B
                for a, _b in loader:
                    ids.append(a[:, 0].numpy().astype(np.int64).copy())
                    loader.mark(ddl.Marker.END_OF_BATCH)
                loader.mark(ddl.Marker.END_OF_EPOCH)
                seen.append(np.concatenate(ids))

    main()
    assert np.array_equal(np.stack(seen), expected_epochs(1, P, n, B, epochs, 0.5)[0])
