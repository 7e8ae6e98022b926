"""BASELINE config 1 on ddl_b200: synthetic TensorDataset 10k x 3x32x32 fp32, 1 DataPusher process + 1 consumer
on the CPU -- the same loop as baseline/ref_config1.py (every batch is read: summed).

    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/config1_cpu.py [--declare]

--declare: the producer declares its shuffle (shuffle_seed), so it gets two windows and the native out-of-place
shuffle one iteration ahead of the consumer; without it the user's numpy shuffle runs in place, as in the reference.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ddl_b200 as ddl  # noqa: E402

N, V, B, EPOCHS, WARM = 10000, 3 * 32 * 32, 250, 12, 2
ap = argparse.ArgumentParser()
ap.add_argument("--declare", action="store_true")
ARGS = ap.parse_args()


class Producer(ddl.ProducerFunctionSkeleton):
    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(self.rank_global)
        return ddl.DataProducerOnInitReturn(N, V, (N, V), (V,), shuffle_seed=self.rank_global if ARGS.declare else None)

    def post_init(self, **kw):
        self.my_ary = kw["my_ary"]
        self.my_ary[...] = np.random.default_rng(7).random((N, V), dtype=np.float32)

    def execute_function(self, **kw):
        self.rng.shuffle(self.my_ary)


def main(env, conn):
    loader = ddl.DistributedDataLoader(Producer(), B, conn, EPOCHS, 0.0, "sendrecv_replace", 0, 1)
    acc, t0, seen = 0.0, None, 0
    for e in range(EPOCHS):
        if e == WARM:
            t0, seen = time.perf_counter(), 0
        for (x,) in loader:
            acc += float(x.sum())
            seen += x.shape[0]
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": "tensor10k", "impl": "ddl_b200", "transport": env.transport, "ranks": 2,
                      "declared_shuffle": ARGS.declare, "samples_per_s": round(seen / dt, 1),
                      "epochs_timed": EPOCHS - WARM, "checksum": acc}), flush=True)


if __name__ == "__main__":
    ddl.distributed_dataloader(main, n_instances=1, device="cpu", config=ddl.LoaderConfig.from_env(metrics=False))()
