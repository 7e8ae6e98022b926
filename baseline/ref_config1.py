"""BASELINE config 1 on the reference: synthetic TensorDataset 10k x 3x32x32 fp32, `mpirun -n 2`
(1 DataPusher + 1 consumer) on the CPU.  Runs the UNMODIFIED reference (baseline/_ref) over the in-box MPI
stand-in and prints samples/s at the consumer (the batch is touched: summed, so the zero-copy views are read).

    python baseline/ref_config1.py            # launches the 2 ranks itself
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))

if os.environ.get("BDDL_MPI_RANK") is None:  # launcher
    sys.path.insert(0, HERE)
    import mpirun

    rc = mpirun.launch(2, [os.path.abspath(__file__)], extra_env={"OMP_NUM_THREADS": "2"}, timeout=300)
    sys.exit(rc)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import ddl  # noqa: E402  (the reference)

N, V, B, EPOCHS, WARM = 10000, 3 * 32 * 32, 250, 12, 2


class Producer(ddl.ProducerFunctionSkeleton):
    def on_init(self, *a, **kw):
        super().on_init(*a, **kw)
        self.rng = np.random.default_rng(self.rank_global)
        return ddl.DataProducerOnInitReturn(N, V, (N, V), (V,))

    def post_init(self, *a, **kw):
        self.my_ary = kw["my_ary"]
        self.my_ary[...] = np.random.default_rng(7).random((N, V), dtype=np.float32)

    def execute_function(self, *a, **kw):
        self.rng.shuffle(self.my_ary)


@ddl.distributed_dataloader
def main(mpi_env, conn):
    loader = ddl.DistributedDataLoader(producer_function=Producer(), batch_size=B, connection=conn, n_epochs=EPOCHS,
                                       fraction_exchange=0.0, exchange_method="sendrecv_replace", instance_idx=0,
                                       n_instances=1)
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
    print(json.dumps({"config": "tensor10k", "impl": "reference", "mpi": "in-box MPI stand-in", "ranks": 2,
                      "samples_per_s": round(seen / dt, 1), "epochs_timed": EPOCHS - WARM, "checksum": acc}), flush=True)


if __name__ == "__main__":
    main()
