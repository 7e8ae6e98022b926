"""The reference's demo program (reference tests/run_ddl.py) on ddl_b200.

Synthetic point-wise data: `nTimesteps * 10052` rows of 9 float32 values, min-max normalised,
delivered as three fields (3 coordinates, 5 targets, 1 sample weight) in batches of 4096 for 3
epochs -- the workload of the reference's only test (reference tests/run_ddl.py:301-316).

Run it the reference's way (1 consumer + 3 producer processes, host windows, numpy shuffle):

    torchrun --nproc-per-node 4 --master-addr 127.0.0.1 examples/run_ddl_b200.py

or the B200 way (one process per GPU, shards in HBM, shuffles and batch assembly on the GPU):

    DDL_B200_INSTANCES=world torchrun --nproc-per-node <gpus> --master-addr 127.0.0.1 examples/run_ddl_b200.py
    python examples/run_ddl_b200.py            # one GPU (or the CPU execution path)
"""
import os
import sys
from dataclasses import dataclass

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ddl_b200 as ddl  # noqa: E402


@dataclass
class Params:
    nepoch: int = 3
    batch_size: int = 64 * 64
    nData: int = 10


def make_points(n_rows: int, seed: int) -> np.ndarray:
    """[n_rows, 9] float32: parameter, x(2), targets(5), sample weight -- min-max normalised."""
    raw = np.random.default_rng(seed).random((n_rows, 9)).astype(np.float32)
    lo, hi = raw[:, :3].min(axis=0), raw[:, :3].max(axis=0)
    raw[:, :3] = (raw[:, :3] - 0.5 * (lo + hi)) / (0.5 * (hi - lo))
    raw[:, 3:8] /= np.abs(raw[:, 3:8]).max(axis=0)
    raw[:, 8] /= raw[:, 8].mean()
    return raw


class PointProducer(ddl.ProducerFunctionSkeleton):
    def __init__(self, params: Params, idx: int, n_instances: int):
        super().__init__()
        self.params, self.idx, self.n_instances = params, idx, n_instances
        self.rng = None
        self.data = None

    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(seed=self.rank_global)
        n_rows = self.params.nData * 10052 // self.n_instances
        self.data = make_points(n_rows, seed=1000 + self.rank_global)
        # shuffle_seed declares execute_function below, so that on a GPU it runs as perm_gen
        return ddl.DataProducerOnInitReturn(n_rows, 9, (n_rows, 9), (3, 5, 1), shuffle_seed=self.rank_global)

    def post_init(self, **kw):
        self.my_ary = kw["my_ary"]
        self.my_ary[...] = self.data
        self.data = None

    def execute_function(self, **kw):
        self.rng.shuffle(self.my_ary)


@ddl.distributed_dataloader
def main(params, env, conn):
    print(f"[{env.comm_global.rank:03d}] consumer of instance {env.color}/{env.n_instances}: "
          f"{env.transport} transport, {env.n_producers} producer(s), device {env.device}", flush=True)
    with ddl.DistributedDataLoader(PointProducer(params, env.color, env.n_instances), params.batch_size, conn,
                                   params.nepoch, 0.5, "sendrecv_replace", env.color, env.n_instances) as loader:
        net = torch.nn.Sequential(torch.nn.Linear(3, 30), torch.nn.SiLU(), torch.nn.Linear(30, 5)).to(env.device)
        opt = torch.optim.Adam(net.parameters(), lr=7e-4)
        for epoch in range(params.nepoch):
            total, n = 0.0, 0
            for pos, target, weight in loader:
                loss = (weight * (net(pos) - target) ** 2).mean()
                opt.zero_grad(set_to_none=True)
                loss.backward()
                opt.step()
                total, n = total + float(loss), n + 1
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            print(f"[{env.comm_global.rank:03d}] epoch {epoch + 1}/{params.nepoch}: {n} batches, loss {total / n:.4f}",
                  flush=True)
    print("Training finished")


if __name__ == "__main__":
    main(Params())
