"""Program run under baseline/mpirun.py against the UNMODIFIED reference (`import ddl` resolves to
baseline/_ref/ddl).  It only uses the reference's public API: ProducerFunctionSkeleton,
DataProducerOnInitReturn, distributed_dataloader, DistributedDataLoader, Marker.

Modes (BDDL_REF_CFG json):
  order  rows carry their global id; every consumer dumps the ids it received per epoch
         (parity oracle for ddl_b200: same seed -> same permutation and batch order).
  bench  ImageNet-shaped uint8 rows stored as float32 by byte reinterpretation (the reference
         only has float32 windows, reference ddl/connection.py:105); the consumer does what a user
         of the reference must do to get GPU batches: view as uint8, .to(device), normalise, cast.

`wiring_fix`: the reference never calls its global shuffler.  With the fix the
harness's producer function -- which is callbacks[0] -- defines global_shuffle() itself and forwards
to the reference's own SendRecvReplaceGlobalShuffler.  No reference file is edited either way.
"""
import json
import os
import time

import numpy as np
import torch

import ddl  # the reference
from mpi4py import MPI

CFG = json.loads(os.environ["BDDL_REF_CFG"])
MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


class RefProducer(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, v, fill, fraction):
        super().__init__()
        self.n, self.v, self.fill, self.fraction = n, v, fill, fraction
        self.rng = None
        self._gs = None

    def on_init(self, *a, **kw):
        super().on_init(*a, **kw)
        self.rng = np.random.default_rng(seed=self.rank_global)
        splits = (1, self.v - 1) if self.fill == "ids" else (self.v,)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), splits)

    def post_init(self, *a, **kw):
        self.my_ary = kw["my_ary"]
        if self.fill == "ids":
            gid = self.rank_global * 1_000_000 + np.arange(self.n)
            self.my_ary[:, 0] = gid.astype(np.float32)
            self.my_ary[:, 1:] = (gid[:, None] % 97).astype(np.float32)
        else:
            g = np.random.default_rng(1234 + self.rank_global)
            flat = self.my_ary.reshape(-1).view(np.uint8)
            step = 1 << 26
            for o in range(0, flat.size, step):
                m = min(step, flat.size - o)
                flat[o:o + m] = np.frombuffer(g.bytes(m), dtype=np.uint8)
        self._post_init_extra(kw)

    def _post_init_extra(self, kw):
        pass

    def execute_function(self, *a, **kw):
        self.rng.shuffle(self.my_ary)


class RefProducerWired(RefProducer):
    """Harness-side wiring fix (option i)."""

    def _post_init_extra(self, kw):
        n_instance = kw.get("n_instance", 1)
        ne = int(self.n * self.fraction)
        if n_instance > 1 and ne > 0:
            from ddl.shuffle import SendRecvReplaceGlobalShuffler

            world = MPI.COMM_WORLD
            group = world.Get_size() // n_instance
            k = world.Get_rank() % group
            comm = world.Create_group([k + i * group for i in range(n_instance)])  # members only
            self._gs = SendRecvReplaceGlobalShuffler(self.my_ary, ne, comm, world)

    def global_shuffle(self, *a, **kw):
        if self._gs is not None:
            self._gs.global_shuffle()


@ddl.distributed_dataloader
def main(cfg, mpi_env, conn):
    inst = mpi_env.comm_nth_pusher.Get_rank()
    n_inst = mpi_env.comm_nth_pusher.Get_size()
    n, v, B, E = cfg["n"], cfg["v"], cfg["B"], cfg["epochs"]
    cls = RefProducerWired if cfg.get("wiring_fix") else RefProducer
    loader = ddl.DistributedDataLoader(producer_function=cls(n, v, cfg["fill"], cfg.get("fraction", 0.5)),
                                       batch_size=B, connection=conn, n_epochs=E,
                                       fraction_exchange=cfg.get("fraction", 0.5),
                                       exchange_method="sendrecv_replace", instance_idx=inst, n_instances=n_inst)
    if cfg["mode"] == "order":
        seen = []
        for _ in range(E):
            ids = []
            for i, (a, b) in enumerate(loader):
                ids.append(a.numpy()[:, 0].astype(np.int64).copy())
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            seen.append(np.concatenate(ids))
        np.save(os.path.join(cfg["out"], f"ref_seen_{inst}.npy"), np.stack(seen))
        return

    # bench: deliver normalised bf16 batches to this consumer's GPU
    use_cuda = torch.cuda.is_available() and not cfg.get("cpu")
    dev = torch.device("cuda", inst % torch.cuda.device_count()) if use_cuda else torch.device("cpu")
    hw = v * 4 // 3
    scale = torch.tensor([1.0 / (255.0 * s) for s in STD], device=dev).view(1, 3, 1)
    bias = torch.tensor([-m / s for m, s in zip(MEAN, STD)], device=dev).view(1, 3, 1)
    steps, warm, e2e = cfg["steps"], cfg["warmup"], cfg.get("e2e", False)
    done, t0, ev0, ev1, checks = 0, None, None, None, 0.0
    host_out = torch.empty(1).pin_memory() if use_cuda else torch.empty(1)
    finished = False
    for _ in range(E):
        for i, (x,) in enumerate(loader):
            if not finished:
                if done == warm:
                    if use_cuda:
                        torch.cuda.synchronize(dev)
                        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        ev0.record()
                    t0 = time.perf_counter()
                u8 = x.view(torch.uint8)                       # [B, 3*H*W] bytes of the shared window
                d = u8.to(dev)                                 # pageable host -> device
                y = (d.view(B, 3, hw).float() * scale + bias).to(torch.bfloat16)
                if e2e:
                    host_out.copy_(y[0, 0, :64].float().sum().reshape(1), non_blocking=True)
                    if use_cuda:
                        torch.cuda.current_stream(dev).synchronize()
                    checks += float(host_out[0])
                done += 1
                if done == warm + steps:
                    if use_cuda:
                        ev1.record()
                        torch.cuda.synchronize(dev)
                    wall = time.perf_counter() - t0
                    ms_dev = ev0.elapsed_time(ev1) if use_cuda else wall * 1e3
                    with open(os.path.join(cfg["out"], f"ref_bench_{inst}.json"), "w") as fh:
                        json.dump({"ms_device": ms_dev, "ms_wall": wall * 1e3, "steps": steps, "device": str(dev)}, fh)
                    finished = True
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)


if __name__ == "__main__":
    main(CFG)
