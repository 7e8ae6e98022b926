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
import faulthandler
import json
import os
import sys
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

    # bench: deliver normalised bf16 batches to this consumer's GPU.  One run measures both numbers:
    # phase 1 = `timed` steps device-timed (CUDA events, no per-step synchronisation), phase 2 = the
    # same number of steps end to end (per step: pageable H2D of the batch, normalise, a 4-byte D2H
    # read of a result and a stream synchronise; wall clock).
    use_cuda = torch.cuda.is_available() and not cfg.get("cpu")
    dev = torch.device("cuda", inst % torch.cuda.device_count()) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)  # pinned allocations and events belong to this consumer's GPU
    hw = v * 4 // 3
    scale = torch.tensor([1.0 / (255.0 * s) for s in STD], device=dev).view(1, 3, 1)
    bias = torch.tensor([-m / s for m, s in zip(MEAN, STD)], device=dev).view(1, 3, 1)
    timed, warm = cfg["steps"], cfg["warmup"]
    host_out = torch.empty(1).pin_memory() if use_cuda else torch.empty(1)
    res_path = os.path.join(cfg["out"], f"ref_bench_{inst}.json")
    res = {"steps": timed, "device": str(dev)}
    done, checks = 0, 0.0
    ev0 = ev1 = t0 = None

    def sync():
        if use_cuda:
            torch.cuda.synchronize(dev)

    def save():
        with open(res_path + ".tmp", "w") as fh:
            json.dump(res, fh)
        os.replace(res_path + ".tmp", res_path)

    for _ in range(E):
        for i, (x,) in enumerate(loader):
            phase = 0 if done < warm else (1 if done < warm + timed else (2 if done < warm + 2 * timed else 3))
            if phase < 3:
                if done == warm:  # phase 1 starts
                    sync()
                    if use_cuda:
                        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        ev0.record()
                    t0 = time.perf_counter()
                elif done == warm + timed:  # phase 1 ends, phase 2 starts
                    if use_cuda:
                        ev1.record()
                    sync()
                    res["ms_device"] = ev0.elapsed_time(ev1) if use_cuda else (time.perf_counter() - t0) * 1e3
                    save()
                    t0 = time.perf_counter()
                u8 = x.view(torch.uint8)                       # [B, 3*H*W] bytes of the shared window
                d = u8.to(dev)                                 # pageable host -> device
                y = (d.view(B, 3, hw).float() * scale + bias).to(torch.bfloat16)
                if phase == 2:
                    host_out.copy_(y[0, 0, :64].float().sum().reshape(1), non_blocking=True)
                    if use_cuda:
                        torch.cuda.current_stream(dev).synchronize()
                    checks += float(host_out[0])
                done += 1
                if done == warm + 2 * timed:
                    sync()
                    res["ms_e2e"] = (time.perf_counter() - t0) * 1e3
                    res["check"] = checks
                    save()
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)


if __name__ == "__main__":
    # a rank still running after `watchdog` seconds dumps every thread's stack into its stderr log,
    # so a hang names its own line (the harness surfaces the log tail in the bench JSON)
    faulthandler.dump_traceback_later(float(CFG.get("watchdog", 240)), exit=False, file=sys.stderr)
    main(CFG)
