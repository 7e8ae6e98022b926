"""The other BASELINE.json configs through the public API (loader-only throughput, device-timed,
max over ranks; the timed region lasts >= --min-seconds, spans >= 2 window turnovers and ends after
every fetch stream; one batch per rank is checked bit for bit against an independent fetch).  Launch with torch.distributed.run for N > 1.

  --config tensor10k   config 1: TensorDataset-like 10k x 3x32x32 fp32.  With DDL_B200_INSTANCES=1 and
                       2 processes this is the reference layout (1 DataPusher process + 1 consumer).
  --config gpt2        config 3: int32 token rows of length 1024, widened to int64 while collating.
  --config volume      config 4: 4 x 128^3 fp32 volumes (32 MiB per sample), producers:consumers 1:1.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ddl_b200 as ddl  # noqa: E402
from ddl_b200.utils.benchmarking import check_numerics, timed_loop  # noqa: E402

CONFIGS = {
    # name: (rows, nValues, dtype, batch, convert factory, splits)
    "tensor10k": (10000, 3 * 32 * 32, np.float32, 250, lambda: None, (3 * 32 * 32,)),
    "gpt2": (65536, 1024, np.int32, 512, lambda: ddl.Convert.widen_tokens(), (1024,)),
    "volume": (64, 4 * 128 ** 3, np.float32, 4, lambda: None, (4 * 128 ** 3,)),
    "volume_bf16": (64, 4 * 128 ** 3, np.float32, 4, lambda: ddl.Convert.cast_f32_bf16(), (4 * 128 ** 3,)),
    # ImageNet-shaped uint8 -> bf16; with --replicate every consumer receives instance 0's window (push form:
    # DDL_B200_EXCHANGE=push; DDL_B200_MULTICAST=auto|off selects multimem.st vs one store per consumer)
    "imagenet": (4096, 3 * 224 * 224, np.uint8, 256,
                 lambda: ddl.Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), 224 * 224),
                 (3 * 224 * 224,)),
}


class Synthetic(ddl.ProducerFunctionSkeleton):
    def __init__(self, rows, nv, dtype, convert, splits, replicate=False):
        super().__init__()
        self.rows, self.nv, self.dtype, self.convert, self.splits = rows, nv, np.dtype(dtype).str, convert, splits
        self.replicate = replicate
        self.rng = None

    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(self.rank_global)
        return ddl.DataProducerOnInitReturn(self.rows, self.nv, (self.rows, self.nv), self.splits,
                                            dtype=np.dtype(self.dtype), convert=self.convert,
                                            shuffle_seed=self.rank_global, replicate=self.replicate)

    def post_init(self, **kw):
        a = kw["my_ary"]
        self.my_ary = a
        flat = a.reshape(-1).view(np.uint8)
        g = np.random.default_rng(99 + self.rank_global)
        step = 1 << 26
        for o in range(0, flat.size, step):
            m = min(step, flat.size - o)
            flat[o:o + m] = np.frombuffer(g.bytes(m), dtype=np.uint8) & 0x3F  # keeps fp32 patterns finite

    def execute_function(self, **kw):
        self.rng.shuffle(self.my_ary)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=sorted(CONFIGS))
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--instances", default="world")
    ap.add_argument("--storage", type=int, default=0, help="dedicated storage GPUs (producers:consumers = storage:rest)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--timeout-s", type=float, default=None)
    ap.add_argument("--replicate", action="store_true", help="replicated window: everybody receives instance 0's batches")
    ap.add_argument("--min-seconds", type=float, default=0.5)
    args = ap.parse_args()
    rows, nv, dtype, B, cvf, splits = CONFIGS[args.config]
    out = {}
    eng_ref = []
    loader_cfg = ddl.LoaderConfig.from_env(metrics=False, ring_depth=4, prefetch_depth=2, fetch_streams=2,
                                           timeout_s=args.timeout_s)

    def body(env, conn):
        loader = ddl.DistributedDataLoader(Synthetic(rows, nv, dtype, cvf(), splits, args.replicate), B, conn, 10**9, 0.5,
                                           "sendrecv_replace", env.color, env.n_instances)
        eng_ref.append(loader.engine)
        t = timed_loop(loader, env, args.steps, args.warmup, False, args.min_seconds, int(2.5 * (rows // B)),
                       sample_clocks=env.comm_global.rank == 0)
        nums = env.comm_global.allgather_obj(check_numerics(loader, t.pop("it")))
        x = loader[0]
        ms, steps = t["ms"], t["total_steps"]
        n = env.n_instances
        sample_bytes = nv * np.dtype(dtype).itemsize
        out.update(config=args.config, transport=env.transport, device=str(env.device.type), n_instances=n,
                   producers_per_instance=env.n_producers, storage_gpus=env.n_storage, batch_per_instance=B, sample_bytes=sample_bytes,
                   out_dtype=str(x[0].dtype), ms_per_step=round(ms / steps, 5),
                   samples_per_s=round(n * B * steps / ms * 1e3, 1),
                   input_gbs=round(n * B * steps * sample_bytes / ms / 1e6, 1), rank=env.comm_global.rank,
                   exchange=loader_cfg.exchange, replicate=args.replicate,
                   multicast=bool(getattr(eng_ref[0], "mc_ring", 0)) if eng_ref else None,
                   steps=args.steps, repeats=t["repeats"], timed_region_s=round(ms * 1e-3, 4),
                   windows_crossed=t["windows_crossed"], clocks=t["clocks"],
                   mismatched_rows=sum(m["mismatched_rows"] for m in nums),
                   rows_checked=sum(m["rows"] for m in nums))
        if args.config == "gpt2":
            out["tokens_per_s"] = round(out["samples_per_s"] * 1024)
        loader.close()

    n_inst = args.instances if args.instances == "world" else int(args.instances)
    ddl.distributed_dataloader(body, n_instances=n_inst, n_producers=None if n_inst != "world" else 1, config=loader_cfg,
                               n_storage=args.storage)()
    if out.get("rank") == 0:
        out.pop("rank")
        print(json.dumps(out), flush=True)
        if args.out:
            with open(args.out, "a") as fh:
                fh.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    main()
