"""BASELINE config 2: loader stall while training ResNet-50 (random init, synthetic ImageNet-shaped
uint8 shards, per-epoch global shuffle).  Launch with torch.distributed.run on N GPUs (or plain
python for one).

A/B on the same model and step: (A) every step trains on a fresh batch delivered by the loader,
(B) every step trains on one fixed resident batch (no loader work at all).  Both timed on the
device with CUDA events, max over ranks.  stall % = (A - B) / A.  Target <= 5 % at 8 GPUs.
"""
import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ddl_b200 as ddl  # noqa: E402
from bench import make_producer  # noqa: E402
from ddl_b200.models import resnet50  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--rows", type=int, default=8192)
    ap.add_argument("--max-ctas", type=int, default=0)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    res = {}

    def body(env, conn):
        dev, B = env.device, args.batch
        loader = ddl.DistributedDataLoader(make_producer(args.rows), B, conn, 10**9, 0.5, "sendrecv_replace",
                                           env.color, env.n_instances)
        model = resnet50().to(dev).to(memory_format=torch.channels_last)
        if env.n_instances > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index],
                                                              process_group=env.consumer_group)
        opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
        labels = torch.randint(0, 1000, (B,), device=dev)

        def step(x):
            x = x.view(B, 3, 224, 224).contiguous(memory_format=torch.channels_last)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(model(x), labels)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
            return loss

        def batches():
            while True:
                for (x,) in loader:
                    yield x
                    loader.mark(ddl.Marker.END_OF_BATCH)
                loader.mark(ddl.Marker.END_OF_EPOCH)

        it = batches()
        fixed = next(it).clone()

        def timed(feed):
            for _ in range(args.warmup):
                step(feed())
            torch.cuda.synchronize(dev)
            env.comm_global.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(args.steps):
                step(feed())
            b.record()
            torch.cuda.synchronize(dev)
            return max(env.comm_global.allgather_obj(a.elapsed_time(b) / args.steps))

        ms_fixed = timed(lambda: fixed)
        ms_loader = timed(lambda: next(it))
        ms_fixed2 = timed(lambda: fixed)
        base = min(ms_fixed, ms_fixed2)
        res.update(n_gpus=env.n_instances, batch_per_gpu=B, ms_step_with_loader=round(ms_loader, 3),
                   ms_step_fixed_batch=round(base, 3), stall_pct=round(100 * (ms_loader - base) / ms_loader, 2),
                   images_per_s=round(env.n_instances * B / ms_loader * 1e3, 1), model="resnet50 bf16 channels_last",
                   gather_max_ctas=args.max_ctas, rank=env.comm_global.rank)
        loader.close()

    cfg = ddl.LoaderConfig.from_env(metrics=False, ring_depth=3, gather_max_ctas=args.max_ctas)
    ddl.distributed_dataloader(body, n_instances="world", n_producers=1, config=cfg)()
    if res.get("rank") == 0:
        res.pop("rank")
        print(json.dumps(res), flush=True)
        if args.out:
            with open(args.out, "a") as fh:
                fh.write(json.dumps(res) + "\n")


if __name__ == "__main__":
    main()
