"""BASELINE config 5: sample-size sweep of the fused exchange against an NCCL all-to-all.

Launch with torch.distributed.run on N GPUs.  Every rank owns a shard; each step every rank
assembles a batch whose samples are drawn uniformly from all shards, so (N-1)/N of the bytes cross
NVLink.  Timed on the device with CUDA events, max over ranks.  Reported per GPU: inbound NVLink
GB/s and the fraction of the 900 GB/s nominal / 770 GB/s measured peer-copy rate.

  pull  exchange_gather_collate: the consumer's kernel loads from peer shards (TMA bulk); --mode copy moves
        the bytes unchanged, --mode u8bf16 normalises uint8 to bf16 on the way (the NVLink side still carries
        uint8, the local write side twice the bytes)
  nccl  dist.all_to_all_single of the same byte volume (contiguous blocks -- a real loader would
        additionally need gather and scatter kernels around it)
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200 import _native as N  # noqa: E402
from ddl_b200.ops import Convert, GatherPlan, SourceTable  # noqa: E402
from ddl_b200.parallel import World  # noqa: E402
from ddl_b200.parallel.heap import SymmetricHeap  # noqa: E402
from ddl_b200.utils.benchmarking import ClockSampler, nvlink_counters  # noqa: E402


def timed(fn, iters, warmup, dev):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(dev)
    ms = torch.tensor([a.elapsed_time(b) / iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shard-mb", type=int, default=1024)
    ap.add_argument("--batch-mb", type=int, default=256)
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--mode", default="copy", choices=["copy", "u8bf16"])
    ap.add_argument("--sizes", default="4096,16384,65536,262144,1048576,4194304,16777216,67108864")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    world = World.from_environment()
    rank, R = world.rank, world.size
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    heap = SymmetricHeap(world.comm, dev)
    shard_bytes = args.shard_mb << 20
    shard = heap.alloc("shard", shard_bytes, zero=False)
    shard.bytes().random_(0, 255)
    heap.publish()
    torch.cuda.synchronize(dev)
    world.comm.barrier()
    table = SourceTable([heap.peer_ptr(r, "shard") for r in range(R)], dev)
    results = []
    cv = Convert.copy() if args.mode == "copy" else Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), 0)
    for sb in [int(x) for x in args.sizes.split(",")]:
        rows = shard_bytes // sb
        n = max(R, min((args.batch_mb << 20) // sb, rows))
        n -= n % R
        g = torch.Generator(device="cpu").manual_seed(1000 * rank + 7)
        lists = []
        for _ in range(4):
            src = torch.arange(n) % R                      # uniform over owners: (R-1)/R remote
            row = torch.randint(0, rows, (n,), generator=g)
            perm = torch.randperm(n, generator=g)
            lists.append(((src << N.SLOT_ROW_BITS) | row)[perm].to(torch.int64).to(dev))
        plan = GatherPlan(table, sb, sb, cv)
        outs = [plan.alloc_batch(n) for _ in range(2)]
        it = [0]

        def pull():
            i = it[0]
            it[0] += 1
            plan.run(lists[i % 4], outs[i % 2])

        sampler = ClockSampler(dev.index or 0)
        c0 = nvlink_counters(dev.index or 0) if rank == 0 else None
        if rank == 0:
            sampler.start()
        ms_pull = timed(pull, args.iters, 3, dev)
        clocks = sampler.stop() if rank == 0 else None
        c1 = nvlink_counters(dev.index or 0) if rank == 0 else None
        send = torch.empty(n * sb, dtype=torch.uint8, device=dev)
        recv = torch.empty_like(send)

        def a2a():
            dist.all_to_all_single(recv, send)

        ms_nccl = timed(a2a, args.iters, 3, dev)
        remote = n * sb * (R - 1) / R
        rec = {"n_gpus": R, "sample_bytes": sb, "samples_per_step": n, "remote_bytes_per_gpu": int(remote),
               "pull_ms": round(ms_pull, 4), "pull_nvlink_gbs": round(remote / ms_pull / 1e6, 1),
               "pull_frac_of_900": round(remote / ms_pull / 1e6 / 900, 3),
               "pull_frac_of_770_measured": round(remote / ms_pull / 1e6 / 770, 3),
               "nccl_a2a_ms": round(ms_nccl, 4), "nccl_a2a_gbs": round(remote / ms_nccl / 1e6, 1),
               "speedup_vs_nccl": round(ms_nccl / ms_pull, 2), "mode": args.mode, "iters": args.iters,
               "clocks": clocks}
        if c0 is not None and c1 is not None:
            # counters of rank 0's GPU over warm-up + timed launches vs the bytes the slot lists ask for
            rec["nvlink_rx_bytes_counted"] = c1[1] - c0[1]
            rec["nvlink_rx_bytes_algorithmic"] = int(remote * (args.iters + 3))
        results.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
        del send, recv, outs, lists
    if rank == 0 and args.out:
        with open(args.out, "w") as fh:
            for r in results:
                fh.write(json.dumps(r) + "\n")
    torch.cuda.synchronize(dev)
    world.comm.barrier()
    heap.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
