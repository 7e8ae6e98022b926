"""Where does end-to-end delivery from host-resident shards saturate?  

Launch with torch.distributed.run on N GPUs.  Every rank owns a pinned host buffer and measures, with
CUDA events, the host->device rate of
  * cudaMemcpyAsync on a copy engine (what `host_stage="gather"/"memcpy"` uses), and
  * the loader's zero-copy pull: exchange_gather_collate reading 147 KiB rows straight out of mapped
    pinned memory (what `host_stage="pull"` -- the e2e arm of bench.py -- uses),
first ALONE (ranks take turns, the others idle) and then with ALL ranks at once.  If the concurrent
per-GPU rate of the plain copy-engine memcpy drops the same way the loader's does, the limit is the
platform's host memory / root complex, not the loader.  Also prints the NUMA placement the box reports.
"""
import glob
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200.ops import Convert, GatherPlan, SourceTable  # noqa: E402

ROW = 3 * 224 * 224


def rate(fn, nbytes, dev, seconds=0.4):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(dev)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn()
    b.record()
    torch.cuda.synchronize(dev)
    iters = max(3, int(seconds * 1e3 / max(a.elapsed_time(b), 1e-3)))
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize(dev)
    return nbytes * iters / (a.elapsed_time(b) * 1e-3) / 1e9


def main():
    dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
    rank, R = dist.get_rank(), dist.get_world_size()
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    rows = 2048
    host = torch.empty((rows, ROW), dtype=torch.uint8).pin_memory()
    host.random_(0, 255)
    dst = torch.empty_like(host, device=dev)
    nbytes = host.numel()

    def memcpy():
        dst.copy_(host, non_blocking=True)

    # zero-copy pull by the loader's kernel: source 0 = the pinned buffer as the device sees it
    # (under unified addressing a pinned allocation's host address is also its device address)
    table = SourceTable([host.data_ptr()], dev)
    plan = GatherPlan(table, ROW, ROW, Convert.copy())
    slots = torch.randperm(rows).to(torch.int64).to(dev)
    out = plan.alloc_batch(rows)

    def pull():
        plan.run(slots, out)

    res = {"rank": rank, "n_gpus": R}
    for name, fn in (("memcpy", memcpy), ("zero_copy_pull", pull)):
        solo = 0.0
        for turn in range(R):  # alone: everybody else waits at the barrier
            if turn == rank:
                solo = rate(fn, nbytes, dev)
            torch.cuda.synchronize(dev)
            dist.barrier()
        dist.barrier()
        together = rate(fn, nbytes, dev, seconds=0.8)  # all ranks at once
        dist.barrier()
        res[name] = {"alone_gbs": round(solo, 2), "all_ranks_gbs": round(together, 2)}
    everyone = [None] * R
    dist.all_gather_object(everyone, res)
    if rank == 0:
        numa = {}
        for f in glob.glob("/sys/bus/pci/devices/*/numa_node"):
            try:
                vendor = open(os.path.join(os.path.dirname(f), "vendor")).read().strip()
                cls = open(os.path.join(os.path.dirname(f), "class")).read().strip()
                if vendor == "0x10de" and cls.startswith("0x0302"):
                    numa[os.path.basename(os.path.dirname(f))] = int(open(f).read())
            except OSError:
                pass
        summary = {"n_gpus": R, "host_nodes": sorted(os.path.basename(p) for p in glob.glob("/sys/devices/system/node/node[0-9]*")),
                   "gpu_numa_node": numa, "cpus": os.cpu_count()}
        for name in ("memcpy", "zero_copy_pull"):
            a = [e[name]["alone_gbs"] for e in everyone]
            t = [e[name]["all_ranks_gbs"] for e in everyone]
            summary[name] = {"alone_gbs_per_gpu": a, "all_ranks_gbs_per_gpu": t, "alone_mean": round(sum(a) / R, 2),
                             "all_ranks_mean": round(sum(t) / R, 2), "all_ranks_aggregate": round(sum(t), 1)}
        print(json.dumps(summary), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
