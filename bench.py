"""Headline benchmark: GPU-delivered samples/s of the distributed loader.

Metric and config follow BASELINE.json: ImageNet-shaped synthetic uint8 3x224x224 samples,
per-epoch global shuffle (fraction 0.5) across all GPUs, delivered as normalised bf16 batches
in HBM.  One process per GPU (torchrun), weak scaling (fixed shard and batch per GPU).

  value   device-timed (CUDA events, max over ranks) samples/s over all GPUs with the shards
          resident in HBM -- the product's main path: one exchange_gather_collate launch per
          batch pulls rows from every peer over NVLink, converts and collates them.  The timed
          region repeats the requested --steps until it lasts >= --min-seconds and spans several
          window turnovers (the per-epoch index_exchange + perm_gen + barrier are inside it); its
          end event waits for every fetch stream; one batch per rank is compared bit for bit with
          an independent fetch + fp32 PyTorch expression; a value that breaks a physical limit
          (NVLink 900 GB/s, HBM 7.7 TB/s) is not printed.
  e2e     the same metric through the same public API with the dataset resident in pinned
          host memory: every step copies that step's raw samples host->device and reads a
          checksum of the delivered batch back to the host.

`--impl reference` runs the unmodified reference (baseline/_ref) through its own API on the same
config; see baseline/run_reference.py.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)
C_, H_, W_ = 3, 224, 224
SAMPLE = C_ * H_ * W_


def make_producer(rows: int, seed_base: int = 1234):
    import numpy as np

    import ddl_b200 as ddl

    class ImageShard(ddl.ProducerFunctionSkeleton):
        """Synthetic ImageNet-shaped shard: `rows` random uint8 CHW images per producer."""

        def __init__(self):
            super().__init__()

        def on_init(self, **kw):
            super().on_init(**kw)
            cv = ddl.Convert.normalize_u8(MEAN, STD, H_ * W_)
            return ddl.DataProducerOnInitReturn(rows, SAMPLE, (rows, SAMPLE), (SAMPLE,), dtype=np.uint8, convert=cv,
                                                shuffle_seed=self.rank_global)

        def post_init(self, **kw):
            self.my_ary = kw["my_ary"]
            g = np.random.default_rng(seed_base + self.rank_global)
            flat = self.my_ary.reshape(-1)
            step = 1 << 26
            for o in range(0, flat.size, step):  # random bytes, generated in slices to bound memory
                n = min(step, flat.size - o)
                flat[o:o + n] = np.frombuffer(g.bytes(n), dtype=np.uint8)

        def execute_function(self, **kw):  # only used by host-shuffle mode
            np.random.default_rng(self.rank_global).shuffle(self.my_ary)

    return ImageShard()


def run_ours(args) -> dict:
    import ddl_b200 as ddl
    from ddl_b200.utils.benchmarking import check_numerics, timed_loop

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    result = {"rank": int(os.environ.get("RANK", "0"))}

    def body(env, conn):
        B, rows = args.batch, args.rows
        loader = ddl.DistributedDataLoader(make_producer(rows), B, conn, 10**9, args.fraction, "sendrecv_replace",
                                           env.color, env.n_instances)
        t = timed_loop(loader, env, args.steps, args.warmup, False, args.min_seconds, int(2.5 * (rows // B)),
                       sample_clocks=env.comm_global.rank == 0)
        remote = max(env.comm_global.allgather_obj(loader.remote_fraction()))
        nums = env.comm_global.allgather_obj(check_numerics(loader, t.pop("it")))
        loader.close()
        result.update(main=t, remote=remote, numerics=nums, n=env.n_instances)

    cfg = ddl.LoaderConfig.from_env(metrics=False, ring_depth=args.ring, gather_variant=args.variant,
                                    gather_tune=args.tune, prefetch_depth=args.prefetch_depth,
                                    fetch_streams=args.fetch_streams, exchange=args.exchange,
                                    gather_max_ctas=args.max_ctas, timeout_s=args.timeout_s)
    if not args.e2e_only:
        ddl.distributed_dataloader(body, n_instances="world", n_producers=1, config=cfg)()

    e2e = None
    if not args.no_e2e:
        def body_e2e(env, conn):
            B, rows = args.batch, min(args.rows, args.e2e_rows)
            loader = ddl.DistributedDataLoader(make_producer(rows), B, conn, 10**9, args.fraction,
                                               "sendrecv_replace", env.color, env.n_instances)
            t = timed_loop(loader, env, args.steps, args.warmup, True, args.e2e_min_seconds, int(2.5 * (rows // B)),
                           sample_clocks=False)
            nums = env.comm_global.allgather_obj(check_numerics(loader, t.pop("it")))
            loader.close()
            result.update(e2e=t, e2e_numerics=nums)

        cfg2 = ddl.LoaderConfig.from_env(metrics=False, ring_depth=3, residency="host",
                                         host_stage=args.host_stage, host_threads=args.host_threads, gather_variant=args.e2e_variant,
                                         gather_tune=args.e2e_tune)
        ddl.distributed_dataloader(body_e2e, n_instances="world", n_producers=1, config=cfg2)()
        t = result["e2e"]
        v = args.gpus * args.batch * t["total_steps"] / (t["ms"] * 1e-3)
        bad = sum(n["mismatched_rows"] for n in result["e2e_numerics"])
        e2e = {"value": round(v, 1) if bad == 0 else None, "unit": "samples/s", "h2d_bytes_per_step": args.batch * SAMPLE,
               "d2h_bytes_per_step": 8, "steps": args.steps, "repeats": t["repeats"],
               "ms_per_step": round(t["ms"] / t["total_steps"], 5), "timed_region_s": round(t["ms"] * 1e-3, 3),
               "pcie_gbs_per_gpu": round(args.batch * SAMPLE * t["total_steps"] / (t["ms"] * 1e6), 2),
               "numerics": {"mismatched_rows": bad, "rows_checked": sum(n["rows"] for n in result["e2e_numerics"])},
               "path": {"gather": "pinned host shard -> threaded CPU gather -> one cudaMemcpyAsync per step -> convert kernel -> batch checksum kernel -> 8-byte D2H",
                        "memcpy": "pinned host shard -> cudaMemcpyAsync per row -> convert kernel -> batch checksum kernel -> 8-byte D2H",
                        "pull": "pinned host shard -> kernel pulls rows over PCIe (zero-copy) + convert -> batch checksum kernel -> 8-byte D2H"}[args.host_stage]}

    if result["rank"] != 0:
        return {}
    config = {"model": "imagenet_u8_3x224x224_loader", "global_batch": args.batch * args.gpus, "seq_len": SAMPLE,
              "parallelism": f"dp{args.gpus}", "per_gpu_batch": args.batch, "fraction_exchange": args.fraction,
              "l2": "inputs larger than L2 (every row is read once per epoch)"}
    out = {"metric": "gpu_delivered_samples_per_s", "value": None, "unit": "samples/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config,
           "impl": "ddl_b200", "e2e": e2e, "gpu_launches": 0,
           "details": {"rows_per_gpu": args.rows, "exchange": args.exchange, "ring": args.ring,
                       "prefetch_depth": args.prefetch_depth, "fetch_streams": args.fetch_streams}}
    if args.e2e_only:
        out["gpu_launches"] = int(result["e2e"]["launches"])
        return out
    t = result["main"]
    ms_per_step = t["ms"] / t["total_steps"]
    value = args.gpus * args.batch / (ms_per_step * 1e-3)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            hbm_peak, how = float(json.load(fh)["hbm_gbs"]), "measured"
    except Exception:
        hbm_peak, how = 6650.0, "fallback"
    hbm_gbs = args.batch * SAMPLE * 3 / (ms_per_step * 1e6)  # uint8 read + bf16 write per GPU
    nvl = result["remote"] * args.batch * SAMPLE / (ms_per_step * 1e6)
    bad = sum(n["mismatched_rows"] for n in result["numerics"])
    finite = all(n["finite"] for n in result["numerics"])
    # a value that breaks physics or numerics is not printed
    problems = []
    if nvl > 900.0:
        problems.append(f"NVLink inbound {nvl:.0f} GB/s per GPU exceeds the 900 GB/s link rate")
    if hbm_gbs > 7700.0:
        problems.append(f"HBM traffic {hbm_gbs:.0f} GB/s exceeds the 7700 GB/s nominal rate")
    if bad or not finite:
        problems.append(f"{bad} delivered rows differ from the fp32 PyTorch reference")
    out.update({
        "value": None if problems else round(value, 1), "ms_per_step": round(ms_per_step, 5),
        "repeats": t["repeats"], "timed_steps_total": t["total_steps"], "timed_region_s": round(t["ms"] * 1e-3, 4),
        "windows_crossed": t["windows_crossed"],
        "timing": "CUDA events on the training stream, end event ordered after every fetch stream; "
                  "the driver's --steps is repeated `repeats` times back to back and the mean is reported, so the "
                  "region lasts >= %.2f s and spans >= 2 window turnovers (index_exchange + perm_gen + "
                  "cross-rank barrier included)" % args.min_seconds,
        "exchange": {"remote_fraction": round(result["remote"], 4), "nvlink_in_gbs_per_gpu": round(nvl, 1),
                     "frac_of_900_nominal": round(nvl / 900, 3), "frac_of_770_measured": round(nvl / 770, 3),
                     # NVLink data counters of rank 0's GPU around the timed region next to the bytes the
                     # slot tables ask for (prefetches in flight at either end included in the count)
                     "nvlink_rx_bytes_counted": t.get("nvlink_rx_bytes"),
                     "nvlink_rx_bytes_algorithmic": int(result["remote"] * args.batch * SAMPLE * t["total_steps"])},
        "roofline": {"hbm_gbs_per_gpu": round(hbm_gbs, 1), "hbm_peak_gbs": hbm_peak, "peak": how,
                     "frac_of_hbm_copy": round(hbm_gbs / hbm_peak, 3),
                     "bound": "HBM" if args.gpus == 1 else "NVLink (see exchange)"},
        "clocks": t["clocks"], "gpu_launches": int(t["launches"]),
        "gpu_launches_per_step": round(t["launches"] / t["total_steps"], 3),
        "numerics": {"check": "one timed-path batch per rank == cudaMemcpyAsync row fetch + fp32 PyTorch normalise/cast, bit for bit",
                     "rows_checked": sum(n["rows"] for n in result["numerics"]), "mismatched_rows": bad,
                     "max_abs_err": max(n["max_abs_err"] for n in result["numerics"])},
        "numerics_ok": bad == 0 and finite,
    })
    if problems:
        out["rejected"] = problems
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ddl_b200", choices=["ddl_b200", "reference"])
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--rows", type=int, default=16384)
    ap.add_argument("--e2e-rows", type=int, default=4096)
    ap.add_argument("--fraction", type=float, default=0.5)
    ap.add_argument("--ring", type=int, default=6)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--tune", type=int, default=0)
    ap.add_argument("--prefetch-depth", type=int, default=3)
    ap.add_argument("--fetch-streams", type=int, default=3)
    ap.add_argument("--max-ctas", type=int, default=0)
    ap.add_argument("--exchange", default="pull", choices=["pull", "push"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-only", action="store_true")
    ap.add_argument("--e2e-variant", type=int, default=0)
    ap.add_argument("--e2e-tune", type=int, default=0)
    ap.add_argument("--host-threads", type=int, default=4)
    ap.add_argument("--host-stage", default="pull", choices=["gather", "memcpy", "pull"])
    ap.add_argument("--timeout-s", type=float, default=None, help="bound of every device/host wait of the loader")
    ap.add_argument("--watchdog", type=float, default=0.0, help="dump all Python stacks to stderr after this many seconds")
    ap.add_argument("--min-seconds", type=float, default=1.0, help="minimum length of the device-timed region")
    ap.add_argument("--e2e-min-seconds", type=float, default=0.5, help="minimum length of the e2e timed region")
    args = ap.parse_args()
    if args.watchdog > 0:
        import faulthandler

        faulthandler.dump_traceback_later(args.watchdog, exit=False, file=sys.stderr)
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        from baseline.run_reference import run_reference

        out = run_reference(args)
    else:
        out = run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
