"""Per-kernel roofline bench for exchange_gather_collate and perm_gen on one GPU.

Times each launch with CUDA events on the launching stream (warm-up first, synchronize on both
sides), rotates slot lists and output buffers so consecutive launches never reuse L2 lines, and
reports achieved bytes/s against MEASURED_PEAKS.json (hbm_gbs).  Results go to stdout as JSON
lines and, with --out, to a file for profiles/.

The reference has no kernels; the numbers here are the product's own roofline
fractions.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddl_b200 import _native as N  # noqa: E402
from ddl_b200.ops import Convert, GatherPlan, Pcg64, SourceTable, perm_gen, rng_states_to_tensor, round_up  # noqa: E402


def peaks():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        with open(p) as fh:
            return json.load(fh)["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def bench_gather(name, convert, sample_bytes, batch, shard_bytes, variant, tune, max_ctas=0, iters=20, warmup=3):
    dev = torch.device("cuda:0")
    row_bytes = round_up(sample_bytes, 16)
    rows = max(batch * 4, shard_bytes // row_bytes)
    shard = torch.randint(0, 255, (rows * row_bytes,), dtype=torch.uint8, device=dev)
    table = SourceTable([shard.data_ptr()], dev)
    plan = GatherPlan(table, row_bytes, sample_bytes, convert, variant=variant, tune=tune, max_ctas=max_ctas)
    n_lists = iters + warmup
    perm = torch.randperm(rows, device=dev)[: min(rows, n_lists * batch)]
    if perm.numel() < n_lists * batch:
        perm = perm.repeat((n_lists * batch + perm.numel() - 1) // perm.numel())
    slots = perm[: n_lists * batch].to(torch.int64).contiguous().view(n_lists, batch)
    n_out = max(2, min(8, (512 << 20) // max(1, batch * plan.dst_row_bytes)))
    outs = [plan.alloc_batch(batch) for _ in range(n_out)]
    for i in range(warmup):
        plan.run(slots[i], outs[i % n_out])
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for i in range(iters):
        evs[i][0].record()
        plan.run(slots[warmup + i], outs[(warmup + i) % n_out])
        evs[i][1].record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)
    med = ms[len(ms) // 2]
    in_b = batch * sample_bytes
    out_b = batch * plan.out_sample_bytes
    hbm, how = peaks()
    gbs = (in_b + out_b) / med / 1e6
    return {
        "kernel": "exchange_gather_collate", "case": name, "variant": ["auto", "ldg", "tma"][variant], "tune": tune,
        "max_ctas": max_ctas, "batch": batch, "sample_bytes": sample_bytes, "ms_median": round(med, 4),
        "ms_min": round(ms[0], 4), "samples_per_s": round(batch / med * 1e3), "gbs_in_plus_out": round(gbs, 1),
        "frac_of_hbm_copy": round(gbs / hbm, 3), "peak": how,
    }


def bench_gather_overlapped(name, convert, sample_bytes, batch, shard_bytes, variant, tune, streams=3, iters=60,
                            warmup=6):
    """The deployed shape of the loader: `streams` launches in flight on as many streams (ring of output buffers),
    timed as a whole between two events on the default stream -- ramp and tail of one launch hide behind the next."""
    dev = torch.device("cuda:0")
    row_bytes = round_up(sample_bytes, 16)
    rows = max(batch * 4, shard_bytes // row_bytes)
    shard = torch.randint(0, 255, (rows * row_bytes,), dtype=torch.uint8, device=dev)
    plan = GatherPlan(SourceTable([shard.data_ptr()], dev), row_bytes, sample_bytes, convert, variant=variant, tune=tune)
    n_lists = iters + warmup
    perm = torch.randperm(rows, device=dev)
    perm = perm.repeat((n_lists * batch + perm.numel() - 1) // perm.numel())[: n_lists * batch]
    slots = perm.to(torch.int64).contiguous().view(n_lists, batch)
    n_out = 2 * streams
    outs = [plan.alloc_batch(batch) for _ in range(n_out)]
    ss = [torch.cuda.Stream(dev) for _ in range(streams)]
    main = torch.cuda.current_stream(dev)

    def run(first, count):
        for i in range(first, first + count):
            plan.run(slots[i], outs[i % n_out], stream=ss[i % streams])

    run(0, warmup)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(main)
    for st in ss:
        st.wait_event(a)
    run(warmup, iters)
    for st in ss:
        main.wait_stream(st)
    b.record(main)
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / iters
    hbm, how = peaks()
    gbs = batch * (sample_bytes + plan.out_sample_bytes) / ms / 1e6
    return {"kernel": "exchange_gather_collate", "case": name, "variant": ["auto", "ldg", "tma"][variant], "tune": tune,
            "streams": streams, "batch": batch, "sample_bytes": sample_bytes, "ms_per_launch": round(ms, 4),
            "samples_per_s": round(batch / ms * 1e3), "gbs_in_plus_out": round(gbs, 1),
            "frac_of_hbm_copy": round(gbs / hbm, 3), "peak": how}


def bench_perm(n, tables=1, iters=5):
    dev = torch.device("cuda:0")
    t = torch.arange(n, dtype=torch.int64, device=dev).repeat(tables, 1).contiguous()
    st = rng_states_to_tensor([Pcg64(i) for i in range(tables)], dev)
    perm_gen(t, st)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        perm_gen(t, st)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / iters
    host = np.arange(n, dtype=np.int64)
    r = Pcg64(0)
    t0 = time.perf_counter()
    r.shuffle(host)
    host_ms = (time.perf_counter() - t0) * 1e3
    return {"kernel": "perm_gen", "n": n, "tables": tables, "ms": round(ms, 3),
            "msteps_per_s": round(n / ms / 1e3, 2), "host_cxx_ms": round(host_ms, 3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default=None, help="substring filter on case names")
    args = ap.parse_args()
    img = Convert.normalize_u8([0.485, 0.456, 0.406], [0.229, 0.224, 0.225], 224 * 224)
    cases = [
        ("imagenet_u8_bf16", img, 150528, 256, 2 << 30),
        ("imagenet_copy", Convert.copy(), 150528, 256, 2 << 30),
        ("gpt2_tokens_i32_i64", Convert.widen_tokens(), 4096, 4096, 1 << 30),
        ("gpt2_tokens_copy", Convert.copy(), 4096, 4096, 1 << 30),
        ("cifar_f32_bf16", Convert.cast_f32_bf16(), 12288, 4096, 1 << 30),
        ("volume_f32_copy", Convert.copy(), 4 * 128**3 * 4, 8, 2 << 30),
        ("volume_f32_bf16", Convert.cast_f32_bf16(), 4 * 128**3 * 4, 8, 2 << 30),
        ("pointwise_36B_copy", Convert.copy(), 36, 4096, 64 << 20),
    ]
    results = []
    for name, cv, sb, batch, shard in cases:
        if args.only and args.only not in name:
            continue
        combos = [(N.VARIANT_LDG, 0), (N.VARIANT_TMA, 0), (N.VARIANT_TMA, 1), (N.VARIANT_TMA, 2), (N.VARIANT_TMA, 3)]
        if args.quick:
            combos = [(N.VARIANT_LDG, 0), (N.VARIANT_TMA, 0)]
        for variant, tune in combos:
            r = bench_gather(name, cv, sb, batch, shard, variant, tune)
            results.append(r)
            print(json.dumps(r), flush=True)
    if not args.only or "batch_sweep" in args.only:
        for b in (64, 256, 1024, 4096):
            for variant in (N.VARIANT_LDG, N.VARIANT_TMA):
                r = bench_gather("imagenet_u8_bf16_batch_sweep", img, 150528, b, 4 << 30, variant, 0)
                results.append(r)
                print(json.dumps(r), flush=True)
            r = bench_gather("imagenet_copy_batch_sweep", Convert.copy(), 150528, b, 4 << 30, N.VARIANT_TMA, 0)
            results.append(r)
            print(json.dumps(r), flush=True)
    if not args.only or "overlap" in args.only:
        for b in (64, 256, 1024):
            for streams in (1, 3):
                r = bench_gather_overlapped("imagenet_u8_bf16_overlap", img, 150528, b, 4 << 30, N.VARIANT_TMA, 0, streams)
                results.append(r)
                print(json.dumps(r), flush=True)
    if not args.only or "sm_budget" in args.only:
        for ctas in (8, 16, 32, 64):
            r = bench_gather("imagenet_u8_bf16_sm_budget", img, 150528, 256, 2 << 30, N.VARIANT_TMA, 0, max_ctas=ctas)
            results.append(r)
            print(json.dumps(r), flush=True)
    if not args.only or "perm" in args.only:
        for n in (10000, 100520, 1_000_000):
            r = bench_perm(n)
            results.append(r)
            print(json.dumps(r), flush=True)
        r = bench_perm(100520, tables=24)
        results.append(r)
        print(json.dumps(r), flush=True)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as fh:
            for r in results:
                fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
