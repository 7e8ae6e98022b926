"""Shared measurement code of bench.py and benchmarks/*.py: the timed loop, the clock sampler
and the bit-exact numerics check.  The reference ships no benchmark; the rules
followed here are the profiling recipe's: CUDA events on the launching stream after warm-up, max
over ranks, clocks sampled during the timed region."""
from __future__ import annotations

import math
import subprocess
import threading
import time


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.idx, self.rows, self.proc = device_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def batches(loader, ddl, first_field_only=True):
    while True:
        for fields in loader:
            yield fields[0] if first_field_only else fields
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)


def timed_loop(loader, env, steps, warmup, e2e: bool, min_seconds: float, min_total_steps: int, sample_clocks: bool):
    """Warm up, then time `repeats * steps` loader steps between two CUDA events on the training
    stream (max over ranks).  The second event is ordered after every fetch stream, so exactly the
    counted batch assemblies start and finish inside the timed region (batches that were already
    prefetched when it starts are balanced by the prefetches still in flight when it ends).
    `repeats` is chosen so that the region lasts at least `min_seconds` and spans at least
    `min_total_steps` steps (several window turnovers: the per-epoch index_exchange + perm_gen +
    cross-rank barrier are inside it); the reported time per step is the mean."""
    import torch

    import ddl_b200 as ddl
    from ddl_b200 import _native as N

    dev = env.device
    cuda = dev.type == "cuda"  # the CPU engine runs the same loop (dry runs of this file without a GPU)
    main = torch.cuda.current_stream(dev) if cuda else None
    comm = env.comm_global
    host_out = torch.empty(1, dtype=torch.int64).pin_memory() if cuda else None
    it = iter(batches(loader, ddl))

    def run(n):
        for _ in range(n):
            next(it)
            if e2e and cuda:  # device->host read of the step's result: a checksum of the delivered batch
                host_out.copy_(loader.batch_checksum(), non_blocking=True)
                main.synchronize()

    def sync():
        if cuda:
            torch.cuda.synchronize(dev)

    class Span:
        """CUDA events on the training stream; the end event is ordered after every fetch stream."""

        def __enter__(self):
            self.t0 = time.perf_counter()
            if cuda:
                self.a, self.b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                self.a.record(main)
            return self

        def __exit__(self, *exc):
            if cuda:
                for fs in loader.fetch_streams:
                    main.wait_stream(fs)
                self.b.record(main)
                torch.cuda.synchronize(dev)
                eng = getattr(loader, "engine", None)
                if eng is not None and hasattr(eng, "check_waits"):
                    eng.check_waits()  # a bounded device-side wait that gave up raises here
            self.wall = time.perf_counter() - self.t0
            self.ms = self.a.elapsed_time(self.b) if cuda else self.wall * 1e3
            return False

    bpw = len(loader)
    run(max(warmup, bpw + 2))  # untimed: at least the requested steps and one window turnover
    sync()
    comm.barrier()
    with Span() as cal:  # calibration pass (not part of the result): how long do `steps` steps take here?
        run(steps)
    cal_s = max(cal.ms * 1e-3, cal.wall if e2e else 0.0, 1e-7)
    want = max(math.ceil(min_seconds / cal_s), math.ceil(min_total_steps / steps), 1)
    repeats = int(min(max(comm.allgather_obj(want)), 200000))
    total = repeats * steps

    comm.barrier()
    sync()
    sampler = ClockSampler(dev.index or 0)
    if sample_clocks and cuda:
        sampler.start()
    link0 = nvlink_counters(dev.index or 0) if (sample_clocks and cuda) else None
    comm.barrier()
    l0 = N.launch_count()
    with Span() as sp:
        run(total)
    launches = N.launch_count() - l0
    comm.barrier()
    link1 = nvlink_counters(dev.index or 0) if (sample_clocks and cuda) else None
    comm.barrier()
    clocks = sampler.stop() if (sample_clocks and cuda) else None
    ms = max(sp.ms, sp.wall * 1e3) if e2e else sp.ms
    all_ms = comm.allgather_obj(ms)
    return {"ms": max(all_ms), "ms_min_rank": min(all_ms), "launches": launches, "clocks": clocks,
            "repeats": repeats, "total_steps": total, "wall_s": sp.wall, "windows_crossed": total // max(1, bpw), "it": it,
            "nvlink_rx_bytes": (link1[1] - link0[1]) if (link0 and link1) else None,
            "nvlink_tx_bytes": (link1[0] - link0[0]) if (link0 and link1) else None}


def check_numerics(loader, it) -> dict:
    """One more batch, compared bit for bit with ``loader.reference_batch()``: rows fetched by plain
    cudaMemcpyAsync from their owners' shards + the fp32 PyTorch convert expression."""
    import torch

    x = next(it)
    ref = loader.reference_batch()[0]
    if x.is_cuda:
        torch.cuda.synchronize()
    xb = x.contiguous().view(torch.uint8).reshape(x.shape[0], -1)
    rb = ref.contiguous().view(torch.uint8).reshape(ref.shape[0], -1)
    bad_rows = int((~(xb == rb).all(dim=1)).sum().item())
    fl = x.is_floating_point()
    return {"rows": int(x.shape[0]), "mismatched_rows": bad_rows,
            "max_abs_err": float((x.double() - ref.double()).abs().max().item()),
            "finite": bool(torch.isfinite(x.float()).all().item()) if fl else True}


def nvlink_counters(device_index: int):
    """Sum of the NVLink data counters of one GPU, (tx_bytes, rx_bytes), from
    ``nvidia-smi nvlink -gt d`` -- or None where the driver does not report them.  Bracketing a run
    with two readings shows the traffic that really crossed the links next to the algorithmic bytes."""
    import re

    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(device_index)], capture_output=True,
                             text=True, timeout=20).stdout
    except Exception:
        return None
    tx = [int(x) for x in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out)]
    rx = [int(x) for x in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out)]
    if not tx or not rx:
        return None
    return sum(tx) * 1024, sum(rx) * 1024
