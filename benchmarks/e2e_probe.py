import sys, time, torch
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import ddl_b200 as ddl
from bench import make_producer
B, rows = 256, 4096
for stage, sync, prio, ctas in (("pull", True, -1, 0), ("pull", True, 0, 0), ("pull", True, -1, 140), ("pull", True, 0, 140), ("pull", True, 0, 120)):
    cfg = ddl.LoaderConfig(metrics=False, ring_depth=3, residency="host", host_stage=stage, host_threads=8, fetch_priority=prio, gather_max_ctas=ctas)
    @ddl.distributed_dataloader(n_instances=1, n_producers=1, config=cfg)
    def main(env, conn):
        loader = ddl.DistributedDataLoader(make_producer(rows), B, conn, 10**9, 0.5)
        host_out = torch.empty(1).pin_memory()
        def steps():
            while True:
                for (x,) in loader:
                    yield x
                    loader.mark(ddl.Marker.END_OF_BATCH)
                loader.mark(ddl.Marker.END_OF_EPOCH)
        it = steps()
        for _ in range(8): next(it)
        torch.cuda.synchronize()
        t0 = time.perf_counter(); n = 64; tg = 0.0
        for _ in range(n):
            ta = time.perf_counter()
            x = next(it)
            tg += time.perf_counter() - ta
            if sync:
                host_out.copy_(x[0, :64].float().sum().reshape(1), non_blocking=True)
                torch.cuda.current_stream().synchronize()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(stage, "prio", prio, "ctas", ctas, "sync" if sync else "nosync", "ms/step", round(dt / n * 1e3, 3), "samples/s", round(B * n / dt), "host ms in next()", round(tg / n * 1e3, 3))
        loader.close()
    main()
