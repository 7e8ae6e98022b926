"""Worker program for the multi-process loader tests (the analogue of the reference's
tests/run_ddl.py:107-240): every sample's first value is its global id, so the order in which
batches arrive can be compared with an oracle.  Launched under torch.distributed.run."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ddl_b200 as ddl  # noqa: E402

CFG = json.loads(os.environ["BDDL_TEST_CFG"])


class ByteProducer(ddl.ProducerFunctionSkeleton):
    """uint8 rows: bytes 0..3 = global id (little endian), the rest a pattern derived from it;
    delivered as bf16 through the fused convert (scale 1, bias 0: every byte value is exact)."""

    def __init__(self, n, v):
        super().__init__()
        self.n, self.v = n, v

    def on_init(self, **kw):
        super().on_init(**kw)
        from ddl_b200 import _native as N

        cv = ddl.Convert(N.CVT_U8_BF16, (1.0,), (0.0,), 0)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (4, self.v - 4), dtype=np.uint8,
                                            convert=cv, shuffle_seed=self.rank_global)

    def post_init(self, **kw):
        a = kw["my_ary"]
        gid = (self.rank_global * 1_000_000 + np.arange(self.n)).astype(np.int64)
        a[:, :4] = gid.astype("<u4").view(np.uint8).reshape(self.n, 4)
        a[:, 4:] = ((gid[:, None] * 7 + np.arange(self.v - 4)[None, :]) % 251).astype(np.uint8)


class RaggedProducer(ddl.ProducerFunctionSkeleton):
    """Variable-size int32 rows: row of global id g holds g*100 + [0, len) and garbage beyond."""

    def __init__(self, n, v):
        super().__init__()
        self.n, self.v = n, v

    def on_init(self, **kw):
        super().on_init(**kw)
        from ddl_b200 import _native as N

        cv = ddl.Convert(N.CVT_I32_I64, pad_value=-1)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (self.v,), dtype=np.int32, convert=cv,
                                            shuffle_seed=self.rank_global, variable_length=True)

    def post_init(self, **kw):
        a = kw["my_ary"]
        self.lengths = np.random.default_rng(100 + self.rank_global).integers(1, self.v + 1, size=self.n)
        gid = self.rank_global * 1_000_000 + np.arange(self.n)
        a[...] = 77777
        for i, l in enumerate(self.lengths):
            a[i, :l] = gid[i] * 100 + np.arange(l)


class IdProducer(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, v, declare_seed, group_size):
        super().__init__()
        self.n, self.v, self.declare_seed, self.group_size = n, v, declare_seed, group_size
        self.rng = None

    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(seed=self.rank_global)
        cv = ddl.Convert.cast_f32_bf16() if CFG.get("convert") else None
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (1, self.v - 1), convert=cv,
                                            shuffle_seed=self.rank_global if self.declare_seed else None,
                                            replicate=bool(CFG.get("replicate")))

    def post_init(self, **kw):
        self.my_ary = kw["my_ary"]
        # ids travel as float32: they must stay below 2^24, so worlds with many producers use a smaller stride
        gid = (0 if CFG.get("small_ids") else self.rank_global * CFG.get("id_stride", 1_000_000)) + np.arange(self.n)
        self.my_ary[:, 0] = gid.astype(np.float32)
        self.my_ary[:, 1:] = (gid[:, None] % 97).astype(np.float32)

    def execute_function(self, **kw):
        self.calls = getattr(self, "calls", 0) + 1
        if CFG.get("die_after") and self.calls > CFG["die_after"]:
            import time

            time.sleep(12)  # a producer that stops answering mid-run (fault injection)
            os._exit(0)
        self.rng.shuffle(self.my_ary)


def body(env, conn):
    n, v, B, E = CFG["n"], CFG["v"], CFG["B"], CFG["epochs"]
    group_size = env.n_producers + 1
    producer = ByteProducer(n, v) if CFG.get("u8") else IdProducer(n, v, CFG.get("declare_seed", False), group_size)
    if CFG.get("ragged"):
        producer = RaggedProducer(n, v)
    loader = ddl.DistributedDataLoader(producer, B, conn, E, CFG.get("fraction", 0.0), "sendrecv_replace", env.color,
                                       env.n_instances)
    seen = []
    if CFG.get("die_after"):
        try:
            for e in range(E):
                for i, (a, b) in enumerate(loader):
                    loader.mark(ddl.Marker.END_OF_BATCH)
                loader.mark(ddl.Marker.END_OF_EPOCH)
        except ddl.exceptions.LoaderTimeoutError as err:
            with open(os.path.join(CFG["out"], "timeout.txt"), "w") as fh:
                fh.write(str(err))
            os._exit(0)
        os._exit(3)
    if CFG.get("ragged"):
        for e in range(E):
            ids = []
            for (x,) in loader:
                lens = loader.batch_lengths().cpu().numpy()
                x = x.cpu().numpy()
                gid = x[:, 0] // 100
                for j in range(B):
                    want = np.random.default_rng(100 + int(gid[j]) // 1_000_000).integers(1, v + 1, size=n)[gid[j] % 1_000_000]
                    assert lens[j] == want and np.array_equal(x[j, :want], gid[j] * 100 + np.arange(want))
                    assert (x[j, want:] == -1).all(), "padding must replace whatever lies beyond the valid prefix"
                ids.append(gid.astype(np.int64))
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            seen.append(np.concatenate(ids))
        np.save(os.path.join(CFG["out"], f"seen_{env.color}.npy"), np.stack(seen))
        return
    if CFG.get("resume"):
        # a fresh world continues from a checkpointed position (process transport: replayed hand-overs)
        e0, b0 = CFG["resume"]
        loader.load_state_dict({"transport": "process", "epoch": e0, "batch": b0, "target_rank": 0})
        ids = []
        for e in range(e0, E):
            for a, b in loader:
                ids.append(a.detach().cpu().float().numpy()[:, 0].astype(np.int64))
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
        np.save(os.path.join(CFG["out"], f"seen_{env.color}.npy"), np.concatenate(ids))
        return
    if CFG.get("u8"):
        pat = torch.arange(v - 4, device=env.device)[None, :]
        w = torch.tensor([1, 256, 65536, 16777216], device=env.device)
        for e in range(E):
            ids = []
            for a, b in loader:
                assert a.dtype == torch.bfloat16 and a.shape == (B, 4) and b.shape == (B, v - 4)
                gid = (a.long() * w).sum(dim=1)
                assert torch.equal(b.long(), (gid[:, None] * 7 + pat) % 251), "payload does not belong to its id"
                ids.append(gid.cpu().numpy())
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            seen.append(np.concatenate(ids))
        np.save(os.path.join(CFG["out"], f"seen_{env.color}.npy"), np.stack(seen))
        return
    if CFG.get("defer"):
        # no host synchronisation per batch: the fetch streams are held back on the device while the
        # host races ahead through window hand-overs; batches are only looked at after the last epoch
        kept = []
        for e in range(E):
            row = []
            for i, (a, b) in enumerate(loader):
                for fs in loader.fetch_streams:
                    with torch.cuda.stream(fs):
                        torch.cuda._sleep(3_000_000)
                row.append((a.clone(), b[:, 0].clone()))
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            kept.append(row)
        torch.cuda.synchronize()
        for row in kept:
            ids = [a.cpu().float().numpy()[:, 0].astype(np.int64) for a, _ in row]
            for (a, b0), g in zip(row, ids):
                assert np.array_equal(b0.cpu().float().numpy().astype(np.int64), g % 97)
            seen.append(np.concatenate(ids))
        np.save(os.path.join(CFG["out"], f"seen_{env.color}.npy"), np.stack(seen))
        return
    for e in range(E):
        ids = []
        for i, (a, b) in enumerate(loader):
            assert a.shape == (B, 1) and b.shape == (B, v - 1)
            ids.append(a.detach().cpu().float().numpy()[:, 0].astype(np.int64))
            assert np.array_equal(b.cpu().float().numpy()[:, 0].astype(np.int64), ids[-1] % 97)
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)
        seen.append(np.concatenate(ids))
    np.save(os.path.join(CFG["out"], f"seen_{env.color}.npy"), np.stack(seen))


if __name__ == "__main__":
    cfg = ddl.LoaderConfig.from_env(compat=CFG.get("compat"), shuffle_mode=CFG.get("shuffle_mode"),
                                    timeout_s=CFG.get("timeout_s", 120.0), verify=CFG.get("verify"),
                                    exchange=CFG.get("exchange"), barrier=CFG.get("barrier"),
                                    ring_depth=CFG.get("ring_depth"), prefetch_depth=CFG.get("prefetch_depth"),
                                    fetch_streams=CFG.get("fetch_streams"), multicast=CFG.get("multicast"))
    main = ddl.distributed_dataloader(body, n_instances=CFG["n_instances"], n_producers=CFG.get("n_producers"),
                                      config=cfg, device="cpu" if CFG.get("cpu", True) else None,
                                      n_storage=CFG.get("n_storage", 0))
    try:
        main()
    except BaseException:
        import traceback

        # torchrun buries a rank's traceback under its own report: keep a copy the test can show
        with open(os.path.join(CFG["out"], f"error_rank{os.environ.get('RANK', '0')}.txt"), "w") as fh:
            traceback.print_exc(file=fh)
        raise
