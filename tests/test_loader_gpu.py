"""The public API on a real GPU: batches delivered by the sm_100a kernels must be the batches the
reference semantics prescribe (tests/oracle.py), for HBM- and host-resident shards."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

import ddl_b200 as ddl
from oracle import expected_epochs

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


class IdShard(ddl.ProducerFunctionSkeleton):
    def __init__(self, n, v, declare):
        super().__init__()
        self.n, self.v, self.declare, self.rng = n, v, declare, None

    def on_init(self, **kw):
        super().on_init(**kw)
        self.rng = np.random.default_rng(self.rank_global)
        return ddl.DataProducerOnInitReturn(self.n, self.v, (self.n, self.v), (1, self.v - 1),
                                            shuffle_seed=self.rank_global if self.declare else None)

    def post_init(self, **kw):
        a = kw["my_ary"]
        self.my_ary = a
        gid = self.rank_global * 1_000_000 + np.arange(self.n)
        a[:, 0] = gid
        a[:, 1:] = (gid[:, None] % 97)

    def execute_function(self, **kw):
        self.rng.shuffle(self.my_ary)


def _run(cfg, n=3000, v=12, B=256, E=5, P=2, declare=True):
    seen = []

    @ddl.distributed_dataloader(n_instances=1, n_producers=P, config=cfg)
    def main(env, conn):
        assert env.device.type == "cuda"
        loader = ddl.DistributedDataLoader(IdShard(n, v, declare), B, conn, E, 0.5, "sendrecv_replace", 0, 1)
        for _ in range(E):
            ids = []
            for a, b in loader:
                assert a.is_cuda and a.shape == (B, 1) and b.shape == (B, v - 1)
                ids.append(a[:, 0].long().cpu().numpy())
                assert torch.equal(b[:, 0].long().cpu(), torch.from_numpy(ids[-1] % 97))
                loader.mark(ddl.Marker.END_OF_BATCH)
            loader.mark(ddl.Marker.END_OF_EPOCH)
            seen.append(np.concatenate(ids))

    main()
    exp = expected_epochs(1, P, n, B, E, 0.5)[0]
    assert np.array_equal(np.stack(seen), exp)


@pytest.mark.parametrize("mode,declare", [("device", True), ("trace", False), ("host", False)])
def test_single_gpu_matches_reference_order(mode, declare):
    _run(ddl.LoaderConfig(shuffle_mode=mode), declare=declare)


@pytest.mark.parametrize("stage", ["gather", "memcpy", "pull"])
def test_host_resident_shards(stage):
    _run(ddl.LoaderConfig(residency="host", host_stage=stage), v=16)


def test_prefetch_off_and_deep_ring():
    _run(ddl.LoaderConfig(prefetch=False))
    _run(ddl.LoaderConfig(ring_depth=4))
    _run(ddl.LoaderConfig(ring_depth=4, prefetch_depth=2, fetch_streams=2))
    _run(ddl.LoaderConfig(ring_depth=5, prefetch_depth=3, fetch_streams=3, residency="host"), v=16)


def test_checkpoint_resume_on_gpu():
    import test_checkpoint as tc

    tc.run_resume_case(None)


def test_integrity_mode_cross_checks_kernel_forms():
    _run(ddl.LoaderConfig(verify=True, ring_depth=3, prefetch_depth=2), v=64)


def test_canary_mode_on_gpu():
    """(lend, batch) stamps written behind every assembly by cuStreamWriteValue32 and checked on hand-out and mark."""
    _run(ddl.LoaderConfig(canary=True, ring_depth=3, prefetch_depth=2))
    _run(ddl.LoaderConfig(canary=True, residency="host", host_stage="pull"), v=16)


def test_tracing_reports_device_time_per_stage():
    got = {}

    @ddl.distributed_dataloader(n_instances=1, n_producers=1, config=ddl.LoaderConfig(trace=True))
    def main(env, conn):
        loader = ddl.DistributedDataLoader(IdShard(2048, 8, True), 256, conn, 2, 0.5)
        for _ in range(2):
            for a, b in loader:
                loader.mark(ddl.Marker.END_OF_BATCH)
            got.update(loader.trace_summary())
            loader.mark(ddl.Marker.END_OF_EPOCH)

    main()
    assert got["gather"]["count"] >= 16 and got["gather"]["mean_ms"] > 0
    assert got["produce.window1"]["count"] >= 2


def test_push_exchange_single_gpu():
    """Owner-side stores + ready/free counters (one rank pushes to itself)."""
    _run(ddl.LoaderConfig(exchange="push", ring_depth=3, prefetch_depth=2))


def test_uint8_images_normalised_to_bf16():
    n, B = 1024, 64
    hw = 24 * 24
    cv = ddl.Convert.normalize_u8((0.485, 0.456, 0.406), (0.229, 0.224, 0.225), hw)
    data = np.random.default_rng(1).integers(0, 256, size=(n, 3 * hw), dtype=np.uint8)

    class Img(ddl.ProducerFunctionSkeleton):
        def on_init(self, **kw):
            super().on_init(**kw)
            return ddl.DataProducerOnInitReturn(n, 3 * hw, (n, 3 * hw), (3 * hw,), dtype=np.uint8, convert=cv,
                                                shuffle_seed=5)

        def post_init(self, **kw):
            kw["my_ary"][...] = data

    @ddl.distributed_dataloader(n_instances=1, n_producers=1)
    def main(env, conn):
        loader = ddl.DistributedDataLoader(Img(), B, conn, 1)
        perm = np.random.default_rng(5).permutation(n) if False else None
        order = np.arange(n)
        np.random.default_rng(5).shuffle(order)
        sc = torch.tensor(cv.scale).repeat_interleave(hw)
        bi = torch.tensor(cv.bias).repeat_interleave(hw)
        for i, (x,) in enumerate(loader):
            ref = (torch.from_numpy(data[order[i * B:(i + 1) * B]]).float() * sc + bi).to(torch.bfloat16)
            assert x.dtype == torch.bfloat16 and torch.equal(x.cpu().view(torch.int16), ref.view(torch.int16))
            from ddl_b200.ops import checksum_reference
            got = int(loader.batch_checksum().item()) & (2**64 - 1)
            assert got == checksum_reference(ref.view(torch.int16).numpy())
            loader.mark(ddl.Marker.END_OF_BATCH)
        loader.mark(ddl.Marker.END_OF_EPOCH)

    main()


def _torchrun(nproc, cfg, tmp_path, timeout=240):
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cfg = dict(cfg, out=str(tmp_path))
    env = dict(os.environ, BDDL_TEST_CFG=json.dumps(cfg))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "workers", "run_loader.py")]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    if res.returncode != 0:
        import glob

        errs = "".join(f"--- {os.path.basename(f)}\n{open(f).read()[-1500:]}" for f in sorted(glob.glob(os.path.join(str(tmp_path), "error_rank*.txt")))[:2])
        raise AssertionError(errs or (res.stdout[-3000:] + res.stderr[-3000:]))


@pytest.mark.parametrize("declare", [False, True])
def test_process_transport_consumer_on_gpu(tmp_path, declare):
    """Reference layout on one GPU: the consumer stages the borrowed host window to the device
    (with a declared shuffle: from whichever of the producer's two windows is current)."""
    cfg = {"n_instances": 1, "n": 4000, "v": 8, "B": 256, "epochs": 4, "fraction": 0.5, "cpu": False,
           "declare_seed": declare}
    _torchrun(3, cfg, tmp_path)
    exp = expected_epochs(1, 2, 4000, 256, 4, 0.5)[0]
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "seen_0.npy")), exp)


@pytest.mark.parametrize("host_pull", ["0", "1"])
def test_process_transport_window_not_returned_while_gpu_reads_it(tmp_path, host_pull):
    """The consumer never synchronises per batch and its fetch streams lag behind the host: a
    window must not go back to its producer (who reshuffles it at once) before every copy / pull
    that reads it has finished on the device."""
    cfg = {"n_instances": 1, "n": 1024, "v": 16, "B": 128, "epochs": 6, "fraction": 0.0, "cpu": False,
           "defer": True}
    os.environ["DDL_B200_HOST_PULL"] = host_pull
    try:
        _torchrun(2, cfg, tmp_path)
    finally:
        os.environ.pop("DDL_B200_HOST_PULL", None)
    exp = expected_epochs(1, 1, 1024, 128, 6, 0.0)[0]
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "seen_0.npy")), exp)


@pytest.mark.parametrize("host_pull", ["0", "1"])
def test_process_transport_converts_on_the_gpu(tmp_path, host_pull):
    """Producer process + GPU consumer with a fused cast: the borrowed host window is staged
    (cudaMemcpyAsync) or pulled (zero-copy) and converted by the kernel."""
    cfg = {"n_instances": 1, "n": 200, "v": 16, "B": 50, "epochs": 4, "fraction": 0.0, "cpu": False,
           "convert": True, "small_ids": True}
    os.environ["DDL_B200_HOST_PULL"] = host_pull
    try:
        _torchrun(2, cfg, tmp_path)
    finally:
        os.environ.pop("DDL_B200_HOST_PULL", None)
    exp = expected_epochs(1, 1, 200, 50, 4, 0.0)[0] - 1_000_000  # producer 1 has rank_global 1
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "seen_0.npy")), exp)


def _need_gpus(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} CUDA devices")


ID_STRIDE = 100_000  # float32 ids of the id/payload worker stay exact (< 2^24) for up to 160 producers


def _restride(ids):
    """Oracle ids (rank_global * 1e6 + row) in the worker's id_stride encoding."""
    return ids // 1_000_000 * ID_STRIDE + ids % 1_000_000


def _check_order(tmp_path, n_inst, P, n, B, E, fraction, restride=False):
    exp = expected_epochs(n_inst, P, n, B, E, fraction)
    for i in range(n_inst):
        want = _restride(exp[i]) if restride else exp[i]
        assert np.array_equal(np.load(os.path.join(str(tmp_path), f"seen_{i}.npy")), want), f"instance {i}"


@pytest.mark.multigpu
@pytest.mark.parametrize("nproc", [2, 4, 8])
def test_multi_gpu_push_exchange(tmp_path, nproc):
    """Push form, one launch per step and GPU: every GPU stores the samples it owns into its peers'
    batch rings over NVLink and signals `ready` from inside the kernel; order equals the oracle
    (partner tables with real 3-cycles from 3 instances on, reference ddl/shuffle.py:50-79)."""
    _need_gpus(nproc)
    cfg = {"n_instances": nproc, "n_producers": 2, "n": 4096, "v": 32, "B": 256, "epochs": 6, "fraction": 0.5,
           "shuffle_mode": "device", "declare_seed": True, "cpu": False, "exchange": "push", "timeout_s": 60.0,
           "ring_depth": 4, "prefetch_depth": 2, "fetch_streams": 2, "id_stride": ID_STRIDE}
    _torchrun(nproc, cfg, tmp_path)
    _check_order(tmp_path, nproc, 2, 4096, 256, 6, 0.5, restride=True)


@pytest.mark.multigpu
@pytest.mark.parametrize("nproc,barrier", [(2, "nccl"), (2, "flags"), (2, "host"), (4, "flags"), (8, "flags"),
                                           (8, "nccl")])
def test_multi_gpu_global_shuffle_over_nvlink(tmp_path, nproc, barrier):
    """One process per GPU: slot tables exchanged and samples pulled through peer mappings."""
    _need_gpus(nproc)
    cfg = {"n_instances": nproc, "n_producers": 2, "n": 4096, "v": 32, "B": 256, "epochs": 6, "fraction": 0.5,
           "shuffle_mode": "device", "declare_seed": True, "cpu": False, "barrier": barrier, "timeout_s": 60.0,
           "id_stride": ID_STRIDE}
    _torchrun(nproc, cfg, tmp_path)
    _check_order(tmp_path, nproc, 2, 4096, 256, 6, 0.5, restride=True)


@pytest.mark.multigpu
@pytest.mark.parametrize("nproc,exchange", [(2, "pull"), (8, "pull"), (2, "push"), (8, "push")])
def test_multi_gpu_u8_to_bf16_with_integrity_mode(tmp_path, nproc, exchange):
    """uint8 rows of 4 KiB through the TMA/convert form across GPUs: ids and payloads must match,
    and (pull) every batch is re-assembled by the LDG form and fingerprinted (verify=True)."""
    _need_gpus(nproc)
    cfg = {"n_instances": nproc, "n_producers": 1, "n": 2048, "v": 4096, "B": 128, "epochs": 4, "fraction": 0.5,
           "shuffle_mode": "device", "cpu": False, "u8": True, "verify": exchange == "pull", "exchange": exchange,
           "ring_depth": 3, "prefetch_depth": 2, "timeout_s": 60.0}
    _torchrun(nproc, cfg, tmp_path)
    _check_order(tmp_path, nproc, 1, 2048, 128, 4, 0.5)


@pytest.mark.multigpu
@pytest.mark.parametrize("exchange", ["pull", "push"])
def test_two_gpus_variable_length_rows(tmp_path, exchange):
    """Variable-size int32 rows (up to 2400 bytes) exchanged between two GPUs, widened to int64 and padded
    while collating -- by the pull kernel and by the push kernel."""
    _need_gpus(2)
    cfg = {"n_instances": 2, "n_producers": 1, "n": 400, "v": 600, "B": 32, "epochs": 3, "fraction": 0.5,
           "shuffle_mode": "device", "cpu": False, "ragged": True, "exchange": exchange, "timeout_s": 60.0,
           "ring_depth": 3, "prefetch_depth": 2}
    _torchrun(2, cfg, tmp_path)
    _check_order(tmp_path, 2, 1, 400, 32, 3, 0.5)


@pytest.mark.multigpu
@pytest.mark.parametrize("nproc,multicast", [(2, "auto"), (2, "off"), (8, "auto")])
def test_replicated_window_delivered_to_every_consumer(tmp_path, nproc, multicast):
    """replicate=True: every consumer receives instance 0's window.  Each sample is converted once by
    the GPU that owns it and stored once through the NVLS multicast mapping of the batch rings
    (multimem.st inside the push kernel) -- or once per consumer with multicast='off'; both must
    deliver exactly what the unicast oracle prescribes for instance 0."""
    _need_gpus(nproc)
    cfg = {"n_instances": nproc, "n_producers": 2, "n": 4096, "v": 32, "B": 256, "epochs": 6, "fraction": 0.5,
           "shuffle_mode": "device", "declare_seed": True, "cpu": False, "exchange": "push", "timeout_s": 60.0,
           "ring_depth": 4, "prefetch_depth": 2, "fetch_streams": 2, "replicate": True, "multicast": multicast,
           "id_stride": ID_STRIDE}
    _torchrun(nproc, cfg, tmp_path)
    exp = _restride(expected_epochs(nproc, 2, 4096, 256, 6, 0.5)[0])
    for i in range(nproc):
        assert np.array_equal(np.load(os.path.join(str(tmp_path), f"seen_{i}.npy")), exp), f"consumer {i}"


@pytest.mark.multigpu
def test_replicated_delivery_multicast_or_unicast():
    """MulticastBuffer: multimem.st through an NVLS multicast object where the platform has it,
    peer copies otherwise; either way every GPU's copy must hold what the owner wrote."""
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(HERE, "workers", "run_multicast.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "MODE" in res.stdout
    print(res.stdout[res.stdout.index("CAPS"):][:200])


@pytest.mark.multigpu
def test_dedicated_storage_gpu(tmp_path):
    """One GPU only stores the shards, the other trains: every sample is pulled over NVLink."""
    cfg = {"n_instances": 1, "n_storage": 1, "n_producers": 2, "n": 4096, "v": 32, "B": 256, "epochs": 4,
           "fraction": 0.5, "shuffle_mode": "device", "declare_seed": True, "cpu": False}
    _torchrun(2, cfg, tmp_path)
    exp = expected_epochs(1, 2, 4096, 256, 4, 0.5)
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "seen_0.npy")), exp[0])
