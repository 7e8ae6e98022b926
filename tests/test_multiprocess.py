"""Multi-process runs of the whole loader on the CPU (gloo rendezvous, world_size > 1), the
counterpart of the reference's only test (reference tests/test_ddl.py:8-28: mpirun -np 4, exit
code 0) -- but with assertions on *what* arrives: every consumer must see exactly the sample
order the reference semantics prescribe (tests/oracle.py)."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from oracle import expected_epochs

HERE = os.path.dirname(os.path.abspath(__file__))
WORKER = os.path.join(HERE, "workers", "run_loader.py")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_world(nproc, cfg, tmp_path, timeout=150, extra_env=None):
    cfg = dict(cfg, out=str(tmp_path))
    env = dict(os.environ, BDDL_TEST_CFG=json.dumps(cfg), OMP_NUM_THREADS="1")
    env.update(extra_env or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    if res.returncode != 0:
        import glob

        errs = "".join(f"--- {os.path.basename(f)}\n{open(f).read()[-1500:]}" for f in sorted(glob.glob(os.path.join(str(tmp_path), "error_rank*.txt")))[:2])
        raise AssertionError(errs or (res.stdout[-3000:] + res.stderr[-3000:]))
    return res


def check(tmp_path, n_instances, P, cfg, designed=True):
    exp = expected_epochs(n_instances, P, cfg["n"], cfg["B"], cfg["epochs"], cfg.get("fraction", 0.0), designed)
    for i in range(n_instances):
        got = np.load(os.path.join(str(tmp_path), f"seen_{i}.npy"))
        assert got.shape == exp[i].shape
        assert np.array_equal(got, exp[i]), f"instance {i}: order differs from the reference semantics"


@pytest.mark.timeout(200)
def test_process_transport_like_reference_test(tmp_path):
    """1 consumer + 3 producer processes, the reference's own layout (tests/run_ddl.py:301-316)."""
    cfg = {"n_instances": 1, "n": 5000, "v": 9, "B": 512, "epochs": 5, "fraction": 0.5}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 1, 3, cfg)


@pytest.mark.timeout(200)
def test_process_transport_two_instances_global_shuffle(tmp_path):
    """2 instances x (1 consumer + 1 producer): the exchange the reference designed but never runs."""
    cfg = {"n_instances": 2, "n": 2000, "v": 4, "B": 256, "epochs": 4, "fraction": 0.5}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 2, 1, cfg)


@pytest.mark.timeout(200)
def test_process_transport_as_wired_has_no_exchange(tmp_path):
    cfg = {"n_instances": 2, "n": 2000, "v": 4, "B": 256, "epochs": 3, "fraction": 0.5, "compat": "as_wired"}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 2, 1, cfg, designed=False)


@pytest.mark.timeout(200)
@pytest.mark.parametrize("mode,declare", [("device", True), ("trace", False), ("host", False)])
def test_device_transport_three_instances(tmp_path, mode, declare):
    """One process per instance, in-process producers, slot-table exchange over shared memory."""
    cfg = {"n_instances": 3, "n_producers": 2, "n": 1500, "v": 5, "B": 128, "epochs": 6, "fraction": 0.5,
           "shuffle_mode": mode, "declare_seed": declare}
    run_world(3, cfg, tmp_path)
    check(tmp_path, 3, 2, cfg)


@pytest.mark.timeout(200)
def test_device_transport_four_instances_single_producer(tmp_path):
    cfg = {"n_instances": 4, "n_producers": 1, "n": 1000, "v": 3, "B": 100, "epochs": 4, "fraction": 0.3,
           "shuffle_mode": "device", "declare_seed": True}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 4, 1, cfg)


@pytest.mark.timeout(120)
def test_dead_producer_is_detected_not_hung(tmp_path):
    """Failure detection: in the reference a dead producer hangs the consumer in
    Recv forever; here the bounded wait raises LoaderTimeoutError."""
    cfg = dict({"n_instances": 1, "n": 1000, "v": 4, "B": 100, "epochs": 50, "fraction": 0.0,
                "die_after": 2, "timeout_s": 3.0}, out=str(tmp_path))
    env = dict(os.environ, BDDL_TEST_CFG=json.dumps(cfg), OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), WORKER]
    subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=100)
    msg = open(os.path.join(str(tmp_path), "timeout.txt")).read()
    assert "did not answer" in msg


@pytest.mark.timeout(200)
def test_dedicated_storage_ranks_one_to_three(tmp_path):
    """producers:consumers 1:3 -- one storage rank holds the shards of three consumers, which
    still see exactly the reference order (inexpressible in the reference, ddl/connection.py:29-30)."""
    cfg = {"n_instances": 3, "n_storage": 1, "n_producers": 2, "n": 1200, "v": 5, "B": 128, "epochs": 5,
           "fraction": 0.5, "shuffle_mode": "device", "declare_seed": True}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 3, 2, cfg)


@pytest.mark.timeout(200)
def test_process_transport_declared_shuffle_double_windows(tmp_path):
    """Producer processes with a declared shuffle: two windows each, native out-of-place shuffle
    (+ out-of-place exchange) one iteration ahead -- same order as the reference semantics."""
    cfg = {"n_instances": 2, "n": 2000, "v": 4, "B": 256, "epochs": 5, "fraction": 0.5, "declare_seed": True}
    run_world(4, cfg, tmp_path)
    check(tmp_path, 2, 1, cfg)
    cfg = {"n_instances": 1, "n": 3000, "v": 9, "B": 512, "epochs": 7, "fraction": 0.5, "declare_seed": True}
    run_world(3, cfg, tmp_path)
    check(tmp_path, 1, 2, cfg)


@pytest.mark.timeout(200)
@pytest.mark.parametrize("declare", [False, True])
def test_process_transport_resume_replays_to_the_checkpoint(tmp_path, declare):
    """Checkpoint of the process transport: a fresh world (1 consumer + 2 producer processes) loaded with
    a position continues with exactly the batches an uninterrupted run delivers from there on."""
    n, B, E, e0, b0 = 1200, 100, 5, 2, 7
    cfg = {"n_instances": 1, "n": n, "v": 4, "B": B, "epochs": E, "fraction": 0.0, "declare_seed": declare,
           "resume": [e0, b0]}
    run_world(3, cfg, tmp_path)
    exp = expected_epochs(1, 2, n, B, E, 0.0)[0]
    tail = np.concatenate([exp[e0][b0 * B:]] + [exp[e] for e in range(e0 + 1, E)])
    assert np.array_equal(np.load(os.path.join(str(tmp_path), "seen_0.npy")), tail)


@pytest.mark.timeout(200)
@pytest.mark.parametrize("n_storage", [0, 1])
def test_variable_length_windows_across_instances_and_storage_ranks(tmp_path, n_storage):
    """Variable-size samples with the global exchange between two instances, with the shards on the
    consumers' own ranks or on a dedicated storage rank (which then reports the row lengths)."""
    cfg = {"n_instances": 2, "n_storage": n_storage, "n_producers": 1, "n": 200, "v": 24, "B": 16, "epochs": 3,
           "fraction": 0.5, "shuffle_mode": "device", "ragged": True}
    run_world(2 + n_storage, cfg, tmp_path)
    check(tmp_path, 2, 1, cfg)


@pytest.mark.timeout(300)
def test_device_transport_eight_instances_two_producers(tmp_path):
    """8 instances x 2 producers on the CPU engine: partner tables with several cycles (n=8), two
    windows rotating, the order of every consumer equal to the reference semantics.  Ids use a smaller stride so
    that 24 producers' float32 ids stay exact."""
    cfg = {"n_instances": 8, "n_producers": 2, "n": 640, "v": 6, "B": 64, "epochs": 6, "fraction": 0.5,
           "shuffle_mode": "device", "declare_seed": True, "id_stride": 100_000}
    run_world(8, cfg, tmp_path, timeout=280)
    exp = expected_epochs(8, 2, 640, 64, 6, 0.5)
    for i in range(8):
        want = exp[i] // 1_000_000 * 100_000 + exp[i] % 1_000_000
        assert np.array_equal(np.load(os.path.join(str(tmp_path), f"seen_{i}.npy")), want), f"instance {i}"
