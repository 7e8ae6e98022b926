"""The counterpart of the reference's only test (reference tests/test_ddl.py:8-28): run the demo
program under the launcher with 4 ranks -- 1 consumer + 3 producer processes -- and require exit
code 0 within the time limit; plus the one-process form."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "examples", "run_ddl_b200.py")


@pytest.mark.timeout(100)
def test_program_with_four_ranks():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=4", "--master-addr",
           "127.0.0.1", "--master-port", str(port), SCRIPT]
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=95)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "Training finished" in res.stdout and "24 batches" in res.stdout


@pytest.mark.timeout(100)
def test_program_single_process():
    env = dict(os.environ, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    res = subprocess.run([sys.executable, SCRIPT], env=env, capture_output=True, text=True, timeout=95)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "Training finished" in res.stdout


@pytest.mark.timeout(150)
def test_training_and_validation_example_single_process():
    """Two windows per instance, one of them replicated (examples/train_with_validation.py), on the CPU path."""
    env = dict(os.environ, OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "train_with_validation.py")], env=env,
                         capture_output=True, text=True, timeout=140)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "epoch 5 (validate): 16 batches" in res.stdout
