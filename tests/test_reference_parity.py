"""Same seed -> same per-epoch permutation and batch order as the reference.

The UNMODIFIED reference (baseline/_ref/ddl) runs over the baseline-only MPI stand-in and dumps
the sample ids every consumer receives; they must equal tests/oracle.py, which is what the
ddl_b200 loaders are held to in test_multiprocess.py / test_loader_gpu.py.  As wired the
reference never exchanges rows between instances; with the harness-side wiring fix its own
SendRecvReplaceGlobalShuffler runs and must match the designed semantics."""
import json
import os
import sys

import numpy as np
import pytest

from oracle import expected_epochs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
REF = os.path.join(ROOT, "baseline", "_ref", "ddl")
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference not installed under baseline/_ref")


def _run(tmp_path, n_inst, P, cfg, wiring_fix):
    import mpirun

    cfg = dict(cfg, mode="order", fill="ids", out=str(tmp_path), wiring_fix=wiring_fix, cpu=True)
    env = {"BDDL_REF_CFG": json.dumps(cfg), "SLURM_JOBID": "1", "SLURM_GPUS_PER_NODE": str(n_inst),
           "SLURM_NNODES": "1", "OMP_NUM_THREADS": "1"}
    rc = mpirun.launch(n_inst * (P + 1), [os.path.join(ROOT, "baseline", "ref_worker.py")], extra_env=env,
                       timeout=240)
    assert rc == 0
    return [np.load(os.path.join(str(tmp_path), f"ref_seen_{i}.npy")) for i in range(n_inst)]


@pytest.mark.timeout(300)
def test_reference_single_instance_three_producers(tmp_path):
    cfg = {"n": 3000, "v": 6, "B": 256, "epochs": 5, "fraction": 0.5}
    got = _run(tmp_path, 1, 3, cfg, wiring_fix=False)
    exp = expected_epochs(1, 3, 3000, 256, 5, 0.5)
    assert np.array_equal(got[0], exp[0])


@pytest.mark.timeout(300)
def test_reference_as_wired_never_exchanges(tmp_path):
    cfg = {"n": 2000, "v": 4, "B": 256, "epochs": 4, "fraction": 0.5}
    got = _run(tmp_path, 3, 1, cfg, wiring_fix=False)
    exp = expected_epochs(3, 1, 2000, 256, 4, 0.5, designed=False)
    for i in range(3):
        assert np.array_equal(got[i], exp[i])
        assert (got[i] // 1_000_000 == i * 2 + 1).all()  # only its own producer's rows


@pytest.mark.timeout(300)
def test_reference_with_wiring_fix_matches_designed_semantics(tmp_path):
    cfg = {"n": 2000, "v": 4, "B": 256, "epochs": 4, "fraction": 0.5}
    got = _run(tmp_path, 3, 1, cfg, wiring_fix=True)
    exp = expected_epochs(3, 1, 2000, 256, 4, 0.5, designed=True)
    for i in range(3):
        assert np.array_equal(got[i], exp[i])
