"""Golden files: the per-epoch sample order produced by the UNMODIFIED reference (recorded with
baseline/ref_worker.py over the MPI stand-in, see tests/golden/index.json).  They pin the oracle
(tests/oracle.py) -- and through it every ddl_b200 loader test -- to the reference's behaviour even
on machines where baseline/_ref is not installed."""
import json
import os

import numpy as np
import pytest

from oracle import expected_epochs

HERE = os.path.dirname(os.path.abspath(__file__))
INDEX = json.load(open(os.path.join(HERE, "golden", "index.json")))


@pytest.mark.parametrize("name", sorted(INDEX))
def test_oracle_equals_recorded_reference_run(name):
    m = INDEX[name]
    gold = np.load(os.path.join(HERE, "golden", name + ".npy"))
    exp = expected_epochs(m["n_instances"], m["producers"], m["n"], m["B"], m["epochs"], m["fraction"],
                          designed=m["wiring_fix"])
    assert gold.shape == (m["n_instances"],) + exp[0].shape
    for i in range(m["n_instances"]):
        assert np.array_equal(gold[i], exp[i])
