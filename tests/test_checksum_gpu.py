"""Device batch fingerprint against its numpy statement (exact: integer sums mod 2^64)."""
import numpy as np
import pytest
import torch

from ddl_b200 import _native as N
from ddl_b200.ops import BatchChecksum, checksum_reference

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


@pytest.mark.parametrize("rows,payload,stride", [(1, 8, 8), (7, 40, 48), (256, 3072, 3072), (64, 150528, 150528),
                                                  (300, 1000, 1024), (2048, 65536, 65536)])
def test_checksum_matches_numpy(rows, payload, stride):
    g = np.random.default_rng(rows)
    host = g.integers(0, 256, size=(rows, stride), dtype=np.uint8)
    dev = torch.from_numpy(host).to(DEV)
    cs = BatchChecksum(DEV, rows, payload)
    for _ in range(2):  # the ticket re-arms itself: a second launch gives the same answer
        cs.run(dev, payload)
        assert cs.value() == checksum_reference(host, payload)
    host[rows // 2, payload - 1] ^= 1
    cs.run(torch.from_numpy(host).to(DEV), payload)
    assert cs.value() == checksum_reference(host, payload)


def test_checksum_rejects_small_scratch_and_misaligned_rows():
    dev = torch.zeros((64, 4096), dtype=torch.uint8, device=DEV)
    cs = BatchChecksum(DEV, 1, 8)
    with pytest.raises(N.NativeError):
        cs.run(dev)
    cs = BatchChecksum(DEV, 64, 4096)
    with pytest.raises(N.NativeError):
        cs.run(dev, payload_bytes=4092)
