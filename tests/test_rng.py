"""N1/N3: the native RNG must reproduce numpy.random.default_rng bit for bit
(reference ddl/shuffle.py:28-30,53; tests/run_ddl.py:122,167) and the partner tables of
SURVEY.md section 3.5 (reference ddl/shuffle.py:32-79)."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from ddl_b200.ops import Pcg64

SEEDS = [0, 1, 2, 3, 7, 12345, 2**32 + 5, 2**63 + 11]


@pytest.mark.parametrize("seed", SEEDS)
def test_raw_stream_matches_numpy(seed):
    ours, ref = Pcg64(seed), np.random.default_rng(seed)
    raw = ref.bit_generator.random_raw(64)
    assert [ours.next64() for _ in range(64)] == [int(v) for v in raw]


@pytest.mark.parametrize("seed", SEEDS)
@pytest.mark.parametrize("n", [1, 2, 5, 24, 1000, 100520])
def test_shuffle_composes_like_numpy(seed, n):
    ours, ref = Pcg64(seed), np.random.default_rng(seed)
    a = np.arange(n, dtype=np.int64)
    b = a.copy()
    for _ in range(3):  # the stream continues across epochs and shuffles compose in place
        ref.shuffle(a)
        ours.shuffle(b)
        assert np.array_equal(a, b)
    assert ours.numpy_state()["state"] == ref.bit_generator.state["state"]
    assert ours.numpy_state()["has_uint32"] == ref.bit_generator.state["has_uint32"]


def test_row_shuffle_matches_numpy_2d():
    ours, ref = Pcg64(9), np.random.default_rng(9)
    a = np.random.default_rng(0).random((1000, 9)).astype(np.float32)
    b = a.copy()
    ref.shuffle(a)
    ours.shuffle(b)
    assert np.array_equal(a, b)


def test_int32_shuffle_matches_numpy():
    ours, ref = Pcg64(4), np.random.default_rng(4)
    a = np.arange(77, dtype=np.int32)
    b = a.copy()
    ref.shuffle(a)
    ours.shuffle(b)
    assert np.array_equal(a, b)


@given(seed=st.integers(0, 2**64 - 1), n=st.integers(0, 300), k=st.integers(0, 1000))
@settings(max_examples=60, deadline=None)
def test_advance_equals_stepping(seed, n, k):
    a, b = Pcg64(seed), Pcg64(seed)
    for _ in range(k % 50):
        a.next64()
    b.advance(k % 50)
    assert a.next64() == b.next64()
    x = np.arange(n, dtype=np.int64)
    y = x.copy()
    a.shuffle(x)
    np.random.default_rng(0)  # no-op: numpy is only the oracle above
    b.shuffle(y)
    assert np.array_equal(x, y)


def test_state_roundtrip_and_clone():
    r = Pcg64(31)
    r.next32()  # leaves a buffered half behind
    d = r.state_dict()
    c = r.clone()
    q = Pcg64(0)
    q.load_state_dict(d)
    w = Pcg64(0)
    w.from_words(r.to_words())
    vals = [r.next32() for _ in range(9)]
    assert vals == [c.next32() for _ in range(9)] == [q.next32() for _ in range(9)] == [w.next32() for _ in range(9)]


# (send_to, recv_from) per nth-pusher rank, seed 1, three consecutive iterations.
PARTNERS = {
    2: [[(1, 1), (0, 0)]] * 3,
    3: [[(2, 1), (0, 2), (1, 0)], [(1, 2), (2, 0), (0, 1)], [(2, 1), (0, 2), (1, 0)]],
    4: [[(1, 2), (3, 0), (0, 3), (2, 1)], [(2, 3), (3, 2), (1, 0), (0, 1)], [(1, 3), (2, 0), (3, 1), (0, 2)]],
    8: [
        [(1, 7), (2, 0), (5, 1), (4, 6), (7, 3), (6, 2), (3, 5), (0, 4)],
        [(6, 3), (7, 2), (1, 5), (0, 6), (5, 7), (2, 4), (3, 0), (4, 1)],
        [(5, 1), (0, 2), (1, 5), (4, 6), (7, 3), (2, 0), (3, 7), (6, 4)],
    ],
}


@pytest.mark.parametrize("n", sorted(PARTNERS))
def test_partner_tables(n):
    for me in range(n):
        rng = Pcg64(1)
        for it in range(3):
            s, r, perm, _ = rng.partner(n, me)
            assert (s, r) == PARTNERS[n][it][me]
            if n >= 3:  # accepted permutations are derangements without 2-cycles
                assert all(perm[i] != i and perm[perm[i]] != i for i in range(n))


def _reference_partner(rng, n, me):
    """The reference loop restated with numpy (ddl/shuffle.py:32-79)."""
    row, col = np.arange(n, dtype=np.int32), np.arange(n, dtype=np.int32)
    if n == 1:
        return 0, 0
    if n == 2:
        return (1, 1) if me == 0 else (0, 0)
    while True:
        rng.shuffle(col)
        bad = False
        for r in range(n):
            if col[row == r][0] == row[col == r][0] or (row == col).any():
                bad = True
                break
        if not bad:
            return int(col[row == me][0]), int(row[col == me][0])


@pytest.mark.parametrize("n", [3, 5, 6, 7, 16])
@pytest.mark.parametrize("seed", [0, 1, 2, 5])
def test_partner_matches_numpy_restatement(n, seed):
    ours, ref = Pcg64(seed), np.random.default_rng(seed)
    for _ in range(4):
        assert ours.partner(n, n - 1)[:2] == _reference_partner(ref, n, n - 1)
