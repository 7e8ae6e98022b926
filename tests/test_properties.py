"""Property tests: whatever the number of instances, exchange fraction and
iterations, the slot tables of the k-th producers stay a global bijection of all sample ids, and
an epoch delivers every id of its window exactly once minus the dropped tail."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from ddl_b200.ops import Pcg64, index_exchange, make_slot, perm_gen, rng_states_to_tensor, slot_iota


@given(n_inst=st.integers(1, 9), n=st.integers(2, 400), frac=st.floats(0.0, 1.0), iters=st.integers(1, 6),
       seed=st.integers(0, 2**20))
@settings(max_examples=60, deadline=None)
def test_tables_stay_a_global_bijection(n_inst, n, frac, iters, seed):
    ne = int(n * frac)
    tables = [torch.empty(n, dtype=torch.int64) for _ in range(n_inst)]
    for i, t in enumerate(tables):
        slot_iota(t, i, 0)
    everything = sorted(make_slot(i, r) for i in range(n_inst) for r in range(n))
    partner_rngs = [Pcg64(1) for _ in range(n_inst)]            # same seed on every member
    states = rng_states_to_tensor([Pcg64(seed + i) for i in range(n_inst)], torch.device("cpu"))
    for _ in range(iters):
        partners = [partner_rngs[i].partner(n_inst, i)[:2] for i in range(n_inst)]
        if n_inst > 1:  # partners agree: who I send to receives from me
            for i, (send_to, recv_from) in enumerate(partners):
                assert partners[send_to][1] == i and partners[recv_from][0] == i
        new = []
        for i, (send_to, recv_from) in enumerate(partners):
            dst = torch.empty(n, dtype=torch.int64)
            active = n_inst > 1 and ne > 0
            index_exchange(dst, tables[i], tables[recv_from] if active else None,
                           tables[send_to] if active else None, ne if active else 0)
            new.append(dst)
        tables = new
        for i in range(n_inst):
            perm_gen(tables[i].unsqueeze(0), states[i].unsqueeze(0))
        assert sorted(torch.cat(tables).tolist()) == everything


@given(n=st.integers(1, 500), B=st.integers(1, 64), seed=st.integers(0, 2**32 - 1))
@settings(max_examples=40, deadline=None)
def test_epoch_delivers_each_row_once_minus_tail(n, B, seed):
    t = torch.arange(n, dtype=torch.int64).unsqueeze(0)
    perm_gen(t, rng_states_to_tensor([Pcg64(seed)], torch.device("cpu")))
    ref = np.arange(n)
    np.random.default_rng(seed).shuffle(ref)
    assert np.array_equal(t[0].numpy(), ref)
    delivered = t[0, : (n // B) * B].tolist()
    assert len(set(delivered)) == len(delivered) and set(delivered) <= set(range(n))
    assert n - len(delivered) == n % B
