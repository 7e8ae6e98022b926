"""GPU numerics of the native kernels against plain PyTorch / host statements of the same op.

Ranks are emulated as several shard buffers on one device (one kernel over all ranks' data),
never as co-resident spinning kernels.  The oracles are
``gather_reference`` (fp32 PyTorch) and the host Fisher-Yates, itself checked against numpy in
test_rng.py.
"""
import numpy as np
import pytest
import torch

from ddl_b200 import _native as N
from ddl_b200.ops import (Convert, GatherPlan, Pcg64, SourceTable, gather_reference, index_exchange, make_slot,
                          perm_gen, rng_states_from_tensor, rng_states_to_tensor, round_up, slot_iota)

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _shards(n_sources, rows, sample_bytes, seed=0, dtype=np.uint8):
    g = np.random.default_rng(seed)
    row_bytes = round_up(sample_bytes, 16)
    out = []
    for _ in range(n_sources):
        a = np.zeros((rows, row_bytes), dtype=np.uint8)
        if dtype == np.float32:
            vals = g.standard_normal((rows, sample_bytes // 4)).astype(np.float32) * 3
            a[:, :sample_bytes] = vals.view(np.uint8)
        elif dtype == np.int32:
            vals = g.integers(-2**31, 2**31 - 1, size=(rows, sample_bytes // 4), dtype=np.int64).astype(np.int32)
            a[:, :sample_bytes] = vals.view(np.uint8)
        else:
            a[:, :sample_bytes] = g.integers(0, 256, size=(rows, sample_bytes), dtype=np.uint8)
        out.append(a)
    return out, row_bytes


def _slots(n, n_sources, rows, seed=1):
    g = np.random.default_rng(seed)
    return np.array([make_slot(int(g.integers(n_sources)), int(g.integers(rows))) for _ in range(n)], dtype=np.int64)


@pytest.mark.parametrize("n", [1, 2, 33, 1000, 4097, 100520])
def test_perm_gen_matches_host(n):
    seeds = [0, 3, 12345]
    tables = torch.arange(n, dtype=torch.int64).repeat(len(seeds), 1).contiguous().to(DEV)
    host_rngs = [Pcg64(s) for s in seeds]
    host_tables = [np.arange(n, dtype=np.int64) for _ in seeds]
    states = rng_states_to_tensor(host_rngs, torch.device(DEV))
    for _ in range(3):  # composed in place, stream continues
        perm_gen(tables, states)
        for r, t in zip(host_rngs, host_tables):
            r.shuffle(t)
        torch.cuda.synchronize()
        got = tables.cpu().numpy()
        for i, t in enumerate(host_tables):
            assert np.array_equal(got[i], t), f"table {i} differs at n={n}"
        dev_rngs = rng_states_from_tensor(states)
        for a, b in zip(dev_rngs, host_rngs):
            da, db = a.state_dict(), b.state_dict()
            assert da["state"] == db["state"] and da["has_uint32"] == db["has_uint32"]
            assert da["draws32"] == db["draws32"]
            if da["has_uint32"]:
                assert da["uinteger"] == db["uinteger"]


def test_perm_gen_strided_tables():
    n, stride = 500, 512
    buf = torch.full((2, stride), -1, dtype=torch.int64, device=DEV)
    buf[:, :n] = torch.arange(n, device=DEV)
    rngs = [Pcg64(5), Pcg64(6)]
    states = rng_states_to_tensor(rngs, torch.device(DEV))
    perm_gen(buf, states, n=n)
    for i, r in enumerate(rngs):
        t = np.arange(n, dtype=np.int64)
        r.shuffle(t)
        assert np.array_equal(buf[i, :n].cpu().numpy(), t)
    assert bool((buf[:, n:] == -1).all())


def test_slot_iota_and_index_exchange():
    n, ne = 1001, 500
    own = torch.empty(n, dtype=torch.int64, device=DEV)
    a = torch.empty_like(own)
    b = torch.empty_like(own)
    slot_iota(own, 0, 10)
    slot_iota(a, 1)
    slot_iota(b, 2, 7)
    assert own[0].item() == make_slot(0, 10) and b[5].item() == make_slot(2, 12)
    dst = torch.empty_like(own)
    index_exchange(dst, own, a, b, ne)
    ref = own.cpu().clone()
    ref[: ne // 2] = a.cpu()[: ne // 2]
    ref[ne // 2: ne] = b.cpu()[ne // 2: ne]
    assert torch.equal(dst.cpu(), ref)
    index_exchange(dst, own, None, None, 0)
    assert torch.equal(dst, own)


CASES = [
    # (mode, sample_bytes, dtype, channels)
    ("copy", 36, np.uint8, 0),
    ("copy", 4096, np.uint8, 0),
    ("copy", 12288, np.uint8, 0),
    ("copy", 150528, np.uint8, 0),
    ("copy", 10000, np.uint8, 0),
    ("u8_bf16", 150528, np.uint8, 3),
    ("u8_bf16", 3 * 32 * 32, np.uint8, 3),
    ("u8_bf16", 3 * 30 * 30, np.uint8, 3),
    ("u8_bf16", 1000, np.uint8, 0),
    ("u8_f32", 3 * 64 * 64, np.uint8, 3),
    ("f32_bf16", 12288, np.float32, 3),
    ("f32_bf16", 36, np.float32, 0),
    ("f32_bf16", 4 * 4 * 16 * 16 * 16, np.float32, 4),
    ("i32_i64", 4096, np.int32, 0),
]


def _convert(mode, sample_bytes, channels):
    if mode == "copy":
        return Convert.copy()
    if mode == "i32_i64":
        return Convert.widen_tokens()
    mean = [0.485, 0.456, 0.406, 0.5][: max(channels, 1)]
    std = [0.229, 0.224, 0.225, 0.25][: max(channels, 1)]
    if mode in ("u8_bf16", "u8_f32"):
        stride = sample_bytes // channels if channels else 0
        return Convert.normalize_u8(mean, std, stride, out="bf16" if mode == "u8_bf16" else "f32")
    stride = (sample_bytes // 4) // channels if channels else 0
    return Convert.cast_f32_bf16([1.0 / s for s in std], [-m / s for m, s in zip(mean, std)], stride)


@pytest.mark.parametrize("mode,sample_bytes,dtype,channels", CASES)
@pytest.mark.parametrize("variant,tune", [(N.VARIANT_LDG, 0), (N.VARIANT_TMA, 0), (N.VARIANT_TMA, 1),
                                          (N.VARIANT_TMA, 2), (N.VARIANT_TMA, 3), (N.VARIANT_AUTO, 0)])
def test_gather_matches_reference(mode, sample_bytes, dtype, channels, variant, tune):
    n_sources, rows, n = 3, 64, 97
    shards, row_bytes = _shards(n_sources, rows, sample_bytes, dtype=dtype)
    dev_shards = [torch.from_numpy(s).to(DEV) for s in shards]
    cv = _convert(mode, sample_bytes, channels)
    table = SourceTable([t.data_ptr() for t in dev_shards], torch.device(DEV))
    plan = GatherPlan(table, row_bytes, sample_bytes, cv, variant=variant, tune=tune)
    slots = _slots(n, n_sources, rows)
    d_slots = torch.from_numpy(slots).to(DEV)
    raw = plan.alloc_batch(n)
    raw.fill_(0xAB)
    plan.run(d_slots, raw)
    torch.cuda.synchronize()
    got = plan.view(raw).cpu()
    ref = gather_reference(shards, slots, sample_bytes, cv)
    assert got.shape == ref.shape
    if got.dtype == torch.bfloat16:
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16))
    else:
        assert torch.equal(got, ref)


@pytest.mark.parametrize("variant", [N.VARIANT_LDG, N.VARIANT_TMA])
@pytest.mark.parametrize("mode", ["copy", "i32_i64", "f32_bf16"])
def test_gather_variable_length_pads(variant, mode):
    g = np.random.default_rng(3)
    n_sources, rows, max_bytes, n = 2, 40, 20000, 50
    es = 1 if mode == "copy" else 4
    cv = {"copy": Convert(N.CVT_COPY, pad_value=7), "i32_i64": Convert(N.CVT_I32_I64, pad_value=-100),
          "f32_bf16": Convert(N.CVT_F32_BF16, (2.0,), (1.0,), 0, pad_value=-1.0)}[mode]
    shards, offs, lens = [], [], []
    for _ in range(n_sources):
        ln = (g.integers(1, max_bytes // es + 1, size=rows) * es).astype(np.int32)
        ln[0] = max_bytes
        ln[1] = es
        starts = np.zeros(rows + 1, dtype=np.int64)
        starts[1:] = np.cumsum([round_up(int(x), 16) for x in ln])
        flat = np.zeros(int(starts[-1]) + 16, dtype=np.uint8)
        for r in range(rows):
            if mode == "f32_bf16":
                v = g.standard_normal(ln[r] // 4).astype(np.float32).view(np.uint8)
            else:
                v = g.integers(0, 256, size=ln[r], dtype=np.uint8)
            flat[starts[r]: starts[r] + ln[r]] = v
        shards.append(flat)
        offs.append(starts[:-1].copy())
        lens.append(ln)
    d_shards = [torch.from_numpy(s).to(DEV) for s in shards]
    d_offs = [torch.from_numpy(o).to(DEV) for o in offs]
    d_lens = [torch.from_numpy(l).to(DEV) for l in lens]
    table = SourceTable([t.data_ptr() for t in d_shards], torch.device(DEV), d_offs, d_lens)
    plan = GatherPlan(table, 0, max_bytes, cv, variant=variant)
    slots = _slots(n, n_sources, rows, seed=9)
    slots[0], slots[1] = make_slot(0, 0), make_slot(1, 1)
    raw = plan.alloc_batch(n)
    raw.fill_(0xEE)
    plan.run(torch.from_numpy(slots).to(DEV), raw)
    torch.cuda.synchronize()
    got = plan.view(raw).cpu()
    ref = gather_reference(shards, slots, max_bytes, cv, lengths=lens, offsets=offs)
    if got.dtype == torch.bfloat16:
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16))
    else:
        assert torch.equal(got, ref)


def test_gather_sm_budget_and_errors():
    shards, row_bytes = _shards(1, 32, 8192)
    d = torch.from_numpy(shards[0]).to(DEV)
    table = SourceTable([d.data_ptr()], torch.device(DEV))
    plan = GatherPlan(table, row_bytes, 8192, Convert.copy(), max_ctas=4)
    slots = _slots(64, 1, 32)
    raw = plan.alloc_batch(64)
    plan.run(torch.from_numpy(slots).to(DEV), raw)
    torch.cuda.synchronize()
    assert torch.equal(plan.view(raw).cpu(), gather_reference(shards, slots, 8192, Convert.copy()))
    bad = GatherPlan(table, row_bytes, 8192, Convert.copy(), dst_row_bytes=8200)
    with pytest.raises(N.NativeError):
        bad.run(torch.from_numpy(slots).to(DEV), raw)


def test_native_library_is_loaded_in_tree():
    import os

    assert os.path.dirname(N.lib_path()).endswith(os.path.join("ddl_b200", "lib"))
    before = N.launch_count()
    t = torch.empty(8, dtype=torch.int64, device=DEV)
    slot_iota(t, 1)
    assert N.launch_count() == before + 1
