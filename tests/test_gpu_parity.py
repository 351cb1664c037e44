"""GPU: CUDA decode through the C ABI == oracle decode on the same container bytes, bit for bit."""
import numpy as np
import pytest

import flsgpu
import oracle_lib
from fls_cases import column_spec, standard_cases, width_sweep_cases, INT_TYPES

pytestmark = pytest.mark.gpu

_NAME_OF = {np.dtype(v): k for k, v in (("I8", np.int8), ("I16", np.int16), ("I32", np.int32), ("I64", np.int64),
                                        ("U8", np.uint8), ("U16", np.uint16), ("U32", np.uint32), ("U64", np.uint64),
                                        ("F32", np.float32), ("F64", np.float64))}


def gpu_decode(t, col, out_type, np_dtype, rg_begin=0, n_rg=None):
    import torch
    n_rg = t.n_rowgroups - rg_begin if n_rg is None else n_rg
    nbytes = t.out_elems(n_rg) * np.dtype(np_dtype).itemsize
    buf = torch.full((nbytes,), 0xA5, dtype=torch.uint8, device="cuda")
    t.decode_column(col, buf.data_ptr(), nbytes, rg_begin, n_rg, out_type)
    flsgpu.sync()
    torch.cuda.synchronize()
    return buf.cpu().numpy().view(np_dtype)


def check_table(cases, vec_per_rg, widen=False):
    cols = [column_spec(c) for c in cases]
    img = flsgpu.write_table(cols, [c[3] for c in cases], vec_per_rowgroup=vec_per_rg)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for i, c in enumerate(cases):
        dt = np.dtype(t.native_dtype(i))
        targets = [(0, dt)]
        if widen:
            if dt.kind == "f":
                targets += [(flsgpu.OUT["F64"], np.dtype(np.float64))] if dt.itemsize == 4 else []
            else:
                for w in (2, 4, 8):
                    if w > dt.itemsize:
                        wd = np.dtype(f"{dt.kind}{w}")
                        targets.append((flsgpu.OUT[_NAME_OF[wd]], wd))
        for out_type, odt in targets:
            exp = o.decode_column(i, 0, t.n_rowgroups, out_type, odt, t.vec_per_rowgroup)
            got = gpu_decode(t, i, out_type, odt)
            for a, b in t.valid_mask_slices():
                assert got[a:b].tobytes() == exp[a:b].tobytes(), (c[0], out_type)
            # whole vectors are written (padding included): the oracle pads the same way
            n_vec_total = sum(t.rowgroup(r).n_vec for r in range(t.n_rowgroups))
            assert n_vec_total > 0
        # the decoded column also equals the original input
        got = gpu_decode(t, i, 0, dt)
        flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
        assert flat.tobytes() == np.ascontiguousarray(c[3]).tobytes(), c[0]
    t.close()


def test_standard_cases_native():
    check_table(standard_cases(), vec_per_rg=2)


def test_standard_cases_widening():
    check_table(standard_cases(n=2048 + 99, seed=3), vec_per_rg=64, widen=True)


@pytest.mark.parametrize("tname", list(INT_TYPES))
def test_width_sweep(tname):
    check_table(width_sweep_cases(tname), vec_per_rg=64)


def test_many_rowgroups_and_subranges():
    n = 70 * 1024 + 13
    r = np.random.default_rng(11)
    a = np.cumsum(r.integers(0, 50, n)).astype(np.int64)
    b = r.integers(-5000, 5000, n).astype(np.int32)
    img = flsgpu.write_table([("a", "INT64", None), ("b", "INT32", None)], [a, b], vec_per_rowgroup=4)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    t.stage(3, 9)
    for col, dt in ((0, np.int64), (1, np.int32)):
        exp = o.decode_column(col, 3, 9, 0, dt, 4)
        got = gpu_decode(t, col, 0, dt, 3, 9)
        assert got.tobytes() == exp.tobytes()
    with pytest.raises(flsgpu.FlsError):
        gpu_decode(t, 0, 0, np.int64, 0, 2)  # not staged
    t.close()


def test_scan_to_host_matches_device_decode():
    import torch
    n = 300 * 1024 + 5
    a = (np.arange(n, dtype=np.int64) * 10)
    d = np.arange(n) / 1000.0
    img = flsgpu.write_table([("a", "INT64", None), ("d", "DOUBLE", None)], [a, d], vec_per_rowgroup=16)
    t = flsgpu.Table(img)
    for col, src in ((0, a), (1, d)):
        out = torch.empty(t.out_elems(t.n_rowgroups) * 8, dtype=torch.uint8).pin_memory()
        t.scan_column_to_host(col, out.data_ptr(), out.numel())
        got = out.numpy().view(src.dtype)
        flat = np.concatenate([got[x:y] for x, y in t.valid_mask_slices()])
        assert flat.tobytes() == src.tobytes()
        t.evict()
    t.close()


def test_scan_columns_to_host_projection():
    """Projection form: a subset of columns, widened outputs, merged staging copies; result == input."""
    import torch
    n = 200 * 1024 + 321
    r = np.random.default_rng(2)
    a = (np.arange(n, dtype=np.int64) * 10)
    b = r.integers(-100, 100, n).astype(np.int8)
    c = np.round(r.random(n) * 100, 2)
    d = np.repeat(r.integers(0, 9, n // 64 + 1), 64)[:n].astype(np.uint16)
    img = flsgpu.write_table([("a", "INT64", None), ("b", "INT8", None), ("c", "DOUBLE", None), ("d", "UINT16", None)],
                             [a, b, c, d], vec_per_rowgroup=8)
    t = flsgpu.Table(img)
    cols = [3, 1, 0]
    out_types = [flsgpu.OUT["U64"], flsgpu.OUT["I32"], 0]
    dts = [np.uint64, np.int32, np.int64]
    bufs = [torch.empty(t.out_elems(t.n_rowgroups) * np.dtype(x).itemsize, dtype=torch.uint8).pin_memory() for x in dts]
    t.scan_columns_to_host(cols, [x.data_ptr() for x in bufs], [x.numel() for x in bufs], out_types=out_types)
    for buf, dt, src in zip(bufs, dts, (d, b, a)):
        got = buf.numpy().view(dt)
        flat = np.concatenate([got[x:y] for x, y in t.valid_mask_slices()])
        assert np.array_equal(flat, src.astype(dt))
    t.close()


def gpu_aggregate(t, col, rg_begin=0, n_rg=None):
    import ctypes as C
    import torch
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    flsgpu.agg_reset(res.data_ptr())
    t.aggregate_column(col, res.data_ptr(), rg_begin, n_rg)
    torch.cuda.synchronize()
    r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
    s = (r.sum_hi << 64) | r.sum_lo
    if s >= 1 << 127:
        s -= 1 << 128
    return dict(sum=s, count=r.count, min=r.min_val, max=r.max_val, fsum=r.fsum)


def test_fused_aggregates_match_oracle():
    """SUM (128-bit) / COUNT / MIN / MAX of every standard case: fused decode+aggregate kernel == oracle, and the
    padding of a partial last vector is not counted."""
    cases = standard_cases(n=5 * 1024 + 333, seed=9)
    img = flsgpu.write_table([column_spec(c) for c in cases], [c[3] for c in cases], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    t.stage()
    o = oracle_lib.Oracle(img)
    for i, c in enumerate(cases):
        exp = o.aggregate(i, 0, t.n_rowgroups)
        got = gpu_aggregate(t, i)
        assert got["count"] == exp["count"] == len(c[3]), c[0]
        if c[3].dtype.kind == "f":
            fin = c[3][np.isfinite(c[3])].astype(np.float64)
            if np.isfinite(exp["fsum"]):
                assert abs(got["fsum"] - exp["fsum"]) <= 1e-9 * max(1.0, np.abs(fin).sum()), c[0]
            continue
        assert got["sum"] == exp["sum"] == int(c[3].astype(object).sum()), c[0]
        assert (got["min"], got["max"]) == (exp["min"], exp["max"]), c[0]
    t.close()


def test_aggregate_128bit_carry_and_accumulation():
    """Sums that overflow 64 bits (HUGEINT semantics) and accumulation over two calls into one result."""
    import torch
    n = 40 * 1024
    a = np.full(n, np.iinfo(np.int64).max, dtype=np.int64)
    b = np.full(n, np.iinfo(np.int64).min, dtype=np.int64)
    u = np.full(n, np.iinfo(np.uint64).max, dtype=np.uint64)
    img = flsgpu.write_table([("a", "INT64", "FFOR"), ("b", "INT64", "FFOR"), ("u", "UINT64", "UNCOMPRESSED")], [a, b, u],
                             vec_per_rowgroup=4)
    t = flsgpu.Table(img)
    t.stage()
    for col, src in ((0, a), (1, b), (2, u)):
        assert gpu_aggregate(t, col)["sum"] == int(src.astype(object).sum())
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    flsgpu.agg_reset(res.data_ptr())
    t.aggregate_column(0, res.data_ptr(), 0, 3)
    t.aggregate_column(0, res.data_ptr(), 3, t.n_rowgroups - 3)
    torch.cuda.synchronize()
    r = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
    assert ((r.sum_hi << 64) | r.sum_lo) == int(a.astype(object).sum()) and r.count == n
    t.close()


@pytest.mark.parametrize("n", [1024, 1_000_000])
def test_reference_test_tables_on_gpu(n):
    """BASELINE config 1: the reference's round-trip tables (test/all_types_single_threaded.test) at 1024 and 1 M rows:
    CUDA decode of (id, col) == what was written, and == the oracle."""
    from ref_fixtures import reference_test_tables
    ids = np.arange(1, n + 1, dtype=np.int64)
    for name, spec in reference_test_tables(n).items():
        img = flsgpu.write_table([("id", "INT64", None), ("col", spec[0], None)], [ids, spec[1]])
        t = flsgpu.Table(img)
        t.stage()
        o = oracle_lib.Oracle(img)
        for c, src in ((0, ids), (1, spec[1])):
            dt = t.native_dtype(c)
            got = gpu_decode(t, c, 0, dt)
            exp = o.decode_column(c, 0, t.n_rowgroups, 0, dt, t.vec_per_rowgroup, n_threads=8)
            for a, b in t.valid_mask_slices():
                assert got[a:b].tobytes() == exp[a:b].tobytes(), (name, c)
            flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
            assert flat.tobytes() == np.ascontiguousarray(src).tobytes(), (name, c)
        t.close()


def test_alp_golden_vectors_on_gpu():
    """ALP decode on the GPU reproduces the reference implementation's values (tests/golden/alp_golden.npz): the
    golden doubles/floats are written as ALP columns and must come back bit-exact."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "alp_golden.npz"))
    for key, dt, tname in (("f64_bits", np.float64, "DOUBLE"), ("f32_bits", np.float32, "FLOAT")):
        vals = g[key].view(dt)
        img = flsgpu.write_table([("v", tname, "ALP")], [vals], vec_per_rowgroup=2)
        t = flsgpu.Table(img)
        t.stage()
        got = gpu_decode(t, 0, 0, dt)
        flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
        assert flat.tobytes() == vals.tobytes()
        t.close()
    for bits, udt in ((64, np.uint64), (32, np.uint32)):
        t = flsgpu.Table(g[f"rd{bits}_image"])
        t.stage()
        got = gpu_decode(t, 0, 0, udt)
        assert got[:2048].tobytes() == g[f"rd{bits}_expected"].tobytes()
        t.close()


@pytest.mark.parametrize("dtype,tname", [(np.int64, "INT64"), (np.int32, "INT32"), (np.int16, "INT16"), (np.int8, "INT8"), (np.float64, "DOUBLE"),
                                         (np.float32, "FLOAT")])
def test_columns_that_mix_kernel_families(dtype, tname):
    """A column whose row groups fall into several launch kinds (light, 64-bit DELTA, DICT by code width, RLE, ALP, ALP_RD): every
    kind's launch walks the device list of its own vectors and decodes them in place (FlsColumnPlan::kind_map) - whole range,
    sub-ranges, widened output: always the input, always the oracle."""
    vpr, rows = 4, 4 * 1024
    r = np.random.default_rng(3)
    info = np.iinfo(dtype) if np.dtype(dtype).kind == "i" else None
    parts = []
    for rg in range(14):
        k = rg % 7
        if np.dtype(dtype).kind == "f":
            x = [np.round(r.random(rows) * 100, 2), r.random(rows) * 2 - 1, np.repeat(r.random(rows // 64), 64), np.full(rows, 2.5),
                 np.round(r.random(rows) * 1e6, 1), (r.random(rows) - 0.5) * 1e30, np.repeat(np.round(r.random(rows // 256), 3), 256)][k]
        else:
            hi = min(int(info.max), 1 << 40)
            x = [r.integers(0, hi, rows), np.sort(r.integers(0, hi, rows)), np.repeat(r.integers(0, hi, rows // 128), 128),
                 r.integers(0, 3, rows) * (hi // 3), np.full(rows, 7), np.arange(rows) * 3 % 100, r.choice(r.integers(0, hi, 300), rows)][k]
        parts.append(np.asarray(x).astype(dtype))
    col = np.concatenate(parts)[:14 * rows - 700]  # partial last row group
    img = flsgpu.write_table([("x", tname, None)], [col], vec_per_rowgroup=vpr)
    t = flsgpu.Table(img)
    kinds = {flsgpu.ENC_NAME[t.chunk(rg, 0).encoding] for rg in range(t.n_rowgroups)}
    assert len(kinds) >= 3, kinds
    o = oracle_lib.Oracle(img)
    dt = np.dtype(dtype)
    wide = np.dtype(np.float64) if dt.kind == "f" else np.dtype(np.int64)
    t.stage()
    for rg_begin, n_rg in ((0, t.n_rowgroups), (3, 6), (5, 1), (t.n_rowgroups - 2, 2)):
        for out_type, odt in ((0, dt), (flsgpu.OUT[_NAME_OF[wide]], wide)):
            exp = o.decode_column(0, rg_begin, n_rg, out_type, odt, t.vec_per_rowgroup)
            got = gpu_decode(t, 0, out_type, odt, rg_begin, n_rg)
            for a, b in t.valid_mask_slices(rg_begin, n_rg):
                assert got[a:b].tobytes() == exp[a:b].tobytes(), (kinds, rg_begin, n_rg, out_type)
    got = gpu_decode(t, 0, 0, dt)
    flat = np.concatenate([got[a:b] for a, b in t.valid_mask_slices()])
    assert flat.tobytes() == col.tobytes()
    t.close()
