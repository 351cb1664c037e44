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
