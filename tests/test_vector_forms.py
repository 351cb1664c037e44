"""Vector forms (flsgpu_column_vector_forms: CONSTANT / SEQUENCE hand-off without decode) and compact scans.

CPU: every vector the forms call CONSTANT or SEQUENCE must expand to exactly what the oracle decodes for it; FLAT vectors
are numbered in range order.  GPU: a compact scan (only FLAT vectors decoded and copied, packed) + the expanded forms gives
the same column, bit for bit, as the oracle and as the full scan.
"""
import numpy as np
import pytest

import flsgpu
import oracle_lib

VEC = 1024


def columns(n):
    i = np.arange(n, dtype=np.int64)
    step = np.where((i // 5000) % 2 == 0, 3, 7)  # the increment changes every 5000 rows: some vectors are sequences, some not
    return [
        ("seq32", "INT32", None, (i + 1).astype(np.int32)),
        ("seq64", "INT64", None, i * 10 - 7),
        ("useq64", "UINT64", None, (i * 10).astype(np.uint64)),
        ("down16", "INT16", None, (30000 - (i % 20000)).astype(np.int16)),      # decreasing, restarts inside a vector
        ("wrap8", "INT8", None, ((i % 256) - 128).astype(np.int8)),             # wraps inside every vector: never a sequence
        ("udown32", "UINT32", None, (4_000_000_000 - i * 3).astype(np.uint32)),  # unsigned, decreasing: increment -3
        ("big_u64", "UINT64", None, (np.uint64(1 << 63) + (i * 2).astype(np.uint64))),  # beyond int64: FLAT
        ("const", "INT64", None, np.full(n, -42, dtype=np.int64)),
        ("piece", "INT64", None, np.cumsum(step)),
        ("const_run", "INT32", "FFOR", np.repeat(np.arange(n // 3000 + 1), 3000)[:n].astype(np.int32)),
        ("rand", "INT32", None, ((i * 2654435761) % 1000003).astype(np.int32)),
        ("dbl", "DOUBLE", None, np.full(n, 2.5)),
        ("date", "DATE", None, (8000 + i % 10000).astype(np.int32)),
        ("delta_only", "INT64", "DELTA", i * 1000),
    ]


def expand(forms, flat, dtype):
    """forms + packed FLAT vectors -> the column in decode_column layout."""
    out = np.zeros(len(forms) * VEC, dtype=dtype)
    k = np.arange(VEC, dtype=np.int64)
    for v, f in enumerate(forms):
        if f["form"] == flsgpu.FORM_FLAT:
            s = int(f["flat_slot"])
            out[v * VEC:(v + 1) * VEC] = flat[s * VEC:(s + 1) * VEC]
        elif np.dtype(dtype).kind == "f":
            bits = np.array([f["start"]], dtype=np.int64).view(np.uint64)
            val = bits.astype(np.uint32).view(np.float32)[0] if np.dtype(dtype).itemsize == 4 else bits.view(np.float64)[0]
            out[v * VEC:(v + 1) * VEC] = val
        else:
            vals = int(f["start"]) + k * int(f["increment"])  # exact: the forms guarantee no wrap inside int64 / the type
            out[v * VEC:(v + 1) * VEC] = vals.astype(np.int64).view(np.int64).astype(dtype) if np.dtype(dtype).kind == "i" \
                else vals.astype(np.uint64).astype(dtype)
    return out


@pytest.mark.parametrize("n,vpr", [(70000, 4), (1024 * 8 + 5, 8), (300, 2)])
def test_forms_expand_to_what_the_oracle_decodes(n, vpr):
    cols = columns(n)
    img = flsgpu.write_table([(c[0], c[1], c[2]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=vpr)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    seen = {flsgpu.FORM_FLAT: 0, flsgpu.FORM_CONSTANT: 0, flsgpu.FORM_SEQUENCE: 0}
    for c, col in enumerate(cols):
        dt = col[3].dtype
        forms, n_flat = t.vector_forms(c)
        exp = o.decode_column(c, 0, t.n_rowgroups, 0, dt, t.vec_per_rowgroup)
        flat_ids = [v for v, f in enumerate(forms) if f["form"] == flsgpu.FORM_FLAT]
        assert n_flat == len(flat_ids) and [int(forms[v]["flat_slot"]) for v in flat_ids] == list(range(n_flat)), col[0]
        packed = np.concatenate([exp[v * VEC:(v + 1) * VEC] for v in flat_ids]) if flat_ids else np.zeros(0, dtype=dt)
        got = expand(forms, packed, dt)
        for r, (a, b) in enumerate(t.valid_mask_slices()):
            assert got[a:b].tobytes() == exp[a:b].tobytes(), (col[0], r)
        for f in forms:
            seen[int(f["form"])] += 1
        if col[0] in ("wrap8", "rand", "big_u64"):
            assert n_flat == sum(t.rowgroup(r).n_vec for r in range(t.n_rowgroups)), col[0]
        if col[0] in ("seq32", "seq64", "useq64", "udown32", "const", "dbl", "delta_only") and n >= 2048:
            assert n_flat <= (1 if n % 1024 else 0), (col[0], n_flat)  # a tiny tail chunk may be written FFOR
    assert all(seen.values()) or n < 2048
    t.close()


def test_sub_ranges_number_their_own_flat_vectors():
    cols = columns(50000)
    img = flsgpu.write_table([(c[0], c[1], c[2]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=4)
    t = flsgpu.Table(img)
    c = [x[0] for x in cols].index("piece")
    whole, _ = t.vector_forms(c)
    part, n_flat = t.vector_forms(c, 3, 5)
    assert (part["form"] == whole["form"][12:32]).all() and (part["start"] == whole["start"][12:32]).all()
    flat = part["flat_slot"][part["form"] == flsgpu.FORM_FLAT]
    assert list(flat) == list(range(n_flat))
    with pytest.raises(flsgpu.FlsError):
        t.vector_forms(c, 10, 100)
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n,vpr", [(70000, 4), (400000, 64)])
def test_compact_scan_plus_forms_equals_full_scan(n, vpr):
    cols = columns(n)
    img = flsgpu.write_table([(c[0], c[1], c[2]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=vpr)
    t = flsgpu.Table(img)
    o = oracle_lib.Oracle(img)
    import ctypes as C
    rows = t.out_elems(t.n_rowgroups)
    for rg0, nrg in ((0, t.n_rowgroups), (1, max(1, t.n_rowgroups - 2))):
        scan, bufs, meta = [], [], []
        for c, col in enumerate(cols):
            dt = col[3].dtype
            forms, n_flat = t.vector_forms(c, rg0, nrg)
            nbytes = max(16, n_flat * VEC * dt.itemsize)
            p = flsgpu.host_alloc(nbytes)
            bufs.append(p)
            scan.append(flsgpu.ScanColumn(col=c, out_type=0, compact=1, out_host=p, out_capacity_bytes=nbytes))
            meta.append((forms, n_flat, dt))
        t.evict()
        launches = flsgpu.launch_count()
        t.scan_projection_to_host(scan, rg0, nrg)
        assert flsgpu.launch_count() > launches
        for c, col in enumerate(cols):
            forms, n_flat, dt = meta[c]
            packed = np.ctypeslib.as_array((C.c_uint8 * max(1, n_flat * VEC * dt.itemsize)).from_address(bufs[c]))[:n_flat * VEC * dt.itemsize].view(dt)
            got = expand(forms, packed, dt)
            exp = o.decode_column(c, rg0, nrg, 0, dt, t.vec_per_rowgroup)
            for r, (a, b) in enumerate(t.valid_mask_slices(rg0, nrg)):
                assert got[a:b].tobytes() == exp[a:b].tobytes(), (col[0], rg0, r)
        for p in bufs:
            flsgpu.host_free(p)
    # a fully formed column costs no launch and no staging at all
    c = [x[0] for x in cols].index("seq64")
    t.evict()
    before, resident = flsgpu.launch_count(), t.resident_bytes()
    p = flsgpu.host_alloc(16)
    t.scan_projection_to_host([flsgpu.ScanColumn(col=c, out_type=0, compact=1, out_host=p, out_capacity_bytes=16)])
    assert flsgpu.launch_count() == before and t.resident_bytes() == resident
    flsgpu.host_free(p)
    assert rows > 0
    t.close()


@pytest.mark.gpu
@pytest.mark.parametrize("threads", [1, 4])
def test_reader_hands_constant_and_sequence_vectors_on(tmp_path, threads):
    """FastLanesReader::Scan over a compact scan: CONSTANT / SEQUENCE vectors, zero-copy FLAT slices, chunks put together
    cut where the forms of neighbouring vectors differ — the rows a consumer reads are the input, with the forms on and off."""
    import reader_lib
    n = 333333
    cols = columns(n)
    path = str(tmp_path / "forms.fls")
    flsgpu.write_table([(c[0], c[1], c[2]) for c in cols], [c[3] for c in cols], vec_per_rowgroup=8).tofile(path)
    for on in (1, 0):
        r = reader_lib.Reader(path, rowgroups_per_batch=5)
        r.set_option("compressed_vectors", on)
        got, _, stats = r.scan(n_threads=threads)
        for c, col in enumerate(cols):
            exp = col[3].astype(np.uint64) if col[3].dtype.kind == 'u' else col[3]  # every unsigned width reads back as UBIGINT
            assert got[c].view(exp.dtype).tobytes() == exp.tobytes(), (col[0], on)
        assert stats["rows"] == n
        if on:
            assert stats["sequence_vectors"] > 0 and stats["constant_vectors"] > 0 and stats["max_chunk_rows"] == 2048, stats
            assert stats["chunks"] > (n + 2047) // 2048  # chunks are cut where the forms of two neighbouring vectors differ
        else:
            assert stats["sequence_vectors"] == stats["constant_vectors"] == 0, stats
        st2 = r.consume(n_threads=threads)
        assert st2["rows"] == n and st2["chunks"] == stats["chunks"]
        r.close()
