"""CPU: libflsgpu.so loads and exports every symbol include/flsgpu.h declares; error behaviour without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import flsgpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "flsgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(flsgpu_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(flsgpu.lib, n), f"{n} declared in include/flsgpu.h but not exported by libflsgpu.so"
    assert sorted(flsgpu.ABI_SYMBOLS) == names, "flsgpu.ABI_SYMBOLS must list exactly the header's entry points"


def test_struct_sizes_match_header():
    assert C.sizeof(flsgpu.ChunkDesc) == 144 and C.sizeof(flsgpu.ColumnDesc) == 80 and C.sizeof(flsgpu.AggResult) == 64


def test_version_and_error_slot():
    assert b"sm_100a" in flsgpu.lib.flsgpu_version()
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Table(np.zeros(10, dtype=np.uint8))
    assert e.value.status == -1 and "header" in str(e.value)
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Table(path="/nonexistent/x.fls")
    assert e.value.status == -2


def test_malformed_containers_are_rejected():
    a = np.arange(5000, dtype=np.int32)
    img = flsgpu.write_table([("a", "INT32", "FFOR")], [a], vec_per_rowgroup=2)
    flsgpu.Table(img).close()
    bad = img.copy()
    bad[0] ^= 0xFF  # magic
    with pytest.raises(flsgpu.FlsError):
        flsgpu.Table(bad)
    with pytest.raises(flsgpu.FlsError):
        flsgpu.Table(img[: len(img) - 100])  # truncated footer
    t = flsgpu.Table(img)
    ch = t.chunk(0, 0)
    t.close()
    bad = img.copy()
    # vector 0's width -> 77 (> T): fls_vec_meta.width is byte 18 of the 32-byte entry
    bad[ch.offset + ch.seg_off[0] + 18] = 77
    with pytest.raises(flsgpu.FlsError) as e:
        flsgpu.Table(bad)
    assert "width" in str(e.value)


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    a = np.arange(3000, dtype=np.int32)
    t = flsgpu.Table(flsgpu.write_table([("a", "INT32", None)], [a]))
    with pytest.raises(flsgpu.FlsError) as e:
        t.stage()
    assert e.value.status == -3  # FLSGPU_ERR_CUDA: the product never decodes on the CPU
    t.close()
    # string columns: heap sizing is host metadata, every decode entry point needs the device
    t = flsgpu.Table(flsgpu.write_table([("s", "VARCHAR", None)], [[b"Value %d" % (i % 10) for i in range(3000)]]))
    assert t.string_heap_bytes(0) == sum((t.chunk(r, 0).const_val + 15) // 16 * 16 for r in range(t.n_rowgroups))
    rows, heap = np.zeros((t.out_elems(t.n_rowgroups), 16), dtype=np.uint8), np.zeros(4096, dtype=np.uint8)
    with pytest.raises(flsgpu.FlsError) as e:
        t.scan_strings_to_host(0, rows.ctypes.data, rows.nbytes, heap.ctypes.data, heap.nbytes)
    assert e.value.status == -3
    t.close()


def test_writer_rejects_bad_input():
    with pytest.raises(flsgpu.FlsError):
        flsgpu.write_table([("a", "INT32", "CONSTANT")], [np.arange(10, dtype=np.int32)])
    with pytest.raises(flsgpu.FlsError):
        flsgpu.write_table([("a", "INT32", "ALP")], [np.arange(10, dtype=np.int32)])


def test_streaming_writer_writes_the_same_bytes_as_the_image(tmp_path):
    """flsgpu_writer_stream_to: row groups reach the file as they fill, the finished file is byte for byte the in-memory
    container (inline footer), or data file + '<path>.footer' (reference WithInlinedFooter(false), fls_writer.cpp:229-230)."""
    import fls_cases
    import str_cases
    n = 5 * 2048 + 333
    cs = fls_cases.standard_cases(n=n)
    sc = str_cases.cases(n)
    cols = [fls_cases.column_spec(c) for c in cs] + [(k, sc[k][0], sc[k][1]) for k in sc]
    arrs = [c[3] for c in cs] + [sc[k][2] for k in sc]
    img = flsgpu.write_table(cols, arrs, vec_per_rowgroup=2)
    path = str(tmp_path / "s.fls")
    w = flsgpu.Writer(cols, 2)
    w.stream_to(path)
    cut = [0, 100, 2048, 2049, 7000, n]  # uneven pieces: pending rows, several row groups at once, a partial tail
    for a, b in zip(cut, cut[1:]):
        w.append([x[a:b] for x in arrs])
    w.finish_stream()
    w.save(path)  # the path it streams to: accepted, nothing left to do
    with pytest.raises(flsgpu.FlsError):
        w.save(path + ".elsewhere")
    with pytest.raises(flsgpu.FlsError):
        w.finish()  # a streaming writer keeps no image
    w.close()
    assert np.array_equal(np.fromfile(path, dtype=np.uint8), img)
    t = flsgpu.Table(path=path)
    assert t.n_rowgroups == (n + 2047) // 2048
    t.close()

    # footer beside the data file
    a = np.arange(10_000, dtype=np.int64)
    w = flsgpu.Writer([("a", "INT64", None)], 2)
    _check = flsgpu._check
    _check(flsgpu.lib.flsgpu_writer_set_inline_footer(w._h, 0), "set_inline_footer")
    split = str(tmp_path / "split.fls")
    w.stream_to(split)
    w.append([a])
    w.finish_stream()
    w.close()
    ref = flsgpu.Writer([("a", "INT64", None)], 2)
    _check(flsgpu.lib.flsgpu_writer_set_inline_footer(ref._h, 0), "set_inline_footer")
    ref.append([a])
    ref.finish()
    ref.save(str(tmp_path / "ref.fls"))
    ref.close()
    for suffix in ("", ".footer"):
        assert open(split + suffix, "rb").read() == open(str(tmp_path / "ref.fls") + suffix, "rb").read(), suffix

    # misuse: after the first row group, twice, unwritable path
    w = flsgpu.Writer([("a", "INT64", None)], 2)
    with pytest.raises(flsgpu.FlsError) as e:
        w.stream_to(str(tmp_path / "no_such_dir" / "x.fls"))
    assert e.value.status == -2
    w.append([a])
    with pytest.raises(flsgpu.FlsError):
        w.stream_to(str(tmp_path / "late.fls"))
    w.close()
    w = flsgpu.Writer([("a", "INT64", None)], 2)
    w.stream_to(str(tmp_path / "once.fls"))
    with pytest.raises(flsgpu.FlsError):
        w.stream_to(str(tmp_path / "twice.fls"))
    w.close()
