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
