#!/usr/bin/env python
"""Generate tests/golden/alp_golden.npz from the reference tree's own ALP / ALP-RD code (oracle/_ref/libalpref.so =
DuckDB's vendored alp.hpp / alprd.hpp compiled from /root/reference by oracle/Makefile).

Run in the build container (where /root/reference exists):   make -C oracle ref && python oracle/make_golden.py
The .npz travels with the repo; tests/test_golden.py checks the oracle restatement against it everywhere and, where
the .so exists, re-derives the vectors to prove the committed file is not stale.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlanes_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = os.path.join(ROOT, "oracle", "_ref", "libalpref.so")
OUT = os.path.join(ROOT, "tests", "golden", "alp_golden.npz")


def load_ref():
    l = C.CDLL(REF)
    l.ref_alp_decode_f64.restype = C.c_double
    l.ref_alp_decode_f64.argtypes = [C.c_int64, C.c_int, C.c_int]
    l.ref_alp_decode_f32.restype = C.c_float
    l.ref_alp_decode_f32.argtypes = [C.c_int64, C.c_int, C.c_int]
    l.ref_alp_encode_f64.restype = C.c_int64
    l.ref_alp_encode_f64.argtypes = [C.c_double, C.c_int, C.c_int]
    l.ref_alp_encode_f32.restype = C.c_int64
    l.ref_alp_encode_f32.argtypes = [C.c_float, C.c_int, C.c_int]
    l.ref_alp_impossible_f64.argtypes = [C.c_double]
    l.ref_alp_impossible_f32.argtypes = [C.c_float]
    for name in ("ref_alprd_decompress_u64", "ref_alprd_decompress_u32"):
        getattr(l, name).restype = None
        getattr(l, name).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    return l


def alp_vectors(l, seed=20261104, n=4096):
    r = np.random.default_rng(seed)
    g = {}
    for name, max_e, dec, ibits in (("f64", 18, l.ref_alp_decode_f64, 63), ("f32", 10, l.ref_alp_decode_f32, 31)):
        e = r.integers(0, max_e, n, endpoint=True)
        f = (r.random(n) * (e + 1)).astype(np.int64)  # 0 <= f <= e
        mag = r.integers(0, ibits, n, endpoint=True)
        digits = (r.integers(0, 2 ** 62, n) >> (62 - np.minimum(mag, 62))) * r.choice([-1, 1], n)
        digits[:8] = [0, 1, -1, 10, -10, 2 ** (ibits - 1), -(2 ** (ibits - 1)), 123456789]
        fdt, udt = (np.float64, np.uint64) if name == "f64" else (np.float32, np.uint32)
        out = np.array([dec(int(d), int(a), int(b)) for d, a, b in zip(digits, e, f)], dtype=fdt)
        g[f"{name}_digits"] = digits.astype(np.int64)
        g[f"{name}_e"] = e.astype(np.uint8)
        g[f"{name}_f"] = f.astype(np.uint8)
        g[f"{name}_bits"] = out.view(udt)
    # encode side (alp.hpp:108-123): decimal-looking values, specials, their digits and "impossible" flags
    vals = np.round(r.random(2048) * 10.0 ** r.integers(0, 6, 2048), 3)
    vals[:10] = [0.0, -0.0, np.nan, np.inf, -np.inf, 1e300, -1e300, 0.1, 123.456, 9.2e18]
    e = r.integers(0, 18, 2048, endpoint=True)
    f = (r.random(2048) * (e + 1)).astype(np.int64)
    g["enc64_values"] = vals
    g["enc64_e"] = e.astype(np.uint8)
    g["enc64_f"] = f.astype(np.uint8)
    g["enc64_digits"] = np.array([l.ref_alp_encode_f64(float(v), int(a), int(b)) for v, a, b in zip(vals, e, f)], dtype=np.int64)
    g["enc64_impossible"] = np.array([l.ref_alp_impossible_f64(float(v)) for v in vals], dtype=np.uint8)
    return g


META_DT = np.dtype([("base", "<u8"), ("data_off16", "<u4"), ("aux_off", "<u4"), ("aux_cnt", "<u2"), ("width", "u1"),
                    ("width2", "u1"), ("e", "u1"), ("f", "u1"), ("n_valid", "<u2"), ("reserved", "<u8")])


def alprd_container(l, bits, seed):
    """A 2-vector ALP_RD container written by the product encoder + the values the REFERENCE glue arithmetic
    (alprd.hpp:222-236) yields from its packed parts, dictionary and exceptions."""
    import flsgpu
    import oracle_lib
    r = np.random.default_rng(seed)
    fdt, udt, tname = (np.float64, np.uint64, "DOUBLE") if bits == 64 else (np.float32, np.uint32, "FLOAT")
    x = (r.random(2048) * 2 - 1).astype(fdt)
    x[5] = np.nan
    x[700] = fdt(1e30)
    x[1500] = -0.0
    img = flsgpu.write_table([("x", tname, "ALP_RD")], [x], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    ch = t.chunk(0, 0)
    meta = np.frombuffer(img, dtype=META_DT, count=ch.n_vec, offset=ch.offset + ch.seg_off[0])
    data0 = ch.offset + ch.seg_off[1]
    decompress = l.ref_alprd_decompress_u64 if bits == 64 else l.ref_alprd_decompress_u32
    rd_dict = np.array([int(k) for k in ch.rd_dict], dtype=np.uint16)
    expected = np.zeros(2048, dtype=udt)
    for v in range(ch.n_vec):
        m = meta[v]
        blob = data0 + int(m["data_off16"]) * 16
        right = oracle_lib.unpack(np.frombuffer(img, dtype=np.uint8, count=128 * int(m["width"]), offset=blob), bits, int(m["width"]))
        left = oracle_lib.unpack(np.frombuffer(img, dtype=np.uint8, count=128 * int(m["width2"]), offset=blob + 128 * int(m["width"])),
                                 bits, int(m["width2"]))
        n = int(m["aux_cnt"])  # exception tail of the blob: [n whole words, padded to 16 B][n u16 positions]
        tail = blob + 128 * (int(m["width"]) + int(m["width2"]))
        exc_vals = np.frombuffer(img, dtype=udt, count=n, offset=tail)
        exc_pos = np.ascontiguousarray(np.frombuffer(img, dtype=np.uint16, count=n, offset=tail + (n * (bits // 8) + 15) // 16 * 16))
        # the reference keeps an exception's LEFT part (alprd.hpp:229-234); the container keeps the whole word, whose right
        # part is the packed one: the left part is what sits above right_bit_width
        assert np.array_equal(exc_vals & udt((1 << int(m["width"])) - 1), np.asarray(right, dtype=udt)[exc_pos])
        exc_left = np.ascontiguousarray((exc_vals >> udt(int(m["width"]))).astype(np.uint16))
        left16 = np.ascontiguousarray((np.asarray(left) & 7).astype(np.uint16))
        right_u = np.ascontiguousarray(np.asarray(right, dtype=udt))
        out = np.zeros(1024, dtype=udt)
        # AlpRDDecompression<T>::Decompress of the reference tree, unchanged (oracle/ref_alp.cpp packs the parts with DuckDB's
        # BitpackingPrimitives first: that is the layout it unpacks)
        decompress(left16.ctypes.data, right_u.ctypes.data, 1024, rd_dict.ctypes.data, exc_left.ctypes.data, exc_pos.ctypes.data, n,
                   int(m["width2"]), int(m["width"]), out.ctypes.data)
        expected[v * 1024:(v + 1) * 1024] = out
    t.close()
    assert expected.tobytes() == x.tobytes(), "reference glue over the container's parts must reproduce the input"
    return img.copy(), expected


def generate():
    l = load_ref()
    g = alp_vectors(l)
    for bits in (64, 32):
        img, exp = alprd_container(l, bits, seed=bits)
        g[f"rd{bits}_image"] = img
        g[f"rd{bits}_expected"] = exp
    return g


if __name__ == "__main__":
    g = generate()
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, {k: v.shape for k, v in g.items()})
