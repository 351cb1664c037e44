#!/usr/bin/env python
"""Generate tests/golden/fsst_golden.npz from the reference tree's own FSST code (oracle/_ref/libfsstref.so =
duckdb/third_party/fsst compiled from /root/reference by oracle/Makefile).

Two directions pin the string path:
  ref_*   the REFERENCE compressor's symbol table + code streams for three corpora; the oracle's decoder
          (oracle/fls_oracle_str.c) must turn them back into the original bytes;
  prod_*  FSST chunks written by the PRODUCT encoder (csrc/host/fls_fsst.cpp), and what the REFERENCE decoder
          (fsst.h:162-230) makes of their code streams: must equal the original bytes, i.e. the product's streams are valid
          input for the reference's fsst_decompress.
Run in the build container:   make -C oracle ref && python oracle/make_golden_fsst.py
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "duckdb-fastlanes_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = os.path.join(ROOT, "oracle", "_ref", "libfsstref.so")
OUT = os.path.join(ROOT, "tests", "golden", "fsst_golden.npz")
FSST_TABLE_BYTES = 256 * 8 + 256


def load_ref():
    l = C.CDLL(REF)
    l.ref_fsst_compress.restype = C.c_longlong
    l.ref_fsst_compress.argtypes = [C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    l.ref_fsst_decompress.restype = C.c_size_t
    l.ref_fsst_decompress.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p]
    return l


def corpora(n=1500):
    import str_cases
    cs = str_cases.cases(n)
    return {"comment": cs["comment"][2], "blob": cs["blob"][2], "value": cs["value_unique"][2]}


def ref_compress(l, strings):
    import flsgpu
    off, data = flsgpu.pack_strings(strings)
    n = len(strings)
    symbol = np.zeros(256, dtype=np.uint64)
    length = np.zeros(256, dtype=np.uint8)
    cap = 7 + 2 * int(off[-1]) + 16 * n + 1024
    out = np.zeros(cap, dtype=np.uint8)
    comp_end = np.zeros(n, dtype=np.uint64)
    nb = l.ref_fsst_compress(n, off.ctypes.data, data.ctypes.data, symbol.ctypes.data, length.ctypes.data, out.ctypes.data, cap,
                             comp_end.ctypes.data)
    assert nb >= 0, nb
    return off, data, symbol, length, out[:nb].copy(), comp_end


def ref_decompress(l, symbol, length, codes, comp_end, dec_cap):
    out = np.zeros(dec_cap + 8, dtype=np.uint8)
    ends = np.zeros(len(comp_end), dtype=np.uint64)
    pos, prev = 0, 0
    codes = np.ascontiguousarray(codes)
    for i, e in enumerate(comp_end):
        e = int(e)
        tmp = np.zeros(8 * (e - prev) + 8, dtype=np.uint8)
        k = l.ref_fsst_decompress(symbol.ctypes.data, length.ctypes.data, e - prev, codes.ctypes.data + prev, len(tmp), tmp.ctypes.data)
        out[pos:pos + k] = tmp[:k]
        pos += k
        ends[i] = pos
        prev = e
    return out[:pos].copy(), ends


def product_chunk(strings):
    """(table symbols, lens, codes, comp_end, dec_end) of an FSST chunk written by the product encoder."""
    import flsgpu
    img = flsgpu.write_table([("s", "BLOB", "FSST")], [strings], vec_per_rowgroup=2)
    t = flsgpu.Table(img)
    assert t.n_rowgroups == 1
    ch = t.chunk(0, 0)
    a0, a1 = ch.offset + ch.seg_off[2], ch.offset + ch.seg_off[3]
    symbol = np.frombuffer(img, dtype=np.uint64, count=256, offset=a0).copy()
    length = np.frombuffer(img, dtype=np.uint8, count=256, offset=a0 + 2048).copy()
    codes = np.frombuffer(img, dtype=np.uint8, count=ch.seg_size[2] - FSST_TABLE_BYTES, offset=a0 + FSST_TABLE_BYTES).copy()
    n = ch.aux_count
    comp_end = np.frombuffer(img, dtype=np.uint32, count=n, offset=a1).astype(np.uint64)
    dec_end = np.frombuffer(img, dtype=np.uint32, count=n, offset=a1 + 4 * n).astype(np.uint64)
    t.close()
    return symbol, length, codes, comp_end, dec_end


def generate():
    l = load_ref()
    g = {}
    for name, strings in corpora().items():
        off, data, symbol, length, codes, comp_end = ref_compress(l, strings)
        dec, ends = ref_decompress(l, symbol, length, codes, comp_end, int(off[-1]))
        assert dec.tobytes() == data[:int(off[-1])].tobytes() and ends.tolist() == off[1:].tolist()
        g[f"ref_{name}_offsets"], g[f"ref_{name}_bytes"] = off, data[:int(off[-1])]
        g[f"ref_{name}_symbol"], g[f"ref_{name}_len"] = symbol, length
        g[f"ref_{name}_codes"], g[f"ref_{name}_comp_end"] = codes, comp_end
        n = min(len(strings), 2048)
        psym, plen, pcodes, pcend, pdend = product_chunk(strings[:n])
        pdec, pends = ref_decompress(l, psym, plen, pcodes, pcend, int(pdend[-1]))
        assert pdec.tobytes() == b"".join(strings[:n]), "reference decoder over the product's streams must give back the input"
        assert pends.tolist() == pdend.tolist()
        g[f"prod_{name}_symbol"], g[f"prod_{name}_len"] = psym, plen
        g[f"prod_{name}_codes"], g[f"prod_{name}_comp_end"] = pcodes, pcend
        g[f"prod_{name}_ref_decoded"], g[f"prod_{name}_ref_ends"] = pdec, pends
    return g


if __name__ == "__main__":
    g = generate()
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, {k: v.shape for k, v in g.items()})
