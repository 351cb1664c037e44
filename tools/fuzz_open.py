"""Mutation fuzzing of the container validator (CPU only; test infrastructure).

flsgpu_table_open / flsgpu_table_open_memory are the gate in front of the device: the kernels trust every offset, width and
count of a table that was opened (DESIGN.md "validation").  This tool takes containers written by the product writer (every
numeric and string encoding), overwrites bytes of their metadata (footer, chunk descriptors, per-vector entry points,
dictionary offsets, run lengths, exception positions, FSST tables ...) and checks for each mutant:

  1. the open either fails with a status and a message or succeeds - it never crashes or reads outside the image (the image
     ends at a PROT_NONE guard page, so an over-read faults);
  2. when it succeeds, the metadata the kernels would follow is safe to follow: the CPU oracle - which walks the same layout
     with no checks of its own - decodes every column of the mutant with the image still against the guard page.

A crash kills the worker process; the driver (`run`) reports the seed image, iteration and mutation so that
`--replay NAME:ITER` reproduces it.  tests/test_fuzz_open.py runs a bounded number of iterations per seed image.
"""
import argparse
import ctypes as C
import json
import mmap
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

PAGE = mmap.PAGESIZE
libc = C.CDLL(None, use_errno=True)
libc.mprotect.argtypes = [C.c_void_p, C.c_size_t, C.c_int]


class Guarded:
    """A byte buffer whose last byte sits (up to 15 bytes of alignment slack) in front of an inaccessible page."""

    def __init__(self, capacity):
        self.n_pages = (capacity + PAGE - 1) // PAGE + 1
        self.map = mmap.mmap(-1, (self.n_pages + 1) * PAGE)
        self.base = C.addressof(C.c_char.from_buffer(self.map))
        if libc.mprotect(self.base + self.n_pages * PAGE, PAGE, 0) != 0:
            raise OSError(C.get_errno(), "mprotect")
        self.all = np.frombuffer(self.map, dtype=np.uint8, count=self.n_pages * PAGE)

    def tail(self, nbytes):
        """Address of nbytes of writable memory that end (16-byte aligned) at the guard page, or None when too large."""
        if nbytes > (self.n_pages - 1) * PAGE:
            return None
        return self.base + ((self.n_pages * PAGE - nbytes) & ~15)

    def place(self, image):
        start = (self.n_pages * PAGE - image.size) & ~15
        view = self.all[start:start + image.size]
        view[:] = image
        self.all[start + image.size:] = 0xA5
        return view


def seed_images():
    import flsgpu
    import fls_cases
    import str_cases
    out = {}
    n = 2 * 1024 + 300
    cases = fls_cases.standard_cases(n=n)
    for tname in ("INT8", "INT16", "INT32", "INT64", "FLOAT", "DOUBLE"):
        cs = [c for c in cases if c[1] == tname]
        out["num_" + tname] = flsgpu.write_table([fls_cases.column_spec(c) for c in cs], [c[3] for c in cs], vec_per_rowgroup=2)
    sc = str_cases.cases(n)
    names = list(sc)
    for lo in range(0, len(names), 5):
        part = names[lo:lo + 5]
        out["str_%d" % (lo // 5)] = flsgpu.write_table([(k, sc[k][0], sc[k][1]) for k in part], [sc[k][2] for k in part], vec_per_rowgroup=2)
    return out


INTERESTING = [0, 1, 2, 15, 16, 31, 32, 63, 64, 65, 127, 128, 255, 256, 1023, 1024, 1025, 4095, 4096, 65535, 65536, 1 << 20,
               (1 << 31) - 1, 1 << 31, (1 << 32) - 1, 1 << 32, (1 << 63) - 1, 1 << 63, (1 << 64) - 1, (1 << 64) - 16, (1 << 64) - 128]


def regions_of(flsgpu, image):
    """[(offset, size, weight)] of the byte ranges that hold metadata: the tail (header / column / row-group / chunk
    descriptors) and, per chunk, the META segment and the AUX segments (dictionaries, offsets, run lengths, FSST tables)."""
    t = flsgpu.Table(image.copy())
    regs = []
    first_chunk = image.size
    for r in range(t.n_rowgroups):
        for c in range(t.n_columns):
            ch = t.chunk(r, c)
            first_chunk = min(first_chunk, ch.offset)
            for s, w in ((0, 4.0), (2, 1.5), (3, 1.5), (1, 0.5)):
                if ch.seg_size[s]:
                    regs.append((ch.offset + ch.seg_off[s], int(ch.seg_size[s]), w))
    end_of_chunks = max((t.chunk(r, c).offset + t.chunk(r, c).size for r in range(t.n_rowgroups) for c in range(t.n_columns)), default=0)
    t.close()
    if end_of_chunks < image.size:
        regs.append((end_of_chunks, image.size - end_of_chunks, 8.0 * max(1, len(regs)) / 4))
    if first_chunk > 0:
        regs.append((0, first_chunk, 2.0))
    return regs


def mutate(rng, image, regs):
    img = image.copy()
    weights = np.array([w for _, _, w in regs], dtype=np.float64)
    weights /= weights.sum()
    what = []
    for _ in range(int(rng.choice([1, 1, 1, 2, 3]))):
        off, size, _ = regs[int(rng.choice(len(regs), p=weights))]
        kind = int(rng.integers(0, 8))
        if kind == 6:  # a block of metadata from elsewhere lands here (a descriptor of another chunk, another vector's entry)
            src_off, src_size, _ = regs[int(rng.choice(len(regs), p=weights))]
            ln = int(min(rng.choice([8, 16, 32, 64, 144]), size, src_size))
            a = src_off + (int(rng.integers(0, src_size - ln + 1)) & ~7)
            b = off + (int(rng.integers(0, size - ln + 1)) & ~7)
            a, b = min(a, src_off + src_size - ln), min(b, off + size - ln)
            img[b:b + ln] = image[a:a + ln]
            what.append((b, -ln, a))
            continue
        if kind == 7:  # the file ends early (or a few bytes of the tail are gone)
            cut = int(rng.choice([int(rng.integers(0, len(img))), len(img) - int(rng.integers(1, 300))]))
            img = img[:max(0, cut)].copy()
            what.append((len(img), 0, 0))
            break
        if kind <= 2:  # an aligned little-endian field takes an interesting value
            width = int(rng.choice([1, 2, 4, 8]))
            if size < width:
                continue
            pos = off + (int(rng.integers(0, size - width + 1)) & ~(width - 1))
            pos = min(pos, off + size - width)
            val = int(rng.choice(len(INTERESTING)))
            val = INTERESTING[val] & ((1 << (8 * width)) - 1)
            if kind == 2:  # ... or its own value +- a small step
                cur = int.from_bytes(img[pos:pos + width].tobytes(), "little")
                val = (cur + int(rng.choice([-17, -16, -2, -1, 1, 2, 16, 17, 1024]))) & ((1 << (8 * width)) - 1)
            img[pos:pos + width] = np.frombuffer(val.to_bytes(width, "little"), dtype=np.uint8)
            what.append((pos, width, val))
        elif kind == 3:  # bit flip
            pos = off + int(rng.integers(0, size))
            img[pos] ^= 1 << int(rng.integers(0, 8))
            what.append((pos, 1, int(img[pos])))
        elif kind == 4:  # random byte
            pos = off + int(rng.integers(0, size))
            img[pos] = int(rng.integers(0, 256))
            what.append((pos, 1, int(img[pos])))
        else:  # a short run of 0x00 / 0xFF
            ln = int(rng.integers(1, 17))
            pos = off + int(rng.integers(0, max(1, size - ln)))
            img[pos:min(pos + ln, off + size)] = int(rng.choice([0, 255]))
            what.append((pos, -ln, int(img[pos])))
    return img, what


def check(flsgpu, oracle_lib, view, out_guard, heap_guard):
    """Open the mutant in place; if the product accepts it, let the oracle walk it, writing into buffers that end at guard
    pages as well (a decoded size that disagrees with the metadata faults).  Returns 'rejected' | 'accepted'."""
    try:
        t = flsgpu.Table(view)
    except flsgpu.FlsError:
        return "rejected"
    try:
        try:
            o = oracle_lib.Oracle(view)
        except RuntimeError:
            return "accepted"  # the oracle's own header check is stricter than needed: nothing to walk
        heap_base = 0x7F0000001000
        n_rg, vpr = t.n_rowgroups, t.vec_per_rowgroup
        for c in range(t.n_columns):
            col = t.column(c)
            if col.phys_bits == 128:
                try:  # host metadata entry points a reader calls right after the open
                    need = t.string_heap_bytes(c)
                    t.string_stats(c)
                except flsgpu.FlsError:
                    continue
                if need != o.string_heap_bytes(c, 0, n_rg):
                    raise AssertionError("product and oracle size the string heap differently")
                rows, heap = out_guard.tail(n_rg * vpr * 1024 * 16), heap_guard.tail(max(16, need))
                if rows and heap:
                    oracle_lib.lib.fls_oracle_decode_strings(C.byref(o.t), c, 0, n_rg, C.c_void_p(rows), C.c_void_p(heap), C.c_uint64(heap_base))
            else:
                out = out_guard.tail(n_rg * vpr * 1024 * (col.phys_bits // 8))
                if out:
                    oracle_lib.lib.fls_oracle_decode_column(C.byref(o.t), c, 0, n_rg, 0, C.c_void_p(out), 1)
    finally:
        t.close()
    return "accepted"


def worker(name, begin, end, report):
    import flsgpu
    import oracle_lib
    image = seed_images()[name]
    regs = regions_of(flsgpu, image)
    guard, out_guard, heap_guard = Guarded(image.size), Guarded(8 << 20), Guarded(8 << 20)
    counts = {"rejected": 0, "accepted": 0}
    for it in range(begin, end):
        rng = np.random.default_rng([sum(name.encode()), it])
        img, what = mutate(rng, image, regs)
        if report:
            with open(report, "w") as f:  # the driver reads this after a crash
                json.dump({"seed": name, "iter": it, "mutation": what}, f)
        counts[check(flsgpu, oracle_lib, guard.place(img), out_guard, heap_guard)] += 1
    print(json.dumps({"seed": name, "begin": begin, "end": end, **counts}), flush=True)


def run(names, iters, report_dir, chunk=250):
    """Drives workers; returns (summary, crashes)."""
    os.makedirs(report_dir, exist_ok=True)
    summary, crashes = {}, []
    for name in names:
        acc = rej = 0
        for begin in range(0, iters, chunk):
            rep = os.path.join(report_dir, f"fuzz_{name}.json")
            cur = begin
            while cur < min(iters, begin + chunk):
                p = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", name, "--begin", str(cur),
                                    "--end", str(min(iters, begin + chunk)), "--report", rep], capture_output=True, text=True)
                if p.returncode == 0:
                    line = json.loads(p.stdout.strip().splitlines()[-1])
                    acc += line["accepted"]
                    rej += line["rejected"]
                    break
                last = json.load(open(rep)) if os.path.exists(rep) else {"seed": name, "iter": cur, "mutation": None}
                last["returncode"] = p.returncode
                last["stderr"] = p.stderr[-400:]
                crashes.append(last)
                cur = last["iter"] + 1  # carry on behind the crashing mutant
        summary[name] = {"accepted": acc, "rejected": rej}
    return summary, crashes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=2000)
    ap.add_argument("--seeds", default=None, help="comma list of seed image names (default: all)")
    ap.add_argument("--worker", default=None)
    ap.add_argument("--begin", type=int, default=0)
    ap.add_argument("--end", type=int, default=0)
    ap.add_argument("--report", default=None)
    ap.add_argument("--replay", default=None, help="NAME:ITER - run one mutant in this process")
    ap.add_argument("--out", default="/tmp/fls_fuzz")
    args = ap.parse_args()
    if args.worker:
        worker(args.worker, args.begin, args.end, args.report)
        return
    if args.replay:
        name, it = args.replay.split(":")
        worker(name, int(it), int(it) + 1, None)
        return
    names = args.seeds.split(",") if args.seeds else list(seed_images())
    summary, crashes = run(names, args.iters, args.out)
    print(json.dumps({"summary": summary, "crashes": crashes}, indent=1))
    sys.exit(1 if crashes else 0)


if __name__ == "__main__":
    main()
