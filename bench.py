#!/usr/bin/env python
"""bench.py — read_fls scan throughput of the B200 FastLanes decode engine (BASELINE.json metric).

A step = one full scan of the all_types table (BASELINE.json configs[1]: 100 M rows; TINYINT .. TIMESTAMP, DOUBLE via
ALP, VARCHAR via dictionary codes): every column of every row group is decoded through the C ABI
(flsgpu_decode_column), one fused kernel launch per column.

  value : decoded GB/s with the compressed container resident in HBM (CUDA events, max over ranks)
  e2e   : the same scan through flsgpu_scan_column_to_host: pinned host container -> H2D -> decode -> D2H into pinned
          host column buffers, every step
  roofline : the kernel template that takes the largest share of the step, algorithmic bytes / its
          CUDA-event time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference : oracle/liboracle.so's reference-shaped scan (2048-row chunks, one row group per
          thread) on the host cores.  The reference's own binary cannot be built here (its FastLanes library is an
          un-vendored dependency), so kind = "port".

Multi-GPU: one process per GPU (torchrun); every rank scans its own shard of row groups (an all_types table of
--rows rows per rank: weak scaling, no data-path collective); rank 0 gathers the per-rank times and the reduced
scan results (row counts + checksums) with one NCCL all_reduce.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "duckdb-fastlanes_b200"), os.path.join(ROOT, "tests"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "read_fls_decoded_GBps"
UNIT = "GB/s"
WORKLOAD = "all_types_100M"

# name, fls type, forced encoding (None: the writer's own selection), numpy generator over row ids i = first..first+n
COLUMNS = [
    ("tiny", "INT8", None, lambda i: ((i % 256) - 128).astype(np.int8)),
    ("small", "INT16", None, lambda i: ((i % 65536) - 32768).astype(np.int16)),
    ("int", "INT32", None, lambda i: i.astype(np.int32)),
    ("big", "INT64", None, lambda i: i * 10),
    ("ubig", "UINT64", None, lambda i: (i * 10).astype(np.uint64)),
    ("flt", "FLOAT", None, lambda i: i.astype(np.float32) / np.float32(100)),
    ("dbl", "DOUBLE", None, lambda i: i / 1000.0),
    ("dec", "DECIMAL", None, lambda i: (i % 10_000_000) * 100, 10, 2),
    ("date", "DATE", None, lambda i: (8116 + (i % 10000)).astype(np.int32)),
    ("ts", "TIMESTAMP", None, lambda i: 1_735_689_600_000_000 + i * 1_000_000),
    # VARCHAR 'Value ' || (i % 1000): a true string column, 1000-key dictionary per row group (DICT_STR).  The scan emits
    # what the reference hands DuckDB for such a chunk: a DICTIONARY_VECTOR = the keys' string_t + dictionary bytes once
    # per row group and a u32 selection vector per row (dictionary_kernel.hpp:60-101)
    ("varchar", "VARCHAR", None, lambda i: value_strings(i % 1000)),
]

_VALUE_KEYS = [b"Value %d" % k for k in range(1000)]
_VALUE_LEN = np.array([len(k) for k in _VALUE_KEYS], dtype=np.int64)
_VALUE_MAT = np.zeros((1000, 9), dtype=np.uint8)
for _k, _v in enumerate(_VALUE_KEYS):
    _VALUE_MAT[_k, :len(_v)] = np.frombuffer(_v, dtype=np.uint8)


def value_strings(k):
    """(offsets, bytes) of 'Value <k>' for every element of k (vectorised: no Python string objects)."""
    lens = _VALUE_LEN[k]
    off = np.zeros(len(k) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:])
    mat = _VALUE_MAT[k]
    data = mat[np.arange(9)[None, :] < lens[:, None]]
    return off, np.ascontiguousarray(data)


def build_table_image(n_rows, threads, batch_rows=64 * 65536):
    """Encode the all_types table with the product writer (host C++), batch by batch so the raw columns never exist
    in full.  Returns the container image as a numpy uint8 array."""
    import flsgpu
    specs = []
    for c in COLUMNS:
        specs.append((c[0], c[1], c[2]) + tuple(c[4:6]))
    w = flsgpu.Writer(specs, 64)
    done = 0
    while done < n_rows:
        n = min(batch_rows, n_rows - done)
        i = np.arange(done + 1, done + n + 1, dtype=np.int64)
        w.append([c[3](i) for c in COLUMNS], threads)
        done += n
    img = w.finish()
    return w, img


META_DT = np.dtype([("base", "<u8"), ("data_off16", "<u4"), ("aux_off", "<u4"), ("aux_cnt", "<u2"), ("width", "u1"),
                    ("width2", "u1"), ("e", "u1"), ("f", "u1"), ("n_valid", "<u2"), ("reserved", "<u8")])
assert META_DT.itemsize == 32


def algorithmic_bytes(t, img, col):
    """SURVEY.md 8(d), per 1024-value vector: C = compressed bytes read, D = 1024*sizeof(T_out) written.
    FFOR-like: 128*W + sizeof(T) + 1; DELTA: + 128 (lane bases); DICT: 128*W_code + 1 (+ |dict|*sizeof(T) once per row
    group); RLE: 128*W_idx + 128 + 1 + n_runs*sizeof(T); ALP: 128*W + 8 + 1 + 2 + 2 + n_exc*(sizeof(T)+2);
    ALP_RD: 128*(W_r + W_l) + 2 + 16 + 2 + n_exc*4; CONSTANT: sizeof(T); UNCOMPRESSED: 1024*sizeof(T);
    + 4 B per vector for the entry point."""
    cd = t.column(col)
    sz = cd.phys_bits // 8
    read = 0
    written = 0
    if cd.phys_bits == 128:
        # dictionary form: packed codes (128*W + 1 + 4 per vector) + dictionary bytes and end offsets read; u32 selection
        # entries + the keys' string_t + dictionary bytes written
        for r in range(t.n_rowgroups):
            ch = t.chunk(r, col)
            m = np.frombuffer(img, dtype=META_DT, count=ch.n_vec, offset=ch.offset + ch.seg_off[0])
            read += int((128 * m["width"].astype(np.int64)).sum()) + ch.n_vec * 5 + ch.const_val + 4 * ch.aux_count
            written += ch.n_vec * 4096 + 16 * ch.aux_count + ch.const_val
        return read, written
    for r in range(t.n_rowgroups):
        ch = t.chunk(r, col)
        off = ch.offset + ch.seg_off[0]
        m = np.frombuffer(img, dtype=META_DT, count=ch.n_vec, offset=off)
        w = m["width"].astype(np.int64)
        w2 = m["width2"].astype(np.int64)
        nx = m["aux_cnt"].astype(np.int64)
        enc = ch.encoding
        nv = ch.n_vec
        if enc == 0:      # UNCOMPRESSED
            c = nv * 1024 * sz
        elif enc == 1:    # CONSTANT
            c = nv * sz
        elif enc == 3:    # DELTA
            c = int((128 * w).sum()) + nv * (128 + sz + 1)
        elif enc == 4:    # DICT
            c = int((128 * w).sum()) + nv * 1 + ch.aux_count * sz
        elif enc == 5:    # RLE
            c = int((128 * w).sum()) + nv * (128 + 1) + int(nx.sum()) * sz
        elif enc == 6:    # ALP
            c = int((128 * w).sum()) + nv * (8 + 1 + 2 + 2) + int(nx.sum()) * (sz + 2)
        elif enc == 7:    # ALP_RD
            c = int((128 * (w + w2)).sum()) + nv * (2 + 16 + 2) + int(nx.sum()) * 4
        elif enc == 8:    # CROSS_RLE
            c = ch.aux_count * (sz + 4)
        else:             # FFOR, FREQUENCY, SLPATCH
            c = int((128 * w).sum()) + nv * (sz + 1) + int(nx.sum()) * (sz + 2)
        read += c + 4 * nv
        written += nv * 1024 * sz
    return read, written


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region, polled through NVML every few ms (the timed region of a
    resident scan is tens of ms, shorter than nvidia-smi's start-up)."""

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.sm, self.reasons = [], set()
        self.stop_flag = threading.Event()
        self.thread = None
        self.max_mhz = None
        self.err = None

    def _run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
            while not self.stop_flag.is_set():
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for name, bit in bits.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.002)
        except Exception as e:  # NVML missing: report it, never fake a clock
            self.err = f"{type(e).__name__}: {e}"

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        time.sleep(0.05)
        self.sm.clear()

    def stop(self):
        self.stop_flag.set()
        self.thread.join(timeout=2)
        out = {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
               "reasons": sorted(self.reasons), "samples": len(self.sm)}
        if self.err:
            out["error"] = self.err
        return out


def bench_config(rows, n_cols):
    """The workload, and nothing but the workload: both arms print exactly this dict."""
    return {"workload": WORKLOAD, "rows_per_gpu": rows, "columns": n_cols, "rowgroup_rows": 65536,
            "sharding": "row groups, one table shard per rank",
            "l2": "inputs + outputs per step exceed L2 (126 MB); no explicit flush"}


def build_table_file(path, rows, threads):
    """The all_types container written to `path` by a child process (the reference arm must not load the product library)."""
    subprocess.check_call([sys.executable, os.path.abspath(__file__), "--impl", "build-table", "--rows", str(rows), "--out", path,
                           "--threads", str(threads)], stdout=subprocess.DEVNULL)


# ---- parity: order-sensitive checksums of a decoded column -------------------------------------------------------------
def words_of(arr):
    """numpy column -> int64 words per row: the physical value sign-extended (floats: their bits); (n, 16) uint8 string_t
    rows -> two words per row."""
    if arr.dtype == np.uint8 and arr.ndim == 2:
        return np.ascontiguousarray(arr).view(np.int64)
    kind = arr.dtype.kind
    if kind == "f":
        arr = arr.view(np.int32 if arr.dtype.itemsize == 4 else np.int64)
    elif kind == "u":
        arr = arr.view(np.dtype(f"i{arr.dtype.itemsize}"))
    return arr.astype(np.int64, copy=False).reshape(len(arr), 1)


def checksum_np(arr):
    """(sum of words, sum of words weighted by 1-based row number), both modulo 2^64."""
    w = words_of(arr).view(np.uint64)
    idx = np.arange(1, w.shape[0] + 1, dtype=np.uint64)
    s1 = s2 = 0
    for k in range(w.shape[1]):
        col = np.ascontiguousarray(w[:, k])
        s1 = (s1 + int(col.sum(dtype=np.uint64)) * (k + 1)) & (2 ** 64 - 1)
        s2 = (s2 + int((col * idx).sum(dtype=np.uint64)) * (k + 1)) & (2 ** 64 - 1)
    return s1, s2


def checksum_torch(words):
    """The same for an int64 tensor [n, k] on the device."""
    import torch
    idx = torch.arange(1, words.shape[0] + 1, dtype=torch.int64, device=words.device)
    s1 = s2 = 0
    for k in range(words.shape[1]):
        col = words[:, k]
        s1 = (s1 + (int(col.sum().item()) & (2 ** 64 - 1)) * (k + 1)) & (2 ** 64 - 1)
        s2 = (s2 + (int((col * idx).sum().item()) & (2 ** 64 - 1)) * (k + 1)) & (2 ** 64 - 1)
    return s1, s2


def oracle_checksums(img, t, n_cols, threads):
    """Every column of the table decoded by the CPU oracle (untimed) -> its checksum."""
    import oracle_lib
    o = oracle_lib.Oracle(img)
    out = []
    for c in range(n_cols):
        cd = t.column(c)
        if cd.phys_bits == 128:
            rows, _ = o.decode_strings(c, 0, t.n_rowgroups, t.vec_per_rowgroup, 0)
            lens = rows[:t.n_tuples].view(np.uint32)[:, 0]
            assert int(lens.max()) <= 12, "checksum of string_t rows assumes inlined strings"
            out.append(checksum_np(rows[:t.n_tuples]))
        else:
            dt = t.native_dtype(c)
            vals = o.decode_column(c, 0, t.n_rowgroups, 0, dt, t.vec_per_rowgroup, threads)
            out.append(checksum_np(vals[:t.n_tuples]))
    return out


def timed_cpu_scan(o, cols, n_rg, threads, target_seconds):
    """Repeat the bounded oracle scan until ~target_seconds of CPU work have been timed."""
    total_b, total_t, passes = 0, 0.0, 0
    while total_t < target_seconds and passes < 1000:
        t0 = time.perf_counter()
        nb, _ = o.scan(cols, 0, n_rg, threads)
        total_t += time.perf_counter() - t0
        total_b += nb
        passes += 1
    return total_b, total_t, passes


def cpu_scan(img, n_rowgroups, n_cols, threads, target_seconds):
    """Oracle scan (reference-shaped: 2048-row chunks, a row group per thread) over a bounded prefix of the row
    groups.  Returns (decoded GB/s, sample description, seconds per pass, rg count)."""
    import oracle_lib
    o = oracle_lib.Oracle(img)
    cols = list(range(n_cols))
    probe = min(n_rowgroups, max(threads * 2, 32))
    t0 = time.perf_counter()
    nb, _ = o.scan(cols, 0, probe, threads)
    dt = time.perf_counter() - t0
    rate = nb / dt
    per_rg = nb / probe
    n_rg = int(min(n_rowgroups, max(probe, target_seconds * rate / per_rg)))
    return o, cols, n_rg


def run_reference(args, rank, world):
    """--impl reference: the CPU path on the host cores (oracle port; the reference binary is unbuildable here)."""
    if rank != 0:
        return
    import tempfile
    threads = os.cpu_count() or 1
    # the input container comes from a child process (the writer is product code); this process loads the oracle only
    tmp = tempfile.mkdtemp(prefix="fls_bench_")
    path = os.path.join(tmp, "all_types.fls")
    build_table_file(path, args.rows, threads)
    img = np.fromfile(path, dtype=np.uint8)
    os.remove(path)
    os.rmdir(tmp)
    hdr = np.frombuffer(img[:64].tobytes(), dtype=np.dtype([("magic", "S8"), ("version", "<u4"), ("n_columns", "<u4"), ("n_rowgroups", "<u4"),
                                                           ("vec_per_rowgroup", "<u4"), ("n_tuples", "<u8"), ("footer_offset", "<u8"),
                                                           ("footer_size", "<u8"), ("reserved", "S16")]))[0]

    class _T:  # what the code below reads off the table
        n_rowgroups, n_columns = int(hdr["n_rowgroups"]), int(hdr["n_columns"])

        @staticmethod
        def close():
            pass
    t = _T
    o, cols, n_rg = cpu_scan(img, t.n_rowgroups, t.n_columns, threads, args.cpu_seconds / max(1, args.steps))
    for _ in range(args.warmup):
        o.scan(cols, 0, min(n_rg, 64), threads)
    total_b, total_t = 0, 0.0
    per_step = args.cpu_seconds / max(1, args.steps)
    passes = 0
    for _ in range(args.steps):
        b_, t_, p_ = timed_cpu_scan(o, cols, n_rg, threads, per_step)
        total_b += b_
        total_t += t_
        passes += p_
    gbps = total_b / total_t / 1e9
    sample = (f"{passes} passes over the first {n_rg} of {t.n_rowgroups} row groups x {t.n_columns} columns, "
              f"{total_t:.1f} s, oracle scan (gcc -O3 -march={__import__('oracle_lib').MARCH}), one row group per thread")
    line = {
        "impl": "reference", "metric": METRIC, "value": gbps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": bench_config(args.rows, t.n_columns),
        "note": "CPU restatement of the reference scan (oracle port, %d threads); the reference's own FastLanes library is not vendored" % threads,
        "cpu_baseline": {"value": gbps, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": gbps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def kernel_name(t, col):
    """Which kernel instantiation(s) decode this column: the launcher (flsgpu_device.cu decode_column_impl) issues one
    launch per family present among the column's chunks: light (integer encodings), 64-bit DELTA, 8-bit FFOR/DELTA (SWAR),
    ALP, ALP_RD (one launch for both when T = 32), DICT per code width, RLE.  A column whose chunks all fall in one family is one launch of one instantiation (the roofline groups those);
    a column with several families is reported as "mixed(...)"."""
    import flsgpu
    fams, encs = set(), set()
    for r in range(t.n_rowgroups):
        ch = t.chunk(r, col)
        e = flsgpu.ENC_NAME[ch.encoding]
        encs.add(e)
        fams.add("str_dict" if e in ("DICT_STR", "FSST_DICT") else "dict%d" % ch.idx_bits if e == "DICT" else "rle" if e == "RLE" else "alp" if e == "ALP" else "alprd" if e == "ALP_RD"
                 else "delta" if e == "DELTA" and ch.phys_bits == 64 else "swar8" if e in ("FFOR", "DELTA") and ch.phys_bits == 8 else "light")
    T = t.column(col).phys_bits
    fams = sorted(fams)  # "delta": 64-bit DELTA chunks have their own instantiation
    if fams == ["str_dict"]:
        # heap fill (+ keys' string_t in the same launch for plain DICT_STR dictionaries, else a keys launch) + codes
        return "fls_str_kernels<dictionary form>", sorted(encs), 2 if encs == {"DICT_STR"} else 3
    if len(fams) == 1:
        return f"fls_decode_kernel<{fams[0]},T={T}>", sorted(encs), 1
    fused = T == 32 and "alp" in fams and "alprd" in fams  # FLOAT: ALP and ALP_RD chunks share one launch
    return f"mixed({'+'.join(fams)}),T={T}", sorted(encs), len(fams) - (1 if fused else 0)


def run_gpu(args, rank, world, local_rank):
    import torch
    import flsgpu
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # NCCL would print its version line on stdout, next to the JSON line
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    flsgpu.init((local_rank,))
    threads = max(1, (os.cpu_count() or 1) // max(1, world))
    t_build = time.perf_counter()
    w, img = build_table_image(args.rows, threads)
    t = flsgpu.Table(img)  # zero-copy view of the writer's image; page-locked on first stage
    build_s = time.perf_counter() - t_build
    n_cols = t.n_columns
    stream = torch.cuda.current_stream().cuda_stream
    t.stage(stream=stream)
    torch.cuda.synchronize()

    # string columns are scanned in dictionary form: 4-byte selection entries per row (+ keys and dictionary bytes per row group)
    is_str = [t.column(c).phys_bits == 128 for c in range(n_cols)]
    widths = [4 if is_str[c] else t.column(c).phys_bits // 8 for c in range(n_cols)]
    out_elems = t.out_elems(t.n_rowgroups)
    outs = [torch.empty(out_elems * wd, dtype=torch.uint8, device="cuda") for wd in widths]
    decoded_bytes = sum(t.n_tuples * wd for wd in widths)  # real tuples only (padding of the last vector excluded)
    compressed_bytes = t.resident_bytes()
    str_keys, str_heap = {}, {}
    for c in range(n_cols):
        if is_str[c]:
            str_keys[c] = torch.empty(max(16, t.string_dict_keys(c) * 16), dtype=torch.uint8, device="cuda")
            str_heap[c] = torch.empty(max(16, t.string_heap_bytes(c)), dtype=torch.uint8, device="cuda")

    def scan_step(events=None):
        for c in range(n_cols):
            if events is not None:
                events[c][0].record()
            if is_str[c]:
                t.decode_string_dict(c, str_keys[c].data_ptr(), str_keys[c].numel(), str_heap[c].data_ptr(), str_heap[c].numel(),
                                     str_heap[c].data_ptr(), stream=stream)
                t.decode_string_codes(c, outs[c].data_ptr(), outs[c].numel(), stream=stream)
            else:
                t.decode_column(c, outs[c].data_ptr(), outs[c].numel(), stream=stream)
            if events is not None:
                events[c][1].record()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        scan_step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = flsgpu.launch_count()
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_cols)]
          for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(args.steps):
        scan_step()
    e1.record()
    barrier()
    launches = flsgpu.launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    # Per-launch durations (roofline, per_column): the same K steps once more with a CUDA event on either side of every
    # launch.  They are not taken in the timed pass because an event record between two launches is a stream operation of
    # its own: it would undo the programmatic dependent launch that lets column k+1 start as column k's CTAs retire
    # (measured on B200: 1.17 ms per step without, 1.30 ms with the events in place; 1.26 / 1.32 ms with FLSGPU_PDL=0).
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for s in range(args.steps):
        scan_step(ev[s])
    e3.record()
    torch.cuda.synchronize()
    instrumented_ms = e2.elapsed_time(e3) / args.steps
    col_ms = [float(np.mean([ev[s][c][0].elapsed_time(ev[s][c][1]) for s in range(args.steps)])) for c in range(n_cols)]

    # ---- parity of what was just timed: every decoded column still sitting in HBM against the CPU oracle's decode
    import oracle_lib
    t_par = time.perf_counter()
    want = oracle_checksums(img, t, n_cols, threads)
    key_base = {}
    parity_bytes = 0
    for c in range(n_cols):
        wd = widths[c]
        if is_str[c]:  # dictionary form -> the string_t of every row: keys of the row's row group, picked by its selection entry
            counts = torch.tensor([t.chunk(r, c).aux_count for r in range(t.n_rowgroups)], dtype=torch.int64)
            key_base[c] = (torch.cumsum(counts, 0) - counts).cuda()
            codes = outs[c].view(torch.int32)[:t.n_tuples].to(torch.int64)
            rg = torch.arange(t.n_tuples, device="cuda") // (t.vec_per_rowgroup * 1024)
            keys = str_keys[c][:t.string_dict_keys(c) * 16].view(torch.int64).view(-1, 2)
            words = keys[key_base[c][rg] + codes]
            parity_bytes += t.n_tuples * 16
        else:
            tdt = {1: torch.int8, 2: torch.int16, 4: torch.int32, 8: torch.int64}[wd]
            words = outs[c].view(tdt)[:t.n_tuples].to(torch.int64).view(-1, 1)
            parity_bytes += t.n_tuples * wd
        got = checksum_torch(words)
        assert got == want[c], f"resident scan: column {COLUMNS[c][0]} differs from the oracle decode ({got} != {want[c]})"
        del words

    # ---- end to end through the reader (FastLanesReader::TryInitializeScan / Scan on worker threads, what DuckDB's
    # MultiFileScan drives): container in host memory -> H2D of the compressed chunks -> decode -> D2H into pinned staging
    # buffers -> 2048-row chunks handed to the consumer, every step.  compressed_vectors: vectors the file marks CONSTANT /
    # SEQUENCE reach the consumer as such (no decode, no PCIe); off = every vector flat.
    import reader_lib
    import tempfile
    tmpdir = tempfile.mkdtemp(prefix="fls_bench_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    path = os.path.join(tmpdir, f"all_types_{rank}.fls")
    np.asarray(img).tofile(path)
    e2e_threads = int(os.environ.get("FLS_E2E_THREADS", 0)) or max(1, min(8, threads))
    e2e_batch = int(os.environ.get("FLS_E2E_BATCH", 0)) or 32
    e2e = {}
    for mode, on in (("compressed_vectors", 1), ("flat", 0)):
        rd = reader_lib.Reader(path, rowgroups_per_batch=e2e_batch)
        rd.set_option("compressed_vectors", on)
        times, stats = [], None
        for s in range(2 + args.steps):  # two untimed passes: pinned staging buffers and device slabs get allocated
            if os.environ.get("FLS_E2E_NOEVICT") != "1":  # experiments only: keep the compressed chunks resident
                rd.evict()
            barrier()
            if s == 2:
                x0 = flsgpu.transfer_bytes()
            t0 = time.perf_counter()
            stats = rd.consume(n_threads=e2e_threads)
            dt = time.perf_counter() - t0
            if s >= 2:
                times.append(dt)
        x1 = flsgpu.transfer_bytes()
        assert stats["rows"] == t.n_tuples
        e2e[mode] = {"s": float(np.sum(times)), "h2d": (x1[0] - x0[0]) // args.steps, "d2h": (x1[1] - x0[1]) // args.steps, "stats": stats}
        if on:  # untimed: the rows a consumer of this scan reads, every column, against the oracle
            raw = rd.scan_raw(n_threads=e2e_threads)
            for c in range(n_cols):
                if is_str[c]:
                    data, lens = raw[c]
                    rows16 = np.zeros((t.n_tuples, 16), dtype=np.uint8)
                    rows16[:, :4] = lens.view(np.uint8).reshape(-1, 4)
                    rows16[:, 4:] = data[:, :12]
                    got = checksum_np(rows16)
                    del rows16
                else:
                    got = checksum_np(raw[c].view(t.native_dtype(c)) if raw[c].dtype.itemsize == t.column(c).phys_bits // 8 else raw[c])
                assert got == want[c], f"end-to-end scan: column {COLUMNS[c][0]} differs from the oracle decode ({got} != {want[c]})"
            del raw
        rd.close()
    os.remove(path)
    os.rmdir(tmpdir)
    parity_s = time.perf_counter() - t_par
    e2e_s = e2e["compressed_vectors"]["s"]
    e2e_flat_s = e2e["flat"]["s"]
    # PCIe limit of this box, measured the plain way: one 1 GiB pinned<->device copy each direction
    probe_dev = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    probe_host = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
    pcie = {}
    for name, (dst, src) in {"d2h": (probe_host, probe_dev), "h2d": (probe_dev, probe_host)}.items():
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        dst.copy_(src, non_blocking=True)
        p1.record()
        torch.cuda.synchronize()
        pcie[name + "_GBps"] = round((1 << 30) / p0.elapsed_time(p1) / 1e6, 1)
    # the box's ceiling when every rank copies at once: all ranks start the same D2H copy behind a barrier
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(4):
        probe_host.copy_(probe_dev, non_blocking=True)
    p1.record()
    torch.cuda.synchronize()
    pcie["d2h_concurrent_GBps_per_gpu"] = round(4 * (1 << 30) / p0.elapsed_time(p1) / 1e6, 1)
    del probe_dev, probe_host

    # ---- reduce over ranks: time = max, bytes = sum; reduced scan result (COUNT, SUM(int)) via NCCL
    tm = torch.tensor([total_ms, e2e_s, e2e_flat_s], dtype=torch.float64, device="cuda")
    by = torch.tensor([float(decoded_bytes), float(t.n_tuples)], dtype=torch.float64, device="cuda")
    # reduced scan result: fused SUM/COUNT of the BIGINT column on every rank, combined with one NCCL all-reduce
    from flsgpu import sharding
    res = torch.zeros(64, dtype=torch.uint8, device="cuda")
    ea0, ea1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):  # warm-up: first launch of a kernel pays lazy module loading
        flsgpu.agg_reset(res.data_ptr(), stream)
        t.aggregate_column(3, res.data_ptr(), stream=stream)
    n_agg = 10
    ea0.record()
    for _ in range(n_agg):
        flsgpu.agg_reset(res.data_ptr(), stream)
        t.aggregate_column(3, res.data_ptr(), stream=stream)
    ea1.record()
    torch.cuda.synchronize()
    agg_ms = ea0.elapsed_time(ea1) / n_agg
    r_ = flsgpu.AggResult.from_buffer_copy(res.cpu().numpy().tobytes())
    s128 = (r_.sum_hi << 64) | r_.sum_lo
    s128 -= (1 << 128) if s128 >= (1 << 127) else 0
    assert s128 == 10 * args.rows * (args.rows + 1) // 2 and r_.count == args.rows, "fused aggregate returned a wrong SUM/COUNT"
    red_sum, red_count = sharding.all_reduce_partial(s128, r_.count, dist, device="cuda")
    # the reduced scan as ONE timed region: fused aggregates of the column on every rank + the NCCL all-reduce of the limbs
    # (the only collective of the design), max over ranks; the all-reduce alone is timed by events around it
    eb0, eb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ec = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_agg)]
    for it in range(2):  # first round warms NCCL's channels up
        barrier()
        eb0.record()
        for k in range(n_agg):
            flsgpu.agg_reset(res.data_ptr(), stream)
            t.aggregate_column(3, res.data_ptr(), stream=stream)
            limbs = sharding.limbs_of_results(res)
            ec[k][0].record()
            if dist is not None:
                dist.all_reduce(limbs, op=dist.ReduceOp.SUM)
            ec[k][1].record()
        eb1.record()
        torch.cuda.synchronize()
    agg_red_ms = eb0.elapsed_time(eb1) / n_agg
    allreduce_us = float(np.median([a.elapsed_time(b) for a, b in ec])) * 1e3
    got_sum, got_count = sharding.unpack_partial(limbs.cpu()[0])
    assert (got_sum, got_count) == (red_sum, red_count), "reduced scan: all-reduced limbs differ"
    tr = torch.tensor([agg_red_ms, allreduce_us], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tr, op=dist.ReduceOp.MAX)
    agg_red_ms, allreduce_us = tr.tolist()
    # the same kind of query end to end from pinned host memory: H2D of the compressed row groups -> fused SUM/COUNT/MIN/MAX of
    # every numeric column -> D2H of the 64-byte results (no decoded byte crosses PCIe or reaches HBM)
    num_cols = [c for c in range(n_cols) if not is_str[c]]
    res_all = torch.zeros(64 * len(num_cols), dtype=torch.uint8, device="cuda")
    res_host = torch.empty(64 * len(num_cols), dtype=torch.uint8).pin_memory()
    agg_e2e = []
    for s in range(4):
        t.evict()
        barrier()
        t0 = time.perf_counter()
        t.stage(stream=stream)
        for i, c in enumerate(num_cols):
            flsgpu.agg_reset(res_all.data_ptr() + 64 * i, stream)
            t.aggregate_column(c, res_all.data_ptr() + 64 * i, stream=stream)
        res_host.copy_(res_all, non_blocking=True)
        torch.cuda.synchronize()
        if s > 0:
            agg_e2e.append(time.perf_counter() - t0)
    agg_e2e_s = float(np.mean(agg_e2e))
    r3 = flsgpu.AggResult.from_buffer_copy(res_host.numpy()[64 * num_cols.index(3):64 * num_cols.index(3) + 64].tobytes())
    assert r3.count == args.rows and ((r3.sum_hi << 64) | r3.sum_lo) == (s128 % (1 << 128)), "end-to-end aggregate differs"
    agg_e2e_bytes = float(sum(t.n_tuples * widths[c] for c in num_cols))
    if dist is not None:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(by, op=dist.ReduceOp.SUM)
    total_ms, e2e_s, e2e_flat_s = tm.tolist()
    all_bytes, all_rows = by.tolist()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    value = all_bytes * args.steps / (total_ms / 1e3) / 1e9
    e2e_value = all_bytes * args.steps / e2e_s / 1e9
    e2e_flat_value = all_bytes * args.steps / e2e_flat_s / 1e9

    # ---- roofline of the dominant kernel template
    groups = {}
    per_column = []
    for c in range(n_cols):
        name, encs, n_launch = kernel_name(t, c)
        rd, wr = algorithmic_bytes(t, img, c)
        g = groups.setdefault(name, {"ms": 0.0, "bytes": 0, "launches": 0})
        g["ms"] += col_ms[c]
        g["bytes"] += rd + wr
        g["launches"] += n_launch
        per_column.append({"column": COLUMNS[c][0], "kernel": name, "launches": n_launch, "encodings": encs, "ms": round(col_ms[c], 4),
                           "alg_read_MB": round(rd / 1e6, 2), "alg_write_MB": round(wr / 1e6, 2),
                           "alg_GBps": round((rd + wr) / col_ms[c] / 1e6, 1)})
    # dominant kernel = the single-instantiation group with the largest share of the step
    dom = max((k for k in groups if not k.startswith("mixed")), key=lambda k: groups[k]["ms"])
    g = groups[dom]
    achieved = g["bytes"] / g["ms"] / 1e6  # GB/s
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
            peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        pass
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(dom)
    except (OSError, ValueError):
        pass
    roofline = {"bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                "launches_per_step": g["launches"], "alg_bytes_per_launch": g["bytes"] // g["launches"],
                "avg_launch_ms": round(g["ms"] / g["launches"], 4), "share_of_step": round(g["ms"] / sum(col_ms), 3),
                "timing": "CUDA events around every launch in a second pass of the same K steps right after the timed pass "
                          "(events between launches would serialise the timed pass's programmatic dependent launches); "
                          f"that pass ran at {instrumented_ms:.4f} ms per step"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        o, cols, n_rg = cpu_scan(img, t.n_rowgroups, n_cols, cores, args.cpu_seconds)
        nb, dt, passes = timed_cpu_scan(o, cols, n_rg, cores, args.cpu_seconds)
        cpu = {"value": round(nb / dt / 1e9, 3), "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{passes} passes over the first {n_rg} of {t.n_rowgroups} row groups x {n_cols} columns, {dt:.1f} s, oracle scan (gcc -O3 -march={oracle_lib.MARCH}), one row group per thread"}

    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": bench_config(args.rows, n_cols),
        "details": {"row_groups_per_gpu": t.n_rowgroups, "compressed_MB_per_gpu": round(compressed_bytes / 1e6, 1),
                    "decoded_MB_per_gpu": round(decoded_bytes / 1e6, 1), "table_build_s": round(build_s, 1),
                    "parity_check_s": round(parity_s, 1)},
        "parity_checked": f"all {n_cols} columns x {t.n_tuples} rows, {parity_bytes} decoded bytes: resident scan (checksums computed on the "
                          "device) and the end-to-end reader scan (rows as a consumer reads them) == CPU oracle decode, order-sensitive "
                          "64-bit checksums, untimed",
        "values_per_s": round(all_rows * n_cols * args.steps / (total_ms / 1e3), 1),
        "e2e": {"value": round(e2e_value, 3), "unit": UNIT, "h2d_bytes_per_step": int(e2e["compressed_vectors"]["h2d"]),
                "d2h_bytes_per_step": int(e2e["compressed_vectors"]["d2h"]), "ms_per_step": round(e2e_s / args.steps * 1e3, 2),
                "path": "FastLanesReader::TryInitializeScan/Scan on %d worker threads (libflsreader), 2048-row chunks handed to a consumer "
                        "that takes the vectors as they are; vectors the file marks CONSTANT / arithmetic SEQUENCE are handed on in that "
                        "form (not decoded, not copied), everything else is decoded on the device and crosses PCIe" % e2e_threads,
                "vectors": {k: e2e["compressed_vectors"]["stats"][k] for k in ("chunks", "constant_vectors", "sequence_vectors",
                                                                              "dictionary_vectors", "assembled_vectors")},
                "flat": {"value": round(e2e_flat_value, 3), "unit": UNIT, "h2d_bytes_per_step": int(e2e["flat"]["h2d"]),
                         "d2h_bytes_per_step": int(e2e["flat"]["d2h"]), "ms_per_step": round(e2e_flat_s / args.steps * 1e3, 2),
                         "note": "the same scan with every vector decoded and copied (compressed_vectors off)"},
                "pcie_measured": pcie},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
        "per_column": per_column, "reduced_scan": {"query": "SELECT SUM(big), COUNT(*) over all shards (fused decode+aggregate, NCCL all-reduce of limbs)",
                                                     "count": red_count, "sum": str(red_sum), "fused_agg_ms_per_gpu": round(agg_ms, 4),
                                                     "fused_agg_values_per_s": round(args.rows / (agg_ms / 1e3), 1),
                                                     "fused_agg_plus_allreduce_ms": round(agg_red_ms, 4), "nccl_allreduce_us": round(allreduce_us, 1),
                                                     "e2e_all_numeric_columns": {
                                                         "query": "SUM/COUNT/MIN/MAX of the %d numeric columns from pinned host memory (H2D compressed, fused aggregates, D2H 64 B per column)" % len(num_cols),
                                                         "ms_per_gpu": round(agg_e2e_s * 1e3, 2),
                                                         "decoded_equivalent_GBps_per_gpu": round(agg_e2e_bytes / agg_e2e_s / 1e9, 1)}},
    }
    emit(line)
    t.close()
    w.close()
    if dist is not None:
        dist.destroy_process_group()


_JSON_FD = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Native libraries write there too (NCCL's version banner, whatever the box's
    NCCL_DEBUG says), so the real stdout is kept on a private descriptor for the JSON line and fd 1 is pointed at stderr."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        while data:
            data = data[os.write(_JSON_FD, data):]


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flsgpu", choices=["flsgpu", "reference", "build-table"])
    ap.add_argument("--out", default=None, help="--impl build-table: where to write the container")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--rows", type=int, default=100_000_000, help="rows of the all_types table per GPU")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU baseline sample budget")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("FLS_ORACLE_NATIVE", "1")  # timed CPU baselines: the oracle compiled for this box's cores
    if args.impl == "build-table":
        w, img = build_table_image(args.rows, args.threads or os.cpu_count() or 1)
        np.asarray(img).tofile(args.out)
        w.close()
    elif args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
