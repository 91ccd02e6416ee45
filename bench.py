#!/usr/bin/env python3
"""bench.py -- cudasp_scan rows/s on N B200s (one process per GPU), with roofline and CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--rows R] [--labels L]

Workload (BASELINE.json configs[1]): 1 M-row synthetic tweak table (1-week shape), no labels,
batch_size = 300000, per GPU (weak scaling: every rank scans its own table, no collective on the
data path).  A "step" is one pass of the whole table through the hot path.
  value  rows/s with the table already resident in HBM (cudasp_scan_device on torch's current stream,
         CUDA events, L2 flushed between steps, max over ranks)
  e2e    rows/s through the public host call cudasp_scan_batch with HOST buffers: memcpy into pinned
         staging, async H2D, kernels, D2H of the match ids -- all inside the timed region
  roofline  integer-multiply bound: achieved IMAD32 lane-ops/s of the dominant kernel
         (k_ladder) = rows * 168786 / its CUDA-event time, against the IMAD32 peak measured by
         tools/imad_peak on this box (fallback: profiles/r01n_imad_peak.json)
  cpu_baseline  the oracle's C port of the reference algorithm on the host cores, bounded sample (the fastest correct CPU scan
         in the repo); cpu_openssl = baseline/cpu_scan_openssl.cpp (OpenSSL 3 EC_POINT_mul, BASELINE.md B-cpu) beside it
  python_baseline  the reference's own CPU code for the path is a pure-Python verifier: oracle/ref_scan.py, 1 core, ~4 s
  one_handle (N > 1, rank 0 alone after the ranks' timed region)  ONE query handle over all N devices, ONE host table of
         N x rows rows, staged by several host threads through the C ABI -- the in-process form of the reference's
         thread->GPU round-robin (src/cudasp_extension.cu:516-532), the case its README names as its scaling wall
  duckdb  the same through the real DuckDB engine + the `cudasp` extension from a Parquet file (duckdb_ext/bin/duckdb):
         N = 1: 10 M rows with the bench's label count; N > 1: config C5's shape (12.5 M rows per GPU, 16 label keys,
         multi-output rows; 100 M rows at N = 8), all GPUs of the box in one SQL query
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

W32_BASE, W32_PER_LABEL = 231526, 674          # IMAD32 lane-ops per row
W32_LADDER = 2 * (666 * 73 + 795 * 45)         # stage A alone = 168786
SEED = 352002          # C2 352002 (L=0), C3 352003 (L=1), C4 352004 (L=8)
REF_SAMPLE_ROWS = 1200000   # reference arm: rows per GPU and step (4 batches), a bounded sample of the workload


def table_params(L):
    """(seed, planted base matches, planted matches per label) of the BASELINE config with L labels."""
    return {0: (352002, 100, 0), 1: (352003, 100, 100), 8: (352004, 100, 50)}.get(L, (SEED + L, 100, 50))


def metric_name(L):
    return "cudasp_scan rows/sec (%d label%s)" % (L, "" if L == 1 else "s")     # BASELINE.json: "cudasp_scan rows/sec (1 label)"


def shape_name(rows):
    return {1000000: "1-week", 9400000: "8-week", 32700000: "32-week"}.get(rows, "custom")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    # default workload = BASELINE.json configs[2] (C3): the configuration the metric "rows/sec (1 label)" is quoted on
    ap.add_argument("--rows", type=int, default=9400000)
    ap.add_argument("--labels", type=int, default=1)
    ap.add_argument("--batch", type=int, default=300000)
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--host-threads", type=int, default=0, help="staging threads of the e2e call (0 = the library's choice)")
    ap.add_argument("--no-legs", action="store_true", help="skip the one_handle and duckdb legs (quick runs)")
    ap.add_argument("--duckdb-rows", type=int, default=0, help="rows of the duckdb leg's Parquet file (0 = 10 M at N = 1, 12.5 M per GPU at N > 1)")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.stop_flag, self.proc = gpu, [], False, None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.samples.append([x.strip() for x in line.split(",")])
                if self.stop_flag:
                    break
        except Exception:
            pass

    def finish(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = max([int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm)}


def measured_imad_peak():
    """(IMAD32 lane-ops/s, source, IMAD.WIDE products/s): tools/imad_peak on this GPU if the binary travelled, else the committed profile."""
    exe = os.path.join(ROOT, "tools", "imad_peak")
    try:
        j = json.loads(subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout)
        src = "measured live by tools/imad_peak (mad.lo.u32; wide: 64-bit multiply-add, distinct operands)"
    except Exception:
        with open(os.path.join(ROOT, "profiles", "r01n_imad_peak.json")) as f:
            j = json.load(f)
        src = "profiles/r01n_imad_peak.json (tools/imad_peak, earlier call)"
    return float(j["imad_lo"]["lane_ops_per_s"]), src, float(j["imad_wide_distinct_operands"]["lane_ops_per_s"])


# 64-bit multiply-adds (IMAD.WIDE) K1 EXECUTES per row: 73 per multiplication (64 + 9 fold), 44 per squaring (36 + 8), so
# 3M + 4S = 395 per doubling, 8M + 3S = 716 per mixed addition, 4 106 for the odd-multiples table and the final Z scale
# (rowfn.cuh ladder_row); the op counts are the query's own (cudasp_query_info).  ncu's count for the C3 key: 86 988
# (profiles/traffic.json, profiles/r02a_ncu_summary.md).
def ladder_wide_per_row(info):
    return 4106 + 395 * info["ladder_dbl"] + 716 * info["ladder_add"]


def cpu_sample(T, rows_sample, threads):
    from oracle import c_oracle as CO
    n = min(rows_sample, T["tweaks"].shape[0])
    off = T["offsets"][: n + 1]
    t0 = time.perf_counter()
    flags, _ = CO.scan(T["sk"], T["spend"], T["labels"], T["tweaks"][:n], T["outputs"], off, threads=threads)
    dt = time.perf_counter() - t0
    return n / dt, n, dt, flags


def gecc_reference_run(T, rows, batch, n_gpus, steps, warmup):
    """rows/s of the unmodified reference GPU pipeline; one host thread per GPU (the reference maps DuckDB
    worker threads round-robin onto devices, src/cudasp_extension.cu:516-532), contiguous row ranges."""
    import ctypes as C
    from oracle import gecc_ref as GR
    GR.load()
    rt = GR._rt
    rt.cudaSetDevice.argtypes = [C.c_int]
    rt.cudaSetDevice.restype = C.c_int
    share = (rows + n_gpus - 1) // n_gpus
    flags_all = np.zeros(rows, dtype=np.uint8)

    def worker(g):
        rt.cudaSetDevice(g)
        lo, hi = g * share, min(rows, (g + 1) * share)
        for base in range(lo, hi, batch):
            cnt = min(batch, hi - base)
            f, _ = GR.scan_batch(T["sk"], T["spend"], T["labels"], T["tweaks"][base: base + cnt], T["outputs"],
                                 T["offsets"][base: base + cnt + 1])
            flags_all[base: base + cnt] = f

    def one_pass():
        th = [threading.Thread(target=worker, args=(g,)) for g in range(n_gpus)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        return time.perf_counter() - t0

    for _ in range(warmup):
        one_pass()
    ts = [one_pass() for _ in range(steps)]
    assert np.array_equal(np.nonzero(flags_all)[0].astype(np.uint32), T["expected"]), "reference flags != planted matches"
    return rows / (sum(ts) / len(ts)), 1e3 * sum(ts) / len(ts)


def reference_arm(args, config, cores, synth):
    L = args.labels
    from oracle import gecc_ref as GR
    use_gecc = GR.available()
    if use_gecc:
        try:
            import torch
            use_gecc = torch.cuda.is_available() and torch.cuda.device_count() >= args.gpus
        except Exception:
            use_gecc = False
    base = {"impl": "reference", "metric": metric_name(L), "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic",
            "config": config}
    # CPU port, bounded sample (always reported)
    sample = cores * 6000
    seed, n_base, per_label = table_params(L)
    Tc = synth.make_table(max(sample, 4000), seed, n_labels=L, n_base=20, n_per_label=2 if L else 0, batch_rows=args.batch)
    cpu_sample(Tc, min(sample, 1000), cores)
    cpu_rate, n, dt, flags = cpu_sample(Tc, sample, cores)
    assert np.array_equal(np.nonzero(flags)[0], Tc["expected"][Tc["expected"] < n]), "oracle disagrees with planted matches"
    cpu_port = {"value": cpu_rate, "unit": "rows/s", "cores": cores, "kind": "port",
                "sample": "%d rows, oracle/cudasp_oracle.c, %d pthreads, %.1f s" % (n, cores, dt)}
    ref_exe = os.path.join(ROOT, "oracle", "_ref", "duckdb_ref")
    gpu_ok = False
    try:
        import torch
        gpu_ok = torch.cuda.is_available() and torch.cuda.device_count() >= args.gpus
    except Exception:
        pass
    if gpu_ok and os.path.exists(ref_exe) and not os.environ.get("CUDASP_REFERENCE_ARM_FFI"):
        # First choice: the reference's own DuckDB extension through its own SQL surface, all N GPUs from ONE DuckDB process (its
        # own multi-GPU mode).  One table of rows_per_gpu rows (the full C3 table at N = 1; a bounded sample of the N x larger
        # weak-scaling workload at N > 1 -- the reference needs ~4.5 s for it on one B200).
        rows = args.rows
        T = synth.make_table(rows, seed, n_labels=L, n_base=n_base, n_per_label=per_label, batch_rows=args.batch)
        os.environ["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(args.gpus))
        value, ms, ts = reference_sql_run(ref_exe, synth, T, rows, L, args.batch, args.steps, max(args.warmup, 1))
        base.update({"value": value, "ms_per_step": ms, "dtype": "u32",
                     "cpu_baseline": {"value": value, "unit": "rows/s", "cores": cores, "kind": "reference",
                                      "sample": "the UNMODIFIED reference DuckDB extension (src/cudasp_extension.cu + src/cudasp_gpu.cu + gECC, compiled for "
                                                "sm_100 into oracle/_ref/duckdb_ref by `make -C oracle ref_ext`) running its README's SQL on a %d-row Parquet "
                                                "table (%d labels, batch_size %d) with %d DuckDB threads on %d B200; match count equal to the planted matches; "
                                                "the reference has no CPU implementation of this path" % (rows, L, args.batch, cores, args.gpus)},
                     "cpu_port": cpu_port, "step_seconds": ts,
                     "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    elif use_gecc:
        per_gpu = min(args.rows, REF_SAMPLE_ROWS)   # bounded sample: the reference needs ~1.3 s per million rows
        rows = per_gpu * args.gpus                  # weak scaling: the same sample on each of N GPUs
        T = synth.make_table(rows, seed, n_labels=L, n_base=n_base, n_per_label=per_label, batch_rows=args.batch)
        value, ms = gecc_reference_run(T, rows, args.batch, args.gpus, args.steps, max(args.warmup, 1))
        base.update({"value": value, "ms_per_step": ms, "dtype": "u32",
                     "cpu_baseline": {"value": value, "unit": "rows/s", "cores": args.gpus, "kind": "reference",
                                      "sample": "the UNMODIFIED reference GPU pipeline (src/cudasp_gpu.cu + gECC, compiled for sm_100 into "
                                                "oracle/_ref by oracle/Makefile) on %d B200, a %d-row sample of the workload per step (%d per GPU, %d labels), "
                                                "%d host thread(s); the reference has no CPU implementation of this path"
                                                % (args.gpus, rows, per_gpu, L, args.gpus)},
                     "cpu_port": cpu_port,
                     "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    else:
        base.update({"value": cpu_rate, "ms_per_step": 1e3 * dt, "dtype": "u64", "cpu_baseline": cpu_port,
                     "e2e": {"value": cpu_rate, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                     "note": "oracle/_ref (the compiled reference GPU pipeline) or a GPU is not available: CPU port timed instead"})
    return base


def cpu_openssl_sample(T, rows_sample, threads):
    """baseline B-cpu: OpenSSL 3 EC_POINT_mul / EC_POINT_add / SHA256, all host cores, bounded sample."""
    try:
        from baseline import cpu_scan
        if not cpu_scan.available():
            return {"unavailable": "OpenSSL development files not present"}
        n = min(rows_sample, T["tweaks"].shape[0])
        t0 = time.perf_counter()
        flags, _ = cpu_scan.scan(T["sk"], T["spend"], T["labels"], T["tweaks"][:n], T["outputs"], T["offsets"][: n + 1], threads=threads)
        dt = time.perf_counter() - t0
        ok = bool(np.array_equal(np.nonzero(flags)[0].astype(np.uint32), T["expected"][T["expected"] < n]))
        return {"value": n / dt, "unit": "rows/s", "cores": threads, "rows_per_s_per_core": n / dt / threads, "kind": "port",
                "library": "OpenSSL 3 (EC_POINT_mul, EC_POINT_add, SHA256) -- baseline/cpu_scan_openssl.cpp", "flags_equal_planted": ok,
                "sample": "%d rows of the same table, %d threads, %.1f s" % (n, threads, dt)}
    except Exception as ex:      # noqa: BLE001  (the baseline is optional; the product line must still print)
        return {"unavailable": repr(ex)[:200]}


def sql_blob(b):
    return "'" + "".join("\\x%02x" % c for c in b) + "'::BLOB"


def write_parquet_table(synth, rows_file, labels, geom_p, batch, table=None):
    """A synthetic table (tests/synth.py: valid curve points, planted matches) as an uncompressed Parquet file with the reference's
    four columns.  -> (table dict, path, directory, seconds spent)"""
    import tempfile
    import pyarrow as pa
    import pyarrow.parquet as pq
    t0 = time.perf_counter()
    T = table
    if T is None:
        seed = 352005 if labels == 16 else 352003
        T = synth.make_table(rows_file, seed, n_labels=labels, n_base=1000 if labels == 16 else 100, n_per_label=100 if labels else 0,
                             batch_rows=batch, geom_p=geom_p, max_extra=63 if geom_p < 0.5 else 15)
    off = T["offsets"].astype(np.int64)
    tw = pa.FixedSizeBinaryArray.from_buffers(pa.binary(64), rows_file, [None, pa.py_buffer(T["tweaks"].tobytes())]).cast(pa.binary())
    txid = np.zeros((rows_file, 32), np.uint8)
    txid[:, :8] = np.arange(rows_file, dtype=np.uint64).view(np.uint8).reshape(rows_file, 8)
    tx = pa.FixedSizeBinaryArray.from_buffers(pa.binary(32), rows_file, [None, pa.py_buffer(txid.tobytes())]).cast(pa.binary())
    outs = pa.ListArray.from_arrays(pa.array(off.astype(np.int32), pa.int32()), pa.array(T["outputs"], pa.int64()))
    tbl = pa.table({"txid": tx, "height": pa.array((800000 + np.arange(rows_file) // 3000).astype(np.int32)), "tweak_key": tw, "outputs": outs})
    d = tempfile.mkdtemp(prefix="cudasp_bench_")
    path = os.path.join(d, "t.parquet")
    pq.write_table(tbl, path, row_group_size=122880, compression="NONE")
    return T, path, d, time.perf_counter() - t0


def reference_sql_run(exe, synth, T, rows, L, batch, steps, warmup):
    """The reference's OWN DuckDB extension (oracle/_ref/duckdb_ref: src/cudasp_extension.cu + src/cudasp_gpu.cu + gECC, unmodified,
    in the vendored DuckDB v1.4.1) through its own public API: the SQL of its README (README.md:100-124) on a Parquet file of the
    table, PRAGMA threads = host cores; it spreads its worker threads over every visible GPU itself (src/cudasp_extension.cu:516-532).
    -> (rows/s, ms per step, seconds per step list)"""
    _, path, d, _ = write_parquet_table(synth, rows, L, 0.5, batch, table=T)
    try:
        lab = "[" + ", ".join(sql_blob(l) for l in T["labels"]) + "]" + ("" if T["labels"] else "::BLOB[]")
        scan = "SELECT count(*) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM read_parquet('%s')), %s, %s, %s, batch_size := %d);" % (
            path, sql_blob(T["sk"]), sql_blob(T["spend"]), lab, batch)
        sql = "PRAGMA threads=%d;\n.timer on\n" % (os.cpu_count() or 1) + (scan + "\n") * (warmup + steps)
        p = subprocess.run([exe, "-csv", "-noheader"], input=sql, capture_output=True, text=True, timeout=3000)
        lines = [l for l in p.stdout.strip().splitlines() if l]
        times = [float(l.split("real")[1].split()[0]) for l in lines if l.startswith("Run Time")]
        counts = [int(l) for l in lines if not l.startswith("Run Time")]
        assert p.returncode == 0 and len(times) == warmup + steps, (p.returncode, p.stderr[-400:])
        assert all(c == len(T["expected"]) for c in counts), "the reference's match count differs from the planted matches: %r" % counts[:3]
        ts = times[warmup:]
        return rows / (sum(ts) / len(ts)), 1e3 * sum(ts) / len(ts), ts
    finally:
        try:
            os.remove(path)
            os.rmdir(d)
        except OSError:
            pass


def one_handle_leg(M, T, rows, n_dev, batch, steps):
    """ONE query handle over n_dev devices, ONE host table of n_dev x rows rows (rank 0's table repeated), several host staging
    threads through cudasp_scan_batch_mt.  Rank 0 alone; the other ranks have left."""
    cores = os.cpu_count() or 1
    threads = max(2, min(4 * n_dev, cores // 2))
    tw = np.tile(T["tweaks"], (n_dev, 1))
    outs = np.tile(T["outputs"], n_dev)
    n_out = int(T["offsets"][-1])
    off = np.concatenate([T["offsets"][:-1].astype(np.uint64) + np.uint64(k * n_out) for k in range(n_dev)] + [np.array([n_dev * n_out], np.uint64)])
    assert int(off[-1]) < 2 ** 32
    off = off.astype(np.uint32)
    expected = np.concatenate([T["expected"].astype(np.uint64) + np.uint64(k * rows) for k in range(n_dev)]).astype(np.uint32)
    q = M.Query(T["sk"], T["spend"], T["labels"], devices=list(range(n_dev)), batch_rows=batch, max_outputs_per_batch=17 * batch)
    try:
        ok = True
        for _ in range(2):
            ok &= bool(np.array_equal(q.scan_batch(tw, outs, off, host_threads=threads), expected))
        h0, d0 = q.transfer_bytes()
        t0 = time.perf_counter()
        for _ in range(steps):
            ids = q.scan_batch(tw, outs, off, host_threads=threads)
        dt = (time.perf_counter() - t0) / steps
        ok &= bool(np.array_equal(ids, expected))
        h1, d1 = q.transfer_bytes()
        return {"value": n_dev * rows / dt, "unit": "rows/s", "ms_per_step": 1e3 * dt, "rows": n_dev * rows, "devices": n_dev,
                "host_threads": threads, "host_cores": cores, "steps": steps, "match_set_ok": ok,
                "h2d_bytes_per_step": (h1 - h0) // steps, "d2h_bytes_per_step": (d1 - d0) // steps,
                "api": "one cudasp_query over %d devices, cudasp_scan_batch_mt from ONE host table, %d staging threads" % (n_dev, threads)}
    finally:
        q.close()


def duckdb_leg(synth, rows_file, replicate, labels, geom_p, batch, n_dev):
    """SELECT count(*) FROM cudasp_scan((SELECT ... FROM read_parquet(...)), ...) through duckdb_ext/bin/duckdb: all visible GPUs in
    one query, Parquet decode + AccumulateInput + staging + kernels + result inside the timed query.  The file is read `replicate`
    times by the one query (config C5: 100 M rows, 16 label keys, multi-output rows on 8 GPUs)."""
    exe = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
    if not os.path.exists(exe):
        return {"unavailable": "duckdb_ext/bin/duckdb not built (tools/build_duckdb.sh needs the vendored DuckDB source)"}
    T, path, d, gen_s = write_parquet_table(synth, rows_file, labels, geom_p, batch)

    blob = sql_blob
    lab = "[" + ", ".join(blob(l) for l in T["labels"]) + "]" + ("" if T["labels"] else "::BLOB[]")
    src = "read_parquet([%s])" % ", ".join(["'%s'" % path] * replicate)
    threads = os.cpu_count() or 1
    scan = "SELECT count(*) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM %s), %s, %s, %s, batch_size := %d);" % (
        src, blob(T["sk"]), blob(T["spend"]), lab, batch)
    plain = "SELECT count(*), sum(height), sum(octet_length(tweak_key)), sum(octet_length(txid)), sum(list_sum(outputs)) FROM %s;" % src
    # second form: the reference README's own example (README.md:100-110) first loads the Parquet file into a DuckDB table
    # (CREATE TABLE tweak AS SELECT ... FROM read_parquet) and scans THAT; the load is outside the timed queries
    scan_t = scan.replace(src, "tweak")
    sql = ("PRAGMA threads=%d;\n.timer on\n%s\n%s\n%s\n%s\n.timer off\nCREATE TABLE tweak AS SELECT txid, height, tweak_key, outputs FROM %s;\n"
           ".timer on\n%s\n%s\n%s\n" % (threads, scan, scan, scan, plain, src, scan_t, scan_t, plain.replace(src, "tweak")))
    gen_s = round(gen_s, 1)
    try:
        p = subprocess.run([exe, "-csv", "-noheader"], input=sql, capture_output=True, text=True, timeout=1200)
        lines = [l for l in p.stdout.strip().splitlines() if l]
        times = [float(l.split("real")[1].split()[0]) for l in lines if l.startswith("Run Time")]
        vals = [l for l in lines if not l.startswith("Run Time")]
        want = len(T["expected"]) * replicate
        counts = [int(v) for v in vals if "," not in v]
        ok = p.returncode == 0 and len(counts) >= 5 and all(c == want for c in counts[:5]) and len(times) >= 7
        if not ok:
            return {"unavailable": "duckdb query failed: rc %d, %s" % (p.returncode, (p.stderr or p.stdout)[-300:])}
        warm = min(times[1:3])
        warm_t = min(times[4:6])
        rows = rows_file * replicate
        # third form: the library's own page decoders (include/cudasp_parquet.h) behind the table function
        # cudasp_scan_parquet -- the same files, no read_parquet.  Its own process: a failure here leaves the two forms above standing.
        native = {"unavailable": "not run"}
        try:
            nq = "SELECT count(*) FROM cudasp_scan_parquet([%s], %s, %s, %s, batch_size := %d);" % (
                ", ".join(["'%s'" % path] * replicate), blob(T["sk"]), blob(T["spend"]), lab, batch)
            pn = subprocess.run([exe, "-csv", "-noheader"], input="PRAGMA threads=%d;\n.timer on\n%s\n%s\n%s\n" % (threads, nq, nq, nq),
                                capture_output=True, text=True, timeout=600)
            nl = [l for l in pn.stdout.strip().splitlines() if l]
            nt = [float(l.split("real")[1].split()[0]) for l in nl if l.startswith("Run Time")]
            nc = [int(v) for v in nl if not v.startswith("Run Time") and "," not in v]
            if pn.returncode == 0 and len(nt) >= 3 and len(nc) >= 3 and all(c == want for c in nc[:3]):
                native = {"value": rows / min(nt[1:3]), "unit": "rows/s", "source": "cudasp_scan_parquet(paths, ...): pages decoded by the library, staged into the pinned slots",
                          "scan_s_warm": min(nt[1:3]), "scan_s_first": nt[0], "match_count_ok": True}
            else:
                native = {"unavailable": "query failed: rc %d, %s" % (pn.returncode, (pn.stderr or pn.stdout)[-200:])}
        except Exception as e:  # noqa: BLE001 -- a reported leg, never the bench's exit status
            native = {"unavailable": "%s: %s" % (type(e).__name__, e)}
        return {"value": rows / warm_t, "unit": "rows/s", "source": "DuckDB table loaded from the Parquet file (the reference README's example)",
                "scan_s_warm": warm_t, "plain_read_all_columns_s": times[6],
                "from_parquet": {"value": rows / warm, "unit": "rows/s", "source": "read_parquet() inside the query", "scan_s_warm": warm,
                                 "scan_s_first": times[0], "plain_read_all_columns_s": times[3]},
                "from_parquet_native": native,
                "rows": rows, "labels": labels, "mean_outputs_per_row": float(len(T["outputs"])) / rows_file,
                "devices": n_dev, "duckdb_threads": threads, "matches_expected": want, "match_count_ok": True, "file_rows": rows_file, "read_times": replicate, "table_gen_s": gen_s,
                "query": "SELECT count(*) FROM cudasp_scan((SELECT txid,height,tweak_key,outputs FROM read_parquet(...)), sk, spend, labels, batch_size := %d)" % batch}
    finally:
        try:
            os.remove(path)
            os.rmdir(d)
        except OSError:
            pass


_REAL_STDOUT = None


def emit(line):
    """The ONE JSON line of the contract goes to the process's real stdout; everything else libraries print to fd 1
    during the run (NCCL's version banner, for one) has been sent to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    L = args.labels
    w32 = W32_BASE + W32_PER_LABEL * L
    seed, n_base, per_label = table_params(L)
    workload = "%d-row synthetic tweak table (%s shape), %d label key(s), planted matches, batch_size=%d, per GPU" % (
        args.rows, shape_name(args.rows), L, args.batch)
    config = {"workload": workload, "rows_per_gpu": args.rows, "labels": L, "batch_size": args.batch, "seed": seed,
              "l2": "flushed (256 MiB write) between timed steps; inputs per step (%.0f MB) also exceed L2" % (args.rows * 84e-6),
              "planted_matches": n_base + per_label * L}
    from tests import synth

    # ------------------------------------------------------------------ reference arm
    # The reference has no CPU implementation of the path (its only CPU code is the one-vector Python
    # verifier scripts).  When oracle/_ref/libcudasp_gecc_ref.so exists -- the UNMODIFIED reference GPU
    # pipeline (src/cudasp_gpu.cu + gECC) compiled for sm_100 by `make -C oracle ref` -- and a GPU is
    # visible, this arm times THAT, driven exactly like the reference's ProcessBatch (kind "reference").
    # Otherwise it times the CPU port oracle/cudasp_oracle.c on all host cores (kind "port").
    if args.impl == "reference":
        if rank != 0:
            return
        cores = os.cpu_count() or 1
        line = reference_arm(args, config, cores, synth)
        emit(line)
        return

    # ------------------------------------------------------------------ our arm (GPU)
    import torch
    import cudasp_b200_loader
    M = cudasp_b200_loader.load()
    from importlib import import_module
    S = import_module("cudasp_b200.host.sharding")      # batch ranges, CSR rebasing, max-over-ranks (the N > 1 host logic)
    if not torch.cuda.is_available() or M.lib().cudasp_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank
    torch.cuda.set_device(dev)
    T = synth.make_table(args.rows, seed + 1000 * rank, n_labels=L, n_base=n_base, n_per_label=per_label, batch_rows=args.batch)
    rows = args.rows
    q = M.Query(T["sk"], T["spend"], T["labels"], devices=[dev], batch_rows=args.batch,
                max_outputs_per_batch=int(np.diff(T["offsets"][:: args.batch]).max(initial=0)) + 17 * args.batch)
    batches = []
    d_tw = torch.from_numpy(T["tweaks"]).cuda()
    d_out = torch.from_numpy(T["outputs"]).cuda()
    for base, end in S.batch_ranges(0, rows, args.batch):
        cnt = end - base
        off, o0, _ = S.local_csr(T["offsets"], base, end)
        batches.append(dict(base=base, cnt=cnt, tw=d_tw[base: base + cnt], out=d_out[o0:], off=torch.from_numpy(off.astype(np.int64)).to(torch.int32).cuda(),
                            ids=torch.zeros(cnt, dtype=torch.int32, device="cuda"), n=torch.zeros(1, dtype=torch.int32, device="cuda")))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    main_s = torch.cuda.current_stream()
    side = [torch.cuda.Stream() for _ in range(max(args.streams, 1))]

    def step_device():
        start = torch.cuda.Event(enable_timing=True)
        end = torch.cuda.Event(enable_timing=True)
        start.record(main_s)
        for s in side:
            s.wait_event(start)
        for i, b in enumerate(batches):
            s = side[i % len(side)]
            q.scan_device(dev, b["tw"].data_ptr(), b["out"].data_ptr(), b["off"].data_ptr(), b["cnt"], b["ids"].data_ptr(),
                          b["n"].data_ptr(), s.cuda_stream)
        for s in side:
            e = torch.cuda.Event()
            e.record(s)
            main_s.wait_event(e)
        end.record(main_s)
        return start, end

    def gathered_ids():
        return S.merge_matches([(b["base"], b["ids"][: int(b["n"].item())].cpu().numpy().astype(np.uint32)) for b in batches])

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()

    for _ in range(args.warmup):
        step_device()
    torch.cuda.synchronize()
    assert np.array_equal(gathered_ids(), T["expected"]), "GPU match set differs from the planted matches"
    sampler = ClockSampler(dev)
    sampler.start()
    launches0 = q.launch_count()
    barrier()
    evs = []
    for _ in range(args.steps):
        flush.fill_(1)
        evs.append(step_device())
    barrier()
    dev_ms = sum(s.elapsed_time(e) for s, e in evs) / args.steps
    launches = q.launch_count() - launches0

    # e2e: public host API, host buffers, copies inside the timed region (bytes counted by the library where it issues the copies)
    ht = args.host_threads if args.host_threads > 0 else None
    for _ in range(2):
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"], host_threads=ht)
    assert np.array_equal(ids, T["expected"]), "e2e match set differs from the planted matches"
    h2d0, d2h0 = q.transfer_bytes()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"], host_threads=ht)
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / args.steps
    barrier()
    h2d1, d2h1 = q.transfer_bytes()
    h2d, d2h = (h2d1 - h2d0) // args.steps, (d2h1 - d2h0) // args.steps
    assert np.array_equal(ids, T["expected"]), "e2e match set differs from the planted matches"
    clocks = sampler.finish()

    # per-kernel times of one full batch (CUDA events on the launching stream, warm)
    b0 = batches[0]
    st = None
    for _ in range(3):
        st = M.profile_stages(q, dev, b0["tw"].data_ptr(), b0["out"].data_ptr(), b0["off"].data_ptr(), b0["cnt"], b0["ids"].data_ptr(),
                              b0["n"].data_ptr())
    q_info = q.info()
    dev_ms, e2e_ms = S.max_over_ranks(dev_ms / 1e3, dist) * 1e3, S.max_over_ranks(e2e_ms / 1e3, dist) * 1e3
    # The ranks' part ends here.  Everything below is rank 0 alone: the other ranks leave (their GPUs must be idle for the
    # one-handle and DuckDB legs, which use every device of the box from ONE process).
    q.close()
    del batches, d_tw, d_out, flush
    torch.cuda.empty_cache()
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        peak, peak_src, wide_peak = measured_imad_peak()
        # frac_wide: the ladder's EXECUTED 64-bit multiply-adds (profiles/traffic.json, ncu instruction counts) against the
        # measured issue rate of IMAD.WIDE itself -- the pipe the kernel is bound by
        value = world * rows / (dev_ms / 1e3)
        e2e = world * rows / (e2e_ms / 1e3)
        ladder_rate = b0["cnt"] * W32_LADDER / (st["ladder"] / 1e3)
        wide_per_row = ladder_wide_per_row(q_info)
        frac_wide = b0["cnt"] * wide_per_row / (st["ladder"] / 1e3) / wide_peak
        traffic = fmaheavy = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            traffic, fmaheavy = tj.get("k_ladder_dram_bytes_per_launch"), tj.get("k_ladder_fmaheavy_pipe_busy")
        cores = os.cpu_count() or 1
        cpu_rate, cpu_n, cpu_dt, _ = cpu_sample(T, min(rows, cores * 6000), cores)
        cpu_ossl = cpu_openssl_sample(T, min(rows, cores * 1500), cores)
        one_handle = duck = None
        if not args.no_legs:
            try:
                if world > 1:
                    one_handle = one_handle_leg(M, T, rows, world, args.batch, min(args.steps, 3))
            except Exception as ex:      # noqa: BLE001  (a leg must not take the contract line down with it)
                one_handle = {"unavailable": repr(ex)[:300]}
            try:
                if world > 1:      # config C5's shape: 12.5 M rows per GPU (100 M at N = 8), 16 labels, multi-output rows
                    duck = duckdb_leg(synth, args.duckdb_rows or 12500000, world, 16, 0.25, args.batch, world)
                else:
                    duck = duckdb_leg(synth, args.duckdb_rows or 10000000, 1, L, 0.5, args.batch, 1)
            except Exception as ex:      # noqa: BLE001
                duck = {"unavailable": repr(ex)[:300]}
        py_base = None
        if world == 1:
            # the reference's only CPU code for this path is its Python verifier (scripts/verify_bip352_test_vector.py);
            # oracle/ref_scan.py restates it row by row -- 1 core, a few seconds
            try:
                from oracle import ref_scan as RS
                sk_i, spend_p = RS.parse_scalar(T["sk"]), RS.parse_point(T["spend"])
                labels_p = [RS.parse_point(l) for l in T["labels"]]
                tw = T["tweaks"].reshape(-1, 64)
                n_py, t0p = 0, time.perf_counter()
                while n_py < 400 and time.perf_counter() - t0p < 4.0:
                    RS.row_candidates(sk_i, RS.parse_point(tw[n_py].tobytes()), spend_p, labels_p)
                    n_py += 1
                dtp = time.perf_counter() - t0p
                py_base = {"value": n_py / dtp, "unit": "rows/s", "cores": 1, "kind": "port",
                           "sample": "%d rows of the same table, oracle/ref_scan.py (pure Python, as the reference's verifier scripts), %.1f s" % (n_py, dtp)}
            except Exception as ex:
                py_base = {"unavailable": repr(ex)[:200]}
        gecc = gecc_sql = None
        if world == 1:
            try:
                from oracle import gecc_ref as GR
                if GR.available():
                    nref = min(rows, args.batch)
                    GR.scan_batch(T["sk"], T["spend"], T["labels"], T["tweaks"][:nref], T["outputs"], T["offsets"][: nref + 1])
                    fl, dt = GR.scan_batch(T["sk"], T["spend"], T["labels"], T["tweaks"][:nref], T["outputs"], T["offsets"][: nref + 1])
                    ok = bool(np.array_equal(np.nonzero(fl)[0].astype(np.uint32), T["expected"][T["expected"] < nref]))
                    gecc = {"value": nref / dt, "unit": "rows/s", "kind": "reference",
                            "sample": "one %d-row batch through the unmodified reference GPU pipeline (oracle/_ref, gECC, sm_100) on the same "
                                      "B200, host buffers in / flags out, %.3f s" % (nref, dt), "flags_equal_planted": ok}
            except Exception as ex:   # the baseline is optional; the product line must still print
                gecc = {"unavailable": repr(ex)[:200]}
            # ... and through the reference's own DuckDB extension: the same SQL as the `duckdb` leg, the same table
            ref_exe = os.path.join(ROOT, "oracle", "_ref", "duckdb_ref")
            if os.path.exists(ref_exe) and not args.no_legs:
                try:
                    v, ms, _ = reference_sql_run(ref_exe, synth, T, rows, L, args.batch, 1, 1)
                    gecc_sql = {"value": v, "unit": "rows/s", "kind": "reference", "ms_per_query": ms,
                                "sample": "the unmodified reference DuckDB extension (oracle/_ref/duckdb_ref) on the same %d-row table as Parquet, its README's "
                                          "SQL, %d DuckDB threads, same B200; match count equal to the planted matches" % (rows, cores)}
                except Exception as ex:      # noqa: BLE001
                    gecc_sql = {"unavailable": repr(ex)[:200]}
        line = {"metric": metric_name(L), "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u32", "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e, "unit": "rows/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                        "api": "cudasp_scan_batch (C ABI, host buffers -> pinned staging by the call's host threads -> async H2D -> kernels -> D2H ids)"},
                "gpu_launches": int(launches),
                "roofline": {"bound": "int-mul", "kernel": "k_ladder", "achieved": ladder_rate / 1e12, "peak": peak / 1e12,
                             "unit": "T IMAD32 lane-ops/s", "frac": ladder_rate / peak, "frac_wide": frac_wide, "wide_products_per_row": wide_per_row,
                             "wide_peak_products_per_s": wide_peak, "fmaheavy_pipe_busy_ncu": fmaheavy, "traffic": traffic,
                             "peak_source": peak_src, "algorithmic_ops_per_row": W32_LADDER, "rows_per_launch": b0["cnt"],
                             "kernel_ms": st["ladder"], "stage_ms": st,
                             "pipeline_frac_per_gpu": (value / world) * w32 / peak, "pipeline_ops_per_row": w32,
                             "hbm_gbs_per_gpu": (value / world) * (64 + 8 * float(T["outputs"].shape[0]) / rows + 4) / 1e9},
                "cpu_baseline": {"value": cpu_rate, "unit": "rows/s", "cores": cores, "kind": "port",
                                 "sample": "%d rows of the same table, oracle/cudasp_oracle.c, %d pthreads, %.1f s" % (cpu_n, cores, cpu_dt)},
                "cpu_openssl": cpu_ossl, "python_baseline": py_base, "gecc_baseline": gecc, "gecc_sql_baseline": gecc_sql,
                "one_handle": one_handle, "duckdb": duck}
        emit(line)


if __name__ == "__main__":
    main()
