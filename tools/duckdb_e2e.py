#!/usr/bin/env python3
"""End-to-end timing through the real DuckDB engine + the C++ `cudasp` extension (duckdb_ext/bin/duckdb):
writes a synthetic Parquet table (planted matches), then times
    SELECT count(*) FROM cudasp_scan((SELECT txid,height,tweak_key,outputs FROM read_parquet(...)), sk, spend, [labels])
and, for comparison, the same scan of the table without the function (the cost of reading + decoding alone).

    python tools/duckdb_e2e.py [--rows N] [--labels L] [--threads T] [--in-memory]       (GPU box)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
DUCKDB = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
# --impl reference: the REFERENCE's own extension in the same DuckDB (oracle/_ref/duckdb_ref, `make -C oracle ref_ext`)
DUCKDB_REF = os.path.join(ROOT, "oracle", "_ref", "duckdb_ref")


def blob(b: bytes) -> str:
    return "'" + "".join("\\x%02x" % c for c in b) + "'::BLOB"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4000000)
    ap.add_argument("--labels", type=int, default=1)
    ap.add_argument("--threads", type=str, default="0", help="PRAGMA threads; a comma list runs the queries once per value on the same file")
    ap.add_argument("--batch", type=int, default=300000)
    ap.add_argument("--replicate", type=int, default=1, help="scan the same Parquet file N times in one query (N x rows rows)")
    ap.add_argument("--geom-p", type=float, default=0.5, help="outputs per row = 1 + min(Geom(p), 15|63); C5 uses 0.25")
    ap.add_argument("--compressed", action="store_true", help="store tweak_key as 33-byte SEC1 (decompressed on the GPU)")
    ap.add_argument("--details", action="store_true", help="details := true (one result row per found output)")
    ap.add_argument("--in-memory", action="store_true", help="load the table into a DuckDB table first (no Parquet decode in the timed query)")
    ap.add_argument("--native", action="store_true", help="also time cudasp_scan_parquet(paths, ...): the library's own page decoders instead of DuckDB's read_parquet")
    ap.add_argument("--native-threads", type=str, default="", help="with --native: also time cudasp_scan_parquet with threads := N for every N of this comma list")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "both"], help="which extension runs the SQL (same file, same query)")
    a = ap.parse_args()
    import numpy as np
    import pyarrow as pa
    import pyarrow.parquet as pq
    from tests import synth
    t0 = time.perf_counter()
    T = synth.make_table(a.rows, 352003, n_labels=a.labels, n_base=100, n_per_label=100 if a.labels else 0, batch_rows=a.batch,
                         geom_p=a.geom_p, max_extra=63 if a.geom_p < 0.5 else 15)
    rows = a.rows
    off = T["offsets"].astype(np.int64)
    if a.compressed:
        from tests.test_gpu_compressed import compress
        tw = pa.FixedSizeBinaryArray.from_buffers(pa.binary(33), rows, [None, pa.py_buffer(compress(T["tweaks"]).tobytes())]).cast(pa.binary())
    else:
        tw = pa.FixedSizeBinaryArray.from_buffers(pa.binary(64), rows, [None, pa.py_buffer(T["tweaks"].tobytes())]).cast(pa.binary())
    txid = np.zeros((rows, 32), np.uint8)
    txid[:, :8] = np.arange(rows, dtype=np.uint64).view(np.uint8).reshape(rows, 8)
    tx = pa.FixedSizeBinaryArray.from_buffers(pa.binary(32), rows, [None, pa.py_buffer(txid.tobytes())]).cast(pa.binary())
    outs = pa.ListArray.from_arrays(pa.array(off.astype(np.int32), pa.int32()), pa.array(T["outputs"], pa.int64()))
    tbl = pa.table({"txid": tx, "height": pa.array((800000 + np.arange(rows) // 3000).astype(np.int32)), "tweak_key": tw, "outputs": outs})
    d = tempfile.mkdtemp(prefix="cudasp_e2e_")
    path = os.path.join(d, "t.parquet")
    pq.write_table(tbl, path, row_group_size=122880, compression="NONE")
    gen_s = time.perf_counter() - t0
    labels = "[" + ", ".join(blob(l) for l in T["labels"]) + "]" + ("" if T["labels"] else "::BLOB[]")
    src = "read_parquet('%s')" % path if a.replicate == 1 else "read_parquet([%s])" % ", ".join(["'%s'" % path] * a.replicate)
    impls = ["ours", "reference"] if a.impl == "both" else [a.impl]
    for threads, impl in [(int(x), i) for x in a.threads.split(',') for i in impls]:
        exe = DUCKDB if impl == "ours" else DUCKDB_REF
        if impl == "reference" and (a.compressed or a.details):
            continue                      # the reference has neither
        pre = "PRAGMA threads=%d;\n" % threads if threads else ""
        cur = src
        if a.in_memory:
            pre += "CREATE TABLE t AS SELECT * FROM %s;\n" % src
            cur = "t"
        scan = "SELECT count(*) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM %s), %s, %s, %s, batch_size := %d%s);" % (
            cur, blob(T["sk"]), blob(T["spend"]), labels, a.batch, ", details := true" if a.details else "")
        plain = "SELECT count(*), sum(height), sum(octet_length(tweak_key)), sum(len(outputs)) FROM %s;" % cur
        native = ""
        if a.native and impl == "ours" and not a.details:
            paths = "[%s]" % ", ".join(["'%s'" % path] * a.replicate)
            nq = "SELECT count(*) FROM cudasp_scan_parquet(%s, %s, %s, %s, batch_size := %d);" % (paths, blob(T["sk"]), blob(T["spend"]), labels, a.batch)
            native = nq + "\n" + nq + "\n" + nq + "\n"
            for nt in [int(x) for x in a.native_threads.split(",") if x]:
                nq2 = nq.replace("batch_size :=", "threads := %d, batch_size :=" % nt)
                native += nq2 + "\n" + nq2 + "\n" + nq2 + "\n"
        sql = pre + ".timer on\n" + scan + "\n" + scan + "\n" + plain + "\n" + native + "SELECT current_setting('threads');\n"
        p = subprocess.run([exe, "-csv", "-noheader"], input=sql, capture_output=True, text=True, timeout=1800)
        if os.environ.get("CUDASP_TRACE"):
            sys.stderr.write(p.stderr)
        lines = [l for l in p.stdout.strip().splitlines() if l]
        times = [float(l.split("real")[1].split()[0]) for l in lines if l.startswith("Run Time")]
        vals = [l for l in lines if not l.startswith("Run Time")]
        want = len(T["expected"]) * a.replicate
        rows = a.rows * a.replicate
        ok = p.returncode == 0 and len(vals) >= 2 and int(vals[0]) == want and int(vals[1]) == want
        res = {"impl": impl, "rows": rows, "replicate": a.replicate, "geom_p": a.geom_p, "mean_outputs": float(len(T["outputs"])) / a.rows, "labels": a.labels, "compressed": a.compressed, "details": a.details, "threads": vals[-1] if vals else None, "in_memory": a.in_memory, "matches_expected": want,
               "matches_ok": ok, "scan_s_cold": times[0] if times else None, "scan_s_warm": times[1] if len(times) > 1 else None,
               "plain_read_s": times[2] if len(times) > 2 else None, "gen_s": round(gen_s, 1),
               "rows_per_s_warm": rows / times[1] if len(times) > 1 else None, "stderr": p.stderr[-300:]}
        if native and len(times) >= 6 and len(vals) >= 6:
            res["native_parquet"] = {"scan_s_first": times[3], "scan_s_warm": min(times[4:6]), "rows_per_s_warm": rows / min(times[4:6]),
                                     "matches_ok": all(int(v) == want for v in vals[3:6])}
            for j, nt in enumerate([int(x) for x in a.native_threads.split(",") if x]):
                lo = 6 + 3 * j
                if len(times) >= lo + 3 and len(vals) >= lo + 3:
                    res["native_parquet_threads_%d" % nt] = {"scan_s_warm": min(times[lo + 1:lo + 3]), "rows_per_s_warm": rows / min(times[lo + 1:lo + 3]),
                                                            "matches_ok": all(int(v) == want for v in vals[lo:lo + 3])}
        print(json.dumps(res))
    try:
        os.remove(path)
        os.rmdir(d)
    except OSError:
        pass


if __name__ == "__main__":
    main()
