#!/usr/bin/env python3
"""Host-side cost of the DuckDB extension per input row, with no GPU in the loop (CPU box): the `duckdb` shell runs against the
test double of the C ABI (tests/hostcheck/mock_cabi, LD_LIBRARY_PATH) with its stand-in scan switched off, so the query time is
DuckDB's Parquet decode + AccumulateInput + slot handling.  Compared with the same SELECT without the table function.

    python tools/host_feed_bench.py [--rows N] [--threads 1,8] [--geom-p 0.25]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUCKDB = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
MOCK = os.path.join(ROOT, "tests", "hostcheck", "mock_cabi")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4000000)
    ap.add_argument("--threads", default="1,8")
    ap.add_argument("--geom-p", type=float, default=0.25)
    ap.add_argument("--path", default="/tmp/hf/t.parquet")
    a = ap.parse_args()
    import numpy as np
    import pyarrow as pa
    import pyarrow.parquet as pq
    if not os.path.exists(a.path):
        rng = np.random.default_rng(1)
        rows = a.rows
        n_out = 1 + np.minimum(rng.geometric(a.geom_p, rows) - 1, 63)
        off = np.zeros(rows + 1, np.int32)
        off[1:] = np.cumsum(n_out)
        tw = pa.FixedSizeBinaryArray.from_buffers(pa.binary(64), rows, [None, pa.py_buffer(rng.bytes(64 * rows))]).cast(pa.binary())
        tx = pa.FixedSizeBinaryArray.from_buffers(pa.binary(32), rows, [None, pa.py_buffer(rng.bytes(32 * rows))]).cast(pa.binary())
        outs = pa.ListArray.from_arrays(pa.array(off, pa.int32()), pa.array(rng.integers(-2**63, 2**63 - 1, int(off[-1]), dtype=np.int64)))
        tbl = pa.table({"txid": tx, "height": pa.array((800000 + np.arange(rows) // 3000).astype(np.int32)), "tweak_key": tw, "outputs": outs})
        os.makedirs(os.path.dirname(a.path), exist_ok=True)
        pq.write_table(tbl, a.path, row_group_size=122880, compression="NONE")
    rows = pq.ParquetFile(a.path).metadata.num_rows
    env = dict(os.environ, LD_LIBRARY_PATH=MOCK, MOCK_CUDASP_NO_SCAN="1", MOCK_CUDASP_DEVICES="1")
    for th in [int(x) for x in a.threads.split(",")]:
        scan = ("SELECT count(*) FROM cudasp_scan((SELECT txid, height, tweak_key, outputs FROM read_parquet('%s')), "
                "repeat('\\x01',32)::BLOB, repeat('\\x02',64)::BLOB, []::BLOB[]);" % a.path)
        plain = ("SELECT count(*), sum(height), sum(octet_length(tweak_key)), sum(octet_length(txid)), sum(list_sum(outputs)) "
                 "FROM read_parquet('%s');" % a.path)
        light = "SELECT count(*), sum(height), sum(octet_length(tweak_key)), sum(len(outputs)) FROM read_parquet('%s');" % a.path
        sql = "PRAGMA threads=%d;\n.timer on\n%s\n%s\n%s\n%s\n%s\n" % (th, scan, scan, scan, plain, light)
        p = subprocess.run([DUCKDB, "-csv", "-noheader"], input=sql, capture_output=True, text=True, env=env, timeout=1800)
        times = [float(l.split("real")[1].split()[0]) for l in p.stdout.splitlines() if l.startswith("Run Time")]
        if len(times) < 5:
            print(p.stdout, p.stderr)
            continue
        best = min(times[1:3])
        print(json.dumps({"rows": rows, "threads": th, "scan_s": best, "plain_all_columns_s": times[3], "plain_light_s": times[4],
                          "ns_per_row_thread_scan": 1e9 * best * th / rows, "ns_per_row_thread_plain": 1e9 * times[3] * th / rows,
                          "ns_per_row_thread_light": 1e9 * times[4] * th / rows}))


if __name__ == "__main__":
    main()
