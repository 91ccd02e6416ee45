#!/usr/bin/env python3
"""End-to-end timing of the columnar entry (include/cudasp_arrow.h) on a synthetic table with planted matches:
  table   : an in-memory pyarrow Table -> cudasp_scan_arrow_stream            (host Arrow buffers -> pinned -> GPU)
  parquet : ParquetFile.iter_batches -> cudasp_scan_arrow_stream              (decode + scan, overlapped)
for 64-byte keys and for 33-byte compressed keys.

    python tools/arrow_e2e.py [--rows N] [--labels L] [--reps R]       (GPU box)
"""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4000000)
    ap.add_argument("--labels", type=int, default=0)
    ap.add_argument("--batch", type=int, default=300000)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--devices", type=int, default=1, help="GPUs owned by the one query handle (one feeding thread)")
    ap.add_argument("--chunk", type=int, default=122880, help="rows per record batch / Parquet row group")
    a = ap.parse_args()
    import numpy as np
    import pyarrow as pa
    import pyarrow.parquet as pq
    import cudasp_b200_loader
    from tests import synth
    from tests.test_gpu_arrow import arrow_table
    from tests.test_gpu_compressed import compress
    M = cudasp_b200_loader.load()
    T = synth.make_table(a.rows, 352004, n_labels=a.labels, n_base=100, n_per_label=100 if a.labels else 0, batch_rows=a.batch)
    want = T["expected"].astype(np.uint64)
    res = {"rows": a.rows, "labels": a.labels, "chunk": a.chunk, "devices": a.devices}
    d = tempfile.mkdtemp(prefix="cudasp_arrow_")
    with M.Query(T["sk"], T["spend"], T["labels"], devices=list(range(a.devices)), batch_rows=a.batch) as q:
        for name, tbl in (("xy64", arrow_table(T)), ("sec1", arrow_table(T, compress(T["tweaks"]), 33))):
            tbl = pa.Table.from_batches(tbl.to_batches(max_chunksize=a.chunk))
            best = None
            for _ in range(a.reps + 1):
                t0 = time.perf_counter()
                ids = M.scan_table(q, tbl)
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            res[name + "_table_rows_per_s"] = a.rows / best
            res[name + "_table_ok"] = bool(np.array_equal(ids, want))
            path = os.path.join(d, name + ".parquet")
            pq.write_table(tbl.select(["tweak_key", "outputs"]), path, row_group_size=a.chunk, compression="NONE")
            best = None
            for _ in range(a.reps):
                pf = pq.ParquetFile(path)
                reader = pa.RecordBatchReader.from_batches(pf.schema_arrow, pf.iter_batches(batch_size=a.chunk))
                t0 = time.perf_counter()
                ids, seen = M.scan_stream(q, reader)
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            res[name + "_parquet_rows_per_s"] = a.rows / best
            res[name + "_parquet_ok"] = bool(seen == a.rows and np.array_equal(ids, want))
            for th in (2, 4, 8):
                best = None
                for _ in range(a.reps):
                    t0 = time.perf_counter()
                    ids, seen = M.scan_parquet(q, path, threads=th)
                    dt = time.perf_counter() - t0
                    best = dt if best is None else min(best, dt)
                res[name + "_parquet_t%d_rows_per_s" % th] = a.rows / best
                res[name + "_parquet_t%d_ok" % th] = bool(seen == a.rows and np.array_equal(ids, want))
            t0 = time.perf_counter()
            n = sum(b.num_rows for b in pq.ParquetFile(path).iter_batches(batch_size=a.chunk))
            res[name + "_parquet_decode_only_rows_per_s"] = n / (time.perf_counter() - t0)
            os.remove(path)
    os.rmdir(d)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
