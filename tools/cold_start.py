#!/usr/bin/env python3
"""Cold-start breakdown of the C ABI in a fresh process: CUDA initialisation, first query (fixed-base table build +
upload, first slot), second query (slot cache), first small batch (module load of the kernels).   (GPU box)"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import cudasp_b200_loader
    from tests import synth
    t0 = time.perf_counter()
    M = cudasp_b200_loader.load()
    L = M.lib()
    t1 = time.perf_counter()
    n = L.cudasp_device_count()
    t2 = time.perf_counter()
    T = synth.make_table(20000, 1, n_labels=0, n_base=5, batch_rows=300000, specials=False)
    t2b = time.perf_counter()
    all_devices = len(sys.argv) > 1 and sys.argv[1] == "--all-devices"     # one handle over every visible GPU
    if not all_devices:
        M.debug_gtab(0, 0, 1)                # forces the device-side build of the fixed-base table (+ CUDA context)
    t3 = time.perf_counter()
    gtab_s = t3 - t2b
    q = M.Query(T["sk"], T["spend"], [], devices=list(range(n)) if all_devices else [0], batch_rows=300000)
    t4 = time.perf_counter()
    ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
    t5 = time.perf_counter()
    ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
    t6 = time.perf_counter()
    q.close()
    q2 = M.Query(T["sk"], T["spend"], [], devices=list(range(n)) if all_devices else [0], batch_rows=300000)
    t7 = time.perf_counter()
    q2.close()
    print(json.dumps({"devices_visible": n, "load_lib_s": t1 - t0, "device_count_s": t2 - t1, "context_and_fixed_base_table_s": gtab_s, "first_query_create_s": t4 - t3,
                      "first_batch_s": t5 - t4, "second_batch_s": t6 - t5, "second_query_create_s": t7 - t6,
                      "ok": bool(np.array_equal(ids, T["expected"]))}))


if __name__ == "__main__":
    main()
