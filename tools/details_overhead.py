#!/usr/bin/env python3
"""Cost of the opt-in match-details pass (K8): the same 1 M-row table through cudasp_scan_batch_fmt and through
cudasp_scan_batch_details (host buffers, end to end), best of R runs each.     (GPU box)"""
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
    M = cudasp_b200_loader.load()
    rows = int(os.environ.get("ROWS", "1000000"))
    L = int(os.environ.get("LABELS", "0"))
    planted = int(os.environ.get("PLANTED", "100"))
    T = synth.make_table(rows, 352005, n_labels=L, n_base=planted, n_per_label=0, batch_rows=300000)
    res = {"rows": rows, "labels": L, "planted": planted}
    for name, details in (("plain", False), ("details", True)):
        with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
            if details:
                q.enable_details()
            best, n = None, 0
            for _ in range(6):
                t0 = time.perf_counter()
                if details:
                    ids, recs = q.scan_batch_details(T["tweaks"], T["outputs"], T["offsets"])
                    n = len(recs)
                else:
                    ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
            res[name + "_rows_per_s"] = rows / best
            res[name + "_ok"] = bool(np.array_equal(ids, T["expected"]))
            if details:
                res["records"] = n
    res["overhead_pct"] = 100.0 * (res["plain_rows_per_s"] / res["details_rows_per_s"] - 1.0)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
