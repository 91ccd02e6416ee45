#!/usr/bin/env python3
"""BASELINE.json configs C2..C4 at full size on this GPU (SURVEY.md section 8d / BASELINE.md section 3):
seeded synthetic tables with planted base/label matches, scanned through the public C-ABI call with HOST buffers
(cudasp_scan_batch).  Checks, per config: the match set equals the planted set exactly, and the candidates of a
random row sample equal the CPU oracle's.  Prints one JSON line per config.

    python tools/run_configs.py [C2 C3 C4]        (GPU box)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

CONFIGS = {"C2": dict(rows=1000000, labels=0, n_base=100, per_label=0, seed=352002),
           "C3": dict(rows=9400000, labels=1, n_base=100, per_label=100, seed=352003),
           "C4": dict(rows=32700000, labels=8, n_base=100, per_label=50, seed=352004)}


def main():
    import cudasp_b200_loader
    from oracle import c_oracle as CO
    from tests import synth
    M = cudasp_b200_loader.load()
    for name in (sys.argv[1:] or ["C2", "C3", "C4"]):
        c = CONFIGS[name]
        t0 = time.perf_counter()
        T = synth.make_table(c["rows"], c["seed"], n_labels=c["labels"], n_base=c["n_base"], n_per_label=c["per_label"], batch_rows=300000)
        gen_s = time.perf_counter() - t0
        with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
            ids = q.scan_batch(T["tweaks"][:300000], T["outputs"], T["offsets"][:300001])     # warm-up (allocations, table upload)
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
                ts.append(time.perf_counter() - t0)
            ok_set = bool(np.array_equal(ids, T["expected"]))
            rng = np.random.default_rng(c["seed"])
            sel = np.sort(rng.choice(c["rows"], 20000, replace=False))
            got = q.candidates(np.ascontiguousarray(T["tweaks"][sel]))
        _, want = CO.scan(T["sk"], T["spend"], T["labels"], np.ascontiguousarray(T["tweaks"][sel]), None, None, threads=os.cpu_count() or 1,
                          want_cands=True)
        best = min(ts)
        print(json.dumps({"config": name, "rows": c["rows"], "labels": c["labels"], "planted": int(len(T["expected"])),
                          "match_set_equals_planted": ok_set, "sample_candidates_equal_oracle": bool(np.array_equal(got, want)),
                          "sample_rows": int(len(sel)), "e2e_s_best": round(best, 4), "e2e_s_all": [round(x, 4) for x in ts],
                          "e2e_rows_per_s": round(c["rows"] / best), "table_gen_s": round(gen_s, 1), "api": "cudasp_scan_batch, host buffers, 1 GPU"}),
              flush=True)


if __name__ == "__main__":
    main()
