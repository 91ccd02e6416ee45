#!/usr/bin/env python3
"""Tuning aid: e2e rows/s of cudasp_scan_batch (host buffers) for several piece sizes / slot counts."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if os.environ.get("_CHILD"):
    sys.path.insert(0, ROOT)
    import numpy as np, cudasp_b200_loader
    from tests import synth
    M = cudasp_b200_loader.load()
    rows = int(os.environ.get("ROWS", "1000000"))
    T = synth.make_table(rows, 352002, n_labels=0, n_base=100, batch_rows=300000)
    q = M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000)
    for _ in range(3):
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
    ts = []
    for _ in range(7):
        t0 = time.perf_counter(); ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"]); ts.append(time.perf_counter() - t0)
    ok = bool(np.array_equal(ids, T["expected"]))
    print(json.dumps({"waves": os.environ.get("CUDASP_PIECE_WAVES"), "ok": ok, "ms_best": round(1e3 * min(ts), 3), "ms_med": round(1e3 * sorted(ts)[3], 3),
                      "Mrows_s": round(rows / min(ts) / 1e6, 2)}))
    q.close()
else:
    for w in (sys.argv[1:] or ["1", "2", "3", "4", "8"]):
        p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=dict(os.environ, CUDASP_PIECE_WAVES=w, _CHILD="1"), capture_output=True, text=True)
        print(p.stdout.strip() or p.stderr[-300:], flush=True)
        if os.environ.get("CUDASP_TRACE"): print(p.stderr[-700:], flush=True)
