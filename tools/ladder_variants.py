#!/usr/bin/env python3
"""Tuning aid: time the pipeline stages of one 300 K-row batch for each K1 launch variant
(CUDASP_LADDER_VARIANT, see csrc/cudasp_capi.cu launch_ladder) and check the match set each time.

    python tools/ladder_variants.py [variants...]        (GPU box)

Variants 1-3 exist only in a library built with VARIANT_DEFS=-DCSP_LADDER_TUNING (csrc/Makefile); the production
build has variant 0 alone and maps every other number to it.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child():
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import cudasp_b200_loader
    from tests import synth
    M = cudasp_b200_loader.load()
    rows = int(os.environ.get("ROWS", "300000"))
    L = int(os.environ.get("LABELS", "0"))
    T = synth.make_table(rows, 352002, n_labels=L, n_base=100, n_per_label=20 if L else 0, batch_rows=rows)
    q = M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=rows, max_outputs_per_batch=int(T["offsets"][-1]) + 16)
    fmt = int(os.environ.get("KEYFMT", "0"))     # 1: 33-byte SEC1 keys
    if fmt:
        from tests.test_gpu_compressed import compress
        os.environ["CUDASP_PROFILE_KEYFMT"] = "1"
        T["tweaks"] = compress(T["tweaks"])
    d_tw = torch.from_numpy(T["tweaks"]).cuda()
    d_out = torch.from_numpy(T["outputs"]).cuda()
    d_off = torch.from_numpy(T["offsets"].astype(np.int64)).to(torch.int32).cuda()
    ids = torch.zeros(rows, dtype=torch.int32, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    best = None
    for _ in range(6):
        st = M.profile_stages(q, 0, d_tw.data_ptr(), d_out.data_ptr(), d_off.data_ptr(), rows, ids.data_ptr(), n.data_ptr())
        if best is None or st["ladder"] < best["ladder"]:
            best = st
    got = np.sort(ids[: int(n.item())].cpu().numpy().astype(np.uint32))
    ok = bool(np.array_equal(got, T["expected"]))
    tot = sum(best.values())
    print(json.dumps({"variant": os.environ.get("CUDASP_LADDER_VARIANT", "0"), "keyfmt": fmt, "ok": ok, "rows": rows, "total_ms": round(tot, 4),
                      "Mrows_s": round(rows / tot / 1e3, 2), **{k: round(v, 4) for k, v in best.items()}}))
    q.close()


if __name__ == "__main__":
    if os.environ.get("_LV_CHILD"):
        child()
    else:
        variants = sys.argv[1:] or ["0", "1", "2", "3"]
        for v in variants:
            env = dict(os.environ, CUDASP_LADDER_VARIANT=v, _LV_CHILD="1")
            p = subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, capture_output=True, text=True, timeout=300)
            print(p.stdout.strip() or ("variant %s failed: %s" % (v, p.stderr[-400:])), flush=True)
