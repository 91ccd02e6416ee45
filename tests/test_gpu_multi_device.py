"""GPU (-m gpu, needs >= 2 devices; skipped otherwise): ONE query handle that owns several GPUs -- the in-process form of
SURVEY.md 8e (the reference maps DuckDB worker threads round-robin onto devices, src/cudasp_extension.cu:129-132,
516-532).  Slots are handed out device-interleaved; every entry point must give the single-device answer."""
import ctypes as C
import threading

import numpy as np
import pytest

import cudasp_b200_loader
from tests import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    n = m.lib().cudasp_device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2)")
    return m


def test_one_handle_two_devices_all_entries(M):
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    from tests.test_gpu_arrow import arrow_table
    ndev = min(M.lib().cudasp_device_count(), 4)
    devs = list(range(ndev))
    rows = 400000
    T = synth.make_table(rows, 8088, n_labels=1, n_base=80, n_per_label=40, batch_rows=50000, specials=False)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=devs, batch_rows=50000) as q:
        assert q.info()["n_devices"] == ndev
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
        assert np.array_equal(ids, T["expected"])
        # slots held at the same time spread over every device of the handle (devices come up on demand: the first two
        # slots may sit on the device that is already up, cudasp_capi.cu try_acquire)
        L = M.lib()
        seen, held = set(), []
        for _ in range(3 * ndev):
            s = B.Slot()
            assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
            seen.add(s.device)
            held.append(s)
            if seen == set(devs):
                break
        assert seen == set(devs)
        for s in held:      # hand them back through an empty batch
            assert L.cudasp_submit_slot(q.handle, C.byref(s), 0, 0, 0) == 0
        r = B.Result()
        while L.cudasp_wait(q.handle, C.byref(r)) == 1:
            assert r.n_matches == 0
            L.cudasp_release_result(q.handle, C.byref(r))
        # columnar entry
        assert np.array_equal(M.scan_table(q, arrow_table(T)), T["expected"].astype(np.uint64))
        # several host threads on the one handle (the DuckDB worker model)
        out, errs = {}, []

        def work(t, lo, hi):
            try:
                o0, o1 = int(T["offsets"][lo]), int(T["offsets"][hi])
                off = (T["offsets"][lo: hi + 1].astype(np.int64) - o0).astype(np.uint32)
                out[t] = q.scan_batch(T["tweaks"][lo:hi], T["outputs"][o0:o1], off) + lo
            except Exception as e:      # noqa: BLE001
                errs.append(e)
        cuts = np.linspace(0, rows, 9).astype(int)
        th = [threading.Thread(target=work, args=(i, int(cuts[i]), int(cuts[i + 1]))) for i in range(8)]
        [t.start() for t in th]
        [t.join() for t in th]
        assert not errs, errs
        assert np.array_equal(np.sort(np.concatenate([out[i] for i in range(8)])), T["expected"])
