"""GPU (-m gpu): the opt-in match-details pass (K8, include/cudasp.h cudasp_match_detail) -- BIP-352's full receive loop
on the rows that matched -- against the oracle's receive_scan_row on transactions built by a sender-side generator,
plus the guarantee that the default results (match ids) do not change when details are enabled."""
import random

import numpy as np
import pytest

import cudasp_b200_loader
from oracle import ref_scan as R
from tests import synth
from tests.test_gpu_compressed import compress

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def build_table(rows, L, seed, n_paid):
    """Random table in which n_paid transactions pay the recipient 1..4 times (labels mixed in), outputs shuffled among
    foreign ones.  Returns T (synth layout) and {row: [(k, output_index, label_index, t_k)]}."""
    rng = random.Random(seed)
    T = synth.make_table(rows, seed, n_labels=L, n_base=0, n_per_label=0, batch_rows=rows, specials=False)
    assert len(T["expected"]) == 0
    off = T["offsets"].astype(np.int64)
    lists = [T["outputs"][off[i]: off[i + 1]].tolist() for i in range(rows)]
    want = {}
    for row in sorted(rng.sample(range(rows), n_paid)):
        recipe = [rng.randrange(-1, L) for _ in range(rng.randint(1, 4))]
        outs, recs = synth.sender_outputs(T["sk"], T["spend"], T["labels"], T["tweaks"][row].tobytes(), recipe)
        mixed = [(o, k) for k, o in enumerate(outs)] + [(v, None) for v in lists[row]]
        rng.shuffle(mixed)
        lists[row] = [o for o, _ in mixed]
        pos = {k: j for j, (_, k) in enumerate(mixed) if k is not None}
        want[row] = [(k, pos[k], m, t) for k, m, t in recs]
    T["outputs"] = np.array([o for l in lists for o in l], np.int64)
    T["offsets"] = np.concatenate([[0], np.cumsum([len(l) for l in lists])]).astype(np.uint32)
    T["lists"] = lists
    return T, want


@pytest.mark.parametrize("rows,L,batch,fmt", [(3000, 0, 300000, 0), (5000, 3, 1024, 0), (4000, 2, 300000, 1)])
def test_details_equal_oracle_and_sender(M, rows, L, batch, fmt):
    T, want = build_table(rows, L, 900 + rows, 60)
    keys = compress(T["tweaks"]) if fmt else T["tweaks"]
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=batch) as q:
        plain = q.scan_batch(keys, T["outputs"], T["offsets"], key_format=fmt)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=batch) as q:
        q.enable_details()
        ids, recs = q.scan_batch_details(keys, T["outputs"], T["offsets"], key_format=fmt)
    assert np.array_equal(ids, plain)                                   # details never change the default result
    assert ids.tolist() == sorted(want)
    got = {}
    for row, k, oi, li, t in recs:
        got.setdefault(row, []).append((k, oi, li, t))
    assert got == want                                                  # what the sender built
    sk, spend, labels = R.parse_scalar(T["sk"]), R.parse_point(T["spend"]), [R.parse_point(l) for l in T["labels"]]
    for row in list(want)[:12]:                                         # and what the oracle's receiver finds
        tw = R.parse_point(T["tweaks"][row].tobytes())
        assert got[row] == R.receive_scan_row(sk, tw, spend, labels, T["lists"][row])


def test_details_through_streaming_and_table_function(M):
    import ctypes as C
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    T, want = build_table(2000, 1, 4321, 25)
    L = M.lib()
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=1000) as q:
        with pytest.raises(M.CudaspError, match="not enabled"):
            q.scan_batch_details(T["tweaks"], T["outputs"], T["offsets"])
        q.enable_details(4096)
        got = {}
        for b in range(2):
            lo, hi = b * 1000, (b + 1) * 1000
            s = B.Slot()
            assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
            o0, o1 = int(T["offsets"][lo]), int(T["offsets"][hi])
            tw, oo = np.ascontiguousarray(T["tweaks"][lo:hi]), np.ascontiguousarray(T["outputs"][o0:o1])
            off = (T["offsets"][lo: hi + 1].astype(np.int64) - o0).astype(np.uint32)      # named: must outlive the memmove
            C.memmove(s.tweak_keys, tw.ctypes.data, 64000)
            C.memmove(s.outputs, oo.ctypes.data, 8 * (o1 - o0))
            C.memmove(s.out_offsets, off.ctypes.data, 4004)
            assert L.cudasp_submit_slot(q.handle, C.byref(s), 1000, o1 - o0, lo) == 0
        r = B.Result()
        while L.cudasp_wait(q.handle, C.byref(r)) == 1:
            d, n = C.POINTER(B.MatchDetail)(), C.c_uint32(0)
            assert L.cudasp_result_details(q.handle, C.byref(r), C.byref(d), C.byref(n)) == 0
            for i in range(n.value):
                got.setdefault(int(r.tag) + d[i].row, []).append((d[i].k, d[i].output_index, d[i].label_index,
                                                                  int.from_bytes(bytes(d[i].t_k), "big")))
            assert sorted({int(r.tag) + d[i].row for i in range(n.value)}) == [int(r.tag) + r.match_ids[i] for i in range(r.n_matches)]
            assert L.cudasp_release_result(q.handle, C.byref(r)) == 0
    assert got == want
    # the table function with details := true: one result row per found output, richer schema
    rows = [(b"tx%05d" % i, 800000 + i, T["tweaks"][i].tobytes(), T["lists"][i]) for i in range(2000)]
    out = M.cudasp_scan(rows, T["sk"], T["spend"], T["labels"], batch_size=700, details=True)
    want_rows = sorted((b"tx%05d" % r, 800000 + r, T["tweaks"][r].tobytes(), k, oi, li, t.to_bytes(32, "big"))
                       for r, recs in want.items() for k, oi, li, t in recs)
    assert sorted(out) == want_rows
    base = M.cudasp_scan(rows, T["sk"], T["spend"], T["labels"], batch_size=700)
    assert sorted(set(o[:3] for o in out)) == sorted(base)


def test_details_capacity_overflow_is_reported(M):
    T, want = build_table(600, 0, 55, 40)
    with M.Query(T["sk"], T["spend"], [], batch_rows=600) as q:
        q.enable_details(8)                                   # fewer than the batch will produce
        with pytest.raises(M.CudaspError, match="max_details_per_batch"):
            q.scan_batch_details(T["tweaks"], T["outputs"], T["offsets"])
    with M.Query(T["sk"], T["spend"], [], batch_rows=600) as q:
        q.enable_details()
        with pytest.raises(M.CudaspError, match="capacity too small"):
            q.scan_batch_details(T["tweaks"], T["outputs"], T["offsets"], capacity=3)
        ids, recs = q.scan_batch_details(T["tweaks"], T["outputs"], T["offsets"])
        assert len(recs) == sum(len(v) for v in want.values())


def test_streaming_details_overflow_is_flagged_not_lost(M):
    """A batch with more detail records than max_details_per_batch still delivers its (complete) matches through the streaming
    API, flagged CUDASP_RESULT_DETAILS_TRUNCATED, and its slot can be released (it used to fail without naming the slot)."""
    import ctypes as C
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    T, want = build_table(600, 0, 55, 40)
    L = M.lib()
    with M.Query(T["sk"], T["spend"], [], batch_rows=600) as q:
        q.enable_details(8)
        s = B.Slot()
        assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
        # enabling details again while a slot is held is a state error, not a self-deadlock on the query mutex
        assert L.cudasp_query_enable_details(q.handle, 16) == B.ERR_STATE
        assert b"before the first batch" in L.cudasp_last_error(q.handle)
        n_out = int(T["offsets"][600])
        C.memmove(s.tweak_keys, T["tweaks"].ctypes.data, 600 * 64)
        C.memmove(s.outputs, T["outputs"].ctypes.data, 8 * n_out)
        C.memmove(s.out_offsets, T["offsets"].ctypes.data, 601 * 4)
        assert L.cudasp_submit_slot(q.handle, C.byref(s), 600, n_out, 7) == 0
        r = B.Result()
        assert L.cudasp_wait(q.handle, C.byref(r)) == 1
        assert r.tag == 7 and r.flags & 1
        assert [r.match_ids[i] for i in range(r.n_matches)] == sorted(want)      # the matches are complete
        d, n = C.POINTER(B.MatchDetail)(), C.c_uint32(0)
        assert L.cudasp_result_details(q.handle, C.byref(r), C.byref(d), C.byref(n)) == 0
        assert n.value == 8
        assert L.cudasp_release_result(q.handle, C.byref(r)) == 0
        assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0          # the slot came back
        assert L.cudasp_submit_slot(q.handle, C.byref(s), 0, 0, 8) == 0
        assert L.cudasp_wait(q.handle, C.byref(r)) == 1 and r.flags == 0
        assert L.cudasp_release_result(q.handle, C.byref(r)) == 0


def test_details_against_reference_generated_chain(M, golden):
    """Pins K8 for k > 0: golden.json "receive_loop" holds, for 20 transactions, t_k / P_k / P_k + L_i for k = 0..4 computed with
    the REFERENCE's own point_multiply / point_add / compress_point / tagged_hash (tests/golden/make_golden.py section 7) and
    the records a receiver must find in an outputs list paying it 1..5 times.  One query per transaction (own keys)."""
    for tx in golden["receive_loop"]:
        sk = int(tx["sk"], 16).to_bytes(32, "little")
        enc = lambda xy: R.encode_point((int(xy[0], 16), int(xy[1], 16)))      # noqa: E731
        tweak = np.frombuffer(enc(tx["tweak"]), np.uint8).reshape(1, 64)
        outputs = np.array(tx["outputs"], np.int64)
        offsets = np.array([0, len(outputs)], np.uint32)
        with M.Query(sk, enc(tx["spend"]), [enc(l) for l in tx["labels"]], batch_rows=16) as q:
            # the candidates of k = 0 are the default path's
            assert q.candidates(tweak)[0].tolist() == tx["chain"][0]["cands"]
            q.enable_details()
            ids, recs = q.scan_batch_details(tweak, outputs, offsets)
        want = [(0, r["k"], r["output_index"], r["label_index"], int(r["t"], 16)) for r in tx["records"]]
        assert ids.tolist() == ([0] if any(r["k"] == 0 for r in tx["records"]) else [])
        assert recs == want
