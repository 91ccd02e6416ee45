"""GPU (-m gpu): the CUDA path, called through the C ABI, against the oracle and the golden vectors."""
import random

import numpy as np
import pytest

import cudasp_b200_loader
from oracle import c_oracle as CO
from oracle import ref_scan as R
from tests import synth

pytestmark = pytest.mark.gpu
P, N = R.P, R.N


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def I(h):
    return int(h, 16)


def pb(xy):
    return R.encode_point((I(xy[0]), I(xy[1])))


def test_field_kats_device(M, golden):
    ks = golden["field"]
    a, b = [I(k["a"]) for k in ks], [I(k["b"]) for k in ks]
    assert M.test_field(0, a, b) == [I(k["mul"]) for k in ks]
    assert M.test_field(2, a, b) == [I(k["add"]) for k in ks]
    assert M.test_field(3, a, b) == [I(k["sub"]) for k in ks]
    nz = [(x, I(k["inv"])) for x, k in zip(a, ks) if x]
    assert M.test_field(4, [x for x, _ in nz], [0] * len(nz)) == [i for _, i in nz]


def test_field_random_and_edges_device(M):
    rng = random.Random(7)
    top = (1 << 256) - 1
    edge = [0, 1, P - 1, P, P + 1, top, top - 1, (1 << 256) - (1 << 32), 1 << 255, 0x1000003D0, 0x1000003D1, 0x1000003D2,
            top - 0x1000003D1, (1 << 224) - 1, top ^ ((1 << 128) - 1), 0xFFFFFFFF, (1 << 32), (1 << 64) - 1]
    vals = edge + [rng.randrange(1 << 256) for _ in range(110)]
    a = [x for x in vals for _ in vals]
    b = [y for _ in vals for y in vals]
    assert M.test_field(0, a, b) == [x * y % P for x, y in zip(a, b)]
    assert M.test_field(1, a, b) == [x * x % P for x in a]
    assert M.test_field(2, a, b) == [(x + y) % P for x, y in zip(a, b)]
    assert M.test_field(3, a, b) == [(x - y) % P for x, y in zip(a, b)]
    inv = M.test_field(4, vals, vals)
    assert all((x % P == 0 and i == 0) or (x * i % P == 1) for x, i in zip(vals, inv))


@pytest.mark.parametrize("key", ["pmul_seed42", "pmul_gpu_dumps", "pmul_batch_seed42"])
def test_shared_secret_kats(M, golden, key):
    """S = s * P through the device ladder (intermediates API) vs the reference's own results."""
    for k in golden[key]:
        with M.Query(I(k["s"]).to_bytes(32, "little"), R.encode_point(R.G)) as q:
            S, _, _ = q.intermediates(np.frombuffer(pb([k["x"], k["y"]]), np.uint8))
        assert R.parse_point(S[0].tobytes()) == (I(k["rx"]), I(k["ry"]))


def test_pipeline_kats(M, golden):
    for k in golden["pipeline"]:
        with M.Query(I(k["sk"]).to_bytes(32, "little"), pb(k["spend"]), [pb(l) for l in k["labels"]], batch_rows=64) as q:
            tw = np.frombuffer(pb(k["tweak"]), np.uint8)
            assert q.candidates(tw)[0].tolist() == k["cands"]
            S, t, P0 = q.intermediates(tw)
        assert R.parse_point(S[0].tobytes()) == (I(k["S"][0]), I(k["S"][1]))
        assert t[0].tobytes().hex() == k["t"]
        assert R.parse_point(P0[0].tobytes()) == (I(k["P0"][0]), I(k["P0"][1]))


def _sql_table(golden, q):
    rows = golden["sql"]["test_data"][: q["rows_visible"]]
    if q["where_height"] is not None:
        rows = [r for r in rows if r["height"] == q["where_height"]]
    return [(bytes.fromhex(r["txid"]), r["height"], bytes.fromhex(r["tweak"]), r["outputs"]) for r in rows]


def test_sql_logic_cases(M, golden):
    """test/sql/cudasp.test, query by query, through the table-function mirror."""
    rep = golden["sql"]["replicated"]
    for q in golden["sql"]["queries"]:
        if q["table"] == "test_data":
            rows = _sql_table(golden, q)
        else:
            n = rep[q["table"]]
            rows = [(bytes.fromhex(rep["txid"]), rep["height"], bytes.fromhex(rep["tweak"]), rep["outputs"])] * n
        named = {"batch_size": q["batch_size"]} if q["batch_size"] else {}
        got = M.cudasp_scan(rows, bytes.fromhex(q["scan_sk"]), bytes.fromhex(q["spend"]), [bytes.fromhex(l) for l in q["labels"]], **named)
        if q["select"] == "COUNT(*)":
            assert len(got) == q["expect"], q
        else:
            assert [g[1] for g in got] == [q["expect"]], q
        for g in got:
            assert g in [(r[0], r[1], r[2]) for r in rows[:8]]


@pytest.mark.parametrize("L", [0, 1, 8])
def test_candidates_vs_oracle_random(M, L):
    T = synth.make_table(3001, 1000 + L, n_labels=L, n_base=0, specials=False)
    _, want = CO.scan(T["sk"], T["spend"], T["labels"], T["tweaks"], None, None, threads=8, want_cands=True)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=1024) as q:
        got = q.candidates(T["tweaks"])
    assert np.array_equal(got, want)


@pytest.mark.parametrize("rows,L,batch", [(1, 0, 300000), (31, 0, 7), (4097, 2, 1000), (20000, 0, 300000), (30011, 8, 8192)])
def test_match_sets_vs_planted_and_oracle(M, rows, L, batch):
    T = synth.make_table(rows, 77 + rows, n_labels=L, n_base=min(20, rows), n_per_label=3 if L else 0, batch_rows=batch,
                         specials=rows >= 200)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=batch) as q:
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
        launches = q.launch_count()
    assert np.array_equal(ids, T["expected"])
    assert launches > 0
    if rows <= 5000:
        flags, _ = CO.scan(T["sk"], T["spend"], T["labels"], T["tweaks"], T["outputs"], T["offsets"], threads=8)
        assert np.array_equal(np.nonzero(flags)[0].astype(np.uint32), ids)


def test_null_and_ragged_inputs(M):
    sk, spend, labels = synth.make_keys(5, 1)
    g = R.encode_point(R.G)
    tw = [CO.pt_mul((i + 2).to_bytes(32, "little"), g)[0] for i in range(8)]
    c = [CO.row(sk, spend, labels, t)["cands"] for t in tw]
    rows = [(b"a", 1, tw[0], [c[0][0]]), (None, 2, tw[1], [c[1][0]]), (b"c", None, tw[2], [c[2][0]]), (b"d", 4, None, [c[3][0]]),
            (b"e", 5, tw[4], None), (b"f", 6, tw[5], []), (b"g", 7, tw[6], [None, c[6][1], None]), (b"h", 8, tw[7], [c[7][0] ^ 1]),
            (b"i", 9, tw[0][:40], [c[0][0]])]
    got = M.cudasp_scan(rows, sk, spend, labels, batch_size=4)
    assert sorted(got) == sorted([(b"a", 1, tw[0]), (b"g", 7, tw[6])])
    want = R.scan_rows(sk, spend, labels, rows)
    assert sorted(got) == sorted(want)
    assert M.cudasp_scan([], sk, spend, labels) == []


def test_label_edge_cases_and_zero_scalar(M):
    sk, spend = R.encode_scalar(12345), R.encode_point(R.G)
    tw = R.encode_point(R.pt_mul(99, R.G))
    P0 = R.row_intermediates(12345, R.parse_point(tw), R.G)["P0"]
    neg, other = (P0[0], P - P0[1]), R.pt_mul(777, R.G)
    labels = [R.encode_point(neg), R.encode_point(P0), R.encode_point(other)]
    want = R.row_candidates(12345, R.parse_point(tw), R.G, [neg, P0, other])
    with M.Query(sk, spend, labels, batch_rows=16) as q:
        got = q.candidates(np.frombuffer(tw * 3, np.uint8))
    assert got.tolist() == [want] * 3
    for k in (0, N):
        z = R.row_intermediates(0, R.parse_point(tw), R.G)
        with M.Query(k.to_bytes(32, "little"), spend, [], batch_rows=16) as q:
            assert q.candidates(np.frombuffer(tw, np.uint8))[0].tolist() == z["cands"]


def test_bind_errors_match_reference_messages(M):
    ok = (b"\x01" * 32, b"\x02" * 64, [])
    cases = [((b"\x01" * 31, ok[1], ok[2]), {}, "scan_private_key must be exactly 32 bytes, got 31 bytes"),
             ((ok[0], b"\x02" * 65, ok[2]), {}, "spend_public_key must be exactly 64 bytes, got 65 bytes"),
             ((ok[0], ok[1], [b"x"]), {}, "Each label key must be exactly 64 bytes, got 1 bytes"),
             ((ok[0], ok[1], [5]), {}, "All elements in label_keys must be BLOBs"),
             ((ok[0], ok[1], ok[2]), {"batch_size": 0}, "batch_size must be positive, got 0"),
             ((ok[0], ok[1], ok[2]), {"batch_size": 10000001}, "batch_size too large (max 10,000,000), got 10000001"),
             ((ok[0], ok[1], ok[2]), {"batch_size": "7"}, "batch_size parameter must be an INTEGER")]
    for args, named, msg in cases:
        with pytest.raises(M.InvalidInputException) as e:
            M.cudasp_scan([], *args, **named)
        assert str(e.value) == msg


def test_streaming_api_and_device_entry(M):
    import ctypes as C
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    T = synth.make_table(2500, 4242, n_labels=1, n_base=12, n_per_label=4, batch_rows=1000, specials=True)
    L = M.lib()
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=1000) as q:
        got, submitted = {}, 0
        for base in range(0, 2500, 1000):
            cnt = min(1000, 2500 - base)
            s = B.Slot()
            assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
            o0, o1 = int(T["offsets"][base]), int(T["offsets"][base + cnt])
            C.memmove(s.tweak_keys, T["tweaks"][base:base + cnt].ctypes.data, cnt * 64)
            off = (T["offsets"][base:base + cnt + 1] - o0).astype(np.uint32)
            C.memmove(s.out_offsets, off.ctypes.data, (cnt + 1) * 4)
            C.memmove(s.outputs, T["outputs"][o0:o1].ctypes.data, (o1 - o0) * 8)
            assert L.cudasp_submit_slot(q.handle, C.byref(s), cnt, o1 - o0, base) == 0
            submitted += 1
        r = B.Result()
        while L.cudasp_wait(q.handle, C.byref(r)) == 1:
            got[r.tag] = [r.tag + r.match_ids[i] for i in range(r.n_matches)]
            assert L.cudasp_release_result(q.handle, C.byref(r)) == 0
        assert len(got) == submitted
        ids = sorted(x for v in got.values() for x in v)
        assert ids == T["expected"].tolist()
        # device-resident entry with torch-owned buffers
        import torch
        d_tw = torch.from_numpy(T["tweaks"][:1000].copy()).cuda()
        d_out = torch.from_numpy(T["outputs"][: int(T["offsets"][1000])].copy()).cuda()
        d_off = torch.from_numpy(T["offsets"][:1001].astype(np.int64)).to(torch.int32).cuda()
        d_ids = torch.zeros(1000, dtype=torch.int32, device="cuda")
        d_n = torch.zeros(1, dtype=torch.int32, device="cuda")
        st = torch.cuda.current_stream().cuda_stream
        q.scan_device(0, d_tw.data_ptr(), d_out.data_ptr(), d_off.data_ptr(), 1000, d_ids.data_ptr(), d_n.data_ptr(), st)
        torch.cuda.synchronize()
        n = int(d_n.item())
        assert sorted(d_ids[:n].tolist()) == [int(x) for x in T["expected"] if x < 1000]


def test_malformed_offsets_are_refused(M):
    """out_offsets index the outputs array inside K5/K7/K8: a non-monotonic or non-zero-based array is an argument error on
    every entry that takes host offsets, never an out-of-bounds device read."""
    import ctypes as C
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    T = synth.make_table(300, 99, n_labels=0, n_base=3, n_per_label=0, batch_rows=300)
    L = M.lib()
    with M.Query(T["sk"], T["spend"], [], batch_rows=300) as q:
        bad = T["offsets"].copy()
        bad[100], bad[101] = bad[101] + 5, bad[100]
        with pytest.raises(M.CudaspError, match="monoton"):
            q.scan_batch(T["tweaks"], T["outputs"], bad)
        assert np.array_equal(q.scan_batch(T["tweaks"], T["outputs"], T["offsets"]), T["expected"])
        s = B.Slot()
        assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
        n_out = int(T["offsets"][300])
        C.memmove(s.tweak_keys, T["tweaks"].ctypes.data, 300 * 64)
        C.memmove(s.outputs, T["outputs"].ctypes.data, 8 * n_out)
        shifted = (T["offsets"] + 1).astype(np.uint32)
        shifted[300] = n_out
        C.memmove(s.out_offsets, shifted.ctypes.data, 301 * 4)
        assert L.cudasp_submit_slot(q.handle, C.byref(s), 300, n_out, 0) == B.ERR_ARG
        C.memmove(s.out_offsets, T["offsets"].ctypes.data, 301 * 4)
        assert L.cudasp_submit_slot(q.handle, C.byref(s), 300, n_out, 0) == 0
        r = B.Result()
        assert L.cudasp_wait(q.handle, C.byref(r)) == 1
        assert [r.match_ids[i] for i in range(r.n_matches)] == T["expected"].tolist()
        assert L.cudasp_release_result(q.handle, C.byref(r)) == 0


def test_scan_batch_staging_threads_and_counters(M):
    """cudasp_scan_batch_mt: the match set does not depend on the number of staging threads (pieces are taken from one list by
    whichever thread is free, results merged and sorted), and the transfer counters account for every byte handed to the copies."""
    rows = 700000                                    # > 4 pieces of two ladder waves, ragged tail
    T = synth.make_table(rows, 31337, n_labels=1, n_base=60, n_per_label=30, batch_rows=300000)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
        for th in (1, 3, 8, 0):
            h0, d0 = q.transfer_bytes()
            ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"], host_threads=th)
            assert np.array_equal(ids, T["expected"]), th
            h1, d1 = q.transfer_bytes()
            # keys + outputs + one offsets array (rows + 1 entries) per piece
            assert h1 - h0 >= rows * 64 + 8 * int(T["offsets"][-1]) + 4 * (rows + 1)
            assert h1 - h0 <= rows * 64 + 8 * int(T["offsets"][-1]) + 4 * (rows + 64)
            assert 0 < d1 - d0 < 64 * 4 * 1025
