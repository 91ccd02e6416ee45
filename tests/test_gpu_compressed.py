"""GPU (-m gpu): 33-byte compressed SEC1 tweak keys (include/cudasp.h CUDASP_KEYS_SEC1_33), decompressed on
the GPU inside K1, against the oracle's restatement of the reference's CPU recipe (scripts/decompress_tweak_key.py:26-46)
and against the same rows scanned in the reference's own 64-byte format."""
import numpy as np
import pytest

import cudasp_b200_loader
from oracle import ref_scan as R
from tests import synth

pytestmark = pytest.mark.gpu
P = R.P
# the key the reference decompresses by hand for its SQL test (scripts/decompress_tweak_key.py:20)
REF_COMPRESSED = bytes.fromhex("024ac253c216532e961988e2a8ce266a447c894c781e52ef6cee902361db960004")


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def compress(tweaks: np.ndarray) -> np.ndarray:
    """[rows,64] x_LE||y_LE -> [rows,33] (02|03)||x_BE."""
    tw = np.ascontiguousarray(tweaks, dtype=np.uint8).reshape(-1, 64)
    out = np.zeros((tw.shape[0], 33), np.uint8)
    out[:, 0] = 2 + (tw[:, 32] & 1)
    out[:, 1:] = tw[:, 31::-1]
    return out


def test_reference_vector_decompresses_to_the_sql_test_key(M, golden):
    xy = R.decompress_sec1(REF_COMPRESSED)
    assert xy is not None and R.on_curve(xy)
    full = R.encode_point(xy)
    sk, spend = synth.make_keys(11, 0)[:2]
    with M.Query(sk, spend, [], batch_rows=64) as q:
        a = q.candidates(np.frombuffer(full, np.uint8))
        b, valid = q.candidates(np.frombuffer(REF_COMPRESSED, np.uint8), key_format=M.KEYS_SEC1_33, with_valid=True)
    assert valid.tolist() == [True] and np.array_equal(a, b)
    # it is a tweak_key of the reference's own SQL test rows (test/sql/cudasp.test)
    sql_keys = {bytes.fromhex(r["tweak"]) for r in golden["sql"]["test_data"]} | {bytes.fromhex(golden["sql"]["replicated"]["tweak"])}
    assert full in sql_keys


def test_sql_logic_cases_with_compressed_keys(M, golden):
    """test/sql/cudasp.test with every tweak_key stored the way the index stores it (33-byte SEC1)."""
    def ck(h):
        return compress(np.frombuffer(bytes.fromhex(h), np.uint8)).tobytes()
    rep = golden["sql"]["replicated"]
    for q in golden["sql"]["queries"]:
        if q["table"] == "test_data":
            rows = golden["sql"]["test_data"][: q["rows_visible"]]
            if q["where_height"] is not None:
                rows = [r for r in rows if r["height"] == q["where_height"]]
            rows = [(bytes.fromhex(r["txid"]), r["height"], ck(r["tweak"]), r["outputs"]) for r in rows]
        else:
            rows = [(bytes.fromhex(rep["txid"]), rep["height"], ck(rep["tweak"]), rep["outputs"])] * rep[q["table"]]
        named = {"batch_size": q["batch_size"]} if q["batch_size"] else {}
        got = M.cudasp_scan(rows, bytes.fromhex(q["scan_sk"]), bytes.fromhex(q["spend"]), [bytes.fromhex(l) for l in q["labels"]], **named)
        if q["select"] == "COUNT(*)":
            assert len(got) == q["expect"], q
        else:
            assert [g[1] for g in got] == [q["expect"]], q


@pytest.mark.parametrize("L", [0, 3])
def test_candidates_compressed_equal_uncompressed(M, L):
    T = synth.make_table(4099, 4242 + L, n_labels=L, n_base=0, specials=False)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=1500) as q:
        want = q.candidates(T["tweaks"])
        got, valid = q.candidates(compress(T["tweaks"]), key_format=M.KEYS_SEC1_33, with_valid=True)
    assert valid.all()
    assert np.array_equal(got, want)


def test_malformed_compressed_keys_never_match(M):
    rows = 600
    T = synth.make_table(rows, 99, n_labels=1, n_base=rows, n_per_label=0, batch_rows=256, specials=False)   # every row matches
    assert len(T["expected"]) == rows
    ck = compress(T["tweaks"])
    bad = {}
    ck[3, 0] = 4; bad[3] = "prefix 04"
    ck[5, 0] = 0; bad[5] = "prefix 00"
    ck[7, 1:] = np.frombuffer(P.to_bytes(32, "big"), np.uint8); bad[7] = "x = p"
    ck[9, 1:] = 0xFF; bad[9] = "x = 2^256-1"
    x = 1
    while R.decompress_sec1(b"\x02" + x.to_bytes(32, "big")) is not None:
        x += 1
    ck[11, 1:] = np.frombuffer(x.to_bytes(32, "big"), np.uint8); bad[11] = "x not on the curve"
    ck[13, 0] ^= 1      # valid point, wrong parity: -T, decompresses fine but no longer matches the planted output
    for i in bad:
        assert R.decompress_sec1(ck[i].tobytes()) is None, bad[i]
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=256) as q:
        cands, valid = q.candidates(ck, key_format=M.KEYS_SEC1_33, with_valid=True)
        ids = q.scan_batch(ck, T["outputs"], T["offsets"], key_format=M.KEYS_SEC1_33)
    assert sorted(np.nonzero(~valid)[0].tolist()) == sorted(bad)
    want = np.array([i for i in range(rows) if i not in bad and i != 13], np.uint32)
    assert np.array_equal(ids, want)
    # the flipped-parity row is valid and gives the candidates of -T
    neg = (int.from_bytes(T["tweaks"][13, :32].tobytes(), "little"), P - int.from_bytes(T["tweaks"][13, 32:].tobytes(), "little"))
    lab = [R.parse_point(b) for b in T["labels"]]
    assert cands[13].tolist() == R.row_candidates(R.parse_scalar(T["sk"]), neg, R.parse_point(T["spend"]), lab)


@pytest.mark.parametrize("rows,L,batch", [(1, 0, 300000), (5000, 2, 1024), (40000, 0, 300000)])
def test_match_sets_compressed(M, rows, L, batch):
    T = synth.make_table(rows, 31337 + rows, n_labels=L, n_base=min(30, rows), n_per_label=4 if L else 0, batch_rows=batch,
                         specials=False)
    with M.Query(T["sk"], T["spend"], T["labels"], batch_rows=batch) as q:
        ids = q.scan_batch(compress(T["tweaks"]), T["outputs"], T["offsets"], key_format=M.KEYS_SEC1_33)
        ids64 = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])        # same handle, formats interleave freely
    assert np.array_equal(ids, T["expected"]) and np.array_equal(ids64, T["expected"])


def test_table_function_accepts_mixed_key_lengths(M):
    T = synth.make_table(40, 8, n_labels=1, n_base=40, batch_rows=16, specials=False)
    sk, spend, labels = T["sk"], T["spend"], T["labels"]
    ck = compress(T["tweaks"])
    off, outs = T["offsets"], T["outputs"]
    rows = []
    for i in range(40):
        key = ck[i].tobytes() if i % 3 == 0 else T["tweaks"][i].tobytes()
        if i == 6:
            key = b"\x05" + key[1:]          # malformed compressed key
        if i == 10:
            key = key[:50]                   # neither 33 nor 64 bytes
        rows.append((b"tx%02d" % i, 700000 + i, key, [int(v) for v in outs[off[i]: off[i + 1]]]))
    got = M.cudasp_scan(rows, sk, spend, labels, batch_size=16)
    want = R.scan_rows(sk, spend, labels, rows)
    assert sorted(got) == sorted(want)
    assert len(got) == 38


def test_streaming_slots_with_compressed_keys(M):
    import ctypes as C
    from importlib import import_module
    B = import_module("cudasp_b200.host.binding")
    T = synth.make_table(3000, 555, n_labels=0, n_base=25, batch_rows=1000, specials=False)
    ck = compress(T["tweaks"])
    L = M.lib()
    with M.Query(T["sk"], T["spend"], [], batch_rows=1000) as q:
        found = []
        for b in range(3):
            lo, hi = b * 1000, (b + 1) * 1000
            s = B.Slot()
            assert L.cudasp_acquire_slot(q.handle, C.byref(s)) == 0
            C.memmove(s.tweak_keys, ck[lo:hi].ctypes.data, 33 * 1000)
            o0, o1 = int(T["offsets"][lo]), int(T["offsets"][hi])
            C.memmove(s.outputs, T["outputs"][o0:o1].ctypes.data, 8 * (o1 - o0))
            lo_off = (T["offsets"][lo: hi + 1].astype(np.int64) - o0).astype(np.uint32)
            C.memmove(s.out_offsets, lo_off.ctypes.data, 4 * 1001)
            assert L.cudasp_submit_slot_fmt(q.handle, C.byref(s), 1000, o1 - o0, lo, M.KEYS_SEC1_33) == 0
        for _ in range(3):
            r = B.Result()
            assert L.cudasp_wait(q.handle, C.byref(r)) == 1
            found.extend(int(r.tag) + int(r.match_ids[i]) for i in range(r.n_matches))
            L.cudasp_release_result(q.handle, C.byref(r))
    assert sorted(found) == T["expected"].tolist()
