"""GPU (-m gpu): the real DuckDB engine + the C++ `cudasp` extension (duckdb_ext/cudasp_extension.cpp) end to end.

duckdb_ext/bin/duckdb is the vendored DuckDB v1.4.1 shell built HERE from /root/reference/duckdb with the
B200-native `cudasp` extension linked in (recipe: INTEGRATION.md; the binary is git-ignored and travels to the
GPU box with the snapshot).  The SQL below restates the reference's only end-to-end test, test/sql/cudasp.test
(same tables, same queries, same expected values -- extracted into tests/golden/golden.json by
tests/golden/make_golden.py), plus a synthetic table with planted base/label matches and NULL handling.
"""
import os
import subprocess

import numpy as np
import pytest

from tests import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DUCKDB = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin", "duckdb")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(DUCKDB), reason="duckdb_ext/bin/duckdb not built")]


def blob(h: str) -> str:
    return "'" + "".join("\\x%s" % h[i:i + 2] for i in range(0, len(h), 2)) + "'::BLOB"


def run_sql(sql: str, threads=None):
    pre = "PRAGMA threads=%d;\n" % threads if threads else ""
    p = subprocess.run([DUCKDB, "-csv", "-noheader"], input=pre + sql, capture_output=True, text=True, timeout=600)
    return p.returncode, p.stdout.strip().splitlines(), p.stderr


def scan_call(table, q, where=""):
    labels = "[" + ", ".join(blob(l) for l in q["labels"]) + "]" + ("" if q["labels"] else "::BLOB[]")
    named = ", batch_size := %d" % q["batch_size"] if q["batch_size"] else ""
    return "cudasp_scan((SELECT txid, height, tweak_key, outputs FROM %s%s), %s, %s, %s%s)" % (
        table, where, blob(q["scan_sk"]), blob(q["spend"]), labels, named)


def test_sql_logic_file_through_duckdb(golden):
    s = golden["sql"]
    ddl = ["CREATE TABLE test_data (txid BLOB, height INTEGER, tweak_key BLOB, outputs BIGINT[]);"]
    inserted = 0
    rep = s["replicated"]
    out_lines = []
    script = list(ddl)
    expects = []
    for q in s["queries"]:
        if q["table"] == "test_data":
            while inserted < q["rows_visible"]:
                r = s["test_data"][inserted]
                script.append("INSERT INTO test_data VALUES (%s, %d, %s, [%s]);" % (
                    blob(r["txid"]), r["height"], blob(r["tweak"]), ", ".join("%d::BIGINT" % o for o in r["outputs"])))
                inserted += 1
            table = "test_data"
        else:
            table = q["table"]
            if ("CREATE TABLE %s" % table) not in "\n".join(script):
                script.append("CREATE TABLE %s AS SELECT %s AS txid, %d AS height, %s AS tweak_key, [%s] AS outputs FROM range(%d);" % (
                    table, blob(rep["txid"]), rep["height"], blob(rep["tweak"]),
                    ", ".join("%d::BIGINT" % o for o in rep["outputs"]), rep[table]))
        where = " WHERE height = %d" % q["where_height"] if q["where_height"] is not None else ""
        sel = "COUNT(*)" if q["select"] == "COUNT(*)" else "height"
        script.append("SELECT %s FROM %s;" % (sel, scan_call(table, q, where)))
        expects.append(q["expect"])
    rc, lines, err = run_sql("\n".join(script))
    assert rc == 0, err
    assert [int(x) for x in lines] == [int(e) for e in expects]


def test_bind_errors_through_duckdb():
    ok = dict(scan_sk="01" * 32, spend="02" * 64, labels=[], batch_size=None)
    src = "(SELECT '\\x01'::BLOB AS txid, 1 AS height, '\\x02'::BLOB AS tweak_key, [1::BIGINT] AS outputs)"
    cases = [("cudasp_scan(%s, %s, %s, []::BLOB[])" % (src, blob("01" * 31), blob(ok["spend"])), "scan_private_key must be exactly 32 bytes, got 31 bytes"),
             ("cudasp_scan(%s, %s, %s, []::BLOB[])" % (src, blob(ok["scan_sk"]), blob("02" * 65)), "spend_public_key must be exactly 64 bytes, got 65 bytes"),
             ("cudasp_scan(%s, %s, %s, ['\\x01'::BLOB])" % (src, blob(ok["scan_sk"]), blob(ok["spend"])), "Each label key must be exactly 64 bytes, got 1 bytes"),
             ("cudasp_scan(%s, %s, %s, []::BLOB[], batch_size := 0)" % (src, blob(ok["scan_sk"]), blob(ok["spend"])), "batch_size must be positive, got 0"),
             ("cudasp_scan(%s, %s, %s, []::BLOB[], batch_size := 10000001)" % (src, blob(ok["scan_sk"]), blob(ok["spend"])),
              "batch_size too large (max 10,000,000), got 10000001")]
    for call, msg in cases:
        rc, _, err = run_sql("SELECT * FROM %s;" % call)
        assert rc != 0 and msg in err, (call, err)


@pytest.mark.parametrize("threads,batch", [(1, 1000), (8, 700), (16, 300000)])
def test_synthetic_table_through_duckdb(tmp_path, threads, batch):
    """Planted base + label matches, NULL/empty/NULL-child outputs, NULL tweak rows, many worker threads, ragged batches."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    rows, L = 40000, 2
    T = synth.make_table(rows, 5150 + threads, n_labels=L, n_base=30, n_per_label=10, batch_rows=batch, specials=True)
    off = T["offsets"].astype(np.int64)
    outs = [T["outputs"][off[i]:off[i + 1]].tolist() for i in range(rows)]
    tweaks = [T["tweaks"][i].tobytes() for i in range(rows)]
    txids = [i.to_bytes(4, "little") * 8 for i in range(rows)]
    heights = [800000 + i // 3000 for i in range(rows)]
    expected = set(int(x) for x in T["expected"])
    # special rows: NULL list (a planted row loses its match), NULL children around a planted value, NULL tweak
    victim = int(T["expected"][3])
    outs[victim] = None
    expected.discard(victim)
    keep = int(T["expected"][5])
    outs[keep] = [None] + outs[keep] + [None]
    gone = int(T["expected"][7])
    tweaks[gone] = None
    expected.discard(gone)
    tbl = pa.table({"txid": pa.array(txids, pa.binary()), "height": pa.array(heights, pa.int32()),
                    "tweak_key": pa.array(tweaks, pa.binary()), "outputs": pa.array(outs, pa.list_(pa.int64()))})
    path = str(tmp_path / "t.parquet")
    pq.write_table(tbl, path, row_group_size=5000)
    q = dict(scan_sk=T["sk"].hex(), spend=T["spend"].hex(), labels=[l.hex() for l in T["labels"]], batch_size=batch)
    sql = "SELECT hex(txid), height, hex(tweak_key) FROM %s ORDER BY 1;" % scan_call("read_parquet('%s')" % path, q)
    rc, lines, err = run_sql(sql, threads=threads)
    assert rc == 0, err
    want = sorted((txids[i].hex().upper(), heights[i], tweaks[i].hex().upper()) for i in expected)
    got = sorted((a, int(b), c) for a, b, c in (ln.split(",") for ln in lines))
    assert got == want


@pytest.mark.parametrize("mode,threads", [("compressed", 8), ("mixed", 4)])
def test_compressed_tweak_keys_through_duckdb(tmp_path, mode, threads):
    """33-byte SEC1 tweak keys (the index's storage format) scanned directly; 'mixed' alternates runs of both key
    lengths (each change cuts a batch) and sprinkles malformed keys that must never match."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    from tests.test_gpu_compressed import compress
    rows, L, batch = 30000, 1, 4096
    T = synth.make_table(rows, 6060, n_labels=L, n_base=40, n_per_label=10, batch_rows=batch, specials=False)
    off = T["offsets"].astype(np.int64)
    outs = [T["outputs"][off[i]:off[i + 1]].tolist() for i in range(rows)]
    ck = compress(T["tweaks"])
    expected = set(int(x) for x in T["expected"])
    tweaks = []
    for i in range(rows):
        use33 = mode == "compressed" or (i // 777) % 2 == 0
        tweaks.append(ck[i].tobytes() if use33 else T["tweaks"][i].tobytes())
    if mode == "mixed":
        for k, i in enumerate(sorted(expected)[:6]):
            if len(tweaks[i]) == 33:
                tweaks[i] = (b"\x04" if k % 2 else b"\x00") + tweaks[i][1:]       # bad prefix
            else:
                tweaks[i] = tweaks[i][:63]                                        # neither length
            expected.discard(i)
    txids = [i.to_bytes(4, "little") * 8 for i in range(rows)]
    heights = [800000 + i // 3000 for i in range(rows)]
    tbl = pa.table({"txid": pa.array(txids, pa.binary()), "height": pa.array(heights, pa.int32()),
                    "tweak_key": pa.array(tweaks, pa.binary()), "outputs": pa.array(outs, pa.list_(pa.int64()))})
    path = str(tmp_path / "c.parquet")
    pq.write_table(tbl, path, row_group_size=5000)
    q = dict(scan_sk=T["sk"].hex(), spend=T["spend"].hex(), labels=[l.hex() for l in T["labels"]], batch_size=batch)
    sql = "SELECT hex(txid), height, hex(tweak_key) FROM %s ORDER BY 1;" % scan_call("read_parquet('%s')" % path, q)
    rc, lines, err = run_sql(sql, threads=threads)
    assert rc == 0, err
    want = sorted((txids[i].hex().upper(), heights[i], tweaks[i].hex().upper()) for i in expected)
    got = sorted((a, int(b), c) for a, b, c in (ln.split(",") for ln in lines))
    assert got == want


def test_details_parameter_through_duckdb(tmp_path):
    """details := true (extension): one result row per found output with k, output_index, label_index and t_k; the
    default call on the same table still returns the reference's three columns, one row per transaction."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    from tests.test_gpu_details import build_table
    rows = 6000
    T, want = build_table(rows, 2, 31415, 40)
    txids = [i.to_bytes(4, "little") * 8 for i in range(rows)]
    tbl = pa.table({"txid": pa.array(txids, pa.binary()), "height": pa.array([800000 + i for i in range(rows)], pa.int32()),
                    "tweak_key": pa.array([T["tweaks"][i].tobytes() for i in range(rows)], pa.binary()),
                    "outputs": pa.array(T["lists"], pa.list_(pa.int64()))})
    path = str(tmp_path / "d.parquet")
    pq.write_table(tbl, path, row_group_size=1500)
    q = dict(scan_sk=T["sk"].hex(), spend=T["spend"].hex(), labels=[l.hex() for l in T["labels"]], batch_size=1000)
    call = scan_call("read_parquet('%s')" % path, q)
    rc, lines, err = run_sql("SELECT height, k, output_index, label_index, hex(t_k) FROM %s ORDER BY 1, 2;" % call.replace(
        "batch_size := 1000", "batch_size := 1000, details := true"), threads=4)
    assert rc == 0, err
    got = [tuple(ln.split(",")) for ln in lines]
    exp = sorted((800000 + r, k, oi, li, "%064X" % t) for r, recs in want.items() for k, oi, li, t in recs)
    assert [(int(a), int(b), int(c), int(d), e) for a, b, c, d, e in got] == exp
    rc, lines, err = run_sql("SELECT height FROM %s ORDER BY 1;" % call, threads=4)
    assert rc == 0, err
    assert [int(x) for x in lines] == sorted(800000 + r for r in want)
    # details := false is the default schema
    rc, lines, err = run_sql("SELECT count(*) FROM %s;" % call.replace("batch_size := 1000", "details := false"), threads=2)
    assert rc == 0 and int(lines[0]) == len(want), err


@pytest.mark.parametrize("threads", [1, 8])
def test_emit_early_returns_the_same_rows(tmp_path, threads):
    """emit_early := true (extension): matches leave during Execute; the result set must not change.  Every row of the
    table matches here, so full output vectors (HAVE_MORE_OUTPUT re-entry with the same input chunk) are exercised."""
    import pyarrow as pa
    import pyarrow.parquet as pq
    rows = 30000
    T = synth.make_table(rows, 2718, n_labels=1, n_base=rows, n_per_label=0, batch_rows=2048, specials=False)
    assert len(T["expected"]) == rows
    off = T["offsets"].astype(np.int64)
    tbl = pa.table({"txid": pa.array([i.to_bytes(4, "little") * 8 for i in range(rows)], pa.binary()),
                    "height": pa.array(np.arange(rows, dtype=np.int32)),
                    "tweak_key": pa.array([T["tweaks"][i].tobytes() for i in range(rows)], pa.binary()),
                    "outputs": pa.array([T["outputs"][off[i]:off[i + 1]].tolist() for i in range(rows)], pa.list_(pa.int64()))})
    path = str(tmp_path / "e.parquet")
    pq.write_table(tbl, path, row_group_size=4096)
    q = dict(scan_sk=T["sk"].hex(), spend=T["spend"].hex(), labels=[l.hex() for l in T["labels"]], batch_size=2048)
    call = scan_call("read_parquet('%s')" % path, q)
    early = call.replace("batch_size := 2048", "batch_size := 2048, emit_early := true")
    for c in (early, early.replace("emit_early := true", "emit_early := true, details := true")):
        rc, lines, err = run_sql("SELECT count(*), count(DISTINCT height), min(height), max(height), sum(height) FROM %s;" % c, threads=threads)
        assert rc == 0, err
        assert [int(x) for x in lines[0].split(",")] == [rows, rows, 0, rows - 1, rows * (rows - 1) // 2]
    # a LIMIT can now be satisfied before the scan has finished (no assertion on time, only on the answer)
    rc, lines, err = run_sql("SELECT height FROM %s LIMIT 5;" % early, threads=threads)
    assert rc == 0 and len(lines) == 5, err
