"""GPU (-m gpu): BASELINE.json's configurations at FULL size through the C ABI, checked by properties that do not depend
on the size (SURVEY.md 8c "oracle plan" iii): the match set equals the planted set, sampled rows' candidates equal the
C oracle's, the result does not depend on how the table is cut into batches, a permuted table gives the permuted
result, and a second scan gives the same answer.  C2 = 1 M rows, no labels, batch_size 300000; C3 = 9.4 M rows, 1 label; C4 = 32.7 M rows, 8 labels."""
import numpy as np
import pytest

import cudasp_b200_loader
from oracle import c_oracle as CO
from tests import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def sampled_candidates_equal_oracle(M, q, T, n=4096, seed=1):
    rng = np.random.default_rng(seed)
    rows = T["tweaks"].reshape(-1, 64).shape[0]
    # planted rows (where a wrong candidate would change the answer) + random rows
    pick = np.unique(np.concatenate([T["expected"].astype(np.int64), rng.integers(0, rows, n)]))
    tw = np.ascontiguousarray(T["tweaks"].reshape(-1, 64)[pick])
    got = q.candidates(tw)
    _, want = CO.scan(T["sk"], T["spend"], T["labels"], tw, None, None, threads=8, want_cands=True)
    assert np.array_equal(got, want)


def test_c2_one_million_rows_no_labels(M):
    rows = 1000000
    T = synth.make_table(rows, 352002, n_labels=0, n_base=100, batch_rows=300000, specials=True)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
        assert np.array_equal(ids, T["expected"])
        assert np.array_equal(q.scan_batch(T["tweaks"], T["outputs"], T["offsets"]), ids)        # idempotent
        sampled_candidates_equal_oracle(M, q, T)
    # the cut into batches is invisible: ragged batch size, same ids
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=77777) as q:
        assert np.array_equal(q.scan_batch(T["tweaks"], T["outputs"], T["offsets"]), T["expected"])
    # a permuted table gives the permuted match set (rows are independent)
    rng = np.random.default_rng(7)
    perm = rng.permutation(rows)
    tw = T["tweaks"].reshape(-1, 64)
    off = T["offsets"].astype(np.int64)
    lens = (off[1:] - off[:-1])[perm]
    poff = np.zeros(rows + 1, np.int64)
    np.cumsum(lens, out=poff[1:])
    idx = np.repeat(off[:-1][perm] - poff[:-1], lens) + np.arange(poff[-1])
    pout = T["outputs"][idx]
    inv = np.empty(rows, np.int64)
    inv[perm] = np.arange(rows)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
        got = q.scan_batch(np.ascontiguousarray(tw[perm]), pout, poff.astype(np.uint32))
    assert np.array_equal(got, np.sort(inv[T["expected"].astype(np.int64)]).astype(got.dtype))


def test_c3_full_table_one_label(M):
    rows = 9400000
    T = synth.make_table(rows, 352003, n_labels=1, n_base=100, n_per_label=100, batch_rows=300000, specials=True)
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
        assert np.array_equal(ids, T["expected"])
        sampled_candidates_equal_oracle(M, q, T, n=2048, seed=3)


def test_c4_full_table_eight_labels(M):
    rows = 32700000
    T = synth.make_table(rows, 352004, n_labels=8, n_base=100, n_per_label=50, batch_rows=300000, specials=True)
    assert len(T["expected"]) >= 100 + 8 * 50
    with M.Query(T["sk"], T["spend"], T["labels"], devices=[0], batch_rows=300000) as q:
        ids = q.scan_batch(T["tweaks"], T["outputs"], T["offsets"])
        assert np.array_equal(ids, T["expected"])
        sampled_candidates_equal_oracle(M, q, T, n=1024, seed=4)
