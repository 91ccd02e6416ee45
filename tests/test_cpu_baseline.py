"""CPU: baseline B-cpu (baseline/cpu_scan_openssl.cpp, OpenSSL 3 EC_POINT_mul / EC_POINT_add / SHA256) gives the oracle's
candidates and flags row for row -- so the rows/s bench.py reports for it is the rate of a CORRECT scan."""
import os

import numpy as np
import pytest

from oracle import c_oracle as CO
from tests import synth

cpu_scan = pytest.importorskip("baseline.cpu_scan")
pytestmark = pytest.mark.skipif(not cpu_scan.available(), reason="OpenSSL development files not present")


@pytest.mark.parametrize("rows,L", [(4000, 0), (3000, 3)])
def test_openssl_scan_equals_oracle(rows, L):
    T = synth.make_table(rows, 9100 + L, n_labels=L, n_base=25, n_per_label=5 if L else 0, batch_rows=rows)
    th = min(os.cpu_count() or 1, 16)
    f, c = cpu_scan.scan(T["sk"], T["spend"], T["labels"], T["tweaks"], T["outputs"], T["offsets"], threads=th, want_cands=True)
    f2, c2 = CO.scan(T["sk"], T["spend"], T["labels"], T["tweaks"], T["outputs"], T["offsets"], threads=th, want_cands=True)
    assert np.array_equal(c, c2)
    assert np.array_equal(f, f2)
    assert np.array_equal(np.nonzero(f)[0], T["expected"])
