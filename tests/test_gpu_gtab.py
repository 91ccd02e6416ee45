"""GPU (-m gpu): the fixed-base table the device builds for itself (csrc/kernels.cuh k_build_gtab; it replaces the
reference's host-uploaded table of 2^i*G, gECC gsv.h:1094-1122) against the oracle: entry (w, d) must be
(d * 2^(bits*w) + r_w) * G with the published offsets r_w = SHA256("cudasp-b200/gtab-offset" || byte(w)) mod n.
The oracle side is a plain scalar multiplication (oracle/ref_scan.py pt_mul), independent of the builder's
double-and-add + addition chain + shared inversion."""
import hashlib
import random

import pytest

import cudasp_b200_loader
from oracle import ref_scan as R
from tests import synth

pytestmark = pytest.mark.gpu
EPT = 16      # rowfn.cuh GTAB_EPT: entries built by one thread (chunk borders are the interesting digits)


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def offset(w):
    return int.from_bytes(hashlib.sha256(b"cudasp-b200/gtab-offset" + bytes([w])).digest(), "big") % R.N


def want(bits, w, d):
    return R.pt_mul(((d << (bits * w)) + offset(w)) % R.N, R.G)


def test_device_table_entries_equal_the_oracle(M):
    sk, spend = synth.make_keys(5, 0)[:2]
    with M.Query(sk, spend, [], batch_rows=64) as q:
        bits = q.info()["fixed_window_bits"]
    nwin = (256 + bits - 1) // bits
    top = (1 << bits) - 1
    last_top = (1 << (256 - bits * (nwin - 1))) - 1      # the last window only holds the digits the scalar's top bits can form
    rng = random.Random(352)
    # window borders, chunk borders, the all-ones digit, random digits
    runs = [(0, 0, 2 * EPT + 1), (0, top - EPT, EPT + 1), (nwin - 1, 0, 3), (nwin - 1, last_top - 1, 2), (nwin // 2, (1 << (bits - 1)) - 2, 5)]
    runs += [(rng.randrange(nwin - 1), rng.randrange(top - 2), 3) for _ in range(24)]
    for w, d0, n in runs:
        got = M.debug_gtab(w, d0, n)
        for i, xy in enumerate(got):
            assert xy == want(bits, w, d0 + i), "table entry (%d, %d)" % (w, d0 + i)
    with pytest.raises(M.CudaspError):
        M.debug_gtab(nwin, 0, 1)
    with pytest.raises(M.CudaspError):
        M.debug_gtab(0, top, 2)
    with pytest.raises(M.CudaspError):
        M.debug_gtab(nwin - 1, last_top, 2)


def test_sum_of_entries_is_t_times_G_plus_offsets(M):
    # what K3 relies on: sum_w table[w][digit_w(t)] = t*G + (sum_w r_w)*G for any 256-bit t (also t >= n)
    sk, spend = synth.make_keys(6, 0)[:2]
    with M.Query(sk, spend, [], batch_rows=64) as q:
        bits = q.info()["fixed_window_bits"]
    nwin = (256 + bits - 1) // bits
    rng = random.Random(353)
    rsum = sum(offset(w) for w in range(nwin)) % R.N
    for t in [0, 1, (1 << 256) - 1, R.N, R.N - 1] + [rng.randrange(1 << 256) for _ in range(6)]:
        acc = None
        for w in range(nwin):
            d = (t >> (bits * w)) & ((1 << bits) - 1)
            acc = R.pt_add(acc, M.debug_gtab(w, d, 1)[0])
        assert acc == R.pt_mul((t + rsum) % R.N, R.G)
