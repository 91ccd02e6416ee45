"""GPU (-m gpu): cudasp_batch_inverse_static -- the allocation-free entry point (include/cudasp.h; csrc/noalloc.cu: static __device__
scratch, address taken with cudaGetSymbolAddress once per device) -- against the existing field-KAT entry cudasp_test_field(op = inv),
which cudaMallocs its buffers.  Integer work: the comparison is bit-exact.  Two calls in a row share the one scratch array; the second
must not be disturbed by what the first left behind.  Named test_zz_* so that pytest collects it last (the round's end runs -x)."""
import ctypes as C
import random

import numpy as np
import pytest

import cudasp_b200_loader

pytestmark = pytest.mark.gpu
P = 2 ** 256 - 2 ** 32 - 977
MAX = 65536        # CUDASP_STATIC_INV_MAX


@pytest.fixture(scope="module")
def M():
    m = cudasp_b200_loader.load()
    assert m.lib().cudasp_device_count() >= 1, "no CUDA device: the product has no CPU path"
    return m


def le_bytes(vals):
    return np.frombuffer(b"".join(v.to_bytes(32, "little") for v in vals), np.uint8).copy()


def test_static_batch_inverse_equals_the_allocating_entry_twice_in_a_row(M):
    import torch
    L = M.lib()
    rng = random.Random(20261105)
    edge = [0, 1, 2, P - 1, P - 2, P, P + 1, 2 ** 256 - 1, 977, 2 ** 32 + 977, 2 ** 255]      # weakly reduced inputs >= p are legal
    calls = [edge + [rng.randrange(2 ** 256) for _ in range(5000 - len(edge))],                 # several elements per thread
             [rng.randrange(P) for _ in range(777)] + [0, 0, 1]]                                # a shorter call right after it
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    for k, vals in enumerate(calls):
        n = len(vals)
        want = [pow(v % P, P - 2, P) for v in vals]
        assert M.test_field(4, vals, [0] * n) == want                                          # the existing (allocating) entry
        d_in = torch.from_numpy(le_bytes(vals)).cuda()
        d_out = torch.full((32 * n,), 0xEE, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        mem0 = torch.cuda.mem_get_info()[0]
        rc = L.cudasp_batch_inverse_static(0, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_out.data_ptr()), n, C.c_void_p(stream.cuda_stream))
        assert rc == 0, L.cudasp_last_error(None)
        stream.synchronize()
        if k:                # (the first call of a process loads the module's static data; after that free memory does not move)
            assert torch.cuda.mem_get_info()[0] == mem0
        got = d_out.cpu().numpy()
        assert [int.from_bytes(got[32 * i: 32 * i + 32].tobytes(), "little") for i in range(n)] == want
        # in place, on the default stream
        assert L.cudasp_batch_inverse_static(0, C.c_void_p(d_in.data_ptr()), C.c_void_p(d_in.data_ptr()), n, None) == 0
        torch.cuda.synchronize()
        assert np.array_equal(d_in.cpu().numpy(), got)


def test_static_batch_inverse_limits(M):
    import torch
    L = M.lib()
    torch.cuda.set_device(0)
    buf = torch.zeros(32 * (MAX + 1), dtype=torch.uint8, device="cuda")
    p = C.c_void_p(buf.data_ptr())
    assert L.cudasp_batch_inverse_static(0, p, p, MAX + 1, None) == -1                          # CUDASP_ERR_ARG
    assert b"CUDASP_STATIC_INV_MAX" in L.cudasp_last_error(None)
    assert L.cudasp_batch_inverse_static(0, None, p, 4, None) == -1
    assert L.cudasp_batch_inverse_static(0, p, p, 0, None) == 0
    # the maximum itself: all ones -> all ones
    ones = np.zeros((MAX, 32), np.uint8)
    ones[:, 0] = 1
    buf[: 32 * MAX] = torch.from_numpy(ones.reshape(-1)).cuda()
    assert L.cudasp_batch_inverse_static(0, p, p, MAX, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(buf[: 32 * MAX].cpu().numpy().reshape(MAX, 32), ones)
