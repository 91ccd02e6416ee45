"""CPU: the N>1 path (row-range sharding, gather of match ids, max-over-ranks timing) with two gloo ranks.

The per-rank scanner here is the C oracle (the checker) standing in for a GPU: what is under test is the host logic in
host/sharding.py that bench.py / multi-process hosts use, not the arithmetic."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import cudasp_b200_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _sharding():
    cudasp_b200_loader.load()
    from importlib import import_module
    return import_module("cudasp_b200.host.sharding")


def test_shard_ranges_cover_exactly_once():
    S = _sharding()
    for rows in (0, 1, 7, 1000, 300001):
        for world in (1, 2, 3, 8):
            for align in (1, 128):
                rs = [S.shard_range(rows, r, world, align) for r in range(world)]
                assert rs[0][0] == 0 and rs[-1][1] == rows
                assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
                assert all(lo % align == 0 for lo, hi in rs if lo < rows)
    with pytest.raises(ValueError):
        S.shard_range(10, 2, 2)
    assert S.batch_ranges(5, 26, 10) == [(5, 15), (15, 25), (25, 26)]


def test_merge_matches_detects_overlap():
    S = _sharding()
    got = S.merge_matches([(100, np.array([5, 1])), (0, np.array([], dtype=np.uint32)), (0, np.array([99]))])
    assert got.tolist() == [99, 101, 105]
    with pytest.raises(ValueError):
        S.merge_matches([(0, np.array([10])), (5, np.array([5]))])


WORKER = r"""
import os, sys, time
sys.path.insert(0, {root!r})
import numpy as np
import torch.distributed as dist
import cudasp_b200_loader
cudasp_b200_loader.load()
from importlib import import_module
S = import_module("cudasp_b200.host.sharding")
from oracle import c_oracle as CO
from tests import synth
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
T = synth.make_table(1201, 4711, n_labels=1, n_base=9, n_per_label=5, batch_rows=250, specials=True)   # same seed on every rank
lo, hi = S.shard_range(1201, rank, world, align=64)
t0 = time.perf_counter()
parts = []
for b0, b1 in S.batch_ranges(lo, hi, 250):
    off, o0, o1 = S.local_csr(T["offsets"], b0, b1)
    flags, _ = CO.scan(T["sk"], T["spend"], T["labels"], T["tweaks"][b0:b1], T["outputs"][o0:o1], off, threads=2)
    parts.append((b0 - lo, np.nonzero(flags)[0]))
mine = S.merge_matches(parts)
dt = time.perf_counter() - t0
slowest = S.max_over_ranks(dt + rank, dist)          # rank 1 pretends to be 1 s slower
merged = S.gather_matches(lo, mine, dist, dst=0)
if rank == 0:
    assert merged is not None and np.array_equal(merged, T["expected"]), (merged, T["expected"])
    assert slowest >= 1.0
    print("OK", len(merged), flush=True)
else:
    assert merged is None
dist.barrier()
dist.destroy_process_group()
"""


def test_two_gloo_ranks_shard_and_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE="2", OMP_NUM_THREADS="1")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r), LOCAL_RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True) for r in range(2)]
    outs = [p.communicate(timeout=300) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert outs[0][0].startswith("OK")
