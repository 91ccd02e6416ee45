"""CPU: damaged Parquet files never crash the library's own page decoders (csrc/parquet_ingest.cpp).  The mutations run in a child
process (tests/hostcheck/parquet_fuzz.py) so that a crash is a failed assertion here, not the end of the test session.

The same runner under the sanitizers (how the decoders were hardened; 35 000 mutations clean):
    g++ -O1 -g -std=c++17 -fPIC -shared -fsanitize=address,undefined -Iinclude -o /tmp/pq_asan.so \
        duckdb-cudasp-extension_b200/csrc/{parquet_ingest,arrow_ingest}.cpp tests/hostcheck/mock_staging.cpp -lpthread
    LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libstdc++.so)" ASAN_OPTIONS=detect_leaks=0 \
        python tests/hostcheck/parquet_fuzz.py /tmp/pq_asan.so 8000 1 /tmp/fz
"""
import os
import subprocess
import sys

import pytest

pa = pytest.importorskip("pyarrow")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("seed", [11, 12])
def test_damaged_files_are_refused_or_scanned_never_crash(tmp_path, seed):
    from tests import test_parquet_ingest as T
    if not os.path.exists(T.SO) or os.path.getmtime(T.SO) < max(os.path.getmtime(d) for d in T.DEPS):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-o", T.SO] + T.SRCS + ["-lpthread"])
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "hostcheck", "parquet_fuzz.py"), T.SO, "400", str(seed), str(tmp_path)],
                       capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, (p.returncode, p.stderr[-2000:])
    assert p.stdout.strip().startswith("ok"), p.stdout[-500:]
    outcomes = eval(p.stdout.strip()[3:])            # {return code: count}
    assert set(outcomes) <= {0, -5, -6} and sum(outcomes.values()) == 400
    assert outcomes.get(-6, 0) > 20 and outcomes.get(0, 0) > 20      # the mutations do reach both the refusals and the decoders
