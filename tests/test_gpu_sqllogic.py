"""GPU (-m gpu): the reference's own end-to-end test, test/sql/cudasp.test, run UNMODIFIED by DuckDB's sqllogictest runner.

duckdb_ext/bin/unittest is the `unittest` target of the vendored DuckDB v1.4.1 (/root/reference/duckdb, BUILD_UNITTESTS=ON)
with the B200-native `cudasp` extension linked in (tools/build_duckdb.sh) -- what `make test` runs in the reference
(extension-ci-tools/makefiles/duckdb_extension.Makefile:182-207).  /root/reference does not exist on the GPU box, so the file
travels as a byte-identical fixture, tests/golden/sqllogic/test/sql/cudasp.test; tests/test_sqllogic_fixture.py (CPU) pins its
SHA-256 and, where the reference is present, compares it with the original."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "duckdb-cudasp-extension_b200", "duckdb_ext", "bin")
UNITTEST = os.path.join(BIN, "unittest")
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(UNITTEST), reason="duckdb_ext/bin/unittest not built")]


def test_reference_sqllogictest_file_unmodified(tmp_path):
    env = dict(os.environ, LD_LIBRARY_PATH=BIN + os.pathsep + os.path.join(ROOT, "duckdb-cudasp-extension_b200") + os.pathsep +
               os.environ.get("LD_LIBRARY_PATH", ""))
    p = subprocess.run([UNITTEST, "--test-dir", os.path.join(ROOT, "tests", "golden", "sqllogic"), "test/sql/cudasp.test"],
                       cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=900)
    out = p.stdout + p.stderr
    assert p.returncode == 0, out[-3000:]
    assert "All tests passed" in out, out[-3000:]
    assert "1 test case" in out, out[-3000:]        # the file was found and run, not skipped (`require cudasp` satisfied)
