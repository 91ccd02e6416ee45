"""CPU: the sqllogictest fixture is the reference's test/sql/cudasp.test, byte for byte."""
import hashlib
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "sqllogic")
REF = "/root/reference/test/sql/cudasp.test"


def test_fixture_is_the_reference_file():
    data = open(os.path.join(FIX, "test", "sql", "cudasp.test"), "rb").read()
    want = open(os.path.join(FIX, "SHA256")).read().split()[0]
    assert hashlib.sha256(data).hexdigest() == want
    if os.path.exists(REF):          # build container only
        assert data == open(REF, "rb").read()
