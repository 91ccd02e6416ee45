"""CPU: the C-ABI library loads and exports every symbol include/cudasp.h declares; the host mirror's
argument validation; no compute without a GPU (fails loudly, no fallback)."""
import os
import re

import pytest

import cudasp_b200_loader

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def M():
    import __graft_entry__ as ge
    if not os.path.exists(os.path.join(ROOT, "duckdb-cudasp-extension_b200", "libcudasp_b200.so")):
        ge.build()
    return cudasp_b200_loader.load()


def test_exports_match_header(M):
    hdr = open(os.path.join(ROOT, "include", "cudasp.h")).read()
    declared = set(re.findall(r"\b(cudasp_[a-z_]+)\s*\(", hdr))
    assert declared == set(M.EXPORTS), declared ^ set(M.EXPORTS)
    L = M.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert b"sm_100a" in L.cudasp_version()


def test_arrow_exports_match_header(M):
    hdr = open(os.path.join(ROOT, "include", "cudasp_arrow.h")).read()
    declared = set(re.findall(r"\b(cudasp_[a-z_]+)\s*\(", hdr))
    assert declared == set(M.ARROW_EXPORTS), declared ^ set(M.ARROW_EXPORTS)
    L = M.lib()
    for name in declared:
        assert getattr(L, name) is not None
    # the internal staging interface between the two translation units stays internal
    import subprocess
    syms = subprocess.run(["nm", "-D", "--defined-only", M.LIB_PATH], capture_output=True, text=True).stdout
    assert "staging" not in syms


def test_parquet_exports_match_header(M):
    hdr = open(os.path.join(ROOT, "include", "cudasp_parquet.h")).read()
    declared = set(re.findall(r"\b(cudasp_[a-z_]+)\s*\(", hdr)) - {"cudasp_last_error", "cudasp_scan"}      # mentioned in comments
    assert declared == set(M.PARQUET_EXPORTS), declared ^ set(M.PARQUET_EXPORTS)
    L = M.lib()
    for name in declared:
        assert getattr(L, name) is not None


def test_no_cpu_fallback(M):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(M.CudaspError) as e:
        M.Query(b"\x01" * 32, b"\x02" * 64)
    assert "no CUDA device" in str(e.value)
    with pytest.raises(M.IOException):
        M.cudasp_scan([], b"\x01" * 32, b"\x02" * 64, [])


def test_bind_validation_messages(M):
    B = M.CudaspScanBind
    with pytest.raises(M.InvalidInputException, match="requires 4 arguments"):
        B([None, b"", b""])
    with pytest.raises(M.InvalidInputException, match="Second argument must be a BLOB"):
        B([None, 5, b"\0" * 64, []])
    with pytest.raises(M.InvalidInputException, match="Third argument must be a BLOB"):
        B([None, b"\0" * 32, "x", []])
    with pytest.raises(M.InvalidInputException, match="Fourth argument must be a LIST"):
        B([None, b"\0" * 32, b"\0" * 64, b"notalist"])
    bd = B([None, b"\0" * 32, b"\0" * 64, [b"\1" * 64]], {"batch_size": 50000})
    assert bd.batch_size == 50000 and bd.names == ["txid", "height", "tweak_key"]
    assert B([None, b"\0" * 32, b"\0" * 64, []]).batch_size == 300000


def test_product_does_not_touch_oracle():
    """The product tree must not import, link or mention the oracle (tier rule 3)."""
    pkg = os.path.join(ROOT, "duckdb-cudasp-extension_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")):
                src = open(os.path.join(d, f), errors="ignore").read()
                assert "oracle" not in src.lower().replace("test-suite", ""), os.path.join(d, f)
