#!/bin/bash
# Builds the vendored DuckDB (v1.4.1, /root/reference/duckdb) with the B200-native `cudasp` extension linked in, plus
# DuckDB's own sqllogictest runner (`unittest`), and copies what travels to the GPU box into duckdb_ext/bin/ (git-ignored):
#   duckdb                   shell with cudasp + parquet built in        (tests/test_gpu_duckdb.py, tools/duckdb_e2e.py)
#   cudasp.duckdb_extension  loadable extension (metadata footer appended by DuckDB's build)
#   unittest, libduckdb.so   sqllogictest runner for the reference's test/sql/cudasp.test, run UNMODIFIED
# Only possible where /root/reference exists (this container).  Recipe documented in INTEGRATION.md section 1.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
PKG="$ROOT/duckdb-cudasp-extension_b200"
B="$ROOT/build/duckdb"
make -s -C "$PKG/csrc"
if [ ! -f "$B/build.ninja" ]; then
  cmake -G Ninja -S /root/reference/duckdb -B "$B" -DCMAKE_BUILD_TYPE=Release \
        -DOVERRIDE_GIT_DESCRIBE=v1.4.1-0-g0000000000 -DBUILD_UNITTESTS=ON -DEXTENSION_STATIC_BUILD=1 \
        -DDUCKDB_EXTENSION_CONFIGS="$PKG/duckdb_ext/extension_config.cmake" -DCMAKE_BUILD_RPATH="$PKG"
fi
ninja -C "$B" duckdb unittest cudasp_loadable_extension
mkdir -p "$PKG/duckdb_ext/bin"
cp "$B/duckdb" "$B/extension/cudasp/cudasp.duckdb_extension" "$B/test/unittest" "$B/src/libduckdb.so" "$PKG/duckdb_ext/bin/"
ls -la "$PKG/duckdb_ext/bin/"
