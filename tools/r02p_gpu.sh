#!/bin/bash
# Round-2 GPU session (N GPUs, default 1): the native Parquet path -- tests, then timing against DuckDB's read_parquet on the same files.
N=${1:-1}
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02p_n$N; mkdir -p $O
T=$(nproc)
timeout 900 python -m pytest tests/test_gpu_parquet.py tests/test_gpu_duckdb.py tests/test_gpu_arrow.py -m gpu -x -q > $O/pytest_parquet.log 2>&1; echo "rc=$?" >> $O/pytest_parquet.log; tail -5 $O/pytest_parquet.log
if [ "$N" = "1" ]; then
  CUDASP_TRACE=1 timeout 900 python tools/duckdb_e2e.py --rows 10000000 --labels 1 --threads $T --native > $O/parquet_native_10m.jsonl 2> $O/err.log; grep scan_parquet $O/err.log | tail -2; cut -c1-200 $O/parquet_native_10m.jsonl; python -c "import json; print(json.loads(open('$O/parquet_native_10m.jsonl').readline())['native_parquet'])"
  timeout 900 python tools/duckdb_e2e.py --rows 1000000 --labels 0 --threads $T --native >> $O/parquet_native_10m.jsonl 2>> $O/err.log; tail -1 $O/parquet_native_10m.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['scan_s_warm'], d['native_parquet'])"
else
  CUDASP_TRACE=1 timeout 900 python tools/duckdb_e2e.py --rows 12500000 --replicate $N --labels 16 --geom-p 0.25 --threads $T --native --native-threads 48,64 > $O/parquet_native_c5.jsonl 2> $O/err.log; python -c "import json; d=json.loads(open('$O/parquet_native_c5.jsonl').readline()); print(d['scan_s_warm'], d['rows_per_s_warm'], d['native_parquet'], d.get('native_parquet_threads_48'), d.get('native_parquet_threads_64'))"; grep -h "scan_parquet" $O/err.log | tail -9 | cut -c1-700
fi
ls -la $O
