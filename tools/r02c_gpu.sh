#!/bin/bash
# Round-2 multi-GPU session (N GPUs, N = $1): one-handle tests, bench under torchrun with the one_handle + duckdb legs, cold start.
N=${1:-2}
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02c_n$N; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1; nproc >> $O/gpus.txt; free -g | head -2 >> $O/gpus.txt
timeout 600 python -m pytest tests/test_gpu_multi_device.py tests/test_gpu_duckdb.py -m gpu -q > $O/pytest_multi.log 2>&1; echo "rc=$?" >> $O/pytest_multi.log; tail -5 $O/pytest_multi.log
python tools/cold_start.py > $O/cold_start_one_device.json 2>&1; python tools/cold_start.py --all-devices > $O/cold_start_all_devices.json 2>&1; cat $O/cold_start_*.json
CUDASP_TRACE=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench.json 2> $O/bench.err; tail -c 2500 $O/bench.json; grep -E "cudasp\]|Error|error" $O/bench.err | tail -20
if [ "$N" = "8" ]; then   # config C5 through DuckDB with the extension's trace on: where the host time goes
  CUDASP_TRACE=1 timeout 900 python tools/duckdb_e2e.py --rows 12500000 --replicate 8 --labels 16 --geom-p 0.25 --threads $(nproc) > $O/duckdb_c5_parquet.jsonl 2> $O/duckdb_c5_parquet.err
  CUDASP_TRACE=1 timeout 900 python tools/duckdb_e2e.py --rows 12500000 --replicate 8 --labels 16 --geom-p 0.25 --threads $(nproc) --in-memory > $O/duckdb_c5_table.jsonl 2> $O/duckdb_c5_table.err
  cut -c1-400 $O/duckdb_c5_parquet.jsonl $O/duckdb_c5_table.jsonl
fi
ls -la $O
