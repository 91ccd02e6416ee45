#!/bin/bash
# Round-2 second GPU session (1 GPU): full suite incl. sqllogictest runner, bench with the new e2e threads + duckdb leg, DuckDB e2e.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02b; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1; nproc >> $O/gpus.txt
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -15 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 1500 $O/bench_n1.json; tail -5 $O/bench_n1.err
for t in 1 2 4 8; do CUDASP_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-legs --host-threads $t 2> $O/e2e_t$t.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'host_threads': $t, 'e2e': d['e2e']['value'], 'value': d['value']}))" >> $O/e2e_threads.jsonl; done
cat $O/e2e_threads.jsonl
CUDASP_TRACE=1 timeout 600 python tools/duckdb_e2e.py --rows 10000000 --labels 1 --threads 8,16,32 > $O/duckdb_e2e_10m.jsonl 2> $O/duckdb_e2e_10m.err; cat $O/duckdb_e2e_10m.jsonl
python tools/cold_start.py > $O/cold_start.json 2>&1; cat $O/cold_start.json
ls -la $O
