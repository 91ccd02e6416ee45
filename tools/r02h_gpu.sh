#!/bin/bash
# Round-2 GPU session (1 GPU), after the machine was replaced: suite, label-stage batching, submit-event effect, SQL-to-SQL against the
# reference's own extension (oracle/_ref/duckdb_ref) on configs C2 / C3 / C4, bench with legs.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02h; mkdir -p $O
T=$(nproc)
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -4 $O/pytest_gpu.log
for L in 1 8 16; do
  LABELS=$L timeout 300 python tools/ladder_variants.py 0 | sed "s/^{/{\"labels\": $L, \"inv_per_thread\": \"adaptive\", /" >> $O/label_inversion.jsonl
  CUDASP_INV_PER_THREAD=16 LABELS=$L timeout 300 python tools/ladder_variants.py 0 | sed "s/^{/{\"labels\": $L, \"inv_per_thread\": 16, /" >> $O/label_inversion.jsonl
done
cat $O/label_inversion.jsonl
for m in 0 1 0 1; do
  CUDASP_SUBMIT_EVENTS=$m timeout 300 python bench.py --steps 5 --warmup 3 --no-legs 2> /dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(json.dumps({'submit_lead_event': $m, 'e2e': d['e2e']['value'], 'e2e_ms': d['e2e']['ms_per_step'], 'value': d['value']}))" >> $O/submit_events.jsonl
done
cat $O/submit_events.jsonl
if [ -x oracle/_ref/duckdb_ref ]; then
  timeout 900 python tools/duckdb_e2e.py --impl both --rows 1000000 --labels 0 --threads $T > $O/sql.jsonl 2> $O/sql.err
  timeout 900 python tools/duckdb_e2e.py --impl both --rows 1000000 --labels 0 --threads $T --in-memory >> $O/sql.jsonl 2>> $O/sql.err
  timeout 1200 python tools/duckdb_e2e.py --impl both --rows 9400000 --labels 1 --threads $T >> $O/sql.jsonl 2>> $O/sql.err
  timeout 1200 python tools/duckdb_e2e.py --impl both --rows 9400000 --labels 1 --threads $T --in-memory >> $O/sql.jsonl 2>> $O/sql.err
  timeout 1800 python tools/duckdb_e2e.py --impl both --rows 32700000 --labels 8 --threads $T --in-memory >> $O/sql.jsonl 2>> $O/sql.err
  cut -c1-330 $O/sql.jsonl
fi
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 900 $O/bench_n1.json
ls -la $O
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; tail -c 700 $O/bench_reference_arm.json
