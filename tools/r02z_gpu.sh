#!/bin/bash
# Round-2 final GPU session (one B200): the whole GPU suite, the bench line with its legs, smoke(), the reference arm.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02z; mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log; tail -3 $O/pytest_gpu.log
timeout 400 python bench.py > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"; cut -c1-600 $O/bench_n1.json
python -c "import json; d=json.loads(open('$O/bench_n1.json').readline()); print('e2e', d['e2e']['value'], 'duckdb', {k: (v.get('value') if isinstance(v, dict) else v) for k, v in d.get('duckdb', {}).items() if k in ('value', 'from_parquet', 'from_parquet_native')})"
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python bench.py --impl reference > $O/bench_reference_arm.json 2> $O/bench_reference_arm.err; echo "ref rc=$?"; cut -c1-300 $O/bench_reference_arm.json
