#!/bin/bash
# Round-2 GPU session: where the native Parquet path's time goes (trace timeline), ramp policies (tuning aid CUDASP_PARQUET_RAMP).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02q; mkdir -p $O
T=$(nproc)
for R in default -1 0; do
  for ROWS in 1000000 10000000; do
    L=$([ $ROWS = 1000000 ] && echo 0 || echo 1)
    if [ $R = default ]; then unset CUDASP_PARQUET_RAMP; else export CUDASP_PARQUET_RAMP=$R; fi
    CUDASP_TRACE=1 timeout 600 python tools/duckdb_e2e.py --rows $ROWS --labels $L --threads $T --native > $O/native_${ROWS}_ramp$R.jsonl 2> $O/native_${ROWS}_ramp$R.err
    echo "ramp=$R rows=$ROWS"; grep -h "scan_parquet" $O/native_${ROWS}_ramp$R.err | tail -2 | cut -c1-600
    python -c "import json; d=json.loads(open('$O/native_${ROWS}_ramp$R.jsonl').readline()); print(d['scan_s_warm'], d['native_parquet'])"
  done
done
