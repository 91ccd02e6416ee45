#!/bin/bash
# Round-2 last ncu pass (one B200): launch list of one bench step of the round's final build (the command runs plain first).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02y; mkdir -p $O
timeout 200 python bench.py --steps 1 --warmup 1 --rows 1200000 --no-legs > $O/bench_plain.json 2> $O/bench_plain.err; rc=$?; echo "plain rc=$rc"; cut -c1-200 $O/bench_plain.json
if [ $rc = 0 ]; then
  timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 1 --warmup 1 --rows 1200000 --no-legs > $O/ncu_bench.log 2>&1; echo "ncu rc=$?"; wc -l $O/launches.csv
fi
