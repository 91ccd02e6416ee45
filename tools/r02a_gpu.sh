#!/bin/bash
# Round-2 first GPU session (1 GPU): suite, bench, K1 formula / shape variants, K3 fetch granularity, inversion batching, ncu.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/r02a; mkdir -p $O
PKG=duckdb-cudasp-extension_b200
nvidia-smi -L > $O/gpus.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 600 python bench.py --steps 3 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; tail -c 600 $O/bench_n1.json
for v in "" _madd _dbl _both; do
  CUDASP_B200_LIB=$PWD/$PKG/libcudasp_b200$v.so LABELS=1 timeout 300 python tools/ladder_variants.py 0 | sed "s/^{/{\"lib\": \"libcudasp_b200$v\", /" >> $O/formula_variants.jsonl
done
CUDASP_B200_LIB=$PWD/$PKG/libcudasp_b200_tune.so LABELS=1 timeout 600 python tools/ladder_variants.py 0 1 2 3 >> $O/shape_variants.jsonl
for g in 32 128; do
  CUDASP_L2_FETCH=$g LABELS=1 timeout 300 python tools/ladder_variants.py 0 | sed "s/^{/{\"l2_fetch\": $g, /" >> $O/l2_fetch.jsonl
done
for p in 8 16 32; do
  CUDASP_INV_PER_THREAD=$p LABELS=1 timeout 300 python tools/ladder_variants.py 0 | sed "s/^{/{\"inv_per_thread\": $p, /" >> $O/inv_per_thread.jsonl
done
cat $O/formula_variants.jsonl $O/shape_variants.jsonl $O/l2_fetch.jsonl $O/inv_per_thread.jsonl
# ncu: the shipped build, one 300 K-row batch with one label (same command ran plain just above)
export _LV_CHILD=1 LABELS=1
ncu --set full --clock-control none --import-source on -k regex:k_ladder -s 2 -c 1 -o $O/prof_k_ladder -f python tools/ladder_variants.py > $O/ncu_ladder.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_hash_fixedbase -s 2 -c 1 -o $O/prof_k_hash -f python tools/ladder_variants.py > $O/ncu_hash.log 2>&1
ncu --metrics sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_uniform.sum,sm__inst_executed.sum,smsp__inst_issued.sum,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmalite_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum \
    --clock-control none -k regex:"k_ladder|k_hash_fixedbase|k_batch_inv|k_final" -s 7 -c 7 --csv --log-file $O/pipes.csv python tools/ladder_variants.py > $O/ncu_pipes.log 2>&1
for g in 32 128; do
  CUDASP_L2_FETCH=$g ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_hash_fixedbase -s 2 -c 1 --csv --log-file $O/k3_dram_fetch$g.csv python tools/ladder_variants.py > $O/ncu_k3_$g.log 2>&1
done
unset _LV_CHILD
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 1 --warmup 1 --rows 1200000 > $O/ncu_bench.log 2>&1
ls -la $O
