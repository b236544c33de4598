#!/bin/bash
# GPU call D of round 2 (1 GPU): whole suite after the sharded-exchange rework, the new bench line with all legs, a
# 200-block slice of the C4 whole genome through `mps cuskss-all`, ncu launch list of one bench step.
set -u
OUT=gpurun_out/r2_d
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/tests_all.log" 2>&1
echo "all gpu tests: exit $?" | tee -a "$OUT/summary.txt"
CUSK_SHARD_MIN_TESTS=200000 timeout 300 python -m pytest tests/test_skeleton_gpu.py -m gpu -x -q -k "sharded" > "$OUT/tests_sharded_mixed.log" 2>&1
echo "sharded tests with mixed levels: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 python bench.py > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "bench.py (all legs): exit $?" | tee -a "$OUT/summary.txt"
timeout 900 python bench.py --workload c4 --blocks 200 --steps 2 --warmup 1 --verify 10 --ref-sample 20 > "$OUT/c4_200_n1.json" 2> "$OUT/c4_200.err"
echo "bench.py --workload c4 (200 blocks): exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python bench.py --workload c4 --blocks 200 --steps 2 --warmup 1 --verify 0 --ref-sample 0 --slots 1 > "$OUT/c4_200_n1_slots1.json" 2>> "$OUT/c4_200.err"
echo "c4 200 blocks, one slot per GPU: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench_steps1.csv" \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-ref-gpu --no-extras --no-pipelined-probe > "$OUT/ncu_launches.log" 2>&1
echo "ncu launch list: exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
