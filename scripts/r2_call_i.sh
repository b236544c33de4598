#!/bin/bash
# GPU call I (1 GPU): Gram kernel after the wave padding was dropped; host team variants of the host-buffer call.
set -u
OUT=gpurun_out/r2_i
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_corr_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -k "not c5_merged and not c2_full" > "$OUT/tests_corr.log" 2>&1
echo "corr tests: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 python -m pytest tests/test_skeleton_gpu.py tests/test_zz_dfs_kernel_gpu.py tests/test_zzzz_c1_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -k "not c3_size" > "$OUT/tests_skeleton.log" 2>&1
echo "skeleton tests (level-1 kernel rewrite): exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_corr.py --steps 3 > "$OUT/corr_default.json" 2> "$OUT/corr.err"
CUSK_GRAM_KSYNC=256 timeout 300 python scripts/bench_corr.py --steps 3 > "$OUT/corr_ksync256.json" 2>> "$OUT/corr.err"
CUSK_GRAM_KSYNC=0 timeout 300 python scripts/bench_corr.py --steps 3 > "$OUT/corr_ksync0.json" 2>> "$OUT/corr.err"
echo "bench_corr: done" | tee -a "$OUT/summary.txt"
for v in pool omp; do
  CUSK_HOST_POOL=$v timeout 300 python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/bench_host_$v.json" 2>> "$OUT/bench.err"
done
CUSK_HOST_THREADS=8 timeout 300 python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/bench_host_pool8.json" 2>> "$OUT/bench.err"
CUSK_FUSE_PLAN=0 timeout 300 python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/bench_nofuseplan.json" 2>> "$OUT/bench.err"
CUSK_FUSE_PLAN=0 timeout 600 python -m pytest tests/test_skeleton_gpu.py -m gpu -q > "$OUT/tests_skeleton_nofuseplan.log" 2>&1
echo "skeleton tests, plan and compaction as two launches: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench_steps1.csv" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_bench.log" 2>&1
echo "bench host variants: done" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
