#!/bin/bash
# GPU call B of round 2: the new lanes-over-neighbours CI kernel (ci_dfs.cu).
#   gpurun --timeout 1500 -- 'bash scripts/r2_call_b.sh'
# 1. parity: the kernel's own tests, the skeleton suite, the config-size cases
# 2. headline bench (no regression on the light C2 block)
# 3. hub-heavy merged skeletons, ci_dfs on (default) against off (CUSK_DFS=0: the round-1 routing), same input, sha1 compared
# 4. ncu --set full of the new kernel (never timed under ncu)
set -u
OUT=gpurun_out/r2_b
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zz_dfs_kernel_gpu.py tests/test_skeleton_gpu.py -m gpu -x -q > "$OUT/tests_dfs.log" 2>&1
echo "dfs + skeleton tests: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python -m pytest tests/test_zzzzz_config_size_gpu.py tests/test_zzzz_c1_gpu.py tests/test_mps_gpu.py -m gpu -x -q -k "not c3_size" > "$OUT/tests_config.log" 2>&1
echo "config-size + mps tests: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python bench.py --no-pipelined-probe > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "bench.py: exit $?" | tee -a "$OUT/summary.txt"
H30="--loci 100 --causal 30 --steps 1 --warmup 1"
H36="--loci 100 --causal 36 --beta-lo 0.02 --beta-hi 0.06 --steps 1 --warmup 1 --maxlevel 9"
timeout 300 python scripts/bench_merged.py $H30 --uniform --tag heavy30_uniform_dfs > "$OUT/heavy30_uniform_dfs.json" 2> "$OUT/heavy30.err"
echo "heavy30 uniform dfs: exit $?" | tee -a "$OUT/summary.txt"
CUSK_DFS=0 timeout 300 python scripts/bench_merged.py $H30 --uniform --tag heavy30_uniform_nodfs > "$OUT/heavy30_uniform_nodfs.json" 2>> "$OUT/heavy30.err"
echo "heavy30 uniform no dfs: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_merged.py $H30 --tag heavy30_het_dfs > "$OUT/heavy30_het_dfs.json" 2>> "$OUT/heavy30.err"
echo "heavy30 het dfs: exit $?" | tee -a "$OUT/summary.txt"
CUSK_DFS=0 timeout 300 python scripts/bench_merged.py $H30 --tag heavy30_het_nodfs > "$OUT/heavy30_het_nodfs.json" 2>> "$OUT/heavy30.err"
echo "heavy30 het no dfs: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_merged.py $H36 --uniform --tag heavy36_l9_uniform_dfs > "$OUT/heavy36_l9_uniform_dfs.json" 2> "$OUT/heavy36.err"
echo "heavy36 (max level 9) uniform dfs: exit $?" | tee -a "$OUT/summary.txt"
CUSK_DFS=0 timeout 400 python scripts/bench_merged.py $H36 --uniform --tag heavy36_l9_uniform_nodfs > "$OUT/heavy36_l9_uniform_nodfs.json" 2>> "$OUT/heavy36.err"
echo "heavy36 (max level 9) uniform no dfs: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_merged.py $H36 --tag heavy36_l9_het_dfs > "$OUT/heavy36_l9_het_dfs.json" 2>> "$OUT/heavy36.err"
echo "heavy36 (max level 9) het dfs: exit $?" | tee -a "$OUT/summary.txt"
for s in 1 2 4 6; do
  CUSK_DFS_MIN_SPAN=$s timeout 300 python scripts/bench_merged.py $H30 --uniform --tag heavy30_uniform_span$s > "$OUT/heavy30_uniform_span$s.json" 2>> "$OUT/heavy30.err"
done
echo "span sweep done" | tee -a "$OUT/summary.txt"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_dfs_kernel -c 8 -o "$OUT/ci_dfs_full" \
    python scripts/bench_merged.py --loci 100 --causal 36 --beta-lo 0.02 --beta-hi 0.06 --steps 1 --warmup 0 --maxlevel 7 --uniform > "$OUT/ncu_dfs.log" 2>&1
echo "ncu --set full ci_dfs_kernel: exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
