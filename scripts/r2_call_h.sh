#!/bin/bash
# GPU call H of round 2 (1 GPU): the K-panel synchronised Gram kernel (parity, then variants), the host thread pool of the
# host-buffer entry points, ncu evidence for the Gram and CI kernels, compute-sanitizer on the smoke case.
set -u
OUT=gpurun_out/r2_h
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_corr_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -k "not c5_merged and not c2_full" > "$OUT/tests_corr.log" 2>&1
echo "corr + C3-size tests: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_corr.py --steps 2 > "$OUT/corr_ksync128_16x9.json" 2> "$OUT/corr.err"
echo "bench_corr default (ksync 128, 16x9): exit $?" | tee -a "$OUT/summary.txt"
CUSK_GRAM_KSYNC=0 timeout 300 python scripts/bench_corr.py --steps 2 > "$OUT/corr_ksync0.json" 2>> "$OUT/corr.err"
echo "bench_corr ksync off: exit $?" | tee -a "$OUT/summary.txt"
for cfg in "64 16 9" "256 16 9" "128 12 12" "128 24 6" "128 37 4" "32 16 9"; do
  set -- $cfg
  CUSK_GRAM_KSYNC=$1 CUSK_GRAM_SB_M=$2 CUSK_GRAM_SB_N=$3 timeout 300 python scripts/bench_corr.py --steps 2 > "$OUT/corr_ksync$1_$2x$3.json" 2>> "$OUT/corr.err"
done
echo "bench_corr variants: done" | tee -a "$OUT/summary.txt"
timeout 900 python bench.py > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "bench.py: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gram_i8_kernel -s 1 -c 1 -o "$OUT/gram_c3" \
    python scripts/bench_corr.py --steps 1 --warmup 1 > "$OUT/ncu_gram.log" 2>&1
echo "ncu gram: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_gram_traffic.py "$OUT/gram_c3.ncu-rep" "$OUT/gram_traffic.json" > "$OUT/gram_traffic.log" 2>&1
python scripts/ncu_summary.py "$OUT/gram_c3.ncu-rep" > "$OUT/gram_c3_summary.txt" 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_ -c 60 -o "$OUT/prof_ci" \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-ref-gpu --no-extras --no-pipelined-probe > "$OUT/ncu_ci.log" 2>&1
echo "ncu ci kernels: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_ci_traffic.py "$OUT/prof_ci.ncu-rep" 0 "$OUT/ci_level_traffic.json" > "$OUT/ci_level_traffic.txt" 2>&1
python scripts/ncu_summary.py "$OUT/prof_ci.ncu-rep" > "$OUT/prof_ci_summary.txt" 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"degree_plan|compact_kernel|level0_tile" -c 6 -o "$OUT/prof_plan" \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-ref-gpu --no-extras --no-pipelined-probe > "$OUT/ncu_plan.log" 2>&1
echo "ncu plan/compact/level0: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_summary.py "$OUT/prof_plan.ncu-rep" > "$OUT/prof_plan_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
timeout 900 compute-sanitizer --tool memcheck python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/sanitizer_memcheck_smoke.log" 2>&1
echo "memcheck smoke: exit $?" | tee -a "$OUT/summary.txt"
timeout 1200 compute-sanitizer --tool racecheck python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/sanitizer_racecheck_smoke.log" 2>&1
echo "racecheck smoke: exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
