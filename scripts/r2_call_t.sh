#!/bin/bash
# GPU call T (1 GPU): multi-level launches v2 (items cut in the row phase, batched bitmap loads), level-2 side stream, plain-launch probe.
set -u
OUT=gpurun_out/r2_t
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zz_multi_level_gpu.py -m gpu -q -x > "$OUT/tests_multi.log" 2>&1
echo "multi-level tests: exit $?" | tee -a "$OUT/summary.txt"
tail -n 15 "$OUT/tests_multi.log"
timeout 1200 python -m pytest tests/test_skeleton_gpu.py tests/test_zz_dfs_kernel_gpu.py tests/test_zzzz_c1_gpu.py tests/test_zzzzz_config_size_gpu.py tests/test_mps_gpu.py -m gpu -q -x -k "not c3_size" > "$OUT/tests_skeleton.log" 2>&1
echo "skeleton tests: exit $?" | tee -a "$OUT/summary.txt"
tail -n 5 "$OUT/tests_skeleton.log"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_default.json" 2>> "$OUT/bench.err"
CUSK_SELF_PLAN=0 timeout 200 $B > "$OUT/bench_noself.json" 2>> "$OUT/bench.err"
CUSK_SELF_COOP=0 timeout 200 $B > "$OUT/bench_nocoop.json" 2>> "$OUT/bench.err"
CUSK_SIDE_STREAM=0 timeout 200 $B > "$OUT/bench_noside.json" 2>> "$OUT/bench.err"
timeout 200 $B > "$OUT/bench_default_b.json" 2>> "$OUT/bench.err"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_t/bench*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], 'ref diff', d.get('ref_gpu',{}).get('adjacency_entries_differing'), d['config']['level_ms'], d['config'].get('ci_kernel_ms'), d['config']['tests_executed_per_level'][3:])
PY
tail -n 5 "$OUT/bench.err"
timeout 300 python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench_steps1.csv" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_bench.log" 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_small_multi_kernel -c 3 -o "$OUT/prof_multi" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_multi.log" 2>&1
python scripts/ncu_summary.py "$OUT/prof_multi.ncu-rep" > "$OUT/ci_small_multi_full_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
