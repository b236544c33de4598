#!/bin/bash
# GPU call P (1 GPU): programmatic dependent launch (CUSK_PDL=1) and per-level timing off (CUSK_LEVEL_TIMING=0) on the
# headline block: parity with PDL on, then the four combinations; launch list and ncu of the level-1 fast path.
set -u
OUT=gpurun_out/r2_p
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
CUSK_PDL=1 CUSK_LEVEL_TIMING=0 timeout 1200 python -m pytest tests/test_skeleton_gpu.py tests/test_zz_dfs_kernel_gpu.py tests/test_zzzz_c1_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -k "not c3_size" > "$OUT/tests_skeleton_pdl.log" 2>&1
echo "skeleton tests with PDL, timing off: exit $?" | tee -a "$OUT/summary.txt"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_default.json" 2>> "$OUT/bench.err"
CUSK_LEVEL_TIMING=0 timeout 200 $B > "$OUT/bench_notiming.json" 2>> "$OUT/bench.err"
CUSK_PDL=1 timeout 200 $B > "$OUT/bench_pdl.json" 2>> "$OUT/bench.err"
CUSK_PDL=1 CUSK_LEVEL_TIMING=0 timeout 200 $B > "$OUT/bench_pdl_notiming.json" 2>> "$OUT/bench.err"
CUSK_PDL=1 CUSK_LEVEL_TIMING=0 CUSK_FUSE_PLAN=0 timeout 200 $B > "$OUT/bench_pdl_notiming_nofuse.json" 2>> "$OUT/bench.err"
timeout 300 python scripts/bench_merged.py --loci 400 --causal 12 --steps 3 --warmup 1 --tag light20k_n1 > "$OUT/light20k_n1.json" 2>> "$OUT/bench.err"
CUSK_PDL=1 CUSK_LEVEL_TIMING=0 timeout 300 python scripts/bench_merged.py --loci 400 --causal 12 --steps 3 --warmup 1 --tag light20k_n1_pdl > "$OUT/light20k_n1_pdl.json" 2>> "$OUT/bench.err"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_p/*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    if 'ms_per_step' in d: print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'])
    else: print(f.split('/')[-1], {k:d[k] for k in ('ms_per_skeleton','adjacency_sha1') if k in d})
PY
timeout 300 python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench_steps1.csv" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_bench.log" 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_mid_kernel -c 4 -o "$OUT/prof_mid" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_mid.log" 2>&1
python scripts/ncu_summary.py "$OUT/prof_mid.ncu-rep" > "$OUT/ci_mid_full_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
