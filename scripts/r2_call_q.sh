#!/bin/bash
# GPU call Q (1 GPU): new defaults (PDL on, plan and compaction as two launches, per-level timing off): full suite + bench.
set -u
OUT=gpurun_out/r2_q
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 1800 python -m pytest tests -m gpu -q -x > "$OUT/tests_gpu.log" 2>&1
echo "gpu suite: exit $?" | tee -a "$OUT/summary.txt"
tail -n 3 "$OUT/tests_gpu.log"
timeout 400 python bench.py > "$OUT/bench_full_default.json" 2> "$OUT/bench_full.err"
echo "full bench: exit $?" | tee -a "$OUT/summary.txt"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
CUSK_FUSE_PLAN=1 timeout 200 $B > "$OUT/bench_fuseplan.json" 2>> "$OUT/bench.err"
CUSK_PDL=0 timeout 200 $B > "$OUT/bench_nopdl.json" 2>> "$OUT/bench.err"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_q/bench*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], 'roofline', round(d['roofline']['frac'],3), d['config']['level_ms'][:4])
PY
