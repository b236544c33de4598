#!/bin/bash
# GPU call K (1 GPU): which rows of <= 32 neighbours should take the depth-first kernel at deep levels when the level is
# light (latency-bound)?  Sweep of the routing knobs on the headline block; then the usual checks.
set -u
OUT=gpurun_out/r2_k
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_default.json" 2>> "$OUT/bench.err"
for lv in 3 5 8; do for sp in 1 3 5; do for dm in 0 12; do
  CUSK_DFS_NARROW_LEVEL=$lv CUSK_DFS_MIN_SPAN=$sp CUSK_DFS_MIN_D=$dm timeout 200 $B > "$OUT/bench_l${lv}_s${sp}_d${dm}.json" 2>> "$OUT/bench.err"
done; done; done
echo "sweep done" | tee -a "$OUT/summary.txt"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_k/bench_*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1]); print(f.split('/')[-1], round(d['ms_per_step'],4), d['gpu_launches'], round(d['value']/1e9,2))
PY
