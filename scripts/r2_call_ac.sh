#!/bin/bash
# GPU call AC (1 GPU): whole genome through the scheduler with 2 / 3 / 4 host slots per GPU.
set -u
OUT=gpurun_out/r2_ac
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
for sl in 3 4; do
  timeout 300 python bench.py --workload c4 --gpus 1 --blocks 1700 --steps 2 --warmup 1 --verify 0 --ref-sample 0 --slots $sl > "$OUT/c4_slots$sl.json" 2> "$OUT/c4_slots$sl.err"
  echo "slots $sl: exit $?"
done
python - <<'PY'
import json
for sl in (2,3,4):
    ls=[l for l in open('gpurun_out/r2_ac/c4_slots%d.json'%sl) if l.startswith('{')]
    if not ls: print(sl,'EMPTY'); continue
    d=json.loads(ls[-1]); s=d['scheduler']
    print('slots',sl,'processing_s',s['processing_s'],'seconds',s['seconds'],'sum_device_s',s['sum_device_s'],'sum_load_s',s['sum_load_s'],'wall',d['wall_s_per_run'])
PY
