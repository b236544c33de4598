#!/bin/bash
# GPU call AD (1 GPU): scheduler default of up to four slots per GPU: mps tests, whole genome with the default and with 6.
set -u
OUT=gpurun_out/r2_ad
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_mps_gpu.py -m gpu -q -x > "$OUT/tests_mps.log" 2>&1
echo "mps tests: exit $?"; tail -n 2 "$OUT/tests_mps.log"
nproc
timeout 300 python bench.py --workload c4 --gpus 1 --blocks 1700 --steps 2 --warmup 1 --verify 8 --ref-sample 0 > "$OUT/c4_default.json" 2> "$OUT/c4_default.err"
echo "c4 default: exit $?"
timeout 300 python bench.py --workload c4 --gpus 1 --blocks 1700 --steps 2 --warmup 1 --verify 0 --ref-sample 0 --slots 6 > "$OUT/c4_slots6.json" 2> "$OUT/c4_slots6.err"
python - <<'PY'
import json
for f in ('c4_default','c4_slots6'):
    ls=[l for l in open('gpurun_out/r2_ad/%s.json'%f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1]); s=d['scheduler']
    print(f,'slots',s['slots_per_gpu'],'processing_s',s['processing_s'],'seconds',s['seconds'],'sum_device_s',s['sum_device_s'],'wall',d['wall_s_per_run'], d.get('per_block_processes'))
PY
