#!/bin/bash
# GPU call AE (1 GPU): speculative start of the host-buffer het call (scans beside the engine): tests, e2e with and without.
set -u
OUT=gpurun_out/r2_ae
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zz_speculation_gpu.py tests/test_skeleton_gpu.py tests/test_zzzz_c1_gpu.py -m gpu -q -x > "$OUT/tests.log" 2>&1
echo "tests: exit $?"; tail -n 8 "$OUT/tests.log"
B="python bench.py --no-extras --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_spec.json" 2> "$OUT/bench.err"
CUSK_SPECULATE=0 timeout 200 $B > "$OUT/bench_nospec.json" 2>> "$OUT/bench.err"
timeout 200 $B > "$OUT/bench_spec_b.json" 2>> "$OUT/bench.err"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ae/bench*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'pipelined', d.get('e2e_pipelined',{}).get('ms_per_step'), d.get('e2e_pipelined',{}).get('same_as_serial_call'))
PY
tail -n 3 "$OUT/bench.err"
