#!/bin/bash
# GPU call U (1 GPU): phase trace of the multi-level launches (CUSK_SELF_TRACE=1) on the headline block + their tests.
set -u
OUT=gpurun_out/r2_u
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zz_multi_level_gpu.py tests/test_skeleton_gpu.py tests/test_corr_gpu.py tests/test_zzzzzz_no_alloc.py -m gpu -q -x > "$OUT/tests_multi.log" 2>&1
echo "tests: exit $?"; tail -n 3 "$OUT/tests_multi.log"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 10 --warmup 3"
CUSK_SELF_TRACE=1 timeout 200 $B > "$OUT/bench_trace.json" 2> "$OUT/trace.err"
grep "self-trace" "$OUT/trace.err" | tail -n 11
CUSK_WARP_SETS=0 timeout 200 $B > "$OUT/bench_nowarp.json" 2> "$OUT/nowarp.err"
timeout 200 $B > "$OUT/bench_default.json" 2> "$OUT/default.err"
python - <<'PY'
import json
for f in ('bench_trace','bench_nowarp','bench_default'):
    d=json.loads([l for l in open('gpurun_out/r2_u/%s.json'%f) if l.startswith('{')][-1])
    print(f, 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), d['config']['level_ms'], d['config']['tests_executed_per_level'][10:])
PY
timeout 300 python scripts/bench_corr.py --steps 3 > "$OUT/corr.json" 2> "$OUT/corr.err"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_u/corr.json') if l.startswith('{')][-1])
print({k:d[k] for k in d if k in ('ms_per_block','bed_gbps','stages_ms')})
PY
timeout 200 python scripts/bench_prep.py > "$OUT/prep.json" 2> "$OUT/prep.err"; tail -n 1 "$OUT/prep.json"; tail -n 2 "$OUT/prep.err"
