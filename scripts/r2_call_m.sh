#!/bin/bash
# GPU call M (1 GPU): whole GPU suite after the binomial table / upload-path reorder; bench lines; `mps block` bench with
# the reference's banded kernels beside it.
set -u
OUT=gpurun_out/r2_m
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 1500 python -m pytest tests -m gpu -q -x > "$OUT/tests_gpu.log" 2>&1
echo "gpu suite: exit $?" | tee -a "$OUT/summary.txt"
CUSK_HOST_SCAN=0 CUSK_HOST_THREADS=2 timeout 600 python -m pytest tests/test_skeleton_gpu.py -m gpu -q > "$OUT/tests_upload_t2.log" 2>&1
echo "skeleton tests, upload path with 2 host threads: exit $?" | tee -a "$OUT/summary.txt"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_default.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=2 timeout 200 $B > "$OUT/bench_t2_upload.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=4 timeout 200 $B > "$OUT/bench_t4_upload.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=4 CUSK_HOST_SCAN=1 timeout 200 $B > "$OUT/bench_t4_scan.json" 2>> "$OUT/bench.err"
timeout 300 python scripts/bench_block.py > "$OUT/bench_block.json" 2> "$OUT/bench_block.err"
echo "bench_block: exit $?" | tee -a "$OUT/summary.txt"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_m/bench_*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    if 'ms_per_step' in d: print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], 'h2d', d['e2e']['h2d_bytes_per_step'])
    else: print(f.split('/')[-1], {k:d[k] for k in ('kernel_ms','e2e_ms','ref_gpu') if k in d})
PY
timeout 300 python bench.py > "$OUT/bench_full_default.json" 2> "$OUT/bench_full.err"
echo "full bench: exit $?" | tee -a "$OUT/summary.txt"
