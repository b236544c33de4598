#!/bin/bash
# GPU call L (1 GPU): host-buffer entry points after the idle-job change (output zero-filled while the device works, scans
# rewritten, upload-instead-of-scan when host threads are scarce), then the routing sweep of call K.
set -u
OUT=gpurun_out/r2_l
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
nproc > "$OUT/nproc.txt"; lscpu | head -20 >> "$OUT/nproc.txt"
timeout 900 python -m pytest tests/test_skeleton_gpu.py tests/test_zzzz_c1_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -k "not c3_size" > "$OUT/tests_default.log" 2>&1
echo "skeleton tests (default host path): exit $?" | tee -a "$OUT/summary.txt"
CUSK_HOST_SCAN=0 timeout 900 python -m pytest tests/test_skeleton_gpu.py tests/test_zzzz_c1_gpu.py -m gpu -q > "$OUT/tests_upload.log" 2>&1
echo "skeleton tests (upload instead of scan): exit $?" | tee -a "$OUT/summary.txt"
B="python bench.py --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu --steps 20 --warmup 3"
timeout 200 $B > "$OUT/bench_default.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=12 timeout 200 $B > "$OUT/bench_t12.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=6 timeout 200 $B > "$OUT/bench_t6.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=2 timeout 200 $B > "$OUT/bench_t2_upload.json" 2>> "$OUT/bench.err"
CUSK_HOST_THREADS=2 CUSK_HOST_SCAN=1 timeout 200 $B > "$OUT/bench_t2_scan.json" 2>> "$OUT/bench.err"
CUSK_HOST_SCAN=0 timeout 200 $B > "$OUT/bench_upload_allthreads.json" 2>> "$OUT/bench.err"
echo "host variants done" | tee -a "$OUT/summary.txt"
for lv in 3 5 8; do for sp in 1 4; do for dm in 0 12; do
  CUSK_DFS_NARROW_LEVEL=$lv CUSK_DFS_MIN_SPAN=$sp CUSK_DFS_MIN_D=$dm timeout 200 $B > "$OUT/sweep_l${lv}_s${sp}_d${dm}.json" 2>> "$OUT/bench.err"
done; done; done
echo "sweep done" | tee -a "$OUT/summary.txt"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_l/*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1]); print(f.split('/')[-1], 'ms', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], 'h2d', d['e2e']['h2d_bytes_per_step'])
PY
timeout 300 python bench.py > "$OUT/bench_full_default.json" 2> "$OUT/bench_full.err"
echo "full bench: exit $?" | tee -a "$OUT/summary.txt"
