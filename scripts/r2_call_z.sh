#!/bin/bash
# GPU call Z (1 GPU): the indicator matrix is no longer cleared as a whole (only the rows no marker owns): correlation and
# banded-correlation tests, C3 BED -> correlations, `mps block` bench.
set -u
OUT=gpurun_out/r2_z
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_corr_gpu.py tests/test_zzz_block_gpu.py tests/test_mps_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -x > "$OUT/tests.log" 2>&1
echo "tests: exit $?"; tail -n 3 "$OUT/tests.log"
timeout 300 python scripts/bench_corr.py --steps 3 > "$OUT/corr.json" 2> "$OUT/corr.err"
timeout 300 python scripts/bench_corr.py --steps 3 --missing 0.002 > "$OUT/corr_missing.json" 2>> "$OUT/corr.err"
timeout 300 python scripts/bench_block.py > "$OUT/block.json" 2> "$OUT/block.err"
python - <<'PY'
import json
for f in ('corr','corr_missing','block'):
    ls=[l for l in open('gpurun_out/r2_z/%s.json'%f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f, {k:d[k] for k in d if k in ('ms_per_block','bed_gbps','stages_ms','device_ms','e2e_ms','ms','pairs','route','same_as_oracle','max_abs_diff')})
PY
tail -n 3 "$OUT/corr.err" "$OUT/block.err"
