#!/bin/bash
# GPU call AA (1 GPU): Gram kernel on CTA pairs (cta_group::2): parity with the one-CTA form, C3 bench both ways.
set -u
OUT=gpurun_out/r2_aa
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 240 python -m pytest tests/test_corr_gpu.py -m gpu -q -x -k "cta_pair" > "$OUT/tests_pair.log" 2>&1
echo "pair tests: exit $?"; tail -n 12 "$OUT/tests_pair.log"
CUSK_GRAM_PAIR=0 timeout 120 python scripts/bench_corr.py --steps 3 > "$OUT/corr_nopair.json" 2> "$OUT/corr.err"
echo "bench no pair: exit $?"
timeout 120 python scripts/bench_corr.py --steps 3 > "$OUT/corr_pair.json" 2>> "$OUT/corr.err"
echo "bench pair: exit $?"
python - <<'PY'
import json
for f in ('corr_nopair','corr_pair'):
    ls=[l for l in open('gpurun_out/r2_aa/%s.json'%f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f, {k:d[k] for k in d if k in ('ms_per_block','bed_gbps','stages_ms','gram_int8_pops_useful','tensor_frac_useful_of_nominal_4.5')})
PY
tail -n 5 "$OUT/corr.err"
