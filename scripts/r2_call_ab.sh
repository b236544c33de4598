#!/bin/bash
# GPU call AB (1 GPU): CTA-pair Gram as the default: correlation / mps / config-size tests, ncu --set full of the kernel,
# SASS excerpt, super-block variants.
set -u
OUT=gpurun_out/r2_ab
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests/test_corr_gpu.py tests/test_mps_gpu.py tests/test_zzz_block_gpu.py tests/test_zzzzz_config_size_gpu.py -m gpu -q -x > "$OUT/tests.log" 2>&1
echo "tests: exit $?"; tail -n 3 "$OUT/tests.log"
for v in "128 16 9" "128 12 12" "128 24 6" "256 16 9" "64 16 9"; do
  set -- $v
  CUSK_GRAM_KSYNC=$1 CUSK_GRAM_SB_M=$2 CUSK_GRAM_SB_N=$3 timeout 120 python scripts/bench_corr.py --steps 3 > "$OUT/corr_pair_k$1_$2x$3.json" 2>> "$OUT/corr.err"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_ab/corr_pair_*.json')):
    ls=[l for l in open(f) if l.startswith('{')]
    if not ls: print(f,'EMPTY'); continue
    d=json.loads(ls[-1])
    print(f.split('/')[-1], d['ms_per_block'], d['stages_ms']['gram_ms'], d.get('gram_int8_pops_useful'))
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram_i8_kernel -s 1 -c 1 -o "$OUT/gram_pair" \
  python scripts/bench_corr.py --steps 1 --warmup 1 > "$OUT/ncu_gram.log" 2>&1
echo "ncu gram: exit $?"
python scripts/ncu_gram_traffic.py "$OUT/gram_pair.ncu-rep" "$OUT/gram_traffic.json" > "$OUT/gram_traffic.log" 2>&1
python scripts/ncu_summary.py "$OUT/gram_pair.ncu-rep" > "$OUT/gram_pair_full_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
grep -E "duration|dram__bytes_read|tensor|hit_rate" "$OUT/gram_pair_full_summary.txt" | head -12
