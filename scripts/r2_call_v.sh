#!/bin/bash
# GPU call V (1 GPU): ncu --set full of the kernels that so far only had a launch-list time: the BED /
# trait helpers of the correlation path, the separating-set search, the bit-plane builder of `mps block`.
set -u
OUT=gpurun_out/r2_v
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 300 python scripts/bench_corr.py --steps 1 --warmup 0 > "$OUT/corr.json" 2> "$OUT/corr.err"
echo "bench_corr: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"decode_indicators|bed_counts|bed_stats|pxp_kernel|trait_|scatter_mxp|pack_upper|mxp_tiled|level0_strip|set_diag" -c 14 -o "$OUT/corr_helpers" \
  python scripts/bench_corr.py --steps 1 --warmup 0 > "$OUT/ncu_corr.log" 2>&1
echo "ncu corr helpers: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_summary.py "$OUT/corr_helpers.ncu-rep" > "$OUT/corr_helpers_full_summary.txt" 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sepselect_pairs -s 1 -c 1 -o "$OUT/sepselect" \
  python scripts/bench_sepselect.py > "$OUT/ncu_sepselect.log" 2>&1
echo "ncu sepselect: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_summary.py "$OUT/sepselect.ncu-rep" > "$OUT/sepselect_full_summary.txt" 2>&1
CUSK_BLOCK_TC=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"bed_planes|band_row_abs" -c 2 -o "$OUT/planes" \
  python scripts/bench_block.py > "$OUT/ncu_planes.log" 2>&1
echo "ncu bed_planes: exit $?" | tee -a "$OUT/summary.txt"
python scripts/ncu_summary.py "$OUT/planes.ncu-rep" > "$OUT/bed_planes_full_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
grep -c "^==" "$OUT"/*_full_summary.txt
