#!/bin/bash
# GPU call Y (1 GPU): DRAM traffic / issue-slot table of the CI launches of one step for bench.py's roofline keys
# (profiles/ci_level_traffic.json), final kernels.
set -u
OUT=gpurun_out/r2_y
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_ -c 20 -o "$OUT/prof_ci" \
  python bench.py --steps 1 --warmup 0 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_ci.log" 2>&1
python scripts/ncu_ci_traffic.py "$OUT/prof_ci.ncu-rep" 0 "$OUT/ci_level_traffic.json" > "$OUT/ci_level_traffic.txt" 2>&1
tail -n 14 "$OUT/ci_level_traffic.txt"
rm -f "$OUT"/*.ncu-rep
