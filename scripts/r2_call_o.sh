#!/bin/bash
# GPU call O (1 GPU): the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (it runs last).
set -u
OUT=gpurun_out/r2_o
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest tests/test_zzzzzz_no_alloc.py -m gpu -q > "$OUT/tests_noalloc_alone.log" 2>&1
echo "no-alloc test alone: exit $?" | tee -a "$OUT/summary.txt"
timeout 1800 python -m pytest tests -m gpu -q -x > "$OUT/tests_gpu.log" 2>&1
echo "gpu suite (no-alloc test last): exit $?" | tee -a "$OUT/summary.txt"
tail -n 3 "$OUT/tests_noalloc_alone.log" "$OUT/tests_gpu.log"
