#!/bin/bash
# GPU call E of round 2 (1 GPU): suite after the device-side square assembly / fused level 0 / scheduler changes; the C4
# whole genome (1 700 blocks) on one GPU; the C3 pipeline (`mps prep` + `mps cusk` at 458 000 x 10 000).
set -u
OUT=gpurun_out/r2_e
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -m gpu -q > "$OUT/tests_all.log" 2>&1
echo "all gpu tests: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python bench.py --workload c4 --blocks 200 --steps 2 --warmup 1 --verify 10 --ref-sample 0 > "$OUT/c4_200_n1.json" 2> "$OUT/c4_200.err"
echo "c4 200 blocks: exit $?" | tee -a "$OUT/summary.txt"
for s in 1 3 4; do
  timeout 600 python bench.py --workload c4 --blocks 200 --steps 2 --warmup 1 --verify 0 --ref-sample 0 --slots $s > "$OUT/c4_200_n1_slots$s.json" 2>> "$OUT/c4_200.err"
done
echo "c4 slot sweep: done" | tee -a "$OUT/summary.txt"
timeout 1500 python bench.py --workload c4 --blocks 1700 --steps 2 --warmup 1 --verify 20 --ref-sample 50 > "$OUT/c4_1700_n1.json" 2> "$OUT/c4_1700.err"
echo "c4 1700 blocks on one GPU: exit $?" | tee -a "$OUT/summary.txt"
timeout 1200 python scripts/bench_c3_pipeline.py > "$OUT/c3_pipeline.json" 2> "$OUT/c3_pipeline.err"
echo "c3 pipeline: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python bench.py --no-extras --no-pipelined-probe > "$OUT/bench_n1_quick.json" 2> "$OUT/bench_n1.err"
echo "bench quick: exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
