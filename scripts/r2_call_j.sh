#!/bin/bash
# GPU call J of round 2 (8 GPUs, charged 8x: keep it short): the bench contract at N=8 and N=4 as the driver launches it,
# the edge-sharded merged skeleton at N=8/4 (strong scaling), the whole-genome block scheduler over 8 GPUs.
set -u
OUT=gpurun_out/r2_j
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "t0 $(date +%s)" > "$OUT/summary.txt"
nvidia-smi --query-gpu=index,name,memory.total --format=csv > "$OUT/gpus.csv" 2>&1
timeout 300 $TR --nproc-per-node 8 --master-port 29531 bench.py --gpus 8 > "$OUT/bench_n8.json" 2> "$OUT/bench_n8.err"
echo "bench N=8: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
timeout 120 $TR --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --impl reference --steps 1 --warmup 0 > "$OUT/bench_n8_reference.json" 2> "$OUT/bench_n8_ref.err"
echo "bench N=8 reference arm: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
timeout 200 $TR --nproc-per-node 4 --master-port 29533 bench.py --gpus 4 --no-cpu-baseline --no-ref-gpu > "$OUT/bench_n4.json" 2> "$OUT/bench_n4.err"
echo "bench N=4: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
H36="--loci 100 --causal 36 --beta-lo 0.02 --beta-hi 0.06 --steps 1 --warmup 1 --maxlevel 10"
H20="--loci 400 --causal 30 --steps 1 --warmup 1"
for n in 8 4; do
  timeout 200 $TR --nproc-per-node $n --master-port 2954$n scripts/bench_merged.py $H36 --tag heavy36_l10_het_n$n > "$OUT/heavy36_l10_het_n$n.json" 2> "$OUT/heavy36_n$n.err"
  echo "heavy36 het N=$n: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
  timeout 200 $TR --nproc-per-node $n --master-port 2955$n scripts/bench_merged.py $H20 --tag heavy20k_het_n$n > "$OUT/heavy20k_het_n$n.json" 2> "$OUT/heavy20k_n$n.err"
  echo "20k heavy het N=$n: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
done
timeout 300 python scripts/bench_c4.py --gpus 8 --blocks 1700 --steps 2 --warmup 1 --verify 0 --ref-sample 0 > "$OUT/c4_1700_n8.json" 2> "$OUT/c4_n8.err"
echo "c4 1700 blocks, 8 GPUs: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
timeout 200 python scripts/bench_c4.py --gpus 4 --blocks 1700 --steps 1 --warmup 1 --verify 0 --ref-sample 0 > "$OUT/c4_1700_n4.json" 2> "$OUT/c4_n4.err"
echo "c4 1700 blocks, 4 GPUs: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
