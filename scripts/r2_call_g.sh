#!/bin/bash
# GPU call G of round 2 (2 GPUs): why did the torchrun runs of call F print nothing?  Then the N=2 runs themselves.
set -u
OUT=gpurun_out/r2_g
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
echo "t0 $(date +%s)" > "$OUT/summary.txt"
$TR --nproc-per-node 2 --master-port 29521 scripts/diag_rank.py > "$OUT/diag_plain.out" 2> "$OUT/diag_plain.err"
echo "diag plain: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
timeout 300 $TR --nproc-per-node 2 --master-port 29522 scripts/diag_rank.py > "$OUT/diag_timeout.out" 2> "$OUT/diag_timeout.err"
echo "diag under timeout: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
$TR --nproc-per-node 2 --master-port 29523 --tee 3 --log-dir "$OUT/tlogs_merged" scripts/bench_merged.py --loci 20 --causal 20 --steps 1 --warmup 1 --tag small_n2 > "$OUT/merged_small_n2.json" 2> "$OUT/merged_small_n2.err"
echo "bench_merged small N=2: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
$TR --nproc-per-node 2 --master-port 29524 --tee 3 --log-dir "$OUT/tlogs_bench" bench.py --gpus 2 > "$OUT/bench_n2.json" 2> "$OUT/bench_n2.err"
echo "bench N=2: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
$TR --nproc-per-node 2 --master-port 29525 bench.py --gpus 2 --same-block --no-extras --no-cpu-baseline --no-ref-gpu > "$OUT/bench_n2_same_block.json" 2> "$OUT/bench_n2_same.err"
echo "bench N=2 same block: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
$TR --nproc-per-node 2 --master-port 29526 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 > "$OUT/bench_n2_reference.json" 2> "$OUT/bench_n2_ref.err"
echo "bench N=2 reference arm: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
H36="--loci 100 --causal 36 --beta-lo 0.02 --beta-hi 0.06 --steps 1 --warmup 1 --maxlevel 10"
$TR --nproc-per-node 2 --master-port 29527 scripts/bench_merged.py $H36 --tag heavy36_l10_het_n2 > "$OUT/heavy36_l10_het_n2.json" 2> "$OUT/heavy36.err"
echo "heavy36 het N=2: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
H20="--loci 400 --causal 30 --steps 1 --warmup 1"
$TR --nproc-per-node 2 --master-port 29528 scripts/bench_merged.py $H20 --tag heavy20k_het_n2 > "$OUT/heavy20k_het_n2.json" 2> "$OUT/heavy20k.err"
echo "20k heavy het N=2: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
timeout 600 python scripts/bench_c3_pipeline.py > "$OUT/c3_pipeline.json" 2> "$OUT/c3_pipeline.err"
echo "c3 pipeline: exit $? $(date +%s)" | tee -a "$OUT/summary.txt"
ls -R "$OUT/tlogs_bench" > "$OUT/tlogs_listing.txt" 2>&1
cat "$OUT/summary.txt"
