#!/bin/bash
# GPU call F of round 2 (2 GPUs): everything that needs a second device.
#   gpurun --gpus 2 --timeout 2400 -- 'bash scripts/r2_call_f.sh'
# 1. the GPU suite on a 2-GPU box (un-skips the NCCL edge-sharded mps test and the distinct-device refusal)
# 2. `mps block` on two devices against one
# 3. bench.py under torchrun at N=2 (block-parallel weak scaling + the edge-sharded merged leg over NCCL), and at N=1
# 4. strong scaling of ONE heavy merged skeleton, 1 vs 2 GPUs (5k and 20k variables)
# 5. the C4 whole genome on 1 and 2 GPUs
set -u
OUT=gpurun_out/r2_f
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi -L > "$OUT/gpus.txt"
timeout 1200 python -m pytest tests -m gpu -q > "$OUT/tests_all_2gpu.log" 2>&1
echo "all gpu tests on 2 GPUs: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python scripts/check_block_multi.py > "$OUT/block_multi.log" 2>&1
echo "mps block on 2 devices == 1 device: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 python bench.py --no-pipelined-probe > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "bench N=1: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 > "$OUT/bench_n2.json" 2> "$OUT/bench_n2.err"
echo "bench N=2 (torchrun): exit $?" | tee -a "$OUT/summary.txt"
timeout 600 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --same-block --no-extras --no-cpu-baseline --no-ref-gpu > "$OUT/bench_n2_same_block.json" 2>> "$OUT/bench_n2.err"
echo "bench N=2 same block: exit $?" | tee -a "$OUT/summary.txt"
H36="--loci 100 --causal 36 --beta-lo 0.02 --beta-hi 0.06 --steps 1 --warmup 1 --maxlevel 10"
timeout 600 python scripts/bench_merged.py $H36 --tag heavy36_l10_het_n1 > "$OUT/heavy36_l10_het_n1.json" 2> "$OUT/heavy36.err"
echo "heavy36 het N=1: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 $TR --nproc-per-node 2 --master-port 29513 scripts/bench_merged.py $H36 --tag heavy36_l10_het_n2 > "$OUT/heavy36_l10_het_n2.json" 2>> "$OUT/heavy36.err"
echo "heavy36 het N=2: exit $?" | tee -a "$OUT/summary.txt"
H20="--loci 400 --causal 30 --steps 1 --warmup 1"
timeout 900 python scripts/bench_merged.py $H20 --tag heavy20k_het_n1 > "$OUT/heavy20k_het_n1.json" 2> "$OUT/heavy20k.err"
echo "20k heavy het N=1: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 $TR --nproc-per-node 2 --master-port 29514 scripts/bench_merged.py $H20 --tag heavy20k_het_n2 > "$OUT/heavy20k_het_n2.json" 2>> "$OUT/heavy20k.err"
echo "20k heavy het N=2: exit $?" | tee -a "$OUT/summary.txt"
timeout 900 python bench.py --workload c4 --gpus 2 --blocks 1700 --steps 2 --warmup 1 --verify 0 --ref-sample 0 > "$OUT/c4_1700_n2.json" 2> "$OUT/c4.err"
echo "c4 1700 blocks on 2 GPUs: exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
