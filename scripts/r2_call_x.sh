#!/bin/bash
# GPU call X (2 GPUs): the round's final state where a second device matters.
#   gpurun --gpus 2 --timeout 1500 -- 'bash scripts/r2_call_x.sh'
# 1. the tests a one-GPU box skips (NCCL edge-sharded mps run, distinct-device refusal) and the sharded skeleton tests
# 2. bench.py under torchrun at N=2 as the driver launches it (own arm and reference arm)
# 3. the C4 whole genome on 2 GPUs through the scheduler (two slots per GPU share a device: cooperative launches side by side)
set -u
OUT=gpurun_out/r2_x
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi -L > "$OUT/gpus.txt"
timeout 900 python -m pytest tests/test_skeleton_gpu.py tests/test_mps_gpu.py tests/test_zz_multi_level_gpu.py -m gpu -q > "$OUT/tests_2gpu.log" 2>&1
echo "skeleton + mps tests on 2 GPUs: exit $?" | tee -a "$OUT/summary.txt"; tail -n 2 "$OUT/tests_2gpu.log"
timeout 600 $TR --nproc-per-node 2 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > "$OUT/bench_n2.json" 2> "$OUT/bench_n2.err"
echo "bench N=2 (torchrun): exit $?" | tee -a "$OUT/summary.txt"
timeout 300 $TR --nproc-per-node 2 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > "$OUT/bench_n2_reference.json" 2> "$OUT/bench_n2_reference.err"
echo "reference arm N=2: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 python bench.py --workload c4 --gpus 2 --blocks 1700 --steps 2 --warmup 1 --verify 8 --ref-sample 0 > "$OUT/c4_1700_n2.json" 2> "$OUT/c4.err"
echo "c4 1700 blocks on 2 GPUs: exit $?" | tee -a "$OUT/summary.txt"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json
def last(f):
    ls=[l for l in open(f) if l.startswith('{')]
    return json.loads(ls[-1]) if ls else {}
d=last('gpurun_out/r2_x/bench_n2.json')
print('N=2 ms', d.get('ms_per_step'), 'value', d.get('value'), 'per rank', d.get('ms_per_step_per_rank'), 'e2e', d.get('e2e',{}).get('ms_per_step'), 'merged', json.dumps(d.get('merged'))[:400])
c=last('gpurun_out/r2_x/c4_1700_n2.json')
print('c4', json.dumps(c)[:900])
PY
