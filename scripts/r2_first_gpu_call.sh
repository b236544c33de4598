#!/bin/bash
# First GPU call of the next round: validates what was written after round 1's GPU budget was spent and captures the
# profiles that are still missing.  Run as
#   gpurun --timeout 1500 -- 'bash scripts/r2_first_gpu_call.sh'
# (add --gpus 2 to cover the multi-device `mps block` step).  Everything lands in gpurun_out/r2_first/.
# 1. the full GPU suite, with the opt-in tensor-core route of the banded correlations (CUSK_BLOCK_TC=1)
# 2. banded bench on both routes (default AND + POPC bit planes, tcgen05 indicator Gram)
# 3. `mps block` on two devices against the one-device run (only when the box has two)
# 4. headline bench, then launch list + one ncu --set full capture of the banded kernels (never timed under ncu)
set -u
OUT=gpurun_out/r2_first
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 900 python -m pytest tests -m gpu -x -q > "$OUT/gpu_tests_default.log" 2>&1
echo "gpu tests (default routes): exit $?" | tee -a "$OUT/summary.txt"
CUSK_TEST_BLOCK_TC=1 timeout 600 python -m pytest tests/test_zzz_block_gpu.py -m gpu -q > "$OUT/gpu_tests_block_tc.log" 2>&1
echo "block tests incl. tensor-core route: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python scripts/bench_block.py > "$OUT/bench_block_default.json" 2> "$OUT/bench_block_default.err"
echo "bench_block default: exit $?" | tee -a "$OUT/summary.txt"
CUSK_BLOCK_TC=1 timeout 300 python scripts/bench_block.py > "$OUT/bench_block_tc.json" 2> "$OUT/bench_block_tc.err"
echo "bench_block tensor-core: exit $?" | tee -a "$OUT/summary.txt"
CUSK_BLOCK_TC=1 timeout 300 python scripts/bench_block.py 8192 65536 512 1 > "$OUT/bench_block_tc_missing.json" 2>> "$OUT/bench_block_tc.err"
echo "bench_block tensor-core, missing calls: exit $?" | tee -a "$OUT/summary.txt"
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
    timeout 600 python scripts/check_block_multi.py > "$OUT/block_multi.log" 2>&1
    echo "mps block on 2 devices == 1 device: exit $?" | tee -a "$OUT/summary.txt"
fi
timeout 600 python bench.py > "$OUT/bench_n1.json" 2> "$OUT/bench_n1.err"
echo "bench.py: exit $?" | tee -a "$OUT/summary.txt"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file "$OUT/launches_block.csv" \
    python scripts/bench_block.py 4096 65536 512 > "$OUT/ncu_block_launches.log" 2>&1
echo "ncu launch list (banded, default): exit $?" | tee -a "$OUT/summary.txt"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:banded_npn_kernel -c 1 -o "$OUT/banded_npn_full" \
    python scripts/bench_block.py 4096 65536 512 > "$OUT/ncu_block_full.log" 2>&1
echo "ncu --set full banded_npn_kernel: exit $?" | tee -a "$OUT/summary.txt"
CUSK_BLOCK_TC=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:gram_i8_kernel -s 1 -c 1 \
    -o "$OUT/banded_gram_full" python scripts/bench_block.py 4096 65536 512 > "$OUT/ncu_block_tc_full.log" 2>&1
echo "ncu --set full gram_i8_kernel (band tiles): exit $?" | tee -a "$OUT/summary.txt"
cat "$OUT/summary.txt"
