#!/bin/bash
# GPU call W (second run, final) (1 GPU): final state of the round: full GPU suite, smoke, the bench as the driver runs it, the reference arm,
# launch list and ncu --set full of the level loop's kernels.
set -u
OUT=gpurun_out/r2_w3
mkdir -p "$OUT"
export PYTHONUNBUFFERED=1
timeout 1800 python -m pytest tests -m gpu -q -x > "$OUT/tests_gpu.log" 2>&1
echo "gpu suite: exit $?" | tee -a "$OUT/summary.txt"
tail -n 3 "$OUT/tests_gpu.log"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/smoke.log" 2>&1
echo "smoke: exit $?" | tee -a "$OUT/summary.txt"; tail -n 1 "$OUT/smoke.log"
timeout 500 python bench.py > "$OUT/bench_full_default.json" 2> "$OUT/bench_full.err"
echo "full bench: exit $?" | tee -a "$OUT/summary.txt"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > "$OUT/bench_reference.json" 2> "$OUT/bench_reference.err"
echo "reference arm: exit $?" | tee -a "$OUT/summary.txt"
python - <<'PY' | tee -a "$OUT/summary.txt"
import json
d=json.loads([l for l in open('gpurun_out/r2_w3/bench_full_default.json') if l.startswith('{')][-1])
print('ms', round(d['ms_per_step'],4), 'value', d['value'], 'e2e', round(d['e2e']['ms_per_step'],3), 'launches', d['gpu_launches'], 'roofline', round(d['roofline']['frac'],3), d['config']['level_ms'])
for k in ('het_n','workload_dense','corr','merged','ref_gpu','e2e_pipelined','cpu_baseline'):
    print(k, json.dumps(d.get(k))[:300])
r=json.loads([l for l in open('gpurun_out/r2_w3/bench_reference.json') if l.startswith('{')][-1])
print('reference', r.get('value'), r.get('ms_per_step'))
PY
timeout 300 python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > /dev/null 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file "$OUT/launches_bench_steps1.csv" \
  python bench.py --steps 1 --warmup 1 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_bench.log" 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ci_ -c 12 -o "$OUT/prof_ci" \
  python bench.py --steps 1 --warmup 0 --no-extras --no-pipelined-probe --no-cpu-baseline --no-ref-gpu > "$OUT/ncu_ci.log" 2>&1
python scripts/ncu_ci_traffic.py "$OUT/prof_ci.ncu-rep" 0 "$OUT/ci_level_traffic.json" > "$OUT/ci_level_traffic.txt" 2>&1
python scripts/ncu_summary.py "$OUT/prof_ci.ncu-rep" > "$OUT/ci_kernels_full_summary.txt" 2>&1
rm -f "$OUT"/*.ncu-rep
