#!/usr/bin/env python
"""profiles/gram_traffic.json from one `ncu --set full` capture of gram_i8_kernel on the C3 block (bench.py's `corr` key
quotes it as dram_bytes_ncu):

    ncu --set full --clock-control none --import-source on -k regex:gram_i8_kernel -s 1 -c 1 -o gpurun_out/gram_c3 \
        python scripts/bench_corr.py --steps 1 --warmup 1
    python scripts/ncu_gram_traffic.py gpurun_out/gram_c3.ncu-rep [profiles/gram_traffic.json]
"""
import csv
import io
import json
import subprocess
import sys

M = {"gpu__time_duration.sum": "time", "dram__bytes_read.sum": "dram_read_bytes", "dram__bytes_write.sum": "dram_write_bytes",
     "lts__t_sector_hit_rate.pct": "l2_hit_pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct_of_peak",
     "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
     "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct_any",
     "launch__registers_per_thread": "registers", "launch__grid_size": "grid"}
MULT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "us": 1e-6, "ns": 1e-9, "ms": 1e-3, "s": 1.0}


def main():
    rep = sys.argv[1]
    out = sys.argv[2] if len(sys.argv) > 2 else "profiles/gram_traffic.json"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(h)}
    row = rows[2]
    d = {"source": f"ncu --set full --clock-control none, one launch of {row[ix['Kernel Name']].split('(')[0]} ({rep})"}
    for k, name in M.items():
        if k in ix and row[ix[k]] != "":
            d[name] = float(row[ix[k]].replace(",", "")) * MULT.get(u[ix[k]], 1.0)
    if "time" in d and "dram_read_bytes" in d:
        d["dram_read_gbps"] = d["dram_read_bytes"] / d["time"] / 1e9
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d, indent=1))


if __name__ == "__main__":
    main()
