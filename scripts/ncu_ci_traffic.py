#!/usr/bin/env python
"""Summarises an `ncu --set full` capture of the level-l CI kernels of ONE bench step into
profiles/ci_level_traffic.json (bytes_per_step = sum of DRAM read + write over the launches; bench.py reports it
as roofline.traffic) and prints a per-launch table.

    ncu --set full --clock-control none --import-source on -k regex:ci_ -c 60 -o gpurun_out/prof_ci3 \
        python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-ref-gpu
    python scripts/ncu_ci_traffic.py gpurun_out/prof_ci3.ncu-rep <launches of one step> [out.json]
"""
import csv
import io
import json
import subprocess
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "sm__warps_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
           "smsp__inst_executed.sum", "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_bytes.sum",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
MULT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ns": 1e-3, "ms": 1e3}


def main():
    rep, count = sys.argv[1], int(sys.argv[2])
    out = sys.argv[3] if len(sys.argv) > 3 else "profiles/ci_level_traffic.json"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", ",".join(METRICS)],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    ix = {n: i for i, n in enumerate(h)}

    def val(row, k):
        if k not in ix or row[ix[k]] == "":  # e.g. lts__t_bytes.sum is not in ncu's full set
            return 0.0
        x = float(row[ix[k]].replace(",", ""))
        return x * MULT.get(u[ix[k]], 1.0)

    if count <= 0:  # one step = up to the second launch of the first kernel seen (the capture may hold several steps)
        first = rows[2][ix["Kernel Name"]]
        again = [k for k, row in enumerate(rows[3:], start=1) if row[ix["Kernel Name"]] == first]
        count = again[0] if again else len(rows) - 2
    launches = []
    for row in rows[2:2 + count]:
        launches.append(dict(kernel=row[ix["Kernel Name"]].split("(")[0], time_us=val(row, METRICS[0]),
                             dram_read_bytes=val(row, METRICS[1]), dram_write_bytes=val(row, METRICS[2]),
                             warps_active_pct=val(row, METRICS[3]), fp64_pipe_pct=val(row, METRICS[4]),
                             l2_hit_pct=val(row, METRICS[5]), warp_instructions=val(row, METRICS[6]),
                             registers=val(row, METRICS[7]), sm_throughput_pct=val(row, METRICS[8]),
                             issue_active_pct=val(row, METRICS[9]), l2_bytes=val(row, METRICS[10]),
                             l2_throughput_pct=val(row, METRICS[11])))
    tot = sum(x["dram_read_bytes"] + x["dram_write_bytes"] for x in launches)
    tt = sum(x["time_us"] for x in launches)
    # time-weighted roofs of the same launches: issue slots (the roof these kernels are closest to) and L2
    issue = sum(x["issue_active_pct"] * x["time_us"] for x in launches) / tt if tt else 0.0
    l2b = sum(x["l2_bytes"] for x in launches)
    l2pct = sum(x["l2_throughput_pct"] * x["time_us"] for x in launches) / tt if tt else 0.0
    fp64 = sum(x["fp64_pipe_pct"] * x["time_us"] for x in launches) / tt if tt else 0.0
    json.dump(dict(source=f"ncu --set full --clock-control none, the {count} level-l CI kernel launches of one C2 step "
                          f"({rep}; cold-cache, serialised)", bytes_per_step=tot,
                   bound_measured="issue slots (instruction issue / FP64 latency); DRAM and L2 are far from their peaks",
                   issue_slot={"achieved_pct_of_peak": issue, "unit": "% of issue slots, time-weighted over the launches",
                               "fp64_pipe_pct": fp64},
                   l2={"bytes_per_step": l2b, "achieved_gbps": l2b / (tt * 1e-6) / 1e9 if tt else None,
                       "pct_of_peak": l2pct},
                   launches=launches),
              open(out, "w"), indent=1)
    print(f"{'kernel':40s} {'us':>8s} {'rd MB':>8s} {'wr KB':>8s} {'warps%':>7s} {'fp64%':>6s} {'L2hit%':>7s} {'Minst':>7s} regs")
    for x in launches:
        print(f"{x['kernel'][:40]:40s} {x['time_us']:8.1f} {x['dram_read_bytes'] / 1e6:8.2f} "
              f"{x['dram_write_bytes'] / 1e3:8.1f} {x['warps_active_pct']:7.1f} {x['fp64_pipe_pct']:6.1f} "
              f"{x['l2_hit_pct']:7.1f} {x['warp_instructions'] / 1e6:7.2f} {int(x['registers'])}")
    print("bytes_per_step", tot, "time_us", sum(x["time_us"] for x in launches))


if __name__ == "__main__":
    main()
