#!/usr/bin/env python
"""Short per-kernel table from an `ncu --set full` report (run where ncu is installed; no GPU needed):

    python scripts/ncu_summary.py gpurun_out/x.ncu-rep [more.ncu-rep ...] > profiles/rN_x_summary.txt

One block per profiled launch with the counters DESIGN.md / bench.py quote: duration, DRAM bytes, L2 bytes and hit rate,
issue-slot utilisation, warps active, registers, shared-memory wavefronts, pipe utilisations (FP64, ALU, XU/POPC, tensor)."""
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor_subpipe_imma.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_barrier_per_warp_active.pct",
    "smsp__warp_issue_stalled_wait_per_warp_active.pct", "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct",
    "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct",
]


def main():
    for path in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(out.splitlines()))
        if len(rows) < 3:
            print(f"== {path}: no launches")
            continue
        hdr, units = rows[0], rows[1]
        ix = {h: i for i, h in enumerate(hdr)}
        for vals in rows[2:]:
            name = vals[ix["Kernel Name"]]
            print(f"== {path} :: {name[:110]}")
            for k in KEYS:
                if k in ix and vals[ix[k]] != "":
                    print(f"   {k:86s} {vals[ix[k]]:>16s} {units[ix[k]]}")
            # any tensor / fp64 pipe counter not listed above but non-zero
            for h, i in ix.items():
                if h in KEYS:
                    continue
                if ("pipe_tensor" in h or "pipe_fp64" in h) and h.endswith("pct_of_peak_sustained_active") and ".avg." in h:
                    try:
                        if float(vals[i]) > 0:
                            print(f"   {h:86s} {vals[i]:>16s} {units[i]}")
                    except ValueError:
                        pass


if __name__ == "__main__":
    main()
