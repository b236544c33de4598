"""Times the prep-bed statistics (bed_counts_kernel + bed_stats_*_kernel) on a device-resident packed
BED of the C3 shape through cusk_bed_marker_stats_noalloc (device buffers, caller's stream, no allocation), CUDA events on
that stream.  Prints one JSON line.  usage: python scripts/bench_prep.py [n] [m]"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cigwas_b200  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 458000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
col_bytes = (n + 3) // 4
ctx = cigwas_b200.Context(0)
lib = ctx.lib
lib.cusk_last_error.restype = C.c_char_p
g = torch.Generator(device="cuda").manual_seed(7)
d_bed = torch.randint(0, 256, (m * col_bytes,), dtype=torch.uint8, device="cuda", generator=g)
d_means = torch.empty(m, dtype=torch.float32, device="cuda")
d_stds = torch.empty(m, dtype=torch.float32, device="cuda")
d_modes = torch.empty(m, dtype=torch.int32, device="cuda")
d_counts = torch.empty((m, 4), dtype=torch.int32, device="cuda")
st = torch.cuda.Stream()


def call():
    rc = lib.cusk_bed_marker_stats_noalloc(C.c_void_p(d_bed.data_ptr()), m, n, C.c_void_p(d_means.data_ptr()),
                                           C.c_void_p(d_stds.data_ptr()), C.c_void_p(d_modes.data_ptr()),
                                           C.c_void_p(d_counts.data_ptr()), C.c_void_p(st.cuda_stream))
    assert rc == 0, lib.cusk_last_error().decode()


call()
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    call()
    e1.record(st)
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[len(ts) // 2]
gb = m * col_bytes / 1e9
print(json.dumps({"what": "prep-bed statistics (counts + means / stds / modes), device-resident packed BED, noalloc entry point",
                  "n": n, "m": m, "bed_gb": round(gb, 3), "ms": round(ms, 3), "bed_gbps_two_passes": round(2 * gb / (ms / 1e3), 1),
                  "counts_sum_ok": bool((d_counts.sum(1) == n).all().item()), "tiled_path": col_bytes % 4 == 0}))
ctx.close()
