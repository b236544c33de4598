"""Times the banded correlations of `mps block` through the C ABI with host buffers (cusk_mcorrk_banded:
H2D of the packed columns, counts + bit planes + banded_npn_kernel + row sums, D2H of the row sums) on seeded
genotypes, next to the oracle on a bounded sample of rows, and checks a sample of rows against the oracle.
Prints one JSON line.  usage: python scripts/bench_block.py [m] [n] [corr_width] [missing 0|1]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cigwas_b200  # noqa: E402
from tests import oracle_lib  # noqa: E402  (checker + CPU baseline only)

m = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
w = int(sys.argv[3]) if len(sys.argv) > 3 else 512
missing = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
rng = np.random.default_rng(20261104)
cb = (n + 3) // 4
bed = rng.integers(0, 256, (m, cb), dtype=np.uint8)
if not missing:  # turn every 01 (missing) code into 11
    lo = bed & 0x55
    hi = (bed >> 1) & 0x55
    bed |= ((lo & ~hi) & 0x55) << 1
# LD: a marker copies 3/4 of its left neighbour's bytes
for i in range(1, m):
    keep = rng.random(cb) < 0.75
    bed[i, keep] = bed[i - 1, keep]
ctx = cigwas_b200.Context(0)
ctx.mcorrk_banded(bed[:64], 64, n, 8)
ts, st = [], None
for _ in range(3):
    t0 = time.perf_counter()
    r = ctx.mcorrk_banded(bed, m, n, w, want_corrs=False)
    ts.append(time.perf_counter() - t0)
    st = r["stats"]
full = ctx.mcorrk_banded(bed, m, n, w)
orc = oracle_lib.load()
rows = 24
t0 = time.perf_counter()
sub = orc.mcorrk_banded(bed[: rows + w], rows + w, n, w)[:rows]
cpu_s = time.perf_counter() - t0
diff = float(np.nanmax(np.abs(sub - full["corrs"][:rows])))
tail = orc.mcorrk_banded(bed[m - 40:], 40, n, min(w, 39))
dtail = float(np.nanmax(np.abs(tail[:, : min(w, 39)] - full["corrs"][m - 40:, : min(w, 39)])))
pairs = st["pairs"]
ref = None
try:
    from tests import ref_lib

    if ref_lib.corr_available() and not os.environ.get("CUSK_BENCH_NO_REF"):
        # the reference's own banded correlations (cu_marker_corr_pearson_npn_batched_sparse, src/corr_host.cu:1199: one
        # kernel launch per marker), unmodified sources built for sm_100, host buffers in / out like ours
        import ctypes as C

        rc = ref_lib.RefCorr()
        ub, fp = C.POINTER(C.c_ubyte), C.POINTER(C.c_float)
        flat = np.ascontiguousarray(bed).reshape(-1)
        out = np.zeros((m, w), np.float32)
        rc.banded(flat.ctypes.data_as(ub), 256, n, 8, 256, out.ctypes.data_as(fp))  # context + module load
        devnull, saved = os.open(os.devnull, os.O_WRONLY), os.dup(1)
        os.dup2(devnull, 1)
        t0 = time.perf_counter()
        rc.banded(flat.ctypes.data_as(ub), m, n, w, m, out.ctypes.data_as(fp))
        ref_s = time.perf_counter() - t0
        os.dup2(saved, 1)
        ok = np.isfinite(out) & np.isfinite(full["corrs"])
        ref = {"what": "the reference's cu_marker_corr_pearson_npn_batched_sparse (unmodified sources, sm_100 build), host "
                       "buffers in / out, wall clock", "ms": 1e3 * ref_s, "pairs_per_s": pairs / ref_s,
               "max_abs_diff_vs_ours": float(np.max(np.abs(out[ok] - full["corrs"][ok])))}
except Exception as e:  # noqa: BLE001
    ref = {"unavailable": repr(e)[:300]}
print(json.dumps({
    "metric": "banded_marker_pairs_per_s", "unit": "pairs/s", "m": m, "n": n, "corr_width": w, "missing": missing,
    "e2e_value": pairs / float(np.median(ts)), "e2e_ms": 1e3 * float(np.median(ts)), "kernel_ms": st["kernel_ms"],
    "h2d_ms": st["h2d_ms"], "kernel_pairs_per_s": pairs / (st["kernel_ms"] * 1e-3),
    "sample_pair_rate_G": pairs * n / (st["kernel_ms"] * 1e-3) / 1e9, "pairs": pairs,
    "cpu_pairs_per_s": rows * w / cpu_s, "cpu_cores": orc.num_threads(), "cpu_sample": f"{rows} rows x {w}",
    "max_abs_diff_head": diff, "max_abs_diff_tail": dtail, "ref_gpu": ref,
    "route": "AND + POPC bit planes (CUSK_BLOCK_TC=0)" if os.environ.get("CUSK_BLOCK_TC", "") == "0"
    else "tcgen05 indicator Gram (default from 512 markers)",
}))
