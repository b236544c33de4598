"""GPU-box exploration: level profile of the C2 generator under a few parameter choices, ours vs the
reference's kernels (oracle/_ref) on the same inputs."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import cigwas_b200
from cigwas_b200 import synth
from tests import ref_lib

ctx = cigwas_b200.Context(0)
ref = ref_lib.RefLib() if ref_lib.available() else None
th = cigwas_b200.hetcor_threshold(1e-4)
cfgs = [(3000, 17, dict(causal_per_trait=c, window=w, max_parents=mp, wlo=a, whi=b)) for c, w, mp, a, b in
        [(12, 8, 3, 0.5, 0.9), (12, 12, 4, 0.6, 0.95), (12, 16, 5, 0.6, 0.95)]]
if len(sys.argv) > 1:
    cfgs = [cfgs[int(sys.argv[1])]]
for m, p, kw in cfgs:
    t0 = time.time()
    Cm = synth.sem_correlation(m, p, stream=2, num_samples=458000, **kw)
    n = m + p
    N = np.full((n, n), 458000.0, np.float32)
    print("cfg", kw, "gen %.1fs" % (time.time() - t0), flush=True)
    for rep in range(2):
        t0 = time.time()
        r = ctx.hetcor_skeleton(Cm, None, N, th, 14, None)
        wall = time.time() - t0
    st = r["stats"]
    print(" ours: wall %.1f ms dev %.1f ms h2d %.2f d2h %.2f l=%d" % (wall * 1e3, st["total_ms"], st["h2d_ms"], st["d2h_ms"], r["l"]))
    print("  canon", st["tests_canonical"])
    print("  exec ", st["tests_executed"])
    print("  ms   ", [round(x, 3) for x in st["level_ms"]])
    print("  ci_ms", [round(x, 3) for x in st["ci_kernel_ms"]])
    print("  maxdeg", st["max_degree"], "edges", st["edges_after"], "sing", st["singular_pivots"])
    tot = sum(st["tests_executed"])
    print("  tests/s (device) %.3e" % (tot / (st["total_ms"] * 1e-3)), flush=True)
    if ref and tot < 3e8:
        for rep in range(2):
            t0 = time.time()
            rr = ref.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32))
            wall_r = time.time() - t0
        nd = int((rr["G"] != r["G"]).sum())
        print(" ref: wall %.1f ms l=%d  adjacency entries differing from ours: %d (float32 fast-math reference)"
              % (wall_r * 1e3, rr["l"], nd), flush=True)
