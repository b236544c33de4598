"""Per-level device times of the C2 block under tuning knobs (time index on/off, CUSK_ITEM_FLOOR)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import cigwas_b200
from cigwas_b200 import synth
import bench

ctx = cigwas_b200.Context(0)
Cm, N = bench.make_block(0)
n = Cm.shape[0]
dC = torch.from_numpy(Cm).cuda(); dN = torch.from_numpy(N).cuda()
ti = np.zeros(n, np.int32); ti[3000:] = 1
dT = torch.from_numpy(ti).cuda()
dbits = torch.zeros((n, ctx.row_words(n)), dtype=torch.int32, device="cuda")
th = cigwas_b200.hetcor_threshold(1e-4)
for use_t in (False, True):
    best = None
    for it in range(8):
        _, st = ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, dN.data_ptr(), 0.0, th, 14, dT.data_ptr() if use_t else None)
        if best is None or st["total_ms"] < best["total_ms"]:
            best = st
    print(json.dumps({"floor": os.environ.get("CUSK_ITEM_FLOOR"), "time_index": use_t, "total_ms": round(best["total_ms"], 3),
                      "level_ms": [round(x, 3) for x in best["level_ms"]], "ci_ms": [round(x, 3) for x in best["ci_kernel_ms"]]}))
