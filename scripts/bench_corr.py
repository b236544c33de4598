"""BED -> correlation throughput of one block (BASELINE.json configs[2]: n = 458k samples x 10k markers
+ 17 traits), device-resident, timed with CUDA events on the context's stream.

    python scripts/bench_corr.py [--n 458000] [--m 10000] [--p 17] [--steps 3] [--missing 0.0]
Prints one JSON line: GB/s of packed BED consumed, int8 tensor-pipe utilisation of the Gram kernel.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cigwas_b200  # noqa: E402


def random_bed_torch(n, m, seed=3, missing=0.0, sub_block=50, copy_prob=0.95, device="cuda"):
    """Same distribution as synth.random_bed, generated on the device (Philox) in marker order."""
    g = torch.Generator(device=device)
    g.manual_seed(20261104 + seed)
    nb = (n + 3) // 4
    out = torch.zeros((m, nb), dtype=torch.uint8, device=device)
    code_of = torch.tensor([3, 2, 0], dtype=torch.uint8, device=device)
    hap = None
    shifts = torch.tensor([0, 2, 4, 6], dtype=torch.uint8, device=device)
    for j in range(m):
        maf = 0.01 + 0.49 * torch.rand((), generator=g, device=device)
        fresh = torch.rand((2, n), generator=g, device=device) < maf
        if hap is None or j % sub_block == 0:
            hap = fresh
        else:
            keep = torch.rand((2, n), generator=g, device=device) < copy_prob
            hap = torch.where(keep, hap, fresh)
        codes = code_of[(hap[0].to(torch.uint8) + hap[1].to(torch.uint8)).long()]
        if missing > 0:
            miss = torch.rand((n,), generator=g, device=device) < missing
            codes = torch.where(miss, torch.ones_like(codes), codes)
        pad = torch.zeros(nb * 4 - n, dtype=torch.uint8, device=device)
        c4 = torch.cat([codes, pad]).view(nb, 4)
        out[j] = (c4 << shifts).sum(dim=1).to(torch.uint8)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=458000)
    ap.add_argument("--m", type=int, default=10000)
    ap.add_argument("--p", type=int, default=17)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--missing", type=float, default=0.0)
    a = ap.parse_args()
    ctx = cigwas_b200.Context(0)
    bed = random_bed_torch(a.n, a.m, missing=a.missing)
    g = torch.Generator(device="cuda")
    g.manual_seed(99)
    phen = torch.randn((a.p, a.n), generator=g, device="cuda")
    phen[:3][torch.rand((3, a.n), generator=g, device="cuda") < 0.01] = float("nan")
    # statistics through the product path (host round trip is outside the timed region)
    means, stds, _, _ = ctx.bed_marker_stats(bed.cpu().numpy(), a.m, a.n)
    d_mean = torch.from_numpy(means).cuda()
    d_std = torch.from_numpy(stds).cuda()
    nv = a.m + a.p
    d_C = torch.zeros((nv, nv), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    res = []
    for i in range(a.warmup + a.steps):
        st = ctx.corr_block_dev(bed.data_ptr(), phen.data_ptr(), a.m, a.n, a.p, d_mean.data_ptr(), d_std.data_ptr(),
                                d_C.data_ptr())
        if i >= a.warmup:
            res.append(st)
    tot = float(np.mean([r["total_ms"] for r in res]))
    gram = float(np.mean([r["gram_ms"] for r in res]))
    ops = res[-1]["gram_int8_ops"]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    bf16 = float(peaks.get("bf16_tflops", 1590.0))
    useful = 8.0 * a.n * a.m * (a.m - 1) / 2.0  # 8 n int8-ops per unordered marker pair
    line = {
        "metric": "bed_to_corr", "workload": f"n={a.n} samples x m={a.m} markers + {a.p} traits, missing={a.missing}",
        "bed_gb": res[-1]["bed_bytes"] / 1e9, "ms_per_block": tot, "bed_gbps": res[-1]["bed_bytes"] / 1e9 / (tot * 1e-3),
        "stages_ms": {k: float(np.mean([r[k] for r in res])) for k in ("counts_ms", "decode_ms", "gram_ms", "mxp_ms", "pxp_ms")},
        "parts": res[-1]["parts"], "gram_tiles": res[-1]["gram_tiles"],
        "gram_int8_tops_issued": ops / (gram * 1e-3) / 1e12,
        "gram_int8_tops_useful": useful / (gram * 1e-3) / 1e12,
        "tensor_peak_int8_tops": {"nominal": 4500.0, "proxy_2x_measured_bf16": 2 * bf16},
        "tensor_frac_issued_of_nominal": ops / (gram * 1e-3) / 1e12 / 4500.0,
        "tensor_frac_issued_of_proxy": ops / (gram * 1e-3) / 1e12 / (2 * bf16),
    }
    print(json.dumps(line), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
