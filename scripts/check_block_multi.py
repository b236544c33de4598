"""`mps block` on two devices (chromosomes dealt to one worker per device, run_block_all in host/pipeline.cu) against
the one-device run on the same file set: the .blocks files must be identical, chromosomes in .bim order.
Needs a box with at least two GPUs.  usage: python scripts/check_block_multi.py"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cigwas_b200  # noqa: E402
from tests.test_block_cpu import random_bed  # noqa: E402
from tests.test_zzz_block_gpu import write_bfiles  # noqa: E402

rng = np.random.default_rng(20261105)
n, w, max_block = 2000, 40, 200
sizes = [900, 300, 1400, 650, 120]  # five chromosomes of uneven size: largest first across the workers
beds = [(str(c + 1), random_bed(rng, m, n, ld=0.97, miss_rate=0.01 if c == 3 else 0.0)) for c, m in enumerate(sizes)]
mps = cigwas_b200.mps_path()
out = {}
for devs in ("0", "0,1"):
    with tempfile.TemporaryDirectory() as d:
        from pathlib import Path

        stem = write_bfiles(Path(d), beds, n)
        r = subprocess.run([mps, "block", stem, str(max_block), "0", str(w)], capture_output=True, text=True,
                           env=dict(os.environ, CUSK_DEVICES=devs))
        if r.returncode != 0:
            print(r.stdout, r.stderr)
            sys.exit(f"mps block failed on devices {devs}")
        used = sorted({ln.split('"device": ')[1].split(",")[0] for ln in r.stdout.splitlines() if '"device": ' in ln})
        print(f"devices {devs}: workers ran on {used}")
        with open(f"{stem}_m{max_block}.blocks") as f:
            out[devs] = f.read()
if out["0"] != out["0,1"]:
    sys.exit("the two-device .blocks file differs from the one-device file")
chrs = [ln.split("\t")[0] for ln in out["0"].splitlines()]
assert chrs == sorted(chrs, key=int), "chromosomes out of .bim order"
print(f"identical: {len(chrs)} blocks over {len(sizes)} chromosomes")
