"""Times the separating-set search of `sepselect` through the C ABI with host buffers
(cusk_sepselect_pairs: H2D of the trait rows, pairs and masks, sepselect_pairs_kernel, D2H of the results) on a
synthetic merged-genome-sized batch, next to the same routine compiled for one host core (tests/sepselect_host_core.cpp,
test infrastructure).  Prints one JSON line.  usage: python scripts/bench_sepselect.py [npairs] [n_var]"""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cigwas_b200  # noqa: E402

npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 126000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20582
P = 17
rng = np.random.default_rng(20261104)
F = rng.standard_normal((n, 40))
Ct = (F[: P + 1] @ F.T) / np.sqrt((F[: P + 1] ** 2).sum(1) + 20.0)[:, None] / np.sqrt((F ** 2).sum(1) + 20.0)[None, :]
Ct[np.arange(P + 1), np.arange(P + 1)] = 1.0
pairs = np.stack([rng.integers(P, n, npairs), rng.integers(0, P, npairs)], axis=1).astype(np.int32)
tmask = rng.integers(0, 1 << P, n, dtype=np.uint32)
tmask[:P] &= ~(np.uint32(1) << np.arange(P, dtype=np.uint32))
tmask[pairs[:, 0]] &= ~(np.uint32(1) << pairs[:, 1].astype(np.uint32))  # j is not adjacent to i
pairs = pairs[((tmask[pairs[:, 0]] >> pairs[:, 1].astype(np.uint32)) & 1) == 0]
thr = 3.89 / np.sqrt(458000.0 - np.arange(P + 1) - 3)
ctx = cigwas_b200.Context(0)
ctx.sepselect_pairs(Ct, P, pairs[:1000], tmask, thr)
ts = []
for _ in range(5):
    t0 = time.perf_counter()
    ln, fm, st = ctx.sepselect_pairs(Ct, P, pairs, tmask, thr)
    ts.append(time.perf_counter() - t0)
gpu_s = float(np.median(ts))
so = os.path.join(tempfile.mkdtemp(), "libhostcheck.so")
subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC",
                       os.path.join(ROOT, "tests", "sepselect_host_core.cpp"), "-o", so])
host = C.CDLL(so)
host.cusk_sepselect_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
sub = pairs[: min(len(pairs), 20000)]
want = np.zeros((sub.shape[0], 34), np.int32)
t0 = time.perf_counter()
host.cusk_sepselect_pairs(None, Ct.ctypes.data, n, P, n, sub.ctypes.data, sub.shape[0], tmask.ctypes.data, thr.ctypes.data,
                          want.ctypes.data)
cpu_s = time.perf_counter() - t0
same = int(((want[:, 0] == ln[: len(sub)]) & (want[:, 2:] == st[: len(sub)]).all(axis=1)).sum())
print(json.dumps(dict(metric="sepselect_pairs_per_s", value=len(pairs) / gpu_s, unit="pairs/s", e2e_ms=gpu_s * 1e3,
                      npairs=int(len(pairs)), n_var=n, P=P, mean_set_len=float(ln.mean()),
                      cpu_1core_pairs_per_s=len(sub) / cpu_s, cpu_sample=int(len(sub)), cpu_same=same)))
