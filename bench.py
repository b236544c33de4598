#!/usr/bin/env python
"""bench.py -- CI tests/s of the skeleton hot path on BASELINE.json configs[1] (cuskss single LD block:
3 000 markers + 17 traits, max level 14, alpha 1e-4, 458 000 samples), one block per GPU per step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c4|merged] [--distinct-blocks]

Contract (see the task statement): ONE JSON line on stdout from rank 0.  `value` = CI tests executed per second with
inputs resident in HBM; `e2e` = the same metric through the reference-facing C-ABI call (`cusk_hetcor_skeleton`: host
buffers in, host adjacency out, copies inside the timed region).  Multi-GPU: LD blocks are independent units, so ranks
process different blocks with no collective (weak scaling); time = max over ranks.

Extra keys of the same line (rank 0; they cost a few seconds each and never touch the two timed legs above):
  workload_dense   SURVEY 8d's C2 generator (AR(1) LD, 30 causal markers per trait: trait rows keep ~30 neighbours at every
                   level, ~10^11 tests), device-resident, with the reference's own kernels beside it at a level cap
  het_n            the headline block with per-pair effective sample sizes (the `--mxp-se/--pxp-se` path)
  corr             BASELINE.json's "BED->corr GB/s": one C3 block (458 000 samples x 10 000 markers + 17 traits) through
                   the correlation front end, tensor-pipe fractions of the int8 Gram kernel
  merged           one cuskss-merged skeleton (C5-like, 5 017 variables, per-pair N) with its level-l work sharded over
                   the N ranks (NCCL exchange between levels): strong scaling, so that the driver's 1/2/4/8 runs record it
  ref_gpu          the reference's own hetcor_skeleton kernels (unmodified sources, sm_100 build) on the same block
--workload c4 runs scripts/bench_c4.py (whole-genome `mps cuskss-all`), --workload merged scripts/bench_merged.py.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALPHA = 1e-4
NUM_SAMPLES = 458000
M_MARKERS, P_TRAITS, MAXLEVEL = 3000, 17, 14
C2_KW = dict(causal_per_trait=12, window=12, max_parents=4, wlo=0.6, whi=0.95)
WORKLOAD = ("cuskss single LD block 3000 markers + 17 traits (n_var 3017), max level 14, alpha 1e-4, 458000 samples, "
            "time index markers 0 / traits 1 as `mps cuskss` sets it (BASELINE.json configs[1]); one block per GPU "
            "per step")
SAMPLE_MARKERS = M_MARKERS  # cpu_baseline / reference arm: the whole block finishes in seconds on the host cores
DENSE_REF_LEVEL = 4         # level cap at which the reference's kernels run the dense block beside ours
C3_N, C3_M = 458000, 10000
MERGED_KW = dict(loci=100, per_locus=50, causal_per_trait=30)  # 5 017 variables, trait rows keep ~25-30 neighbours


def algorithmic_bytes(tests_executed):
    """one CI test at level l reads C[x,y], C[y,s_1..s_l] and the edge flag: 4 (l + 2) bytes."""
    return sum(t * 4 * (l + 2) for l, t in enumerate(tests_executed) if l >= 1)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md); one process (rank 0)
    queries every GPU of the job."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, indices):
        super().__init__(daemon=True)
        self.indices = ",".join(str(i) for i in indices)
        self.rows = []
        self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", self.indices], capture_output=True, text=True, timeout=5).stdout
                for ln in out.strip().splitlines():
                    self.rows.append([f.strip() for f in ln.split(",")])
            except Exception:
                pass
            time.sleep(0.25)

    def summary(self):
        sm = [float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for i, nm in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows), "gpus": self.indices}


def make_block(rank):
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(M_MARKERS, P_TRAITS, stream=2 + rank, num_samples=NUM_SAMPLES, **C2_KW)
    n = Cm.shape[0]
    N = np.full((n, n), float(NUM_SAMPLES), np.float32)
    return Cm, N


def sample_of(Cm, N):
    ix = np.r_[np.arange(SAMPLE_MARKERS), np.arange(M_MARKERS, M_MARKERS + P_TRAITS)]
    return np.ascontiguousarray(Cm[np.ix_(ix, ix)]), np.ascontiguousarray(N[np.ix_(ix, ix)])


def cpu_port_tests_per_s(Cm, N, steps=1):
    """The FP64 CPU oracle (a port: the reference has no CPU skeleton) on all host threads."""
    from tests import oracle_lib

    orc = oracle_lib.load()
    # all the host cores this process may use, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for its workers)
    orc.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    Cs, Ns = sample_of(Cm, N)
    n = Cs.shape[0]
    th = orc.hetcor_threshold(ALPHA)
    G1 = np.ones((n, n), np.int32)
    ti = np.zeros(n, np.int32)
    ti[n - P_TRAITS:] = 1  # cuskss: markers time 0, traits time 1
    times, tests = [], 0
    for _ in range(steps):
        t0 = time.perf_counter()
        r = orc.hetcor_skeleton(Cs, G1, Ns, th, MAXLEVEL, ti)
        times.append(time.perf_counter() - t0)
        tests = int(r["stats"][15:30].sum())
    return tests, times, orc.num_threads(), n


class quiet_stdout:
    """The reference kernels printf progress lines; keep them off our one-JSON-line stdout."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.null)
        os.close(self.saved)


class stdout_to_stderr:
    """NCCL prints its version banner on stdout when its first communicator comes up; the contract is ONE JSON line"""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def run_reference_arm(args, rank, world):
    """--impl reference: the reference has no CPU implementation of the skeleton; the timed
    code is the FP64 CPU port under oracle/ on all host threads, on a bounded sample of the workload."""
    if rank != 0:
        return
    Cm, N = make_block(0)
    cpu_port_tests_per_s(Cm, N, steps=max(1, min(args.warmup, 1)))
    tests, times, cores, ns = cpu_port_tests_per_s(Cm, N, steps=args.steps)
    tot = sum(times)
    val = tests * len(times) / tot
    sample = f"leading {SAMPLE_MARKERS} markers + {P_TRAITS} traits of the block (n_var {ns}), all levels"
    line = {"impl": "reference", "metric": "ci_tests_per_s", "value": val, "unit": "tests/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / len(times),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "generator": dict(C2_KW, seed_stream="2"), "sample": sample},
            "cpu_baseline": {"value": val, "unit": "tests/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "tests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_pipelined_probe(args):
    """Extra, informational leg (key `e2e_pipelined`, N=1 only, run in a child process so that it cannot touch the
    main line): the same reference-facing call with the same pinned host buffers, but issued from TWO host threads
    on two contexts of the one device, so that the copies of one block overlap the level loop of the other -- what
    the block scheduler (`mps cuskss-all`, host/sched.cu) does with a stream of blocks."""
    import ctypes as C

    import torch

    import cigwas_b200

    torch.cuda.set_device(0)
    Cm, N = make_block(0)
    n = Cm.shape[0]
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ti = np.zeros(n, np.int32)
    ti[M_MARKERS:] = 1
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
    hC = torch.from_numpy(Cm).pin_memory()
    hN = torch.from_numpy(N).pin_memory()
    steps = max(2, min(args.steps, 16))
    steps -= steps % 2
    bufs = [torch.ones((n, n), dtype=torch.int32).pin_memory() for _ in range(steps)]
    ctxs = [cigwas_b200.Context(0), cigwas_b200.Context(0)]
    lib = ctxs[0].lib
    moved = [0, 0]

    def call(ctx, hG):
        st = cigwas_b200.Stats()
        l = C.c_int(0)
        rc = lib.cusk_hetcor_skeleton(ctx.h, C.cast(hC.data_ptr(), fp), n, C.cast(hG.data_ptr(), ip), 0,
                                      C.cast(hN.data_ptr(), fp), C.c_float(th), MAXLEVEL, ti.ctypes.data_as(ip),
                                      C.byref(l), C.byref(st))
        if rc != 0:
            raise RuntimeError(lib.cusk_last_error().decode())
        moved[0], moved[1] = int(st.h2d_bytes), int(st.d2h_bytes)
        return sum(st.tests_executed)

    scratch = torch.ones((n, n), dtype=torch.int32).pin_memory()
    call(ctxs[0], scratch)
    expect = scratch.numpy().copy()
    for _ in range(max(1, args.warmup)):  # both contexts warm, one after the other
        for c in ctxs:
            scratch.fill_(1)
            call(c, scratch)
    torch.cuda.synchronize()
    gate = threading.Barrier(3)
    tests = [0, 0]
    errors = []

    def worker(t):
        try:
            gate.wait()
            for k in range(t, steps, 2):
                tests[t] += call(ctxs[t], bufs[k])
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))
        finally:
            gate.wait()

    th_ = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    for t in th_:
        t.start()
    gate.wait()
    t0 = time.perf_counter()
    gate.wait()
    wall = time.perf_counter() - t0
    for t in th_:
        t.join()
    same = not errors and all(np.array_equal(b.numpy(), expect) for b in bufs)
    print(json.dumps({"value": sum(tests) / wall if same else None, "unit": "tests/s", "ms_per_step": 1e3 * wall / steps,
                      "steps": steps, "host_threads": 2, "contexts_on_device": 2, "same_as_serial_call": bool(same),
                      "errors": errors, "h2d_bytes_per_step": moved[0], "d2h_bytes_per_step": moved[1],
                      "l2": "not flushed: every step's inputs arrive from host memory",
                      "api": "cusk_hetcor_skeleton from two host threads, one context each"}), flush=True)
    for c in ctxs:
        c.close()


def pipelined_probe_in_child(args):
    try:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--probe-pipelined", "--steps", str(args.steps),
                            "--warmup", str(args.warmup)], capture_output=True, text=True, timeout=240,
                           env=dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0")))
        for ln in reversed(r.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)
        return {"unavailable": f"probe exited {r.returncode}: {(r.stderr or r.stdout)[-300:]}"}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)[:300]}


# ---- extra legs ------------------------------------------------------------------------------------------------------
def timed_dev_steps(ctx, torch, stream, flush, fn, steps, warmup):
    """fn() -> stats dict; CUDA events on the context's stream around every step, L2 flushed before each"""
    for _ in range(warmup):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    st = None
    torch.cuda.synchronize()
    for i in range(steps):
        with torch.cuda.stream(stream):
            flush.zero_()
            ev[i][0].record(stream)
        st = fn()
        ev[i][1].record(stream)
    torch.cuda.synchronize()
    # per-level times come from one more, instrumented step (the event records cost ~0.2 ms per call)
    ctx.set_level_timing(True)
    st = fn()
    ctx.set_level_timing(False)
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in ev], st


def leg_dense(ctx, torch, stream, flush, th, want_ref):
    """SURVEY 8d's C2 block: AR(1) LD in sub-blocks of 50 (x 0.3 across neighbours), 30 causal markers per trait."""
    from cigwas_b200 import synth

    Cm = synth.c2_dense_block(M_MARKERS, P_TRAITS, stream=2, num_samples=NUM_SAMPLES)
    n = Cm.shape[0]
    ti = np.zeros(n, np.int32)
    ti[M_MARKERS:] = 1
    dC, dT = torch.from_numpy(Cm).cuda(), torch.from_numpy(ti).cuda()
    dbits = torch.zeros((n, ctx.row_words(n)), dtype=torch.int32, device="cuda")

    def step(ml=MAXLEVEL):
        return ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, None, float(NUM_SAMPLES), th, ml,
                                       dT.data_ptr())[1]

    ms, st = timed_dev_steps(ctx, torch, stream, flush, step, steps=2, warmup=1)
    tests = sum(st["tests_executed"])
    out = {"workload": "SURVEY 8d C2 generator (cigwas_b200.synth.c2_dense_block, stream 2): 3000 markers in AR(1) "
                       "sub-blocks of 50 (r = 0.9^|i-j|, x 0.3 across neighbours) + 17 traits with 30 causal markers "
                       "each; max level 14, alpha 1e-4, uniform N = 458000, device-resident, L2 flushed",
           "ms_per_step": float(np.mean(ms)), "value": tests / (float(np.mean(ms)) * 1e-3), "unit": "tests/s",
           "tests_executed_per_level": st["tests_executed"], "tests_canonical_per_level": st["tests_canonical"],
           "max_degree_per_level": st["max_degree"], "level_ms": [round(x, 3) for x in st["level_ms"]],
           "kernel_launches": st["kernel_launches"]}
    if want_ref:
        from tests import ref_lib

        if ref_lib.available():
            ref = ref_lib.RefLib()
            N = np.full((n, n), float(NUM_SAMPLES), np.float32)
            G1 = np.ones((n, n), np.int32)
            with quiet_stdout():
                ref.hetcor_skeleton(Cm, G1, N, th, DENSE_REF_LEVEL, ti)
                t0 = time.perf_counter()
                rr = ref.hetcor_skeleton(Cm, G1, N, th, DENSE_REF_LEVEL, ti)
                ref_ms = 1e3 * (time.perf_counter() - t0)
            t0 = time.perf_counter()
            ro = ctx.hetcor_skeleton(Cm, G1, N, th, DENSE_REF_LEVEL, ti)
            our_ms = 1e3 * (time.perf_counter() - t0)
            msd, _ = timed_dev_steps(ctx, torch, stream, flush, lambda: step(DENSE_REF_LEVEL), steps=2, warmup=1)
            out["ref_gpu"] = {"what": f"the reference's hetcor_skeleton on this block to max level {DENSE_REF_LEVEL} "
                                      "(its level-l kernels run a thread per conditioning set with an SVD per set from "
                                      "level 4 on; the full depth would take it many minutes), host buffers, wall clock",
                              "max_level": DENSE_REF_LEVEL, "ref_ms_per_block": ref_ms, "ours_same_call_ms": our_ms,
                              "ours_device_resident_ms": float(np.mean(msd)),
                              "adjacency_entries_differing": int((rr["G"] != ro["G"]).sum())}
    return out


def leg_het_n(ctx, torch, stream, flush, th, Cm, dC, dT, dbits):
    """the headline block with per-pair effective sample sizes (the uniform-N fast path is not what a
    `--mxp-se/--pxp-se` run takes)"""
    from cigwas_b200 import synth

    n = Cm.shape[0]
    N = synth.het_sample_sizes(n, M_MARKERS, NUM_SAMPLES, stream=5, lo=50000, na_rate=0.005)
    dN = torch.from_numpy(N).cuda()
    ms, st = timed_dev_steps(ctx, torch, stream, flush,
                             lambda: ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, dN.data_ptr(), 0.0,
                                                             th, MAXLEVEL, dT.data_ptr())[1], steps=3, warmup=1)
    tests = sum(st["tests_executed"])
    return {"workload": "the headline block with marker-trait / trait-trait effective sample sizes U(50000, 458000) and "
                        "0.5 % NaN (synth.het_sample_sizes): per-test thresholds from the mean of trunc(N) over the "
                        "pairs of {X, Y} + S", "ms_per_step": float(np.mean(ms)),
            "value": tests / (float(np.mean(ms)) * 1e-3), "unit": "tests/s", "tests_per_block": tests,
            "level_ms": [round(x, 4) for x in st["level_ms"]]}


def fast_bed(n, m, seed):
    """uniform random genotypes without missing calls (BED code 01 never occurs), whole bytes at memory speed: the
    correlation front end's throughput does not depend on the LD structure"""
    rng = np.random.Generator(np.random.PCG64([20261104, seed]))
    nb = (n + 3) // 4
    r = rng.integers(0, 256, size=(m, nb), dtype=np.uint8)
    lo, hi = r & 0x55, (r >> 1) & 0x55
    r |= ((lo & ~hi) & 0x55) << 1  # 01 -> 11
    if n % 4:
        r[:, -1] &= (1 << (2 * (n % 4))) - 1
    return r


def leg_corr(ctx, torch, peaks):
    """BASELINE.json configs[2]: BED -> correlations of one block, device-resident, CUDA events of the context"""
    n, m, p = C3_N, C3_M, P_TRAITS
    bed = fast_bed(n, m, 3)
    rng = np.random.Generator(np.random.PCG64([20261104, 99]))
    ph = rng.standard_normal((p, n), dtype=np.float32)
    for t in range(3):
        ph[t, rng.random(n) < 0.01] = np.nan
    means, stds, _, _ = ctx.bed_marker_stats(bed, m, n)
    d_bed = torch.from_numpy(bed).cuda()
    d_ph = torch.from_numpy(ph).cuda()
    d_mean, d_std = torch.from_numpy(means).cuda(), torch.from_numpy(stds).cuda()
    nv = m + p
    d_C = torch.zeros((nv, nv), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    res = []
    for i in range(3):
        st = ctx.corr_block_dev(d_bed.data_ptr(), d_ph.data_ptr(), m, n, p, d_mean.data_ptr(), d_std.data_ptr(),
                                d_C.data_ptr())
        if i >= 1:
            res.append(st)
    tot = float(np.mean([r["total_ms"] for r in res]))
    gram = float(np.mean([r["gram_ms"] for r in res]))
    issued = res[-1]["gram_int8_ops"]
    useful = 8.0 * n * m * (m - 1) / 2.0  # 8 n int8 ops per unordered marker pair
    bf16 = float(peaks.get("bf16_tflops", 1590.0))
    dram = None
    try:
        dram = json.load(open(os.path.join(ROOT, "profiles", "gram_traffic.json")))
    except Exception:
        pass
    ref = None
    try:
        from tests import ref_lib

        if ref_lib.corr_available():
            # the reference's own correlation front end (unmodified sources, sm_100) on the leading markers of the same
            # block, host buffers in and out like ours below; both timed by wall clock around the call
            mr = 2000
            rc = ref_lib.RefCorr()
            sub = np.ascontiguousarray(bed[:mr])
            with quiet_stdout():
                rc.corr_pearson_npn(sub[:64], ph, 64, n, p, means[:64], stds[:64])  # context + module load
                t0 = time.perf_counter()
                rmxm, rmxp, rpxp = rc.corr_pearson_npn(sub, ph, mr, n, p, means[:mr], stds[:mr])
                ref_ms = 1e3 * (time.perf_counter() - t0)
            ctx.corr_block(sub[:64], ph, 64, n, p, means[:64], stds[:64])
            t0 = time.perf_counter()
            o = ctx.corr_block(sub, ph, mr, n, p, means[:mr], stds[:mr])
            our_ms = 1e3 * (time.perf_counter() - t0)
            pairs = mr * (mr - 1) // 2
            ref = {"what": "the reference's cu_corr_pearson_npn (src/corr_host.cu:1094, unmodified sources, nvcc -arch=sm_100 "
                           "--use_fast_math) on the leading markers of the same block, host buffers in / out, wall clock",
                   "markers": mr, "marker_pairs": pairs, "ref_ms": ref_ms, "ours_same_call_ms": our_ms,
                   "ref_bed_gbps": mr * ((n + 3) // 4) / 1e9 / (ref_ms * 1e-3),
                   "max_abs_diff_mxm": float(np.nanmax(np.abs(rmxm - o["mxm"]))),
                   "max_abs_diff_mxp": float(np.nanmax(np.abs(rmxp - o["mxp"]))),
                   "extrapolated_ref_ms_full_block": ref_ms * (m * (m - 1) / 2) / pairs}
    except Exception as e:  # noqa: BLE001
        ref = {"unavailable": repr(e)[:300]}
    return {"workload": f"BED -> correlations of one block: {n} samples x {m} markers + {p} traits (1 % NA in 3), no "
                        "missing calls, uniform random genotypes; device-resident (BASELINE.json configs[2])",
            "bed_gb": res[-1]["bed_bytes"] / 1e9, "ms_per_block": tot,
            "bed_gbps": res[-1]["bed_bytes"] / 1e9 / (tot * 1e-3),
            "stages_ms": {k: float(np.mean([r[k] for r in res])) for k in ("counts_ms", "decode_ms", "gram_ms", "mxp_ms", "pxp_ms")},
            "gram_tiles": res[-1]["gram_tiles"],
            "gram_int8_pops_useful": useful / (gram * 1e-3) / 1e15, "gram_int8_pops_issued": issued / (gram * 1e-3) / 1e15,
            "tensor_frac_useful_of_nominal_4.5": useful / (gram * 1e-3) / 4.5e15,
            "tensor_frac_useful_of_2x_measured_bf16": useful / (gram * 1e-3) / (2e12 * bf16),
            "tensor_frac_issued_of_nominal_4.5": issued / (gram * 1e-3) / 4.5e15,
            "dram_bytes_ncu": dram, "ref_gpu": ref}


def leg_merged(ctx, torch, dist, rank, world, th):
    """one cuskss-merged skeleton, level-l work items dealt over the ranks, NCCL exchange of the removals per level"""
    import hashlib

    import cigwas_b200
    from cigwas_b200 import synth

    if world > 1:
        ids = [cigwas_b200.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.attach_comm(rank, world, ids[0])
    m = MERGED_KW["loci"] * MERGED_KW["per_locus"]
    n = m + P_TRAITS
    Cm = synth.merged_correlation(MERGED_KW["loci"], MERGED_KW["per_locus"], P_TRAITS, stream=5, num_samples=NUM_SAMPLES,
                                  causal_per_trait=MERGED_KW["causal_per_trait"])
    N = synth.het_sample_sizes(n, m, NUM_SAMPLES, stream=5, lo=50000, na_rate=0.005)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    dC, dN, dT = torch.from_numpy(Cm).cuda(), torch.from_numpy(N).cuda(), torch.from_numpy(ti).cuda()
    dbits = torch.zeros((n, ctx.row_words(n)), dtype=torch.int32, device="cuda")
    stream = torch.cuda.ExternalStream(ctx.stream)

    def step():
        return ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, dN.data_ptr(), 0.0, th, MAXLEVEL,
                                       dT.data_ptr())[1]

    step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms, tests, xms, xbytes, st = 0.0, 0, 0.0, 0, None
    steps = 2
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        st = step()
        e1.record(stream)
        e1.synchronize()
        ms += e0.elapsed_time(e1)
        tests += sum(st["tests_executed"][1:])
        if world > 1:
            cs = ctx.comm_stats()
            xms += cs["exchange_ms"]
            xbytes += cs["exchange_bytes"]
    digest = hashlib.sha1(dbits.cpu().numpy().tobytes()).hexdigest()
    same = True
    if world > 1:
        t = torch.tensor([ms, xms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, xms = t.tolist()
        c = torch.tensor([tests], dtype=torch.int64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        tests = int(c.item())
        ds = [None] * world
        dist.all_gather_object(ds, digest)
        same = all(d == ds[0] for d in ds)
    ctx.set_level_timing(True)  # level times from one more, instrumented skeleton (every rank takes part)
    st = step()
    ctx.set_level_timing(False)
    torch.cuda.synchronize()
    if world > 1:
        ctx.detach_comm()
    return {"workload": f"cuskss-merged, {m} markers + {P_TRAITS} traits (n_var {n}), per-pair N (0.5 % NaN), 30 causal "
                        "markers per trait (trait rows keep ~25-30 neighbours at every level), max level 14; ONE skeleton, "
                        "level-l work items dealt over the ranks, removals exchanged per level (C5-like; the 20 017-variable "
                        "shape: scripts/bench_merged.py)", "scaling": "strong", "n_gpus": world,
            "ms_per_skeleton": ms / steps, "value": tests / (ms * 1e-3), "unit": "tests/s",
            "tests_per_skeleton": tests // steps, "exchange_ms_per_skeleton": xms / steps,
            "exchange_bytes_per_skeleton": xbytes // steps, "exchange_share": (xms / ms) if ms > 0 else None,
            "exchange_gbps": (xbytes / 1e9) / (xms * 1e-3) if xms > 0 else None,
            "adjacency_sha1": digest, "adjacency_identical_on_all_ranks": bool(same),
            "level_ms_rank0": [round(x, 3) for x in st["level_ms"]]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c4", "merged"])
    ap.add_argument("--same-block", action="store_true", default=True,
                    help="every rank times the SAME block, the one the N=1 run times (default): weak scaling then measures "
                    "contention, not which blocks the extra ranks happened to draw")
    ap.add_argument("--distinct-blocks", dest="same_block", action="store_false",
                    help="rank r times its own block (generator stream 2+r), as rounds 1 and early 2 did")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-pipelined-probe", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip workload_dense / het_n / corr / merged")
    ap.add_argument("--probe-pipelined", action="store_true", help=argparse.SUPPRESS)
    args, rest = ap.parse_known_args()
    if args.probe_pipelined:
        run_pipelined_probe(args)
        return
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.workload == "c4":
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        import bench_c4

        bench_c4.main(["--gpus", str(args.gpus), "--steps", str(min(args.steps, 3)), "--warmup", str(min(args.warmup, 1))] + rest)
        return
    if args.workload == "merged":
        os.execv(sys.executable, [sys.executable, os.path.join(ROOT, "scripts", "bench_merged.py")] + rest)

    import torch
    import torch.distributed as dist

    import cigwas_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        with stdout_to_stderr():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
    ctx = cigwas_b200.Context(local_rank)
    lib = ctx.lib
    import ctypes as C

    Cm, N = make_block(0 if args.same_block else rank)
    n = Cm.shape[0]
    th = cigwas_b200.hetcor_threshold(ALPHA)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    # resident inputs for the device-timed leg
    dC = torch.from_numpy(Cm).cuda()
    dN = torch.from_numpy(N).cuda()
    rw = ctx.row_words(n)
    dbits = torch.zeros((n, rw), dtype=torch.int32, device="cuda")
    # cuskss semantics: markers carry time index 0, traits 1 (cli.cpp:37-42, :219), so traits never
    # enter the conditioning set of a marker-marker pair
    ti = np.zeros(n, np.int32)
    ti[M_MARKERS:] = 1
    dT = torch.from_numpy(ti).cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    # pinned host buffers for the end-to-end leg (the reference-facing call: host in, host out)
    hC = torch.from_numpy(Cm).pin_memory()
    hN = torch.from_numpy(N).pin_memory()
    hG = torch.ones((n, n), dtype=torch.int32).pin_memory()
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
    torch.cuda.synchronize()

    def step_dev():
        return ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, dN.data_ptr(), 0.0, th, MAXLEVEL,
                                       dT.data_ptr())

    def step_e2e():
        st = cigwas_b200.Stats()
        l = C.c_int(0)
        rc = lib.cusk_hetcor_skeleton(ctx.h, C.cast(hC.data_ptr(), fp), n, C.cast(hG.data_ptr(), ip), 0,
                                      C.cast(hN.data_ptr(), fp), C.c_float(th), MAXLEVEL,
                                      ti.ctypes.data_as(ip), C.byref(l),
                                      C.byref(st))
        if rc != 0:
            raise RuntimeError(lib.cusk_last_error().decode())
        return st

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_dev()
        hG.fill_(1)
        step_e2e()

    sampler = None
    if rank == 0:
        sampler = ClockSampler(range(world))
        sampler.start()
    # ---- device-resident leg ---------------------------------------------------------------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    tests_total, alg_bytes, ci_ms, launches = 0, 0, 0.0, 0
    last_stats = None
    barrier()
    for i in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()  # L2 flush between timed iterations (inputs are 73 MB < L2)
            ev[i][0].record(stream)
        _, st = step_dev()
        ev[i][1].record(stream)
        tests_total += sum(st["tests_executed"])
        launches += st["kernel_launches"]  # counted by the engine at its launch sites
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    # ---- the same steps once more with the engine's per-level event records on: they give the CI kernels' durations for
    #      the roofline and level_ms; they are kept out of the timed region above because the records themselves cost
    #      ~0.2 ms of a 0.9 ms step ---------------------------------------------------------------------------------
    ctx.set_level_timing(True)
    for i in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()
        _, st = step_dev()
        alg_bytes += algorithmic_bytes(st["tests_executed"])
        ci_ms += sum(st["ci_kernel_ms"])
        last_stats = st
    ctx.set_level_timing(False)
    barrier()
    # ---- end-to-end leg ----------------------------------------------------------------------------
    e2e_ms, e2e_tests, e2e_up, e2e_down = 0.0, 0, 0, 0
    barrier()
    for i in range(args.steps):
        hG.fill_(1)
        with torch.cuda.stream(stream):
            flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        st = step_e2e()
        e2e_ms += 1e3 * (time.perf_counter() - t0)
        e2e_tests += sum(st.tests_executed)
        e2e_up, e2e_down = int(st.h2d_bytes), int(st.d2h_bytes)
    barrier()
    if sampler:
        sampler.stop_flag = True
        sampler.join()

    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
        per_rank = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(per_rank, t)
        per_rank = [x.tolist() for x in per_rank]
        dev_ms_max, e2e_ms_max = max(x[0] for x in per_rank), max(x[1] for x in per_rank)
        c = torch.tensor([tests_total, e2e_tests, launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        tests_all, e2e_tests_all, launches_all = c.tolist()
    else:
        per_rank = [[dev_ms, e2e_ms]]
        dev_ms_max, e2e_ms_max = dev_ms, e2e_ms
        tests_all, e2e_tests_all, launches_all = tests_total, e2e_tests, launches

    line, peaks = None, {}
    if rank == 0:
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = alg_bytes / (ci_ms * 1e-3) / 1e9 if ci_ms > 0 else 0.0
        traffic, ncu_ci = None, {}
        try:
            ncu_ci = json.load(open(os.path.join(ROOT, "profiles", "ci_level_traffic.json")))
            traffic = ncu_ci["bytes_per_step"]
        except Exception:
            pass
        line = {
            "metric": "ci_tests_per_s", "value": tests_all / (dev_ms_max * 1e-3), "unit": "tests/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "generator": dict(C2_KW, seed_stream="2 (every rank)" if args.same_block else "2+rank"),
                       "l2": "flushed between timed iterations (256 MB write)", "tests_per_block": tests_total // args.steps,
                       "tests_canonical_per_level": last_stats["tests_canonical"],
                       "tests_executed_per_level": last_stats["tests_executed"],
                       "max_degree_per_level": last_stats["max_degree"],
                       "level_ms": [round(x, 4) for x in last_stats["level_ms"]]},
            "ms_per_step_per_rank": [round(x[0] / args.steps, 4) for x in per_rank],
            "e2e_ms_per_step_per_rank": [round(x[1] / args.steps, 4) for x in per_rank],
            "e2e": {"value": e2e_tests_all / (e2e_ms_max * 1e-3), "unit": "tests/s",
                    "h2d_bytes_per_step": e2e_up, "d2h_bytes_per_step": e2e_down,
                    "ms_per_step": e2e_ms_max / args.steps,
                    "api": "cusk_hetcor_skeleton (host C, G, N in; host G out)",
                    "note": "bytes are what the call moved over PCIe (cusk_stats.h2d_bytes / d2h_bytes): C always; the "
                            "all-ones candidate matrix G and the constant N of this workload are recognised by a host scan "
                            "and not uploaded (3 x 36.4 MB would be); the adjacency returns as the bitmap and is expanded "
                            "into the caller's int matrix by host threads"},
            "gpu_launches": launches_all,
            "roofline": {"bound": "hbm", "kernel": "level-l CI kernels of the step (ci_mid_kernel<HET,1|2>, "
                         "ci_small_kernel<HET,l>, the test phases of ci_small_multi_kernel<L0,L1> (levels 4..14), "
                         "ci_dfs_kernel<MODE,l,NS>, ci_level_kernel<HET>)", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                         "bound_measured": ncu_ci.get("bound_measured", "issue slots / FP64 latency (not HBM)"),
                         "issue_slot": ncu_ci.get("issue_slot"), "l2": ncu_ci.get("l2"),
                         "note": "algorithmic bytes = sum_l tests_l * 4 (l+2), time = sum of the CI kernels' "
                                 "durations (CUDA events on the engine's stream, %globaltimer stamps for the test phases inside the "
                                 "multi-level launches; taken in a repeat of the timed steps "
                                 "with the engine's per-level event records switched on: the records cost ~0.2 ms per "
                                 "step and are off in the timed region).  The matrix (36 MB) is L2-resident, so "
                                 "DRAM traffic (ncu) is ~3 % of the algorithmic bytes and HBM is not what bounds these "
                                 "kernels: issue_slot / l2 give the roofs ncu measures for the same launches "
                                 "(profiles/ci_level_traffic.json)"},
            "clocks": sampler.summary() if sampler else None,
        }
        if not args.no_cpu_baseline:
            tests, times, cores, ns = cpu_port_tests_per_s(Cm, N, steps=1)
            line["cpu_baseline"] = {"value": tests / times[0], "unit": "tests/s", "cores": cores, "kind": "port",
                                    "sample": f"leading {SAMPLE_MARKERS} markers + {P_TRAITS} traits of the block "
                                              f"(n_var {ns}), all levels, {times[0]:.1f} s"}
        if not args.no_cpu_baseline:
            # what the reference's own kernels would do differently on this block (int32 rank overflow, F10; the
            # block-level NoEdgeFlag early exit, F9): counted level by level on the oracle's graphs (oracle/cusk_oracle.cpp
            # orc_ref_quirks); all zero means the reference enumerates exactly the sets counted in `value`
            try:
                from tests import oracle_lib

                orc = oracle_lib.load()
                Gq = orc.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 0, ti)["G"]
                per_level = []
                for lq in range(1, MAXLEVEL + 1):
                    q = orc.ref_quirks(Gq, ti, lq)
                    per_level.append([q["rows_int32_rank_overflow"], q["rows_cut_short_by_noedgeflag"], q["rows_not_analysed"]])
                    if not orc.hetcor_level(Cm, Gq, N, th, lq, ti):
                        break
                line["ref_quirks"] = {"per_level_[int32_overflow_rows, noedgeflag_cut_rows, rows_not_analysed]": per_level,
                                      "any": bool(any(a or b for a, b, _ in per_level))}
            except Exception as e:  # noqa: BLE001
                line["ref_quirks"] = {"unavailable": repr(e)[:200]}
        if not args.no_ref_gpu:
            from tests import ref_lib

            if ref_lib.available():
                ref = ref_lib.RefLib()
                G1 = np.ones((n, n), np.int32)
                with quiet_stdout():
                    ref.hetcor_skeleton(Cm, G1, N, th, MAXLEVEL, ti)
                    t0 = time.perf_counter()
                    rr = ref.hetcor_skeleton(Cm, G1, N, th, MAXLEVEL, ti)
                    ref_ms = 1e3 * (time.perf_counter() - t0)
                hG.fill_(1)
                step_e2e()
                line["ref_gpu"] = {"what": "the reference's own hetcor_skeleton (unmodified sources, nvcc -arch=sm_100 "
                                           "--use_fast_math) on the same B200 and block, host buffers in/out, wall clock",
                                   "ms_per_block": ref_ms,
                                   "adjacency_entries_differing": int((rr["G"] != hG.numpy()).sum())}
    # ---- extra legs (never inside the timed regions above) ---------------------------------------------------------
    if not args.no_extras:
        try:
            merged = leg_merged(ctx, torch, dist, rank, world, th)
            if rank == 0:
                line["merged"] = merged
        except Exception as e:  # noqa: BLE001
            if rank == 0:
                line["merged"] = {"unavailable": repr(e)[:300]}
        if rank == 0:
            for key, fn in (("het_n", lambda: leg_het_n(ctx, torch, stream, flush, th, Cm, dC, dT, dbits)),
                            ("workload_dense", lambda: leg_dense(ctx, torch, stream, flush, th, not args.no_ref_gpu)),
                            ("corr", lambda: leg_corr(ctx, torch, peaks))):
                try:
                    line[key] = fn()
                except Exception as e:  # noqa: BLE001
                    line[key] = {"unavailable": repr(e)[:300]}
    if rank == 0:
        if world == 1 and not args.no_pipelined_probe:
            line["e2e_pipelined"] = pipelined_probe_in_child(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
