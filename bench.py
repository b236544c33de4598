#!/usr/bin/env python
"""bench.py -- CI tests/s of the skeleton hot path on BASELINE.json configs[1] (cuskss single LD block:
3 000 markers + 17 traits, max level 14, alpha 1e-4, 458 000 samples), one block per GPU per step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Contract (see the task statement): ONE JSON line on stdout from rank 0.  `value` = CI tests executed
per second with inputs resident in HBM; `e2e` = the same metric through the reference-facing C-ABI
call (`cusk_hetcor_skeleton`: host buffers in, host adjacency out, copies inside the timed region).
Multi-GPU: LD blocks are independent units, so ranks process different blocks with no collective
(weak scaling); time = max over ranks.
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


def algorithmic_bytes(tests_executed):
    """one CI test at level l reads C[x,y], C[y,s_1..s_l] and the edge flag: 4 (l + 2) bytes."""
    return sum(t * 4 * (l + 2) for l, t in enumerate(tests_executed) if l >= 1)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                for ln in out.strip().splitlines():
                    self.rows.append([f.strip() for f in ln.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

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
                "reasons": sorted(reasons), "samples": len(self.rows)}


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
            "config": {"workload": WORKLOAD, "generator": dict(C2_KW, seed_stream="2+rank"), "sample": sample},
            "cpu_baseline": {"value": val, "unit": "tests/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "tests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_pipelined_probe(args):
    """Extra, informational leg (key `e2e_pipelined`, N=1 only, run in a child process so that it cannot touch the
    main line): the same reference-facing call with the same pinned host buffers, but issued from TWO host threads
    on two contexts of the one device, so that the copies of one block overlap the level loop of the other -- what
    the block scheduler (`mps cuskss-all`, host/sched.cu) does with a stream of blocks.  Every step still copies
    its 109 MB in and 36 MB out inside the timed region.  First written without GPU time; its results are checked
    against the serial call before anything is reported."""
    import ctypes as C
    import threading

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

    def call(ctx, hG):
        st = cigwas_b200.Stats()
        l = C.c_int(0)
        rc = lib.cusk_hetcor_skeleton(ctx.h, C.cast(hC.data_ptr(), fp), n, C.cast(hG.data_ptr(), ip), 0,
                                      C.cast(hN.data_ptr(), fp), C.c_float(th), MAXLEVEL, ti.ctypes.data_as(ip),
                                      C.byref(l), C.byref(st))
        if rc != 0:
            raise RuntimeError(lib.cusk_last_error().decode())
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
    nn = n * n
    print(json.dumps({"value": sum(tests) / wall if same else None, "unit": "tests/s", "ms_per_step": 1e3 * wall / steps,
                      "steps": steps, "host_threads": 2, "contexts_on_device": 2, "same_as_serial_call": bool(same),
                      "errors": errors, "h2d_bytes_per_step": 3 * nn * 4, "d2h_bytes_per_step": nn * 4,
                      "l2": "not flushed: every step's inputs arrive from host memory",
                      "api": "cusk_hetcor_skeleton from two host threads, one context each"}), flush=True)
    for c in ctxs:
        c.close()


def pipelined_probe_in_child(args):
    import subprocess

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    ap.add_argument("--no-pipelined-probe", action="store_true")
    ap.add_argument("--probe-pipelined", action="store_true", help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.probe_pipelined:
        run_pipelined_probe(args)
        return
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    import cigwas_b200

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = cigwas_b200.Context(local_rank)
    lib = ctx.lib
    import ctypes as C

    Cm, N = make_block(rank)
    n = Cm.shape[0]
    nn = n * n
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

    sampler = ClockSampler(local_rank)
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
        alg_bytes += algorithmic_bytes(st["tests_executed"])
        ci_ms += sum(st["ci_kernel_ms"])
        launches += st["kernel_launches"]  # counted by the engine at its launch sites
        last_stats = st
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
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
    sampler.stop_flag = True
    sampler.join()

    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms_max, e2e_ms_max = t.tolist()
        c = torch.tensor([tests_total, e2e_tests, launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        tests_all, e2e_tests_all, launches_all = c.tolist()
    else:
        dev_ms_max, e2e_ms_max = dev_ms, e2e_ms
        tests_all, e2e_tests_all, launches_all = tests_total, e2e_tests, launches

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = alg_bytes / (ci_ms * 1e-3) / 1e9 if ci_ms > 0 else 0.0
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "ci_level_traffic.json")))["bytes_per_step"]
        except Exception:
            pass
        line = {
            "metric": "ci_tests_per_s", "value": tests_all / (dev_ms_max * 1e-3), "unit": "tests/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "generator": dict(C2_KW, seed_stream="2+rank"), "l2": "flushed between timed iterations "
                       "(256 MB write)", "tests_per_block": tests_total // args.steps,
                       "tests_canonical_per_level": last_stats["tests_canonical"],
                       "tests_executed_per_level": last_stats["tests_executed"],
                       "level_ms": [round(x, 4) for x in last_stats["level_ms"]]},
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
                         "ci_small_kernel<HET,l>, ci_level_kernel<HET>)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                         "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                         "note": "algorithmic bytes = sum_l tests_l * 4 (l+2), time = sum of the CI kernels' "
                                 "durations (CUDA events on the engine's stream); the kernels work out of shared "
                                 "memory / L2 and are FP64-issue and latency bound, not HBM bound; traffic = DRAM "
                                 "bytes of the same launches (ncu, profiles/ci_level_traffic.json)"},
            "clocks": sampler.summary(),
        }
        if not args.no_cpu_baseline:
            tests, times, cores, ns = cpu_port_tests_per_s(Cm, N, steps=1)
            line["cpu_baseline"] = {"value": tests / times[0], "unit": "tests/s", "cores": cores, "kind": "port",
                                    "sample": f"leading {SAMPLE_MARKERS} markers + {P_TRAITS} traits of the block "
                                              f"(n_var {ns}), all levels, {times[0]:.1f} s"}
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
        if world == 1 and not args.no_pipelined_probe:
            line["e2e_pipelined"] = pipelined_probe_in_child(args)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
