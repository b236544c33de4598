"""Parity of the CUDA skeleton engines (through the C ABI) against the CPU oracle.

Bar (BASELINE.json north_star): adjacency and separating sets identical except for edges whose
Fisher-z statistic lies within 1e-9 (relative) of the threshold; those are counted and listed by
the oracle's `near` records.  Canonical test counts are integers and must match exactly.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NEAR_BAND = 1e-9
ALPHA = 1e-4
NS = 458000


@pytest.fixture(scope="module")
def ctx():
    import cigwas_b200

    c = cigwas_b200.Context(0)
    yield c
    c.close()


def _near_pairs(near):
    return {(int(r[0]), int(r[1])) for r in near} | {(int(r[1]), int(r[0])) for r in near}


def _assert_adj_equal(G_gpu, G_orc, near):
    diff = np.argwhere(G_gpu != G_orc)
    allowed = _near_pairs(near)
    bad = [tuple(d) for d in diff if tuple(d) not in allowed]
    assert not bad, f"{len(bad)} adjacency mismatches outside the 1e-9 band, first: {bad[:5]}"
    return len(diff)


def test_golden_c_n10_both_engines(ctx, golden):
    """cusk/tests/cupc_tests.cpp:17-41 and :65-88."""
    import cigwas_b200

    g = golden["cupc_n10"]
    n = g["n"]
    Cm = np.array(g["C"], np.float32).reshape(n, n)
    A = np.array(g["A"], np.int32).reshape(n, n)
    Th = cigwas_b200.threshold_array(g["sample_size"], g["alpha"])
    r = ctx.skeleton(Cm, Th, 14)
    assert (r["G"] == A).all()
    th = cigwas_b200.hetcor_threshold(g["alpha"])
    N = np.full((n, n), float(g["sample_size"]), np.float32)
    r2 = ctx.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32))
    assert (r2["G"] == A).all()


def test_thresholds_match_oracle(oracle):
    import cigwas_b200

    for n, a in [(500000, 1e-8), (10000, 1e-4), (458000, 1e-4), (20, 0.05)]:
        assert (cigwas_b200.threshold_array(n, a) == oracle.threshold_array(n, a)).all()
        assert cigwas_b200.hetcor_threshold(a) == oracle.hetcor_threshold(a)


@pytest.mark.parametrize("m,p,cp,stream", [(150, 5, 8, 11), (300, 6, 10, 12), (90, 3, 14, 13)])
def test_hetcor_uniform_parity(ctx, oracle, m, p, cp, stream):
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(m, p, stream=stream, num_samples=NS, causal_per_trait=cp)
    n = m + p
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32), near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32))
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    st = rg["stats"]
    assert st["tests_canonical"][: ro["l"] + 1] == ro["stats"][: ro["l"] + 1].tolist()
    assert (rg["G"] == rg["G"].T).all() and np.diag(rg["G"]).sum() == 0
    # all-ones shortcut gives the same graph
    rg2 = ctx.hetcor_skeleton(Cm, None, N, th, 14, None)
    assert (rg2["G"] == rg["G"]).all()


def test_hetcor_general_ess_time_and_candidates(ctx, oracle):
    """Non-uniform effective sample sizes with NaN entries, a two-valued time index and a
    sparse candidate matrix: gaps the reference's tests leave open."""
    from cigwas_b200 import synth

    m, p = 140, 6
    n = m + p
    Cm = synth.sem_correlation(m, p, stream=21, num_samples=NS, causal_per_trait=8)
    N = synth.het_sample_sizes(n, m, NS, stream=21, na_rate=0.02)
    assert np.isnan(N).any()
    rng = synth.rng_for(77)
    G0 = (rng.random((n, n)) < 0.9).astype(np.int32)
    G0 = np.triu(G0, 1)
    G0 = G0 + G0.T
    ti = np.zeros(n, np.int32)
    ti[m:] = rng.integers(1, 4, size=p)
    th = oracle.hetcor_threshold(ALPHA)
    for maxlevel in (14, 2, 0):
        ro = oracle.hetcor_skeleton(Cm, G0, N, th, maxlevel, ti, near_band=NEAR_BAND)
        rg = ctx.hetcor_skeleton(Cm, G0, N, th, maxlevel, ti)
        _assert_adj_equal(rg["G"], ro["G"], ro["near"])
        assert rg["l"] == ro["l"]


@pytest.mark.parametrize("m,p,cp,stream", [(150, 5, 8, 31), (260, 6, 10, 32)])
def test_skeleton_parity_with_sepsets(ctx, oracle, m, p, cp, stream):
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(m, p, stream=stream, num_samples=10000, causal_per_trait=cp)
    n = m + p
    Th = oracle.threshold_array(10000, ALPHA)
    ro = oracle.skeleton(Cm, Th, 14, near_band=NEAR_BAND)
    rg = ctx.skeleton(Cm, Th, 14)
    ndiff = _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["l"] == ro["l"]
    assert rg["stats"]["tests_canonical"][: ro["l"] + 1] == ro["stats"][: ro["l"] + 1].tolist()
    if ndiff == 0 and len(ro["near"]) == 0:
        # canonical separating sets: lowest lexicographic rank of row X (oracle header)
        assert (rg["SepSet"] == ro["SepSet"]).all()
        assert np.allclose(rg["pMax"], ro["pMax"], rtol=1e-5, atol=1e-7)
    # sparse list and dense array agree
    for rec in rg["seps"][:200]:
        lvl = rec["level"]
        assert (rg["SepSet"][rec["x"], rec["y"], :lvl] == rec["s"][:lvl]).all()
        assert rg["G"][rec["x"], rec["y"]] == 0


def test_maxlevel_and_last_level(ctx, oracle):
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(120, 4, stream=41, num_samples=NS, causal_per_trait=6)
    n = Cm.shape[0]
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    for ml in (0, 1, 3, 14):
        ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, ml, np.zeros(n, np.int32))
        rg = ctx.hetcor_skeleton(Cm, None, N, th, ml, None)
        assert rg["l"] == ro["l"], ml
        assert (rg["G"] == ro["G"]).all(), ml


def test_reference_signatures(golden):
    """The drop-in symbols with the reference's exact signatures (cuPC-S.h:192, hetcor-cuPC-S.h:46)."""
    import cigwas_b200

    lib = cigwas_b200.load_library()
    g = golden["cupc_n10"]
    n = g["n"]
    Cm = np.array(g["C"], np.float32).reshape(n, n).copy()
    A = np.array(g["A"], np.int32).reshape(n, n)
    Th = cigwas_b200.threshold_array(g["sample_size"], g["alpha"])
    G = np.zeros((n, n), np.int32)
    pMax = np.zeros((n, n), np.float32)
    Sep = np.zeros((n, n, 14), np.int32)
    P = C.c_int(n)
    l = C.c_int(0)
    ml = C.c_int(14)
    fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)
    lib.Skeleton(Cm.ctypes.data_as(fp), C.byref(P), G.ctypes.data_as(ip), Th.ctypes.data_as(fp), C.byref(l),
                 C.byref(ml), pMax.ctypes.data_as(fp), Sep.ctypes.data_as(ip))
    assert (G == A).all()
    assert (Sep[G == 1] == -1).all()
    G2 = np.ones((n, n), np.int32)
    N = np.full((n, n), float(g["sample_size"]), np.float32)
    th = C.c_float(cigwas_b200.hetcor_threshold(g["alpha"]))
    ti = np.zeros(n, np.int32)
    lib.hetcor_skeleton(Cm.ctypes.data_as(fp), C.byref(P), G2.ctypes.data_as(ip), N.ctypes.data_as(fp), C.byref(th),
                        C.byref(l), C.byref(ml), ti.ctypes.data_as(ip))
    assert (G2 == A).all()


def test_device_resident_api(ctx, oracle):
    import torch
    from cigwas_b200 import synth

    Cm = synth.sem_correlation(200, 5, stream=51, num_samples=NS, causal_per_trait=8)
    n = Cm.shape[0]
    th = oracle.hetcor_threshold(ALPHA)
    N = np.full((n, n), float(NS), np.float32)
    ro = oracle.hetcor_skeleton(Cm, np.ones((n, n), np.int32), N, th, 14, np.zeros(n, np.int32))
    dC = torch.from_numpy(Cm).cuda()
    rw = ctx.row_words(n)
    dbits = torch.zeros((n, rw), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    l, st = ctx.hetcor_skeleton_dev(dC.data_ptr(), n, dbits.data_ptr(), True, None, float(NS), th, 14)
    bits = dbits.cpu().numpy().view(np.uint32)
    G = ((bits[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(n, rw * 32)[:, :n].astype(np.int32)
    assert (G == ro["G"]).all()
    assert l == ro["l"]
    assert st["tests_executed"][1] > 0


def _run_group(world, devices, fn):
    """One context + thread per rank of an in-process group; fn(ctx) is the (identical) call each makes."""
    import threading

    import cigwas_b200

    ctxs = [cigwas_b200.Context(devices[r % len(devices)]) for r in range(world)]
    cigwas_b200.local_group(ctxs)
    out, err = [None] * world, [None] * world

    def work(r):
        try:
            out[r] = fn(ctxs[r])
        except Exception as e:  # noqa: BLE001
            err[r] = e

    ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=120)
    for c in ctxs:
        c.detach_comm()
        c.close()
    assert not any(err), err
    return out


@pytest.mark.parametrize("world", [2, 3])
def test_edge_sharded_individual_level_identical(ctx, world):
    """the individual-level engine sharded over `world` ranks (work items dealt round-robin, MIN
    all-reduce of the separating-set ranks per level): graph, separating sets and pMax of EVERY rank are
    those of the single-GPU call, bit for bit.  The ranks share GPU 0 (in-process exchange by device copies);
    the NCCL exchange between GPUs is covered by tests/test_mps_gpu.py and scripts/bench_merged.py."""
    from cigwas_b200 import synth
    import cigwas_b200

    Cm = synth.sem_correlation(240, 6, stream=91, num_samples=10000, causal_per_trait=10)
    Th = cigwas_b200.threshold_array(10000, ALPHA)
    one = ctx.skeleton(Cm, Th, 14)
    res = _run_group(world, [0], lambda c: c.skeleton(Cm, Th, 14))
    for r in res:
        assert r["l"] == one["l"]
        assert (r["G"] == one["G"]).all()
        assert (r["SepSet"] == one["SepSet"]).all()
        assert np.array_equal(r["pMax"], one["pMax"])
        # canonical counts come from the (replicated) level plan: the same on every rank
        assert r["stats"]["tests_canonical"] == one["stats"]["tests_canonical"]


def test_edge_sharded_hetcor_identical(ctx):
    from cigwas_b200 import synth
    import cigwas_b200

    m, p = 300, 6
    n = m + p
    Cm = synth.sem_correlation(m, p, stream=93, num_samples=20000, causal_per_trait=10)
    N = synth.het_sample_sizes(n, m, 20000, stream=94, lo=4000, na_rate=0.01)
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    G0 = np.ones((n, n), np.int32)
    one = ctx.hetcor_skeleton(Cm, G0, N, th, 14, ti)
    res = _run_group(3, [0], lambda c: c.hetcor_skeleton(Cm, G0, N, th, 14, ti))
    for r in res:
        assert r["l"] == one["l"] and (r["G"] == one["G"]).all()


# ---- all levels through all three CI kernels ---------------------------------------------------------------------
# The bench generator (strong LD, 12 causal markers per trait) at 400 markers + 17 traits runs to level 14: level 1-2
# rows have 33..160+ neighbours (ci_mid.cu and ci_level.cu), from level 3 on every row is short (ci_small.cu, register
# variants to l = 10 and the shared-memory variants above).  No test of this case lies within 1e-9 of its threshold.
def _deep_case():
    import bench
    from cigwas_b200 import synth

    m, p = 400, 17
    return m, p, synth.sem_correlation(m, p, stream=8, num_samples=NS, **bench.C2_KW)


def test_all_levels_individual_level_engine(ctx, oracle):
    m, p, Cm = _deep_case()
    Th = oracle.threshold_array(NS, ALPHA)
    ro = oracle.skeleton(Cm, Th, 14, near_band=NEAR_BAND)
    assert ro["l"] == 15 and len(ro["near"]) == 0
    rg = ctx.skeleton(Cm, Th, 14)
    assert rg["l"] == ro["l"]
    assert (rg["G"] == ro["G"]).all()
    assert rg["stats"]["tests_canonical"][:15] == ro["stats"][:15].tolist()
    assert (rg["SepSet"] == ro["SepSet"]).all()  # lowest-rank separating set of every removed edge, at every level
    assert np.allclose(rg["pMax"], ro["pMax"], rtol=1e-5, atol=1e-7)
    by_level = np.bincount([int(r["level"]) for r in rg["seps"]], minlength=15)
    assert by_level[1] > 0 and by_level[2] > 0 and by_level[3:11].sum() > 0  # removals on every kernel's levels


@pytest.mark.parametrize("general_n", [False, True])
def test_all_levels_hetcor_engine(ctx, oracle, general_n):
    import cigwas_b200
    from cigwas_b200 import synth

    m, p, Cm = _deep_case()
    n = m + p
    N = (synth.het_sample_sizes(n, m, NS, stream=9, lo=50000, na_rate=0.005) if general_n
         else np.full((n, n), float(NS), np.float32))
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    G0 = np.ones((n, n), np.int32)
    ro = oracle.hetcor_skeleton(Cm, G0, N, th, 14, ti, near_band=NEAR_BAND)
    rg = ctx.hetcor_skeleton(Cm, None, N, th, 14, ti)
    assert ro["l"] == 15 and rg["l"] == ro["l"]
    _assert_adj_equal(rg["G"], ro["G"], ro["near"])
    assert rg["stats"]["tests_canonical"][:15] == ro["stats"][:15].tolist()


def test_local_group_refuses_distinct_devices():
    """the in-process copy exchange hung with ranks on distinct GPUs of the NVSwitch box; it is a
    single-device test vehicle and says so (distinct GPUs go through NCCL: cusk_comm_create_all / cusk_comm_create)"""
    import torch

    import cigwas_b200

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctxs = [cigwas_b200.Context(0), cigwas_b200.Context(1)]
    with pytest.raises(cigwas_b200.CuskError):
        cigwas_b200.local_group(ctxs)
    for c in ctxs:
        c.close()


def test_edge_sharded_mixed_levels(ctx, monkeypatch):
    """with a CUSK_SHARD_MIN_TESTS between the levels' test counts some levels are dealt over the ranks and
    some run whole on every rank (the decision is taken on the device, LevelCtl.shard_on); same graph either way"""
    from cigwas_b200 import synth
    import cigwas_b200

    m, p = 300, 6
    n = m + p
    Cm = synth.sem_correlation(m, p, stream=93, num_samples=20000, causal_per_trait=10)
    N = np.full((n, n), 20000.0, np.float32)
    th = cigwas_b200.hetcor_threshold(ALPHA)
    ti = np.zeros(n, np.int32)
    ti[m:] = 1
    one = ctx.hetcor_skeleton(Cm, None, N, th, 14, ti)
    # the library reads the knob once per process: this test only checks that whatever threshold is in force (the
    # suite's conftest sets 0; a run with CUSK_SHARD_MIN_TESTS=200000 exercises the mixed case) the result is the same
    res = _run_group(2, [0], lambda c: c.hetcor_skeleton(Cm, None, N, th, 14, ti))
    for r in res:
        assert r["l"] == one["l"] and (r["G"] == one["G"]).all()
