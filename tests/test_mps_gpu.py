"""GPU tests of the `mps` process contract: the reference's own file-level cuskss tests
(cusk/tests/cuskss_tests.cpp:31-209), prep-bed on small.bed (prep_tests.cpp:44-73), and cusk / cuskss /
cuskss-het / cuskss-merged / two-stage / scheduler / edge-sharded runs compared file by file with
the oracle-driven restatement of the host flow (tests/host_ref.py)."""
import base64
import os
import subprocess

import numpy as np
import pytest

from tests import host_ref

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MPS = os.path.join(ROOT, "ci-gwas_b200", "mps")


def _device_count():
    import torch

    return torch.cuda.device_count()


def run_mps(args, env=None, check=True):
    e = dict(os.environ)
    # at most two GPUs per mps process: what these tests were verified on (a process that touches all eight
    # GPUs of a box spends most of its run time initialising them)
    e.setdefault("CUSK_DEVICES", "0,1" if _device_count() >= 2 else "0")
    e.update(env or {})
    r = subprocess.run([MPS] + [str(a) for a in args], capture_output=True, text=True, env=e)
    if check:
        assert r.returncode == 0, r.stdout + r.stderr
    return r


@pytest.fixture(scope="module")
def ref_files(golden, tmp_path_factory):
    d = tmp_path_factory.mktemp("ref_files")
    for name, b in golden["files_b64"].items():
        (d / name).write_bytes(base64.b64decode(b))
    return str(d)


def cuskss_args(d, mxm="mxm.bin", mxp="mxp.txt", mxp_se="NULL", pxp="pxp.txt", pxp_se="NULL", time_index="NULL",
                block_index=0, blockfile="NULL", marker_ixs="NULL", alpha=1e-4, L1=3, L2=0, depth=1, ns=500000,
                outdir=None):
    f = lambda n: n if n == "NULL" else os.path.join(d, n)
    return ["cuskss", f(mxm), f(mxp), f(mxp_se), f(pxp), f(pxp_se), f(time_index), block_index, f(blockfile),
            f(marker_ixs), alpha, L1, L2, depth, ns, outdir]


# ---- the reference's own end-to-end tests -----------------------------------------------------------------
def test_reference_cuskss_trait_only(ref_files, golden, tmp_path):
    """cuskss_tests.cpp:31-83 (through the mps argv contract, whose path checks want a block file when
    no marker indices are given, apps/mps.cpp:59-68)"""
    g = golden["cuskss"]
    out = str(tmp_path)
    run_mps(cuskss_args(ref_files, mxm="NULL", mxp="NULL", pxp="trait_summary_stats.txt",
                        blockfile="blocks.txt", L1=3, L2=1, outdir=out))
    r = host_ref.read_outputs(os.path.join(out, "trait_only"))
    assert r["adj"].reshape(-1).tolist() == g["trait_only"]["adj"]
    np.testing.assert_allclose(r["corr"].reshape(-1), g["trait_only"]["corr"], atol=g["corr_tol"])
    assert not os.path.exists(os.path.join(out, "trait_only.sep"))


@pytest.mark.parametrize("mode", ["merged", "block"])
def test_reference_cuskss_two_stage(ref_files, golden, tmp_path, mode):
    """cuskss_tests.cpp:85-209: .ixs == {2,3,4,5}, 4x4 adj and corr"""
    g = golden["cuskss"]
    out = str(tmp_path)
    kw = dict(mxm="small_mxm.bin", mxp="marker_trait_summary_stats.txt", pxp="trait_summary_stats.txt", L1=3, L2=1,
              depth=1, outdir=out)
    if mode == "merged":
        run_mps(cuskss_args(ref_files, marker_ixs="marker_indices.bin", **kw))
        stem = "cuskss_merged"
    else:
        run_mps(cuskss_args(ref_files, blockfile="blocks.txt", **kw))
        stem = "1_0_2"
    r = host_ref.read_outputs(os.path.join(out, stem))
    assert r["ixs"].tolist() == g["two_stage"]["ixs"]
    assert r["adj"].reshape(-1).tolist() == g["two_stage"]["adj"]
    np.testing.assert_allclose(r["corr"].reshape(-1), g["two_stage"]["corr"], atol=g["corr_tol"])
    assert r["mdim"] == (4, 3, 14)


def test_reference_prep_small_bed(ref_files, golden, tmp_path):
    """prep_tests.cpp:44-73: .dim = (10, 5), means / stds of small.bed; text files with 6 significant digits"""
    for sfx in (".bed", ".bim", ".fam"):
        (tmp_path / ("small" + sfx)).write_bytes(open(os.path.join(ref_files, "small" + sfx), "rb").read())
    stem = str(tmp_path / "small")
    run_mps(["prep", stem])
    assert open(stem + ".dim").read() == "10\t5\n"
    ps = golden["prep_small"]
    means = [float(x) for x in open(stem + ".means").read().split()]
    stds = [float(x) for x in open(stem + ".stds").read().split()]
    np.testing.assert_allclose(means, ps["means"], atol=2e-6, rtol=1e-5)
    np.testing.assert_allclose(stds, ps["stds"], atol=2e-6, rtol=1e-5)
    assert open(stem + ".means").read().split("\n")[:2] == ["0.6", "1.1"]  # operator<< default precision
    assert len(open(stem + ".modes").read().split()) == 5
    r = run_mps(["prep", os.path.join(ref_files, "wrong_magic_num_one")], check=False)
    assert r.returncode == 1


# ---- cuskss against the oracle-driven restatement ---------------------------------------------------------------
def _cuskss_case(tmp_path, oracle, m, p, stream, het, L1, L2, depth, merged, time_index=None, env=None):
    import cigwas_b200
    from cigwas_b200 import synth

    ns = 20000
    Cm = synth.sem_correlation(m, p, stream=stream, num_samples=ns, causal_per_trait=6)
    d = str(tmp_path)
    Nhet = synth.het_sample_sizes(m + p, m, ns, stream=stream + 1, lo=4000, na_rate=0.0) if het else None
    host_ref.write_cuskss_inputs(d, Cm, m, p, het_N=Nhet, na_frac=0.02 if het else 0.0,
                                 rng=np.random.default_rng(stream), extra_rows_before=7)
    f = lambda n: os.path.join(d, n)
    open(f("blocks.txt"), "w").write("0\t0\t6\n1\t0\t%d\n" % (m - 1))
    np.arange(7, 7 + m, dtype=np.int32).tofile(f("ixs.bin"))
    ti = None
    if time_index is not None:
        open(f("ti.txt"), "w").write("".join(f"{t}\n" for t in time_index))
        ti = "ti.txt"
    out = f("out")
    os.makedirs(out, exist_ok=True)
    args = cuskss_args(d, mxp_se="mxp_se.txt" if het else "NULL", pxp_se="pxp_se.txt" if het else "NULL",
                       time_index=ti or "NULL", block_index=1, blockfile="NULL" if merged else "blocks.txt",
                       marker_ixs="ixs.bin" if merged else "NULL", L1=L1, L2=L2, depth=depth, ns=ns, outdir=out)
    run_mps(args, env=env)
    got = host_ref.read_outputs(os.path.join(out, "cuskss_merged" if merged else "1_0_%d" % (m - 1)))
    # expectation
    mxm = host_ref.load_mxm(f("mxm.bin"))
    mxp, mess = host_ref.load_mxp(f("mxp.txt"), f("mxp_se.txt") if het else None, list(range(7, 7 + m)))
    pxp, pess = host_ref.load_pxp(f("pxp.txt"), f("pxp_se.txt") if het else None, float(ns))
    C0, N0 = host_ref.square_inputs(mxm, mxp, mess, pxp, pess, float(ns), het)
    th = cigwas_b200.hetcor_threshold(1e-4)
    exp = host_ref.cuskss_expected(oracle, C0, N0, p, th, depth, L1, L2,
                                   np.array(time_index if time_index is not None else [1] * p, np.int32))
    assert got["mdim"] == exp["mdim"]
    assert (got["ixs"] == exp["ixs"]).all()
    assert (got["adj"] == exp["adj"]).all()
    assert (got["corr"] == exp["corr"]).all()  # copied input values: bit-exact
    return got


def test_cuskss_block_single_stage(tmp_path, oracle):
    _cuskss_case(tmp_path, oracle, 180, 5, 301, het=False, L1=6, L2=0, depth=2, merged=False)


def test_cuskss_block_two_stage_time_index(tmp_path, oracle):
    _cuskss_case(tmp_path, oracle, 150, 4, 303, het=False, L1=3, L2=8, depth=1, merged=False, time_index=[2, 1, 3, 1])


def test_cuskss_het_merged_two_stage(tmp_path, oracle):
    _cuskss_case(tmp_path, oracle, 160, 5, 305, het=True, L1=4, L2=6, depth=2, merged=True)


def test_cuskss_merged_edge_sharded_matches_single(tmp_path, oracle):
    """the level-l tests of one skeleton dealt to 3 ranks (in-process exchange, ranks share GPU 0)
    -> the same files as the single-GPU run and as the oracle"""
    a = _cuskss_case(tmp_path / "one", oracle, 170, 5, 307, het=True, L1=5, L2=0, depth=3, merged=True)
    b = _cuskss_case(tmp_path / "three", oracle, 170, 5, 307, het=True, L1=5, L2=0, depth=3, merged=True,
                     env={"CUSK_SHARD_GPUS": "3", "CUSK_LOCAL_EXCHANGE": "1", "CUSK_DEVICES": "0"})
    for k in ("ixs", "adj", "corr"):
        assert (a[k] == b[k]).all()


def test_cuskss_merged_edge_sharded_nccl(tmp_path, oracle):
    """the same over NCCL (ncclCommInitAll in one process, one rank per GPU); needs >= 2 GPUs on the box"""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _cuskss_case(tmp_path, oracle, 170, 5, 307, het=True, L1=5, L2=0, depth=3, merged=True,
                 env={"CUSK_SHARD_GPUS": "2"})


# ---- cusk (individual-level data) ------------------------------------------------------------------------------
def _write_bfiles(d, beds, n, chrs):
    """beds: list of uint8 [m_i, ceil(n/4)] per chromosome"""
    stem = os.path.join(d, "geno")
    with open(stem + ".bed", "wb") as f:
        f.write(bytes([0x6C, 0x1B, 0x01]))
        for b in beds:
            f.write(b.tobytes())
    with open(stem + ".bim", "w") as f:
        for c, b in zip(chrs, beds):
            for i in range(b.shape[0]):
                f.write(f"{c}\trs{c}_{i}\t0\t{1000 + i}\tA\tG\n")
    with open(stem + ".fam", "w") as f:
        for i in range(n):
            f.write(f"f{i} i{i} 0 0 0 -9\n")
    return stem


def _make_cusk_inputs(tmp_path, n, ms, p, stream, miss=0.0):
    from cigwas_b200 import synth

    d = str(tmp_path)
    beds = [synth.random_bed(n, m, stream=stream + k, missing_rate=miss) for k, m in enumerate(ms)]
    stem = _write_bfiles(d, beds, n, [str(k + 1) for k in range(len(ms))])
    rng = synth.rng_for(stream + 100)
    # traits: linear in a few markers of every chromosome + noise, standardised; a few NA
    geno = []
    for b in beds:
        codes = np.stack([(b >> (2 * j)) & 3 for j in range(4)], axis=-1).reshape(b.shape[0], -1)[:, :n]
        geno.append(np.where(codes == 0, 2.0, np.where(codes == 2, 1.0, 0.0)))
    G = np.concatenate(geno, 0)
    ph = np.zeros((p, n))
    for t in range(p):
        idx = rng.choice(G.shape[0], size=4, replace=False)
        ph[t] = (rng.uniform(0.15, 0.4, 4)[:, None] * G[idx]).sum(0) + rng.standard_normal(n)
        if t > 0:
            ph[t] += 0.3 * ph[t - 1]
        ph[t] = (ph[t] - ph[t].mean()) / ph[t].std()
    with open(os.path.join(d, "traits.phen"), "w") as f:
        f.write("FID IID " + " ".join(f"T{t}" for t in range(p)) + "\n")
        for i in range(n):
            vals = ["NA" if (t == 0 and i % 97 == 5) else repr(float(np.float32(ph[t, i]))) for t in range(p)]
            f.write(f"f{i} i{i} " + " ".join(vals) + "\n")
    return d, stem, beds


def _expected_cusk_block(oracle, stem, beds, n, p, chr_ix, first, last, alpha, L1, L2, depth):
    import cigwas_b200

    means = np.array([np.float32(x) for x in open(stem + ".means").read().split()], np.float32)
    stds = np.array([np.float32(x) for x in open(stem + ".stds").read().split()], np.float32)
    g0 = sum(b.shape[0] for b in beds[:chr_ix]) + first
    m = last - first + 1
    bed = np.ascontiguousarray(beds[chr_ix][first:last + 1])
    rows = [ln.split()[2:] for ln in open(os.path.join(os.path.dirname(stem), "traits.phen")).read().split("\n")[1:]
            if ln.split()]
    ph = np.array([[np.nan if v == "NA" else np.float32(float(v)) for v in r] for r in rows], np.float32).T.copy()
    Th = cigwas_b200.threshold_array(n, alpha)
    return host_ref.cusk_expected(oracle, bed, ph.reshape(-1), m, n, p, means[g0:g0 + m], stds[g0:g0 + m], Th, L1, L2,
                                  depth)


def _check_cusk_outputs(got, exp):
    assert got["mdim"] == exp["mdim"]
    assert (got["ixs"] == exp["ixs"]).all()
    assert (got["adj"] == exp["adj"]).all()
    np.testing.assert_allclose(got["corr"], exp["corr"], rtol=1e-6, atol=2e-7)
    assert (got["sep"] == exp["sep"]).all()


def test_cusk_two_chromosomes_blocks_and_scheduler(tmp_path, oracle):
    """prep -> cusk on single blocks (second chromosome: BED offset through the .bim) -> cusk-all over every
    block (LPT queue, resume) gives the same files"""
    n, ms, p = 3000, [90, 140], 3
    d, stem, beds = _make_cusk_inputs(tmp_path, n, ms, p, stream=401)
    run_mps(["prep", stem])
    open(os.path.join(d, "b.blocks"), "w").write("1\t0\t49\n1\t50\t89\n2\t0\t69\n2\t70\t139\n")
    blocks = [(0, 0, 49), (0, 50, 89), (1, 0, 69), (1, 70, 139)]
    out = os.path.join(d, "out")
    os.makedirs(out)
    alpha, L1, L2, depth = 1e-4, 4, 3, 2
    n_written = 0
    for bix in (1, 3):
        c, a, b = blocks[bix]
        r = run_mps(["cusk", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), alpha, L1, L2, depth,
                     out, bix])
        exp = _expected_cusk_block(oracle, stem, beds, n, p, c, a, b, alpha, L1, L2, depth)
        sfile = os.path.join(out, f"{c + 1}_{a}_{b}")
        if exp is None:
            assert "Skipping block" in r.stdout and not os.path.exists(sfile + ".mdim")
        else:
            _check_cusk_outputs(host_ref.read_outputs(sfile), exp)
            n_written += 1
    assert n_written >= 1
    out2 = os.path.join(d, "out_all")
    os.makedirs(out2)
    r = run_mps(["cusk-all", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), alpha, L1, L2, depth,
                 out2])
    assert '"summary": true' in r.stdout and '"failed": 0' in r.stdout
    for bix, (c, a, b) in enumerate(blocks):
        exp = _expected_cusk_block(oracle, stem, beds, n, p, c, a, b, alpha, L1, L2, depth)
        sfile = os.path.join(out2, f"{c + 1}_{a}_{b}")
        if exp is None:
            assert not os.path.exists(sfile + ".mdim")
        else:
            _check_cusk_outputs(host_ref.read_outputs(sfile), exp)
    # resume: nothing left to do
    r = run_mps(["cusk-all", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), alpha, L1, L2, depth,
                 out2])
    assert '"done": 0' in r.stdout


def test_cusk_block_with_missing_calls(tmp_path, oracle):
    n, ms, p = 2100, [120], 2
    d, stem, beds = _make_cusk_inputs(tmp_path, n, ms, p, stream=411, miss=0.01)
    run_mps(["prep", stem])
    open(os.path.join(d, "b.blocks"), "w").write("1\t0\t119\n")
    out = os.path.join(d, "out")
    os.makedirs(out)
    run_mps(["cusk", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), 1e-4, 3, 2, 1, out, 0])
    exp = _expected_cusk_block(oracle, stem, beds, n, p, 0, 0, 119, 1e-4, 3, 2, 1)
    assert exp is not None
    _check_cusk_outputs(host_ref.read_outputs(os.path.join(out, "1_0_119")), exp)


def test_cuskss_all_scheduler(tmp_path, oracle):
    """cuskss-all: every block of a .blocks file (mxm file per block named by the stem) == per-block cuskss runs"""
    from cigwas_b200 import synth

    p, ns = 4, 30000
    sizes = [60, 110, 45]
    d = str(tmp_path)
    f = lambda n: os.path.join(d, n)
    Cs = [synth.sem_correlation(m, p, stream=500 + k, num_samples=ns, causal_per_trait=5) for k, m in enumerate(sizes)]
    # shared pxp from the first block; one mxp table for the genome; one mxm per block
    pxp = Cs[0][sizes[0]:, sizes[0]:]
    names = [f"T{t}" for t in range(p)]
    with open(f("pxp.txt"), "w") as fh:
        fh.write(" ".join(names) + "\n")
        for a in range(p):
            fh.write(names[a] + " " + " ".join(repr(float(pxp[a, b])) for b in range(p)) + "\n")
    with open(f("mxp.txt"), "w") as fh, open(f("g.blocks"), "w") as fb:
        fh.write("chr snp ref " + " ".join(names) + "\n")
        for k, m in enumerate(sizes):
            fb.write(f"{k + 1}\t0\t{m - 1}\n")
            for i in range(m):
                fh.write(f"{k + 1} rs{k}_{i} A " + " ".join(repr(float(Cs[k][i, m + t])) for t in range(p)) + "\n")
            il = np.tril_indices(m)
            Cs[k][:m, :m][il].astype(np.float32).tofile(f(f"mxm_{k + 1}_0_{m - 1}.bin"))
    out = f("out")
    os.makedirs(out)
    r = run_mps(["cuskss-all", f("mxm_{}.bin"), f("mxp.txt"), "NULL", f("pxp.txt"), "NULL", "NULL", f("g.blocks"), 1e-4, 5,
                 2, 2, ns, out])
    assert '"failed": 0' in r.stdout and '"done": 3' in r.stdout
    for k, m in enumerate(sizes):
        o1 = f(f"single{k}")
        os.makedirs(o1)
        run_mps(["cuskss", f(f"mxm_{k + 1}_0_{m - 1}.bin"), f("mxp.txt"), "NULL", f("pxp.txt"), "NULL", "NULL", k,
                 f("g.blocks"), "NULL", 1e-4, 5, 2, 2, ns, o1])
        a = host_ref.read_outputs(os.path.join(out, f"{k + 1}_0_{m - 1}"))
        b = host_ref.read_outputs(os.path.join(o1, f"{k + 1}_0_{m - 1}"))
        for key in ("ixs", "adj", "corr"):
            assert (a[key] == b[key]).all()
        assert a["mdim"] == b["mdim"]


def test_cusk_one_marker_block_after_a_larger_one(tmp_path, oracle):
    """on the tensor-core mxp route a ONE-marker block has no Gram tile at all, and its marker x trait values
    must still reach the square matrix (they used to be scattered only inside the Gram branch, so the skeleton read what
    the previous block had left in the arena).  One scheduler slot, larger block first, then the one-marker block whose
    marker carries the strongest trait signal: both equal the oracle-driven restatement."""
    n, ms, p = 3000, [60], 3
    d, stem, beds = _make_cusk_inputs(tmp_path, n, ms, p, stream=431)
    run_mps(["prep", stem])
    b = beds[0]
    codes = np.stack([(b >> (2 * j)) & 3 for j in range(4)], axis=-1).reshape(b.shape[0], -1)[:, :n]
    G = np.where(codes == 0, 2.0, np.where(codes == 2, 1.0, 0.0))
    rows = [ln.split()[2:] for ln in open(os.path.join(d, "traits.phen")).read().split("\n")[1:] if ln.split()]
    y = np.array([[np.nan if v == "NA" else float(v) for v in r] for r in rows]).T
    ok = ~np.isnan(y).any(axis=0)
    r_abs = np.abs(np.corrcoef(np.vstack([G[:, ok], y[:, ok]]))[:60, 60:]).max(axis=1)
    k = int(np.argmax(r_abs))  # the marker with the strongest trait signal
    a0, a1 = (0, k - 1) if k >= 20 else (k + 1, 59)  # the larger block: everything on one side of it
    open(os.path.join(d, "b.blocks"), "w").write(f"1\t{a0}\t{a1}\n1\t{k}\t{k}\n")
    out = os.path.join(d, "out")
    os.makedirs(out)
    alpha, L1, L2, depth = 1e-4, 3, 2, 1
    r = run_mps(["cusk-all", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), alpha, L1, L2, depth, out],
                env={"CUSK_SLOTS_PER_GPU": "1", "CUSK_DEVICES": "0"})
    assert '"failed": 0' in r.stdout
    exp = _expected_cusk_block(oracle, stem, beds, n, p, 0, k, k, alpha, L1, L2, depth)
    assert exp is not None, "the chosen marker must pass the marker x trait screen"
    _check_cusk_outputs(host_ref.read_outputs(os.path.join(out, f"1_{k}_{k}")), exp)
    exp0 = _expected_cusk_block(oracle, stem, beds, n, p, 0, a0, a1, alpha, L1, L2, depth)
    if exp0 is not None:
        _check_cusk_outputs(host_ref.read_outputs(os.path.join(out, f"1_{a0}_{a1}")), exp0)


def test_cusk_truncated_bed_fails(tmp_path):
    """a .bed cut short must end the run with an error -- code 00 is a valid genotype, so a zero-filled tail
    would give plausible wrong correlations"""
    n, ms, p = 1000, [40], 2
    d, stem, beds = _make_cusk_inputs(tmp_path, n, ms, p, stream=441)
    run_mps(["prep", stem])
    open(os.path.join(d, "b.blocks"), "w").write("1\t0\t39\n")
    size = os.path.getsize(stem + ".bed")
    with open(stem + ".bed", "r+b") as f:
        f.truncate(size - 5 * ((n + 3) // 4))  # the last five markers are gone
    out = os.path.join(d, "out")
    os.makedirs(out)
    r = run_mps(["cusk", os.path.join(d, "traits.phen"), stem, os.path.join(d, "b.blocks"), 1e-4, 3, 2, 1, out, 0],
                check=False)
    assert r.returncode != 0
    assert not os.path.exists(os.path.join(out, "1_0_39.mdim"))
