"""CPU tests of the command-line front end (ci_gwas_b200.py): the reference's sub-commands, defaults and input
checks (ci-gwas.py:55-62, :95-252, :386-457) mapped onto the `mps` argv contract.  `mps` itself is replaced by
a recorder script here."""
import json
import os
import stat
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "ci_gwas_b200.py")


@pytest.fixture()
def fake_mps(tmp_path):
    rec = tmp_path / "argv.json"
    exe = tmp_path / "mps"
    exe.write_text("#!%s\nimport json, os, sys\np = %r\nprev = json.load(open(p))['calls'] if os.path.exists(p) else []\n"
                   "json.dump({'argv': sys.argv[1:], 'devices': os.environ.get('CUSK_DEVICES'), "
                   "'shard': os.environ.get('CUSK_SHARD_GPUS'), 'calls': prev + [sys.argv[1:]]}, open(p, 'w'))\n"
                   "sys.exit(int(os.environ.get('FAKE_RC', '0')))\n" % (sys.executable, str(rec)))
    exe.chmod(exe.stat().st_mode | stat.S_IEXEC)
    return str(exe), str(rec)


def cli(args, mps, rc_env=None):
    env = dict(os.environ, CI_GWAS_B200_MPS=mps)
    if rc_env is not None:
        env["FAKE_RC"] = str(rc_env)
    return subprocess.run([sys.executable, CLI] + args, capture_output=True, text=True, env=env)


def test_prep_bed_and_cusk(fake_mps):
    mps, rec = fake_mps
    assert cli(["prep-bed", "/data/geno"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["prep", "/data/geno"]
    # positional order of ci-gwas.py cusk: block_index blocks bfiles phen [alpha max_level max_level_two max_depth outdir]
    assert cli(["cusk", "7", "b.blocks", "/data/geno", "t.phen"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["cusk", "t.phen", "/data/geno", "b.blocks", "0.0001", "3", "14", "1", "./", "7"]
    assert cli(["cusk-all", "b.blocks", "/data/geno", "t.phen", "0.001", "2", "6", "2", "out", "--gpus", "0,3"], mps).returncode == 0
    r = json.load(open(rec))
    assert r["argv"] == ["cusk-all", "t.phen", "/data/geno", "b.blocks", "0.001", "2", "6", "2", "out"]
    assert r["devices"] == "0,3"
    assert cli(["cusk", "0", "b", "g", "p", "0.1", "15"], mps).returncode == 2  # max level above 14


def test_block(fake_mps):
    mps, rec = fake_mps
    # ci-gwas.py:64-93, :390-401: bfiles max-block-size device-mem-gb corr-width, all positional
    assert cli(["block", "/data/geno", "11000", "10", "2000", "--gpus", "2"], mps).returncode == 0
    r = json.load(open(rec))
    assert r["argv"] == ["block", "/data/geno", "11000", "10", "2000"] and r["devices"] == "2"
    assert cli(["block", "/data/geno", "1", "10", "2000"], mps).returncode == 2   # max-block-size >= 2
    assert cli(["block", "/data/geno", "100", "10", "1"], mps).returncode == 2    # corr-width >= 2
    assert cli(["block", "/data/geno", "100", "1.5", "20"], mps).returncode == 2  # device-mem-gb is an int
    assert cli(["block", "/data/geno", "100"], mps).returncode == 2


def test_cuskss_modes_and_checks(fake_mps):
    mps, rec = fake_mps
    base = ["cuskss", "--pxp", "pxp.txt", "--alpha", "0.0001", "--num-samples", "458000"]
    r = cli(base, mps)
    assert r.returncode != 0 and "Either blockfile + block index or marker indices" in r.stderr
    r = cli(base + ["--blockfile", "b", "--mxp-se", "se.txt"], mps)
    assert r.returncode != 0 and "no or both pxp and mxp standard error" in r.stderr
    r = cli(base + ["--blockfile", "b", "--mxm", "m.bin"], mps)
    assert r.returncode != 0 and "no or both mxp and mxm" in r.stderr
    assert cli(base + ["--blockfile", "b", "--block-index", "4", "--mxm", "m.bin", "--mxp", "mxp.txt",
                       "--time-index", "ti.txt", "--outdir", "o"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["cuskss", "m.bin", "mxp.txt", "NULL", "pxp.txt", "NULL", "ti.txt", "4", "b",
                                            "NULL", "0.0001", "3", "14", "1", "458000", "o"]
    # merged + het + sharded over 4 GPUs
    assert cli(base + ["--marker-indices", "ix.bin", "--mxm", "m.bin", "--mxp", "mxp.txt", "--mxp-se", "a", "--pxp-se", "b",
                       "--shard-gpus", "4"], mps).returncode == 0
    r = json.load(open(rec))
    assert r["calls"][-2][:10] == ["cuskss", "m.bin", "mxp.txt", "a", "pxp.txt", "b", "NULL", "0", "NULL", "ix.bin"]
    assert r["shard"] == "4"
    # ci-gwas.py:452-456: a merged skeleton's output is then rewritten in the merge-block-outputs format
    assert r["calls"][-1] == ["reformat-merged", "./"]
    # all blocks
    assert cli(["cuskss-all", "--pxp", "pxp.txt", "--alpha", "0.01", "--num-samples", "1000", "--blockfile", "b",
                "--mxm", "mxm/{}.bin", "--mxp", "mxp.txt"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["cuskss-all", "mxm/{}.bin", "mxp.txt", "NULL", "pxp.txt", "NULL", "NULL", "b",
                                            "0.01", "3", "14", "1", "1000", "./"]


def test_merge_block_outputs(fake_mps):
    mps, rec = fake_mps
    # ci-gwas.py:254-270: positional <cusk-output-dir> <blockfile>
    assert cli(["merge-block-outputs", "out/dir", "all.blocks"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["merge", "all.blocks", "out/dir"]


def test_sepselect_and_orient_v_structs(fake_mps):
    mps, rec = fake_mps
    # ci-gwas.py:335-358 / :302-333: positional <cusk-result-stem> <alpha> <num-samples> [--orientation-prior]
    assert cli(["sepselect", "out/merged_blocks", "0.0001", "458000"], mps).returncode == 0
    assert json.load(open(rec))["argv"] == ["sepselect", "out/merged_blocks", "0.0001", "458000", "NULL"]
    assert cli(["orient-v-structs", "out/merged_blocks", "0.001", "1000", "--orientation-prior", "p.bin", "--gpus", "2"], mps).returncode == 0
    r = json.load(open(rec))
    assert r["argv"] == ["sepselect", "out/merged_blocks", "0.001", "1000", "p.bin"] and r["devices"] == "2"
    assert cli(["sepselect", "s", "1.5", "10"], mps).returncode == 2  # alpha above 1


def test_exit_status_propagates(fake_mps):
    mps, _ = fake_mps
    assert cli(["prep-bed", "x"], mps, rc_env=1).returncode == 1
    r = subprocess.run([sys.executable, CLI, "prep-bed", "x"], capture_output=True, text=True,
                       env=dict(os.environ, CI_GWAS_B200_MPS="/nonexistent/mps"))
    assert r.returncode != 0 and "not found" in r.stderr
