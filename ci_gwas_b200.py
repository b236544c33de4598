#!/usr/bin/env python
"""Command-line front end of the B200 skeleton path: the `prep-bed`, `block`, `cusk` and `cuskss` sub-commands of the
reference's `ci-gwas.py` (ci-gwas.py:55-62, :64-93, :95-150, :152-252; the het and merged modes are `cuskss` with
`--mxp-se/--pxp-se` and `--marker-indices`), `merge-block-outputs` (ci-gwas.py:254-270, :459-464), `sepselect` and
`orient-v-structs` (ci-gwas.py:302-358, :467-476), with the same
arguments, defaults and input checks (ci-gwas.py:386-457), driving `ci-gwas_b200/mps` instead of
`cusk/build/apps/mps`; the merge and the reformatting of a merged skeleton's output (ci-gwas.py:452-456) are the
native `mps merge` / `mps reformat-merged` instead of cusk_postprocessing/merge_blocks.py.

Additions: `cusk-all` / `cuskss-all` run every block of a block file on all GPUs of the box (LPT schedule, resume
by existing outputs); `--gpus` picks the devices; `cuskss --shard-gpus N` deals the level-l tests of ONE merged
skeleton to N GPUs (NCCL exchange per level).  Everything else of ci-gwas.py (srfci, mvivw, ...)
is outside this path: run it from the reference on the files written here.
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))


def mps_path():
    p = os.environ.get("CI_GWAS_B200_MPS") or os.path.join(ROOT, "ci-gwas_b200", "mps")
    if not os.path.exists(p):
        sys.exit(f"{p} not found: build it first (python -c 'import __graft_entry__ as g; g.build()')")
    return p


class Checked:
    """argparse type with bounds, as ci-gwas.py:27-41 (TypeCheck)"""

    def __init__(self, fn, name, lo=None, hi=None):
        self.fn, self.name, self.lo, self.hi = fn, name, lo, hi

    def __call__(self, val):
        try:
            v = self.fn(val)
        except ValueError:
            raise argparse.ArgumentTypeError(f"{self.name} has to be of type {self.fn.__name__}")
        if self.lo is not None and v < self.lo:
            raise argparse.ArgumentTypeError(f"{self.name} has to be >= {self.lo}")
        if self.hi is not None and v > self.hi:
            raise argparse.ArgumentTypeError(f"{self.name} has to be <= {self.hi}")
        return v


def run(argv, args):
    env = dict(os.environ)
    if getattr(args, "gpus", None):
        env["CUSK_DEVICES"] = args.gpus
    if getattr(args, "shard_gpus", 0) and args.shard_gpus > 1:
        env["CUSK_SHARD_GPUS"] = str(args.shard_gpus)
    subprocess.run([mps_path()] + [str(a) for a in argv], check=True, env=env)


def prep_bed(args):
    run(["prep", args.bfiles], args)


def block(args):
    # ci-gwas.py:390-401
    run(["block", args.bfiles, args.max_block_size, args.device_mem_gb, args.corr_width], args)


def cusk_argv(args):
    return [args.phen, args.bfiles, args.blocks, args.alpha, args.max_level, args.max_level_two, args.max_depth,
            args.outdir]


def cusk(args):
    run(["cusk"] + cusk_argv(args) + [args.block_index], args)


def cusk_all(args):
    run(["cusk-all"] + cusk_argv(args), args)


def check_cuskss(args, need_selection=True):
    if need_selection and args.blockfile == "NULL" and args.marker_indices == "NULL":
        sys.exit("Either blockfile + block index or marker indices into the mxp file have to be provided for cuskss.")
    if sum([args.mxp_se == "NULL", args.pxp_se == "NULL"]) == 1:
        sys.exit("Please provide no or both pxp and mxp standard error files.")
    if sum([args.mxp == "NULL", args.mxm == "NULL"]) == 1:
        sys.exit("Please provide no or both mxp and mxm correlation files.")


def cuskss_argv(args, block_index, marker_indices):
    return [args.mxm, args.mxp, args.mxp_se, args.pxp, args.pxp_se, args.time_index, block_index, args.blockfile,
            marker_indices, args.alpha, args.max_level_one, args.max_level_two, args.max_depth, args.num_samples,
            args.outdir]


def cuskss(args):
    check_cuskss(args)
    run(["cuskss"] + cuskss_argv(args, args.block_index, args.marker_indices), args)
    if args.marker_indices != "NULL":
        # ci-gwas.py:452-456: rewrite the merged output in the merge-block-outputs format
        run(["reformat-merged", args.outdir], args)


def merge_blocks(args):
    run(["merge", args.blockfile, args.cusk_output_dir], args)  # ci-gwas.py:459-464


def sepselect(args):
    # ci-gwas.py:467-476: `sepselect` and `orient-v-structs` both end in MergedCuskResults.to_file(<dir>/max_sep_min_pc);
    # the reference's `sepselect` stops there on its missing PAG, so both sub-commands run the orient-v-structs flow
    prior = getattr(args, "orientation_prior", None)
    run(["sepselect", args.cusk_result_stem, repr(args.alpha), args.num_samples, prior or "NULL"], args)


def cuskss_all(args):
    if args.blockfile == "NULL":
        sys.exit("cuskss-all needs --blockfile.")
    check_cuskss(args, need_selection=False)
    argv = cuskss_argv(args, 0, "NULL")
    del argv[8]  # no marker indices ...
    del argv[6]  # ... and no block index: every block of the file
    run(["cuskss-all"] + argv, args)


def add_device_args(p, shard=False):
    p.add_argument("--gpus", type=str, default=None, help="comma-separated CUDA device ids (default: all)")
    if shard:
        p.add_argument("--shard-gpus", type=Checked(int, "shard-gpus", 1, 64), default=1,
                       help="with --marker-indices: deal the level-l tests of the one skeleton to this many GPUs")


def add_cusk_args(p, with_index):
    if with_index:
        p.add_argument("block_index", type=Checked(int, "block-index", 0, None),
                       help="0-based index of the row of the block file to run")
    p.add_argument("blocks", type=str, help="file with genomic block definitions (output of ci-gwas.py block)")
    p.add_argument("bfiles", type=str, help="filestem of .bed, .bim, .fam fileset")
    p.add_argument("phen", type=str, help="path to standardized phenotype tsv.")
    p.add_argument("alpha", type=Checked(float, "alpha", 0.0, 1.0), nargs="?", default=10**-4,
                   help="significance level")
    p.add_argument("max_level", type=Checked(int, "max-level", 0, 14), nargs="?", default=3,
                   help="maximal size of separation sets in the first round")
    p.add_argument("max_level_two", type=Checked(int, "max-level", 0, 14), nargs="?", default=14,
                   help="maximal size of separation sets in the second round")
    p.add_argument("max_depth", type=Checked(int, "max-depth", 1, None), nargs="?", default=1,
                   help="max depth at which marker variables are kept as ancestors")
    p.add_argument("outdir", type=str, nargs="?", default="./", help="directory for output")
    add_device_args(p)


def add_cuskss_args(p, with_index):
    p.add_argument("--mxm", type=str, default="NULL",
                   help="marker-marker correlations (binary lower triangle)" +
                        ("" if with_index else "; a pattern, '{}' stands for the block's <chr>_<first>_<last> stem"))
    p.add_argument("--mxp", type=str, default="NULL", help="marker-trait correlations (text, space-separated)")
    p.add_argument("--pxp", type=str, required=True, help="trait-trait correlations (text)")
    p.add_argument("--mxp-se", type=str, default="NULL", help="standard errors of the marker-trait correlations")
    p.add_argument("--pxp-se", type=str, default="NULL", help="standard errors of the trait-trait correlations")
    if with_index:
        p.add_argument("--block-index", type=Checked(int, "block-index", 0, None), default=0,
                       help="0-based index of the row of the block file to run")
        p.add_argument("--marker-indices", type=str, default="NULL",
                       help="row indices of the markers in the mxp file (binary int32); selects the merged mode")
    p.add_argument("--blockfile", type=str, default="NULL", help="file with genomic block definitions")
    p.add_argument("--alpha", type=Checked(float, "alpha", 0.0, 1.0), required=True, help="significance level")
    p.add_argument("--max-level-one", type=Checked(int, "max-level", 0, 14), default=3,
                   help="maximal size of separation sets in the first round")
    p.add_argument("--max-level-two", type=Checked(int, "max-level", 0, 14), default=14,
                   help="maximal size of separation sets in the second round")
    p.add_argument("--max-depth", type=Checked(int, "max-depth", 1, None), default=1,
                   help="max depth at which marker variables are kept as ancestors")
    p.add_argument("--time-index", type=str, default="NULL", help="one time index per trait (text, one per line)")
    p.add_argument("--num-samples", type=Checked(int, "num-samples", 1, None), required=True,
                   help="number of samples used for computing correlations")
    p.add_argument("--outdir", type=str, default="./", help="directory for output")
    add_device_args(p, shard=with_index)


def build_parser():
    ap = argparse.ArgumentParser(description="CI-GWAS skeleton path on B200: prep-bed, cusk, cuskss")
    sub = ap.add_subparsers(title="commands", dest="cmd", required=True)
    p = sub.add_parser("prep-bed", help="Prepare PLINK bed file for cusk")
    p.add_argument("bfiles", type=str, help="filestem of .bed, .bim, .fam fileset")
    add_device_args(p)
    p.set_defaults(func=prep_bed)
    p = sub.add_parser("block", help="Tile whole-genome LD matrix into block diagonal matrix")  # ci-gwas.py:64-93
    p.add_argument("bfiles", type=str, help="filestem of .bed, .bim, .fam fileset")
    p.add_argument("max_block_size", metavar="max-block-size", type=Checked(int, "max-block-size", 2),
                   help="maximum number of markers per block")
    p.add_argument("device_mem_gb", metavar="device-mem-gb", type=Checked(int, "device-mem-gb", 0),
                   help="maximum memory available on GPU in GB (0: all that is free)")
    p.add_argument("corr_width", metavar="corr-width", type=Checked(int, "corr-width", 2),
                   help="width of banded-correlation matrix")
    add_device_args(p)
    p.set_defaults(func=block)
    p = sub.add_parser("cusk", help="Skeleton of one genomic block from individual-level data")
    add_cusk_args(p, True)
    p.set_defaults(func=cusk)
    p = sub.add_parser("cusk-all", help="cusk for every block of the block file, on all GPUs")
    add_cusk_args(p, False)
    p.set_defaults(func=cusk_all)
    p = sub.add_parser("cuskss", help="Skeleton from summary statistics (block, het and merged modes)")
    add_cuskss_args(p, True)
    p.set_defaults(func=cuskss)
    p = sub.add_parser("cuskss-all", help="cuskss for every block of the block file, on all GPUs")
    add_cuskss_args(p, False)
    p.set_defaults(func=cuskss_all)
    p = sub.add_parser("merge-block-outputs", help="Merge outputs from multiple blocks processed with cusk or cuskss")
    p.add_argument("cusk_output_dir", metavar="cusk-output-dir", type=str, help="output directory of cusk or cuskss")
    p.add_argument("blockfile", type=str, help="file with genomic block definitions (output of ci-gwas block)")
    p.set_defaults(func=merge_blocks)
    for name, hlp in [("sepselect", "Compute maximal and partial-correlation-minimizing separation sets on merged cusk skeletons"),
                      ("orient-v-structs", "Orient v-structures using maximal separation sets on merged cusk skeletons.")]:
        p = sub.add_parser(name, help=hlp)  # ci-gwas.py:302-358
        p.add_argument("cusk_result_stem", metavar="cusk-result-stem", type=str, help="outdir + stem of merged cusk results")
        p.add_argument("alpha", type=Checked(float, "alpha", 0.0, 1.0), help="significance level for conditional independence tests")
        p.add_argument("num_samples", metavar="num-samples", type=Checked(int, "num-samples", 1, None),
                       help="number of samples used for computing correlations")
        if name == "orient-v-structs":
            p.add_argument("--orientation-prior", metavar="orientation-prior", type=str, default=None,
                           help="matrix of (0, 1) (32 bit integers, binary) of dims (n_trait, n_trait) indicating directions to be forced.")
        add_device_args(p)
        p.set_defaults(func=sepselect)
    return ap


def main(argv=None):
    args = build_parser().parse_args(argv)
    try:
        args.func(args)
    except subprocess.CalledProcessError as e:
        sys.exit(e.returncode)


if __name__ == "__main__":
    main()
