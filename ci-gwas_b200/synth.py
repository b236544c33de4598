"""Deterministic synthetic inputs of the shapes BASELINE.json names.

All generators are numpy-only and seeded (PCG64, master seed 20261104 + stream id) so that the
parity tests, the bench and the reference arm see bit-identical inputs.
"""
import os

import numpy as np

MASTER_SEED = 20261104


def rng_for(stream):
    return np.random.Generator(np.random.PCG64([MASTER_SEED, int(stream)]))


def sem_correlation(m, p, stream=2, sub_block=50, window=8, max_parents=3, causal_per_trait=30,
                    trait_density=0.3, num_samples=None, dtype=np.float32, wlo=0.3, whi=0.7):
    """Population correlation matrix of a linear SEM: m markers in LD sub-blocks followed by p traits.

    Markers: each marker has 1..max_parents parents among the previous `window` markers of its
    sub-block (weights U(0.3,0.7), random sign, rescaled so that the explained variance stays
    below 0.9); the first marker of a sub-block depends on the last marker of the previous one
    with weight 0.3.  Traits: `causal_per_trait` causal markers (|beta| ~ U(0.01,0.05)) and
    lower-triangular trait-trait effects U(-0.3,0.3) with density `trait_density`.  All variables
    have unit variance, so the covariance A A' is the correlation matrix (exactly PSD).
    num_samples: if given, symmetric N(0, 1/num_samples) sampling noise is added off the diagonal.
    Returns C [m+p, m+p] with unit diagonal, markers first (the order cusk's square matrix uses,
    cli.cpp:597-649).
    """
    rng = rng_for(stream)
    nv = m + p
    A = np.zeros((nv, nv), np.float64)  # x = A e, e iid N(0,1)
    for i in range(m):
        first_of_block = (i % sub_block) == 0
        row = np.zeros(nv, np.float64)
        if i > 0:
            if first_of_block:
                par = np.array([i - 1])
                w = np.array([0.3])
            else:
                lo = max(i - window, i - (i % sub_block))
                k = int(rng.integers(1, max_parents + 1))
                k = min(k, i - lo)
                par = rng.choice(np.arange(lo, i), size=k, replace=False)
                w = rng.uniform(wlo, whi, size=k) * rng.choice([-1.0, 1.0], size=k)
            row = w @ A[par]
            v = float(row @ row)
            if v > 0.9:
                row *= np.sqrt(0.9 / v)
                v = 0.9
        else:
            v = 0.0
        row[i] = np.sqrt(1.0 - v)
        A[i] = row
    for t in range(p):
        i = m + t
        k = min(causal_per_trait, m)
        par = rng.choice(np.arange(m), size=k, replace=False)
        w = rng.uniform(0.01, 0.05, size=k) * rng.choice([-1.0, 1.0], size=k)
        row = w @ A[par]
        for u in range(t):
            if rng.random() < trait_density:
                row = row + rng.uniform(-0.3, 0.3) * A[m + u]
        v = float(row @ row)
        if v > 0.9:
            row *= np.sqrt(0.9 / v)
            v = 0.9
        row[i] = np.sqrt(1.0 - v)
        A[i] = row
    Cm = A @ A.T
    if num_samples:
        E = rng.standard_normal((nv, nv)) / np.sqrt(float(num_samples))
        E = np.triu(E, 1)
        Cm = Cm + E + E.T
    np.fill_diagonal(Cm, 1.0)
    Cm = np.clip(Cm, -0.999999, 0.999999)
    np.fill_diagonal(Cm, 1.0)
    Cm = Cm.astype(dtype)
    # exact symmetry after the cast
    Cm = np.triu(Cm) + np.triu(Cm, 1).T
    return np.ascontiguousarray(Cm)


def c2_single_block(m=3000, p=17, num_samples=458000, stream=2):
    """BASELINE.json configs[1]: cuskss single LD block, 3k markers + 17 traits."""
    return sem_correlation(m, p, stream=stream, num_samples=num_samples)


def c2_dense_block(m=3000, p=17, stream=2, sub_block=50, rho=0.9, cross=0.3, causal_per_trait=30, trait_density=0.3,
                   num_samples=458000, beta_lo=0.01, beta_hi=0.05, dtype=np.float32):
    """SURVEY 8d, config C2 as specified: marker LD is AR(1) inside sub-blocks of `sub_block` markers
    (r_ij = rho^|i-j|), `cross` times that between markers of neighbouring sub-blocks and 0 beyond, projected to
    positive definite (eigenvalues clipped at 1e-3, unit diagonal restored); every trait has `causal_per_trait`
    causal markers (|beta| ~ U(beta_lo, beta_hi), random sign) plus lower-triangular trait-trait effects
    U(-0.3, 0.3) with density `trait_density`; marker x trait and trait x trait correlations follow analytically
    (R beta, beta' R beta).  Sampling noise N(0, 1/num_samples) is added off the diagonal as in sem_correlation.
    Returns C [m+p, m+p] float32, markers first."""
    rng = rng_for(stream + 200)
    idx = np.arange(m)
    dist = np.abs(idx[:, None] - idx[None, :]).astype(np.float64)
    blk = idx // sub_block
    dblk = np.abs(blk[:, None] - blk[None, :])
    R = np.where(dblk == 0, rho ** dist, np.where(dblk == 1, cross * rho ** dist, 0.0))
    w, V = np.linalg.eigh(R)
    R = (V * np.clip(w, 1e-3, None)) @ V.T
    s = 1.0 / np.sqrt(np.diag(R))
    R = R * s[:, None] * s[None, :]
    nv = m + p
    Wm = np.zeros((p, m))  # trait = Wm g + Wn e (e: the traits' own unit noises)
    Wn = np.zeros((p, p))
    for t in range(p):
        k = min(causal_per_trait, m)
        par = rng.choice(m, size=k, replace=False)
        Wm[t, par] += rng.uniform(beta_lo, beta_hi, size=k) * rng.choice([-1.0, 1.0], size=k)
        for u in range(t):
            if rng.random() < trait_density:
                a = rng.uniform(-0.3, 0.3)
                Wm[t] += a * Wm[u]
                Wn[t] += a * Wn[u]
        v = float(Wm[t] @ R @ Wm[t] + Wn[t] @ Wn[t])
        if v > 0.9:
            sc = np.sqrt(0.9 / v)
            Wm[t] *= sc
            Wn[t] *= sc
            v = 0.9
        Wn[t, t] = np.sqrt(1.0 - v)
    Cm = np.zeros((nv, nv))
    Cm[:m, :m] = R
    Cm[:m, m:] = R @ Wm.T
    Cm[m:, :m] = Cm[:m, m:].T
    Cm[m:, m:] = Wm @ R @ Wm.T + Wn @ Wn.T
    if num_samples:
        E = rng.standard_normal((nv, nv)) / np.sqrt(float(num_samples))
        E = np.triu(E, 1)
        Cm = Cm + E + E.T
    np.fill_diagonal(Cm, 1.0)
    Cm = np.clip(Cm, -0.999999, 0.999999)
    np.fill_diagonal(Cm, 1.0)
    Cm = Cm.astype(dtype)
    Cm = np.triu(Cm) + np.triu(Cm, 1).T
    return np.ascontiguousarray(Cm)


def het_sample_sizes(nv, m, num_samples, stream=5, lo=50000, na_rate=0.005):
    """Effective sample sizes of a heterologous-correlation run: marker-marker entries are the
    nominal sample size, marker-trait / trait-trait entries are U(lo, num_samples) and a fraction
    is NaN (missing SE), symmetric."""
    rng = rng_for(stream + 1000)
    N = np.full((nv, nv), float(num_samples), np.float64)
    blk = rng.uniform(lo, num_samples, size=(nv, nv - m))
    N[:, m:] = blk
    N[m:, :] = N[:, m:].T
    N = np.triu(N) + np.triu(N, 1).T
    mask = rng.random((nv, nv)) < na_rate
    mask[:, :m] &= False
    mask = np.triu(mask, 1)
    mask = mask | mask.T
    N[mask] = np.nan
    return N.astype(np.float32)


def random_bed(n_samples, m, stream=3, missing_rate=0.0, sub_block=50, copy_prob=0.95):
    """Packed PLINK BED columns (no magic bytes) of an LD-structured genotype block.

    Haplotype Markov chain along markers within sub-blocks (copy the previous marker's allele with
    prob `copy_prob`), MAF ~ U(0.01, 0.5), two haplotypes summed (HWE).  Returns uint8
    [m, ceil(n/4)].  Coding (bed_lut_gpu.h): 00 -> 2, 01 -> missing, 10 -> 1, 11 -> 0.
    """
    rng = rng_for(stream)
    nb = (n_samples + 3) // 4
    out = np.zeros((m, nb), np.uint8)
    code_of = np.array([3, 2, 0], np.uint8)  # genotype 0,1,2 -> 2-bit code
    hap = None
    for j in range(m):
        maf = rng.uniform(0.01, 0.5)
        fresh = (rng.random((2, n_samples)) < maf)
        if hap is None or (j % sub_block) == 0:
            hap = fresh
        else:
            keep = rng.random((2, n_samples)) < copy_prob
            hap = np.where(keep, hap, fresh)
        g = hap[0].astype(np.uint8) + hap[1].astype(np.uint8)
        codes = code_of[g]
        if missing_rate > 0:
            miss = rng.random(n_samples) < missing_rate
            codes = np.where(miss, np.uint8(1), codes)
        pad = np.full(nb * 4 - n_samples, 0, np.uint8)  # padding bits 00
        c4 = np.concatenate([codes, pad]).reshape(nb, 4)
        out[j] = c4[:, 0] | (c4[:, 1] << 2) | (c4[:, 2] << 4) | (c4[:, 3] << 6)
    return out


def merged_correlation(loci=400, per_locus=50, p=17, stream=5, num_samples=458000, causal_per_trait=12,
                       window=12, max_parents=4, wlo=0.6, whi=0.95, dtype=np.float32, beta_lo=0.01, beta_hi=0.05):
    """BASELINE.json configs[4] (cuskss-merged): `loci` independent LD loci of `per_locus` selected markers
    each, followed by p traits.  Same SEM as sem_correlation inside a locus; markers of different loci
    are independent, so the population matrix is assembled locus by locus (no dense nv x nv product).
    Sampling noise N(0, 1/num_samples) is added off the diagonal.  Returns C [nv, nv] float32."""
    rng = rng_for(stream)
    m = loci * per_locus
    nv = m + p
    A_loc = []
    for _ in range(loci):
        A = np.zeros((per_locus, per_locus), np.float64)
        for i in range(per_locus):
            row = np.zeros(per_locus, np.float64)
            v = 0.0
            if i > 0:
                lo = max(i - window, 0)
                k = min(int(rng.integers(1, max_parents + 1)), i - lo)
                par = rng.choice(np.arange(lo, i), size=k, replace=False)
                w = rng.uniform(wlo, whi, size=k) * rng.choice([-1.0, 1.0], size=k)
                row = w @ A[par]
                v = float(row @ row)
                if v > 0.9:
                    row *= np.sqrt(0.9 / v)
                    v = 0.9
            row[i] = np.sqrt(1.0 - v)
            A[i] = row
        A_loc.append(A)
    # trait rows in the global noise basis (markers' noises, then the traits' own)
    AT = np.zeros((p, nv), np.float64)
    for t in range(p):
        par = rng.choice(np.arange(m), size=min(causal_per_trait, m), replace=False)
        w = rng.uniform(beta_lo, beta_hi, size=par.size) * rng.choice([-1.0, 1.0], size=par.size)
        row = np.zeros(nv, np.float64)
        for i, wi in zip(par, w):
            L, j = divmod(int(i), per_locus)
            row[L * per_locus:(L + 1) * per_locus] += wi * A_loc[L][j]
        for u in range(t):
            if rng.random() < 0.3:
                row = row + rng.uniform(-0.3, 0.3) * AT[u]
        v = float(row @ row)
        if v > 0.9:
            row *= np.sqrt(0.9 / v)
            v = 0.9
        row[m + t] = np.sqrt(1.0 - v)
        AT[t] = row
    Cm = rng.standard_normal((nv, nv), dtype=np.float32)
    Cm *= np.float32(1.0 / np.sqrt(float(num_samples)))
    Cm = np.triu(Cm, 1)
    Cm = Cm + Cm.T
    for L in range(loci):
        s = slice(L * per_locus, (L + 1) * per_locus)
        Cm[s, s] += (A_loc[L] @ A_loc[L].T).astype(np.float32)
        mt = (A_loc[L] @ AT[:, s].T).astype(np.float32)  # [per_locus, p]
        Cm[s, m:] += mt
        Cm[m:, s] += mt.T
    Cm[m:, m:] += (AT @ AT.T).astype(np.float32)
    np.clip(Cm, -0.999999, 0.999999, out=Cm)
    np.fill_diagonal(Cm, 1.0)
    Cm = np.triu(Cm) + np.triu(Cm, 1).T
    return np.ascontiguousarray(Cm.astype(dtype))


def simulate_dag(n=10000, snp=1600, tr=10, n_latent=2, deg=3, prob_pleio=0.2, lo_mp=0.001, hi_mp=0.05, lo_pp=0.001,
                 hi_pp=0.2, stream=1):
    """BASELINE.json configs[0]: the small synthetic DAG of simulation/simulate_dag.R (gen_rand_dag, :3-98, called
    with the constants of :106-116; the file uses n = 16 000, BASELINE.json names n = 10 000).

    Nodes in topological order: `snp` markers, `n_latent` latent variables, `tr` traits.  Edges only point to later
    nodes: a marker reaches each later node with probability deg/snp (:22-28); a marker with exactly one trait child
    gets each OTHER trait with probability prob_pleio (:30-39); a latent variable or trait reaches each later node with
    probability min(1, deg/tr) (:41-47).  |effect| ~ U(lo_mp, hi_mp) from a marker to a latent variable or trait,
    U(lo_pp, hi_pp) otherwise, random sign (:53-78).  A node without parents is N(0,1); otherwise the parents'
    combination plus N(0, 1 - var) noise (:80-90; a negative residual variance, NaN in R, is clipped to 1e-6 here).
    The latent columns are dropped and the sample correlation of the rest is what cuskss gets (:118-124).
    numpy's PCG64 stream, not R's Mersenne twister: the same distributions, not the same draws.
    Returns (C [snp+tr, snp+tr] float32, true adjacency [snp+tr, snp+tr] bool of the DAG without latent variables)."""
    rng = rng_for(1000 + stream)
    pq = snp + n_latent + tr
    first_trait = snp + n_latent
    G = np.zeros((pq, pq), bool)
    for i in range(snp):
        G[i, i + 1:] = rng.random(pq - i - 1) < deg / snp
    for i in range(snp):
        kids = np.flatnonzero(G[i, first_trait:])
        if kids.size == 1:
            others = np.setdiff1d(np.arange(tr), kids)
            G[i, first_trait + others] = rng.random(tr - 1) < prob_pleio
    for j in range(snp, pq):
        G[j, j + 1:] = rng.random(pq - j - 1) < min(1.0, deg / tr)
    A = np.zeros((pq, pq))
    for i in range(pq):
        kids = np.flatnonzero(G[i])
        lo = np.where((i < snp) & (kids >= snp), lo_mp, lo_pp)
        hi = np.where((i < snp) & (kids >= snp), hi_mp, hi_pp)
        A[i, kids] = rng.uniform(lo, hi) * np.where(rng.random(kids.size) < 0.5, -1.0, 1.0)
    x = np.empty((n, pq))
    for i in range(pq):
        anc = np.flatnonzero(G[:, i])
        if anc.size == 0:
            x[:, i] = rng.standard_normal(n)
        else:
            g = x[:, anc] @ A[anc, i]
            x[:, i] = g + rng.standard_normal(n) * np.sqrt(max(1.0 - g.var(ddof=1), 1e-6))
    keep = np.r_[np.arange(snp), np.arange(first_trait, pq)]
    C = np.corrcoef(x[:, keep], rowvar=False).astype(np.float32)
    np.fill_diagonal(C, 1.0)
    truth = G[np.ix_(keep, keep)]
    return C, truth | truth.T


# ---- BASELINE.json configs[3] (C4): a whole genome of LD blocks as cuskss input files ---------------------------------
def c4_block_sizes(nblocks=1700, stream=4, median=600, sigma=0.6, lo=50, hi=3000):
    """sizes clip(round(LogNormal(ln 600, 0.6)), 50, 3 000)"""
    rng = rng_for(stream)
    return np.clip(np.rint(rng.lognormal(np.log(median), sigma, size=nblocks)), lo, hi).astype(np.int64)


def _c4_marker_factor(m, stream, sub_block=50, window=12, max_parents=4, wlo=0.6, whi=0.95):
    """marker part of sem_correlation (the bench's C2 generator): x = A e with unit variances"""
    rng = rng_for(stream)
    A = np.zeros((m, m), np.float64)
    for i in range(m):
        row = np.zeros(m, np.float64)
        v = 0.0
        if i > 0:
            if (i % sub_block) == 0:
                row = 0.3 * A[i - 1]
            else:
                lo = max(i - window, i - (i % sub_block))
                k = min(int(rng.integers(1, max_parents + 1)), i - lo)
                par = rng.choice(np.arange(lo, i), size=k, replace=False)
                w = rng.uniform(wlo, whi, size=k) * rng.choice([-1.0, 1.0], size=k)
                row = w @ A[par]
            v = float(row @ row)
            if v > 0.9:
                row = row * np.sqrt(0.9 / v)
                v = 0.9
        row[i] = np.sqrt(1.0 - v)
        A[i] = row
    return A, rng


def _c4_worker_init():
    """forked generator workers: one BLAS thread each (the pool already uses every core)"""
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(1)
    except Exception:  # noqa: BLE001
        pass


def _c4_write_block(args):
    """worker: one block's mxm file; returns (block, C_b beta [m, p], beta' C_b beta [p, p])"""
    b, m, stream, beta, outdir, stem, num_samples = args
    A, rng = _c4_marker_factor(int(m), stream)
    Cb = A @ A.T
    E = rng.standard_normal((m, m)) / np.sqrt(float(num_samples))
    E = np.triu(E, 1)
    Cs = np.clip(Cb + E + E.T, -0.999999, 0.999999)
    np.fill_diagonal(Cs, 1.0)
    Cs.astype(np.float32)[np.tril_indices(m)].tofile(os.path.join(outdir, "mxm", stem + ".bin"))
    cb = Cb @ beta  # [m, p] population covariance of the block's markers with the traits' genetic parts
    return b, cb, beta.T @ cb


def c4_genome(outdir, nblocks=1700, p=17, stream=4, num_samples=458000, causal_total=300, trait_density=0.3, chromosomes=22,
              procs=None):
    """Writes a synthetic whole genome as `mps cuskss-all` / `ci-gwas.py cuskss` input: <outdir>/genome.blocks
    (chr, first, last per chromosome), <outdir>/mxm/<chr>_<first>_<last>.bin (float32 lower triangle incl. diagonal), one
    <outdir>/mxp.txt for the whole genome and <outdir>/pxp.txt.  Every block's marker LD comes from the bench's C2
    generator with its own stream; each of the p traits has `causal_total` causal markers spread over the genome
    (|beta| ~ U(0.01, 0.05)) and lower-triangular trait-trait effects; marker-trait and trait-trait correlations follow
    analytically (blocks are independent) plus N(0, 1/num_samples) noise.  Returns dict(sizes, stems, markers)."""
    import multiprocessing as mp

    os.makedirs(os.path.join(outdir, "mxm"), exist_ok=True)
    sizes = c4_block_sizes(nblocks, stream)
    rng = rng_for(stream + 7000)
    total = int(sizes.sum())
    starts = np.concatenate([[0], np.cumsum(sizes)[:-1]])
    # chromosomes: consecutive runs of blocks
    chr_of = (np.arange(nblocks) * chromosomes // nblocks) + 1
    stems, first_in_chr = [], {}
    off = {}
    for b in range(nblocks):
        c = int(chr_of[b])
        f = off.get(c, 0)
        stems.append(f"{c}_{f}_{f + int(sizes[b]) - 1}")
        off[c] = f + int(sizes[b])
    # causal markers and effects, per trait, genome-wide
    betas = [np.zeros((int(m), p)) for m in sizes]
    for t in range(p):
        ix = rng.choice(total, size=min(causal_total, total), replace=False)
        w = rng.uniform(0.01, 0.05, size=ix.size) * rng.choice([-1.0, 1.0], size=ix.size)
        bl = np.searchsorted(starts, ix, side="right") - 1
        for i, wi, b in zip(ix, w, bl):
            betas[b][i - starts[b], t] += wi
    jobs = [(b, int(sizes[b]), 100000 * stream + b, betas[b], outdir, stems[b], num_samples) for b in range(nblocks)]
    procs = procs or min(32, os.cpu_count() or 1)
    covG = np.zeros((p, p))
    cbs = [None] * nblocks
    with mp.get_context("fork").Pool(procs, initializer=_c4_worker_init) as pool:
        for b, cb, g in pool.imap_unordered(_c4_write_block, jobs, chunksize=4):
            cbs[b] = cb
            covG += g
    # traits: z = G + e (unit variance), y = D M z with M unit lower triangular (trait -> later trait effects)
    ve = np.clip(1.0 - np.diag(covG), 0.05, None)
    covZ = covG + np.diag(ve)
    M = np.eye(p)
    for t in range(p):
        for u in range(t):
            if rng.random() < trait_density:
                M[t] += rng.uniform(-0.3, 0.3) * M[u]
    covY = M @ covZ @ M.T
    dinv = 1.0 / np.sqrt(np.diag(covY))
    pxp = covY * dinv[:, None] * dinv[None, :]
    Ep = rng.standard_normal((p, p)) / np.sqrt(float(num_samples))
    Ep = np.triu(Ep, 1)
    pxp = np.clip(pxp + Ep + Ep.T, -0.999999, 0.999999)
    np.fill_diagonal(pxp, 1.0)
    names = [f"T{t + 1}" for t in range(p)]
    with open(os.path.join(outdir, "pxp.txt"), "w") as fh:
        fh.write(" ".join(names) + "\n")
        for a in range(p):
            fh.write(names[a] + " " + " ".join("%.9g" % pxp[a, b] for b in range(p)) + "\n")
    mix = (M * dinv[:, None]).T  # cov(marker, y) = cov(marker, z) M' D
    with open(os.path.join(outdir, "mxp.txt"), "w") as fh, open(os.path.join(outdir, "genome.blocks"), "w") as fb:
        fh.write("chr snp ref " + " ".join(names) + "\n")
        for b in range(nblocks):
            c, f, l = stems[b].split("_")
            fb.write(f"{c}\t{f}\t{l}\n")
            mt = cbs[b] @ mix + rng.standard_normal((int(sizes[b]), p)) / np.sqrt(float(num_samples))
            np.clip(mt, -0.999999, 0.999999, out=mt)
            lines = ["%s rs%d_%d A " % (c, b, i) + " ".join("%.7g" % v for v in mt[i]) for i in range(int(sizes[b]))]
            fh.write("\n".join(lines) + "\n")
    return dict(sizes=sizes, stems=stems, markers=total)
