"""Test infrastructure: numpy restatement of the reference's HOST glue around the skeleton engines
(loaders, square-matrix assembly, variable selection, reduction, output files), driven by the CPU
oracle for the arithmetic.  Used to check what `mps cusk` / `mps cuskss` write, file by file.

Citations are into /root/reference/cusk/.  Never imported by the product.
"""
import os

import numpy as np

ML = 14


# ---- loaders -----------------------------------------------------------------------------------------
def read_blocks(path):
    """src/io.cpp:74-101 -> list of (chr, first, last, chr_global_offset)"""
    out, off, on_chr, cur = [], 0, 0, ""
    for line in open(path):
        f = line.split()
        if not f:
            continue
        if f[0] != cur:
            cur = f[0]
            off += on_chr
            on_chr = 0
        out.append((f[0], int(f[1]), int(f[2]), off))
        on_chr += int(f[2]) - int(f[1]) + 1
    return out


def load_mxm(path):
    """src/marker_summary_stats.cpp:8-24"""
    tril = np.fromfile(path, np.float32)
    m = int((np.sqrt(8 * tril.size + 1) - 1) / 2)
    sq = np.ones((m, m), np.float32)
    k = 0
    for i in range(m):
        for j in range(i + 1):
            v = np.float32(0.0) if np.isnan(tril[k]) else tril[k]
            sq[i, j] = sq[j, i] = v
            k += 1
    return sq


def _ess(rho, se):
    """src/marker_trait_summary_stats.cpp:162-164: float ss_sqrt = (1.0 - rho*rho) / se; N = ss_sqrt^2"""
    rho = np.float32(rho)
    ss = np.float32((1.0 - float(np.float32(rho * rho))) / float(np.float32(se)))
    return np.float32(ss * ss)


def load_mxp(path, se_path, rows):
    """src/marker_trait_summary_stats.cpp:40-299; rows = list of 0-based data-line numbers"""
    lines = open(path).read().split("\n")
    hdr = lines[0].split()
    assert hdr[:3] == ["chr", "snp", "ref"]
    p = len(hdr) - 3
    se_lines = open(se_path).read().split("\n") if se_path else None
    corr, ess = [], []
    for r in rows:
        f = lines[1 + r].split()
        fs = se_lines[1 + r].split() if se_lines else None
        for j in range(3, 3 + p):
            na = f[j] in ("NA", "NaN", "nan") or (se_lines is not None and f[j] == "NAN")
            if na:
                corr.append(np.float32(0))
                ess.append(np.float32(np.nan))
            else:
                corr.append(np.float32(f[j]))
                if se_lines:
                    ess.append(_ess(f[j], fs[j]))
    m = len(rows)
    return (np.array(corr, np.float32).reshape(m, p),
            np.array(ess, np.float32).reshape(m, p) if se_lines else None)


def load_pxp(path, se_path, sample_size):
    """src/trait_summary_stats.cpp:5-169"""
    lines = [ln for ln in open(path).read().split("\n")]
    p = len(lines[0].split())
    corr = np.ones((p, p), np.float32)
    ess = np.full((p, p), 0.0 if se_path else sample_size, np.float32)
    se_lines = open(se_path).read().split("\n") if se_path else None
    r = 0
    for li in range(1, len(lines)):
        f = lines[li].split()
        if not f:
            break
        fs = se_lines[li].split() if se_lines else None
        for j in range(1, p + 1):
            c = np.float32(f[j])
            if se_lines:
                if np.isnan(c):
                    c = np.float32(0)
                    ess[r, j - 1] = np.nan
                else:
                    ess[r, j - 1] = _ess(c, fs[j])
            elif np.isnan(c):
                c = np.float32(0)
            corr[r, j - 1] = c
        r += 1
    iu = np.triu_indices(p, 1)
    corr[(iu[1], iu[0])] = corr[iu]
    if se_lines:
        ess[(iu[1], iu[0])] = ess[iu]
    return corr, ess


def square_inputs(mxm, mxp, mxp_ess, pxp, pxp_ess, sample_size, het):
    """src/cli.cpp:89-173"""
    m, p = mxm.shape[0], pxp.shape[0]
    nv = m + p
    C = np.ones((nv, nv), np.float32)
    N = np.full((nv, nv), sample_size, np.float32)
    C[:m, :m] = mxm
    C[:m, m:] = mxp
    C[m:, :m] = mxp.T
    C[m:, m:] = pxp
    if het:
        N[:m, m:] = mxp_ess
        N[m:, :m] = mxp_ess.T
        N[m:, m:] = pxp_ess
    return C, N


# ---- variable selection / reduction ----------------------------------------------------------------
def subset_variables(G, num_markers, max_depth):
    """src/parent_set.cpp:8-53 (one BFS per trait over marker nodes)"""
    nv = G.shape[0]
    keep = set(range(num_markers, nv))
    for start in range(num_markers, nv):
        cur = set(range(num_markers, nv))
        q = [start]
        for _ in range(max_depth):
            nq = []
            for node in q:
                for c in range(num_markers):
                    if G[node, c] == 1 and c not in cur:
                        cur.add(c)
                        nq.append(c)
            q = nq
        keep |= cur
    return sorted(keep)


def reduce_gc(G, C, N, sel, new_to_old):
    """src/parent_set.cpp:177-238"""
    ix = np.ix_(sel, sel)
    return G[ix].copy(), C[ix].copy(), N[ix].copy(), [new_to_old[i] for i in sel]


def reduce_gcs(G, C, S, sel, index_map):
    """src/parent_set.cpp:130-175 with VariableSubsetIndices(P, index_map) (parent_set.h:156-166): the
    old->new map is keyed by index_map[sel[i]] but queried with the un-mapped member; absent key -> 0."""
    ix = np.ix_(sel, sel)
    new_to_old = [index_map[i] for i in sel]
    old_to_new = {o: i for i, o in enumerate(new_to_old)}
    P = set(sel)
    ns = len(sel)
    Sout = np.full((ns, ns, ML), -1, np.int32)
    for a, i in enumerate(sel):
        for b, j in enumerate(sel):
            k = 0
            for l in range(ML):
                e = int(S[i, j, l])
                if e != -1 and e in P:
                    Sout[a, b, k] = old_to_new.get(e, 0)
                    k += 1
    return G[ix].copy(), C[ix].copy(), Sout, new_to_old


# ---- pipelines ------------------------------------------------------------------------------------------
def cuskss_expected(orc, C, N, num_phen, th, depth, L1, L2, time_index_traits):
    """src/cli.cpp:194-346 (run_cusk twice when L2 > 0) -> dict(ixs, adj, corr, mdim)"""
    nv = C.shape[0]
    G = np.ones((nv, nv), np.int32)
    nto = list(range(nv))
    for stage, L in enumerate([L1] + ([L2] if L2 > 0 else [])):
        nv = C.shape[0]
        m = nv - num_phen
        ti = np.zeros(nv, np.int32)
        ti[m:] = time_index_traits
        r = orc.hetcor_skeleton(C, G, N, th, L, ti)
        sel = subset_variables(r["G"], m, depth)
        G, C, N, nto = reduce_gc(r["G"], C, N, sel, nto)
    return dict(ixs=np.array(nto, np.int32), adj=G.astype(np.int32), corr=C.astype(np.float32),
                mdim=(C.shape[0], num_phen, ML))


def cusk_expected(orc, bed, phen, m, n, p, means, stds, Th, L1, L2, depth):
    """src/cli.cpp:432-678 -> None when the block is skipped, else dict(ixs, adj, corr, sep, mdim)"""
    mxm_u, mxp_f, pxp_u = orc.corr_npn(bed, phen, m, n, p, means, stds)
    cr = dict(mxm=mxm_u, mxp=mxp_f, pxp=pxp_u)
    mxp = cr["mxp"].reshape(m, p)
    z = np.abs(0.5 * (np.log(np.abs(1 + mxp.astype(np.float64))) - np.log(np.abs(1 - mxp.astype(np.float64)))))
    if not (z >= Th[0]).any():
        return None
    nv = m + p
    C = np.ones((nv, nv), np.float32)
    iu = np.triu_indices(m, 1)
    C[:m, :m][iu] = cr["mxm"]
    C[:m, :m][(iu[1], iu[0])] = cr["mxm"]
    C[:m, m:] = mxp
    C[m:, :m] = mxp.T
    iu = np.triu_indices(p, 1)
    C[m:, m:][iu] = cr["pxp"]
    C[m:, m:][(iu[1], iu[0])] = cr["pxp"]
    r1 = orc.skeleton(C, Th, L1)
    sel1 = subset_variables(r1["G"], m, depth)
    C1 = C[np.ix_(sel1, sel1)].copy()
    r2 = orc.skeleton(C1, Th, L2)
    sel2 = subset_variables(r2["G"], len(sel1) - p, depth)
    G, Cc, S, nto = reduce_gcs(r2["G"], C1, r2["SepSet"], sel2, sel1)
    return dict(ixs=np.array(nto, np.int32), adj=G.astype(np.int32), corr=Cc.astype(np.float32), sep=S,
                mdim=(len(sel2), p, ML), C_full=C)


def read_outputs(stem):
    """src/parent_set.h:42-52,99-108"""
    nv, p, ml = [int(x) for x in open(stem + ".mdim").read().split()]
    out = dict(mdim=(nv, p, ml), ixs=np.fromfile(stem + ".ixs", np.int32),
               adj=np.fromfile(stem + ".adj", np.int32).reshape(nv, nv),
               corr=np.fromfile(stem + ".corr", np.float32).reshape(nv, nv))
    if os.path.exists(stem + ".sep"):
        out["sep"] = np.fromfile(stem + ".sep", np.int32).reshape(nv, nv, ml)
    return out


# ---- writers of synthetic input files --------------------------------------------------------------------
def write_cuskss_inputs(d, Cm, m, p, het_N=None, blocks=None, na_frac=0.0, rng=None, extra_rows_before=0):
    """Writes mxm (lower triangle incl. diagonal, float32), mxp / pxp text (and SE files when het_N is
    given: se = (1 - r^2)/sqrt(N)) for ONE block of m markers + p traits taken from the dense matrix Cm.
    extra_rows_before: dummy mxp rows in front (the block's global offset)."""
    os.makedirs(d, exist_ok=True)
    il = np.tril_indices(m)
    Cm[:m, :m][il].astype(np.float32).tofile(os.path.join(d, "mxm.bin"))
    names = [f"T{t}" for t in range(p)]

    def se_of(r, N):
        return float((1.0 - float(r) ** 2) / np.sqrt(float(N)))

    with open(os.path.join(d, "mxp.txt"), "w") as f, open(os.path.join(d, "mxp_se.txt"), "w") as fs:
        f.write("chr snp ref " + " ".join(names) + "\n")
        fs.write("chr snp ref " + " ".join(names) + "\n")
        for r in range(extra_rows_before):
            f.write(f"0 pad{r} A " + " ".join("0.5" for _ in names) + "\n")
            fs.write(f"0 pad{r} A " + " ".join("0.001" for _ in names) + "\n")
        for i in range(m):
            vals, ses = [], []
            for t in range(p):
                if na_frac and rng is not None and rng.random() < na_frac:
                    vals.append("NA")
                    ses.append("NA")
                else:
                    vals.append(repr(float(Cm[i, m + t])))
                    ses.append(repr(se_of(Cm[i, m + t], het_N[i, m + t])) if het_N is not None else "0")
            f.write(f"1 rs{i} A " + " ".join(vals) + "\n")
            fs.write(f"1 rs{i} A " + " ".join(ses) + "\n")
    with open(os.path.join(d, "pxp.txt"), "w") as f, open(os.path.join(d, "pxp_se.txt"), "w") as fs:
        f.write(" ".join(names) + "\n")
        fs.write(" ".join(names) + "\n")
        for a in range(p):
            f.write(names[a] + " " + " ".join(repr(float(Cm[m + a, m + b])) for b in range(p)) + "\n")
            fs.write(names[a] + " " + " ".join(
                repr(se_of(Cm[m + a, m + b], het_N[m + a, m + b]) if (het_N is not None and a != b) else 1e-3)
                for b in range(p)) + "\n")
