"""ctypes binding of libcusk_b200.so (see include/cusk_b200.h for the contract of every symbol)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ML = 14

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_i64_p = C.POINTER(C.c_int64)
c_u32_p = C.POINTER(C.c_uint32)
c_ubyte_p = C.POINTER(C.c_ubyte)


class CuskError(RuntimeError):
    pass


class Stats(C.Structure):
    _fields_ = [
        ("tests_canonical", C.c_int64 * (ML + 1)),
        ("tests_executed", C.c_int64 * (ML + 1)),
        ("level_ms", C.c_float * (ML + 1)),
        ("ci_kernel_ms", C.c_float * (ML + 1)),
        ("max_degree", C.c_int32 * (ML + 1)),
        ("edges_after", C.c_int64 * (ML + 1)),
        ("singular_pivots", C.c_int64),
        ("last_level", C.c_int32),
        ("levels_run", C.c_int32),
        ("total_ms", C.c_float),
        ("h2d_ms", C.c_float),
        ("d2h_ms", C.c_float),
        ("h2d_bytes", C.c_int64),
        ("d2h_bytes", C.c_int64),
        ("kernel_launches", C.c_int64),
    ]

    def as_dict(self):
        return dict(
            tests_canonical=list(self.tests_canonical),
            tests_executed=list(self.tests_executed),
            level_ms=[float(x) for x in self.level_ms],
            ci_kernel_ms=[float(x) for x in self.ci_kernel_ms],
            max_degree=list(self.max_degree),
            edges_after=list(self.edges_after),
            singular_pivots=int(self.singular_pivots),
            last_level=int(self.last_level),
            levels_run=int(self.levels_run),
            total_ms=float(self.total_ms),
            h2d_ms=float(self.h2d_ms),
            d2h_ms=float(self.d2h_ms),
            h2d_bytes=int(self.h2d_bytes),
            d2h_bytes=int(self.d2h_bytes),
            kernel_launches=int(self.kernel_launches),
        )


class CorrStats(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("h2d_ms", "counts_ms", "decode_ms", "gram_ms", "mxp_ms", "pxp_ms",
                                         "d2h_ms", "total_ms")] + [
        ("parts", C.c_int32), ("gram_tiles", C.c_int32), ("gram_int8_ops", C.c_int64), ("bed_bytes", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class BandStats(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("h2d_ms", "kernel_ms", "d2h_ms", "total_ms")] + [
        ("chunks", C.c_int32), ("pairs", C.c_int64), ("bed_bytes", C.c_int64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


SEPSET_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("level", "<i4"), ("pmax", "<f4"), ("s", "<i4", (ML,))])

_LIB = None


def library_path():
    return os.path.join(_HERE, "libcusk_b200.so")


def build_library(verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a build of csrc/ (cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(_HERE, "csrc"), "-j8"]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)
    return library_path()


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def load_library():
    """Loads the CUDA library; raises CuskError when it has not been built (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise CuskError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU fallback)")
    lib = C.CDLL(path)
    lib.cusk_last_error.restype = C.c_char_p
    lib.cusk_version.restype = C.c_char_p
    lib.cusk_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.cusk_ctx_destroy.argtypes = [C.c_void_p]
    lib.cusk_ctx_set_level_timing.argtypes = [C.c_void_p, C.c_int]
    lib.cusk_ctx_set_level_timing.restype = None
    lib.cusk_ctx_stream.restype = C.c_void_p
    lib.cusk_ctx_stream.argtypes = [C.c_void_p]
    lib.cusk_ctx_sync.argtypes = [C.c_void_p]
    lib.cusk_threshold_array.argtypes = [C.c_int, C.c_float, c_float_p]
    lib.cusk_hetcor_threshold.restype = C.c_float
    lib.cusk_hetcor_threshold.argtypes = [C.c_float]
    lib.cusk_row_words.restype = C.c_size_t
    lib.cusk_row_words.argtypes = [C.c_int]
    lib.cusk_skeleton.argtypes = [C.c_void_p, c_float_p, C.c_int, c_int_p, c_float_p, C.c_int, c_int_p,
                                  c_float_p, c_int_p, C.c_void_p, C.c_int64, c_i64_p, C.POINTER(Stats)]
    lib.cusk_hetcor_skeleton.argtypes = [C.c_void_p, c_float_p, C.c_int, c_int_p, C.c_int, c_float_p,
                                         C.c_float, C.c_int, c_int_p, c_int_p, C.POINTER(Stats)]
    lib.cusk_hetcor_skeleton_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                             C.c_float, C.c_float, C.c_int, C.c_void_p, c_int_p,
                                             C.POINTER(Stats)]
    lib.cusk_skeleton_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, c_float_p, C.c_int, c_int_p,
                                      C.c_void_p, C.c_void_p, C.c_int64, c_i64_p, C.POINTER(Stats)]
    lib.cusk_bed_marker_stats.argtypes = [C.c_void_p, c_ubyte_p, C.c_size_t, C.c_size_t, c_float_p, c_float_p,
                                          c_int_p, c_i64_p]
    lib.cusk_bed_marker_stats_noalloc.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                                  C.c_void_p, C.c_void_p]
    lib.cusk_bed_marker_stats_noalloc.restype = C.c_int
    lib.cusk_corr_block.argtypes = [C.c_void_p, c_ubyte_p, c_float_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                    c_float_p, c_float_p, C.c_int, c_float_p, c_float_p, c_float_p, c_i64_p,
                                    C.POINTER(CorrStats)]
    lib.cusk_marker_phen_corr.argtypes = [C.c_void_p, c_ubyte_p, c_float_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                          c_float_p, c_float_p, c_float_p, C.POINTER(CorrStats)]
    lib.cusk_corr_block_dev.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t,
                                        C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(CorrStats)]
    lib.cusk_mcorrk_banded.argtypes = [C.c_void_p, c_ubyte_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_float, c_float_p,
                                       c_float_p, C.POINTER(BandStats)]
    lib.cusk_hanning_smoothing.argtypes = [c_float_p, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
    lib.cusk_block_chr.restype = C.c_int64
    lib.cusk_block_chr.argtypes = [c_float_p, C.c_size_t, C.c_int, c_i64_p, c_i64_p, c_int_p]
    lib.cusk_make_blocks.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_float, C.c_size_t]
    lib.cusk_comm_unique_id.argtypes = [C.c_char_p]
    lib.cusk_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]
    lib.cusk_comm_destroy.argtypes = [C.c_void_p]
    lib.cusk_ctx_set_comm.argtypes = [C.c_void_p, C.c_void_p]
    lib.cusk_comm_stats.argtypes = [C.c_void_p, C.POINTER(C.c_double), c_i64_p]
    _LIB = lib
    return lib


def mps_path():
    """the `mps` executable built next to the library (argv contract of the reference's cusk/apps/mps.cpp)"""
    return os.path.join(_HERE, "mps")


def local_group(ctxs):
    """One process, one thread per context: in-process exchange by peer copies (the contexts may share a
    device).  Every context of the group must then make the same skeleton call from its own thread."""
    lib = ctxs[0].lib
    world = len(ctxs)
    hs = (C.c_void_p * world)(*[c.h for c in ctxs])
    out = (C.c_void_p * world)()
    lib.cusk_comm_create_local.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(C.c_void_p)]
    rc = lib.cusk_comm_create_local(hs, world, out)
    if rc != 0:
        raise CuskError(f"rc={rc}: {lib.cusk_last_error().decode()}")
    for c, h in zip(ctxs, out):
        c.adopt_comm(C.c_void_p(h))


def comm_unique_id():
    """128-byte NCCL unique id; create it on one rank and broadcast it to the others."""
    buf = C.create_string_buffer(128)
    if load_library().cusk_comm_unique_id(buf) != 0:
        raise CuskError(load_library().cusk_last_error().decode())
    return buf.raw


def threshold_array(n, alpha):
    """cuPC_call_prep.cpp:13-23"""
    out = np.zeros(15, np.float32)
    load_library().cusk_threshold_array(int(n), C.c_float(alpha), _p(out, c_float_p))
    return out


def hetcor_threshold(alpha):
    """cuPC_call_prep.cpp:25-28"""
    return float(load_library().cusk_hetcor_threshold(C.c_float(alpha)))


def merge_block_outputs(blockfile, outdir, out_base=None):
    """merge_block_outputs(blockfile, outdir).write_mm(out_base) (cusk_postprocessing/merge_blocks.py:361-403, :300-325;
    ci-gwas.py:459-464): <out_base>_sam.mtx, _scm.mtx, .mdim, .ixs; host only"""
    lib = load_library()
    lib.cusk_merge_block_outputs.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
    if out_base is None:
        out_base = os.path.join(outdir, "merged_blocks")
    if lib.cusk_merge_block_outputs(os.fsencode(blockfile), os.fsencode(outdir), os.fsencode(out_base)) != 0:
        raise CuskError(lib.cusk_last_error().decode())


def hanning_smoothing(v, window):
    """hanning_smoothing (src/blocking.cpp:13-36): float64 [len(v)]; host only"""
    lib = load_library()
    v = np.ascontiguousarray(v, np.float32)
    out = np.zeros(len(v), np.float64)
    if lib.cusk_hanning_smoothing(_p(v, c_float_p), len(v), int(window), out.ctypes.data_as(C.POINTER(C.c_double))) != 0:
        raise CuskError(lib.cusk_last_error().decode())
    return out


def block_chr(row_sums, max_block_size):
    """block_chr (src/blocking.cpp:97-136) -> (blocks int64 [nb, 2] = (first, last) inclusive, Hann window used); host only"""
    lib = load_library()
    v = np.ascontiguousarray(row_sums, np.float32)
    first = np.zeros(len(v), np.int64)
    last = np.zeros(len(v), np.int64)
    win = C.c_int(0)
    nb = lib.cusk_block_chr(_p(v, c_float_p), len(v), int(max_block_size), _p(first, c_i64_p), _p(last, c_i64_p),
                            C.byref(win))
    if nb < 0:
        raise CuskError(lib.cusk_last_error().decode())
    return np.stack([first[:nb], last[:nb]], 1), win.value


def reformat_cuskss_merged(cusk_dir):
    """reformat_cuskss_merged_output(cusk_dir).write_mm(cusk_dir/cuskss_merged) (merge_blocks.py:406-425;
    ci-gwas.py:452-456); host only"""
    lib = load_library()
    lib.cusk_reformat_cuskss_merged.argtypes = [C.c_char_p]
    if lib.cusk_reformat_cuskss_merged(os.fsencode(cusk_dir)) != 0:
        raise CuskError(lib.cusk_last_error().decode())


class Context:
    """One device + stream + reusable workspace (cusk_ctx)."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.cusk_ctx_create(int(device), C.byref(h))
        if rc != 0:
            raise CuskError(self.lib.cusk_last_error().decode())
        self.h = h
        self.device = device

    def set_level_timing(self, on):
        """per-level event records (stats level_ms / ci_kernel_ms) on or off; include/cusk_b200.h"""
        self.lib.cusk_ctx_set_level_timing(self.h, 1 if on else 0)

    def close(self):
        self.detach_comm()
        if getattr(self, "h", None):
            self.lib.cusk_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise CuskError(f"rc={rc}: {self.lib.cusk_last_error().decode()}")

    # -- separating-set selection on a merged skeleton (include/cusk_b200.h) ------------------
    def sepselect(self, stem, alpha, num_samples, out_stem, prior_path=None):
        """orient_v_structures_merged(stem, alpha, num_samples, prior).to_file(out_stem)
        (cusk_postprocessing/sepselect.py:576-584, :545-573); the search runs on this context's device"""
        self.lib.cusk_sepselect.argtypes = [C.c_void_p, C.c_char_p, C.c_double, C.c_int64, C.c_char_p, C.c_char_p]
        self._check(self.lib.cusk_sepselect(self.h, os.fsencode(stem), float(alpha), int(num_samples),
                                            None if prior_path is None else os.fsencode(prior_path), os.fsencode(out_stem)))

    def sepselect_pairs(self, Ct, P, pairs, tmask, thr):
        """The greedy search alone (sepselect.py:262-304) for a batch of outer pairs: Ct = rows 0..P of the correlation
        matrix (float64 [(P+1), n_var]); returns (len[npairs], found_minimum[npairs], set[npairs, 32])"""
        Ct = np.ascontiguousarray(Ct, np.float64)
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        tmask = np.ascontiguousarray(tmask, np.uint32)
        thr = np.ascontiguousarray(thr, np.float64)
        out = np.zeros((pairs.shape[0], 34), np.int32)  # cusk_sepselect_result: len, found_minimum, set[32]
        self.lib.cusk_sepselect_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_int64,
                                                  C.c_void_p, C.c_void_p, C.c_void_p]
        self._check(self.lib.cusk_sepselect_pairs(self.h, Ct.ctypes.data, Ct.shape[1], int(P), Ct.shape[1], pairs.ctypes.data,
                                                  pairs.shape[0], tmask.ctypes.data, thr.ctypes.data, out.ctypes.data))
        return out[:, 0].copy(), out[:, 1].copy(), out[:, 2:].copy()

    # -- edge-sharded mode (one skeleton over several GPUs; include/cusk_b200.h "multi-GPU") --------
    def attach_comm(self, rank, world, unique_id):
        """Joins the NCCL communicator of `world` ranks; later skeleton calls on this context take every
        world-th work item of a level and exchange removals with the other ranks after each level."""
        h = C.c_void_p()
        self._check(self.lib.cusk_comm_create(self.h, int(rank), int(world), unique_id, C.byref(h)))
        self.comm = h
        self._check(self.lib.cusk_ctx_set_comm(self.h, h))

    def adopt_comm(self, handle):
        """Attaches a communicator made elsewhere (local_group below)."""
        self.comm = handle
        self._check(self.lib.cusk_ctx_set_comm(self.h, handle))

    def detach_comm(self):
        if getattr(self, "comm", None):
            self.lib.cusk_ctx_set_comm(self.h, None)
            self.lib.cusk_comm_destroy(self.comm)
            self.comm = None

    def comm_stats(self):
        ms, by = C.c_double(0), C.c_int64(0)
        self._check(self.lib.cusk_comm_stats(self.comm, C.byref(ms), C.byref(by)))
        return dict(exchange_ms=ms.value, exchange_bytes=by.value)

    @property
    def stream(self):
        return self.lib.cusk_ctx_stream(self.h)

    def sync(self):
        self._check(self.lib.cusk_ctx_sync(self.h))

    # -- Skeleton(), src/cuPC-S.cu:61 ------------------------------------------------------------
    def skeleton(self, Cm, Th, maxlevel, want_pmax=True, want_dense_sepset=True):
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        G = np.zeros((n, n), np.int32)
        Th = np.ascontiguousarray(Th, np.float32)
        pMax = np.zeros((n, n), np.float32) if want_pmax else None
        Sep = np.zeros((n, n, ML), np.int32) if want_dense_sepset else None
        l = C.c_int(0)
        st = Stats()
        cap = max(1024, n * 64)
        while True:
            recs = np.zeros(cap, SEPSET_DTYPE)
            nsep = C.c_int64(0)
            self._check(self.lib.cusk_skeleton(self.h, _p(Cm, c_float_p), n, _p(G, c_int_p), _p(Th, c_float_p),
                                               int(maxlevel), C.byref(l), _p(pMax, c_float_p), _p(Sep, c_int_p),
                                               recs.ctypes.data_as(C.c_void_p), cap, C.byref(nsep),
                                               C.byref(st)))
            if nsep.value <= cap:
                break
            cap = int(nsep.value)
        return dict(G=G, l=l.value, pMax=pMax, SepSet=Sep, seps=recs[: nsep.value], stats=st.as_dict())

    # -- hetcor_skeleton(), src/hetcor-cuPC-S.cu:75 ---------------------------------------------
    def hetcor_skeleton(self, Cm, G, N, th, maxlevel, time_index=None):
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        all_ones = G is None
        Gw = np.ones((n, n), np.int32) if all_ones else np.array(G, np.int32, copy=True).reshape(n, n)
        N = np.ascontiguousarray(N, np.float32)
        ti = None if time_index is None else np.ascontiguousarray(time_index, np.int32)
        l = C.c_int(0)
        st = Stats()
        self._check(self.lib.cusk_hetcor_skeleton(self.h, _p(Cm, c_float_p), n, _p(Gw, c_int_p), int(all_ones),
                                                  _p(N, c_float_p), C.c_float(th), int(maxlevel),
                                                  _p(ti, c_int_p), C.byref(l), C.byref(st)))
        return dict(G=Gw, l=l.value, stats=st.as_dict())

    # -- correlation front end ---------------------------------------------------------------------
    def bed_marker_stats(self, bed, m, n):
        """prep-bed statistics (prep.cpp:15-77) -> means, stds, modes, counts[m,4]"""
        bed = np.ascontiguousarray(bed, np.uint8).reshape(-1)
        means = np.zeros(m, np.float32)
        stds = np.zeros(m, np.float32)
        modes = np.zeros(m, np.int32)
        counts = np.zeros((m, 4), np.int64)
        self._check(self.lib.cusk_bed_marker_stats(self.h, _p(bed, c_ubyte_p), m, n, _p(means, c_float_p),
                                                   _p(stds, c_float_p), _p(modes, c_int_p), _p(counts, c_i64_p)))
        return means, stds, modes, counts

    def mcorrk_banded(self, bed, m, n, corr_width, device_mem_gb=0.0, want_corrs=True, want_sums=True):
        """cal_mcorrk_banded + marker_corr_banded_mat_row_abs_sums (corr_host.cu:64-128)
        -> dict(corrs [m, corr_width] | None, row_abs_sums [m] | None, stats)"""
        bed = np.ascontiguousarray(bed, np.uint8).reshape(-1)
        corrs = np.zeros((m, corr_width), np.float32) if want_corrs else None
        sums = np.zeros(m, np.float32) if want_sums else None
        st = BandStats()
        self._check(self.lib.cusk_mcorrk_banded(self.h, _p(bed, c_ubyte_p), m, n, corr_width, float(device_mem_gb),
                                                _p(corrs, c_float_p), _p(sums, c_float_p), C.byref(st)))
        return {"corrs": corrs, "row_abs_sums": sums, "stats": st.as_dict()}

    def make_blocks(self, bfiles, max_block_size, corr_width, device_mem_gb=0.0):
        """make_blocks (cli.cpp:362-411): appends to <bfiles>_m<max_block_size>.blocks"""
        self._check(self.lib.cusk_make_blocks(self.h, os.fsencode(bfiles), int(max_block_size), float(device_mem_gb),
                                              int(corr_width)))

    def corr_block(self, bed, phen, m, n, p, mean, std, mode=0, want_mxm=True, want_tables=False):
        """cu_corr_pearson_npn (corr_host.cu:1094) -> dict(mxm, mxp, pxp[, tables], stats)"""
        bed = np.ascontiguousarray(bed, np.uint8).reshape(-1)
        phen = np.ascontiguousarray(phen, np.float32).reshape(-1) if p else None
        mean = np.ascontiguousarray(mean, np.float32)
        std = np.ascontiguousarray(std, np.float32)
        npairs = m * (m - 1) // 2
        mxm = np.zeros(npairs, np.float32) if want_mxm else None
        mxp = np.zeros(m * p, np.float32) if p else None
        pxp = np.zeros(p * (p - 1) // 2, np.float32) if p > 1 else None
        tables = np.zeros((npairs, 9), np.int64) if want_tables else None
        st = CorrStats()
        self._check(self.lib.cusk_corr_block(self.h, _p(bed, c_ubyte_p), _p(phen, c_float_p), m, n, p,
                                             _p(mean, c_float_p), _p(std, c_float_p), int(mode), _p(mxm, c_float_p),
                                             _p(mxp, c_float_p), _p(pxp, c_float_p), _p(tables, c_i64_p),
                                             C.byref(st)))
        return dict(mxm=mxm, mxp=mxp, pxp=pxp, tables=tables, stats=st.as_dict())

    def corr_block_dev(self, d_bed, d_phen, m, n, p, d_mean, d_std, d_Csq, mode=0):
        st = CorrStats()
        self._check(self.lib.cusk_corr_block_dev(self.h, d_bed, d_phen, m, n, p, d_mean, d_std, int(mode), d_Csq,
                                                 C.byref(st)))
        return st.as_dict()

    # -- device-resident variants (pointers are ints, e.g. torch.Tensor.data_ptr()) -----------------
    def row_words(self, n):
        return int(self.lib.cusk_row_words(int(n)))

    def hetcor_skeleton_dev(self, dC, n, dbits, all_ones, dN, n_uniform, th, maxlevel, d_time=None):
        l = C.c_int(0)
        st = Stats()
        self._check(self.lib.cusk_hetcor_skeleton_dev(self.h, dC, int(n), dbits, int(all_ones), dN,
                                                      C.c_float(n_uniform), C.c_float(th), int(maxlevel),
                                                      d_time, C.byref(l), C.byref(st)))
        return l.value, st.as_dict()

    def skeleton_dev(self, dC, n, dbits, Th, maxlevel, d_pmax=None, want_seps=False):
        Th = np.ascontiguousarray(Th, np.float32)
        l = C.c_int(0)
        st = Stats()
        nsep = C.c_int64(0)
        self._check(self.lib.cusk_skeleton_dev(self.h, dC, int(n), dbits, _p(Th, c_float_p), int(maxlevel),
                                               C.byref(l), d_pmax, None, 0,
                                               C.byref(nsep) if want_seps else None, C.byref(st)))
        return l.value, st.as_dict(), nsep.value
