"""ctypes loader for the CPU oracle (oracle/libcusk_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (cigwas_b200) never imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_double_p = C.POINTER(C.c_double)
c_ll_p = C.POINTER(C.c_longlong)
c_ubyte_p = C.POINTER(C.c_ubyte)


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        lib.orc_qnorm.restype = C.c_double
        lib.orc_qnorm.argtypes = [C.c_double]
        lib.orc_hetcor_threshold.restype = C.c_float
        lib.orc_hetcor_threshold.argtypes = [C.c_float]
        lib.orc_threshold_array.argtypes = [C.c_int, C.c_float, c_float_p]
        lib.orc_binom.restype = C.c_longlong
        lib.orc_canonical_tests.restype = C.c_longlong
        lib.orc_npn_from_table.restype = C.c_double
        lib.orc_npn_from_table.argtypes = [c_ll_p]

    def qnorm(self, p):
        return self.lib.orc_qnorm(p)

    def threshold_array(self, n, alpha):
        out = np.zeros(15, np.float32)
        self.lib.orc_threshold_array(int(n), C.c_float(alpha), _p(out, c_float_p))
        return out

    def hetcor_threshold(self, alpha):
        return float(self.lib.orc_hetcor_threshold(C.c_float(alpha)))

    def num_threads(self):
        return int(self.lib.orc_num_threads())

    def set_num_threads(self, t):
        self.lib.orc_set_num_threads(int(t))

    def skeleton(self, Cm, Th, maxlevel, near_band=0.0):
        """-> dict(G, l, pMax, SepSet, stats, near)"""
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        G = np.ones((n, n), np.int32)
        pMax = np.zeros((n, n), np.float32)
        Sep = np.zeros((n, n, 14), np.int32)
        Th = np.ascontiguousarray(Th, np.float32)
        l = C.c_int(0)
        stats = np.zeros(31, np.int64)
        cap = 100000
        near = np.zeros((cap, 5), np.float64)
        near_n = C.c_int(0)
        self.lib.orc_skeleton(_p(Cm, c_float_p), n, _p(G, c_int_p), _p(Th, c_float_p), C.byref(l),
                              int(maxlevel), _p(pMax, c_float_p), _p(Sep, c_int_p), _p(stats, c_ll_p),
                              C.c_double(near_band), _p(near, c_double_p), cap, C.byref(near_n))
        return dict(G=G, l=l.value, pMax=pMax, SepSet=Sep, stats=stats, near=near[: near_n.value])

    def hetcor_skeleton(self, Cm, G, N, th, maxlevel, time_index, near_band=0.0):
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        G = np.array(G, np.int32, copy=True).reshape(n, n)
        N = np.ascontiguousarray(N, np.float32)
        ti = np.ascontiguousarray(time_index, np.int32)
        l = C.c_int(0)
        stats = np.zeros(31, np.int64)
        cap = 100000
        near = np.zeros((cap, 5), np.float64)
        near_n = C.c_int(0)
        self.lib.orc_hetcor_skeleton(_p(Cm, c_float_p), n, _p(G, c_int_p), _p(N, c_float_p), C.c_float(th),
                                     C.byref(l), int(maxlevel), _p(ti, c_int_p), _p(stats, c_ll_p),
                                     C.c_double(near_band), _p(near, c_double_p), cap, C.byref(near_n))
        return dict(G=G, l=l.value, stats=stats, near=near[: near_n.value])

    def hetcor_level(self, Cm, G, N, th, l, time_index, row_rem=0, row_mod=1):
        """one level l >= 1 on the rows X % row_mod == row_rem (sharded emulation); G is updated in place"""
        n = Cm.shape[0]
        ran = self.lib.orc_hetcor_level(_p(Cm, c_float_p), n, _p(G, c_int_p), _p(N, c_float_p), C.c_float(th), int(l),
                                        _p(time_index, c_int_p), int(row_rem), int(row_mod))
        return bool(ran)

    def ref_quirks(self, G, time_index, l, cap=20_000_000):
        """what the reference's level-l kernels would do differently on graph G at the start of level l:
        dict(rows_int32_rank_overflow, rows_cut_short_by_noedgeflag, sets_never_evaluated, rows_not_analysed)"""
        G = np.ascontiguousarray(G, np.int32)
        n = G.shape[0]
        ti = np.ascontiguousarray(time_index, np.int32)
        out = np.zeros(4, np.int64)
        self.lib.orc_ref_quirks(_p(G, c_int_p), n, _p(ti, c_int_p), int(l), C.c_longlong(cap), _p(out, c_ll_p))
        return dict(rows_int32_rank_overflow=int(out[0]), rows_cut_short_by_noedgeflag=int(out[1]),
                    sets_never_evaluated=int(out[2]), rows_not_analysed=int(out[3]))

    def bed_col_stats(self, bed, m, n):
        bed = np.ascontiguousarray(bed, np.uint8)
        means = np.zeros(m, np.float32)
        stds = np.zeros(m, np.float32)
        modes = np.zeros(m, np.int32)
        counts = np.zeros((m, 4), np.int64)
        self.lib.orc_bed_col_stats(_p(bed, c_ubyte_p), C.c_size_t(m), C.c_size_t(n), _p(means, c_float_p),
                                   _p(stds, c_float_p), _p(modes, c_int_p), _p(counts, c_ll_p))
        return means, stds, modes, counts

    def pair_table(self, bed, n, a, b):
        bed = np.ascontiguousarray(bed, np.uint8)
        t = np.zeros(9, np.int64)
        self.lib.orc_bed_pair_table(_p(bed, c_ubyte_p), C.c_size_t(n), C.c_size_t(a), C.c_size_t(b),
                                    _p(t, c_ll_p))
        return t

    def npn_from_table(self, t):
        t = np.ascontiguousarray(t, np.int64)
        return float(self.lib.orc_npn_from_table(_p(t, c_ll_p)))

    def corr_npn(self, bed, phen, m, n, p, mean, std, want_mxm=True):
        bed = np.ascontiguousarray(bed, np.uint8)
        phen = np.ascontiguousarray(phen, np.float32)
        mean = np.ascontiguousarray(mean, np.float32)
        std = np.ascontiguousarray(std, np.float32)
        mxm = np.zeros(m * (m - 1) // 2, np.float64) if want_mxm else None
        mxp = np.zeros(m * p, np.float64)
        pxp = np.zeros(p * (p - 1) // 2, np.float64)
        self.lib.orc_corr_npn(_p(bed, c_ubyte_p), _p(phen, c_float_p), C.c_size_t(m), C.c_size_t(n),
                              C.c_size_t(p), _p(mean, c_float_p), _p(std, c_float_p), _p(mxm, c_double_p),
                              _p(mxp, c_double_p), _p(pxp, c_double_p))
        return mxm, mxp, pxp

    def mcorrk_banded(self, bed, m, n, corr_width):
        """cal_mcorrk_banded: [m, corr_width] float32, entry (row, k) = npn corr of (row, row + 1 + k)"""
        bed = np.ascontiguousarray(bed, np.uint8)
        out = np.zeros((m, corr_width), np.float32)
        self.lib.orc_mcorrk_banded(_p(bed, c_ubyte_p), C.c_size_t(m), C.c_size_t(n), C.c_size_t(corr_width),
                                   _p(out, c_float_p))
        return out

    def banded_row_abs_sums(self, corrs):
        corrs = np.ascontiguousarray(corrs, np.float32)
        m, w = corrs.shape
        out = np.zeros(m, np.float32)
        self.lib.orc_banded_row_abs_sums(_p(corrs, c_float_p), C.c_size_t(m), C.c_size_t(w), _p(out, c_float_p))
        return out

    def hanning_smoothing(self, v, window):
        v = np.ascontiguousarray(v, np.float32)
        out = np.zeros(len(v), np.float64)
        self.lib.orc_hanning_smoothing(_p(v, c_float_p), C.c_size_t(len(v)), int(window), _p(out, c_double_p))
        return out

    def block_chr(self, v, max_block_size):
        """-> (blocks [nb, 2] int64 (first, last), final window size)"""
        v = np.ascontiguousarray(v, np.float32)
        first = np.zeros(len(v), np.int64)
        last = np.zeros(len(v), np.int64)
        win = C.c_int(0)
        self.lib.orc_block_chr.restype = C.c_longlong
        nb = self.lib.orc_block_chr(_p(v, c_float_p), C.c_size_t(len(v)), int(max_block_size), _p(first, c_ll_p),
                                    _p(last, c_ll_p), C.byref(win))
        return np.stack([first[:nb], last[:nb]], 1), win.value

    def subset_variables(self, G, num_var, num_markers, max_depth):
        G = np.ascontiguousarray(G, np.int32)
        sel = np.zeros(num_var, np.int32)
        self.lib.orc_subset_variables(_p(G, c_int_p), int(num_var), int(num_markers), int(max_depth),
                                      _p(sel, c_int_p))
        return np.nonzero(sel)[0]

    def pinv_courrieu(self, A):
        A = np.ascontiguousarray(A, np.float64)
        out = np.zeros_like(A)
        self.lib.orc_pinv_courrieu(_p(A, c_double_p), _p(out, c_double_p), A.shape[0])
        return out

    def inv_svd(self, A):
        A = np.ascontiguousarray(A, np.float64)
        out = np.zeros_like(A)
        self.lib.orc_inv_svd(_p(A, c_double_p), _p(out, c_double_p), A.shape[0])
        return out

    def ith_combination(self, N, P, L):
        out = np.zeros(P, np.int32)
        self.lib.orc_ith_combination(_p(out, c_int_p), int(N), int(P), C.c_longlong(L))
        return out

    def binom(self, n, k):
        return int(self.lib.orc_binom(int(n), int(k)))

    def canonical_tests(self, deg, l):
        deg = np.ascontiguousarray(deg, np.int32)
        return int(self.lib.orc_canonical_tests(_p(deg, c_int_p), len(deg), int(l)))


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def load():
    global _LIB
    if _LIB is None:
        path = os.path.join(ROOT, "oracle", "libcusk_oracle.so")
        src = os.path.join(ROOT, "oracle", "cusk_oracle.cpp")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            build()
        _LIB = Oracle(C.CDLL(path))
    return _LIB
