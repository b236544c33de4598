"""ctypes loader of oracle/_ref/libcusk_ref.so: the UNMODIFIED reference CUDA skeleton sources built
for sm_100 by oracle/build_ref.sh.  Test / baseline infrastructure only (GPU box)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "oracle", "_ref", "libcusk_ref.so")
fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int)


def available():
    return os.path.exists(PATH)


class RefLib:
    def __init__(self):
        self.lib = C.CDLL(PATH)

    def hetcor_skeleton(self, Cm, G, N, th, maxlevel, time_index):
        """src/hetcor-cuPC-S.cu:75 (float32, fast-math, schedule dependent bookkeeping)."""
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        G = np.array(G, np.int32, copy=True).reshape(n, n)
        N = np.ascontiguousarray(N, np.float32)
        ti = np.ascontiguousarray(time_index, np.int32)
        P, l, ml, t = C.c_int(n), C.c_int(0), C.c_int(maxlevel), C.c_float(th)
        self.lib.hetcor_skeleton(Cm.ctypes.data_as(fp), C.byref(P), G.ctypes.data_as(ip), N.ctypes.data_as(fp),
                                 C.byref(t), C.byref(l), C.byref(ml), ti.ctypes.data_as(ip))
        return dict(G=G, l=l.value)

    def skeleton(self, Cm, Th, maxlevel, want_sepset=True):
        """src/cuPC-S.cu:61"""
        Cm = np.ascontiguousarray(Cm, np.float32)
        n = Cm.shape[0]
        G = np.zeros((n, n), np.int32)
        Th = np.ascontiguousarray(Th, np.float32)
        pMax = np.zeros((n, n), np.float32)
        Sep = np.zeros((n, n, 14), np.int32)
        P, l, ml = C.c_int(n), C.c_int(0), C.c_int(maxlevel)
        self.lib.Skeleton(Cm.ctypes.data_as(fp), C.byref(P), G.ctypes.data_as(ip), Th.ctypes.data_as(fp), C.byref(l),
                          C.byref(ml), pMax.ctypes.data_as(fp), Sep.ctypes.data_as(ip))
        return dict(G=G, l=l.value, pMax=pMax, SepSet=Sep)


CORR_PATH = os.path.join(ROOT, "oracle", "_ref", "libcusk_ref_corr.so")


def corr_available():
    return os.path.exists(CORR_PATH)


class RefCorr:
    """the reference's correlation front end (src/corr_host.cu:1094 cu_corr_pearson_npn and :1199
    cu_marker_corr_pearson_npn_batched_sparse; C++ linkage, hence the mangled names), unmodified sources, sm_100"""

    def __init__(self):
        self.lib = C.CDLL(CORR_PATH)
        ub = C.POINTER(C.c_ubyte)
        self.npn = getattr(self.lib, "_Z19cu_corr_pearson_npnPKhPKfmmmS2_S2_PfS3_S3_")
        self.npn.argtypes = [ub, fp, C.c_size_t, C.c_size_t, C.c_size_t, fp, fp, fp, fp, fp]
        self.npn.restype = None
        self.banded = getattr(self.lib, "_Z41cu_marker_corr_pearson_npn_batched_sparsePKhmmmmPf")
        self.banded.argtypes = [ub, C.c_size_t, C.c_size_t, C.c_size_t, C.c_size_t, fp]
        self.banded.restype = None

    def corr_pearson_npn(self, bed, phen, m, n, p, mean, std):
        bed = np.ascontiguousarray(bed, np.uint8).reshape(-1)
        phen = np.ascontiguousarray(phen, np.float32).reshape(-1)
        mean = np.ascontiguousarray(mean, np.float32)
        std = np.ascontiguousarray(std, np.float32)
        mxm = np.zeros(m * (m - 1) // 2, np.float32)
        mxp = np.zeros(m * p, np.float32)
        pxp = np.zeros(max(1, p * (p - 1) // 2), np.float32)
        ub = C.POINTER(C.c_ubyte)
        self.npn(bed.ctypes.data_as(ub), phen.ctypes.data_as(fp), m, n, p, mean.ctypes.data_as(fp), std.ctypes.data_as(fp),
                 mxm.ctypes.data_as(fp), mxp.ctypes.data_as(fp), pxp.ctypes.data_as(fp))
        return mxm, mxp, pxp
