// cusk_oracle.cpp -- CPU FP64 restatement of the CI-GWAS skeleton hot path.
//
// THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this library.  The product
// path (cigwas_b200/csrc -> libcusk_b200.so) never links, imports or calls it.
//
// Parity status: PINNED against the reference's own golden vectors (tests/golden/*, made
// by tests/golden/make_golden.py from /root/reference/cusk/include/test_data and
// cusk/tests/*.cpp): C_N10 -> A_N10 for both engines, threshold 0.0081045, the 10x3x2 and
// 100x7x5 BED correlation sets, small.bed means/stds, parent-set BFS sets; and
// (tests/golden/make_block_golden.py) the banded 100x7 correlations, the Hann smoothing vector and the 17
// blocks of block_chr(TEST_V, 500).
//
// All file:line citations are relative to /root/reference/cusk/.  The reference computes in
// float32 with --use_fast_math on a GPU (src/CMakeLists.txt:27-35); this restatement follows
// the same formulas in IEEE double and with a deterministic (schedule-independent) choice
// where the reference has a benign race (first atomicCAS winner, src/cuPC-S.cu:568-577):
// the separating set recorded for a directed pair (X,Y) is the one with the LOWEST
// lexicographic combination rank in X's compacted adjacency row.
//
// Build: make -C oracle   (g++ -O2 -fopenmp -shared -fPIC)

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <queue>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_ML 14  // include/mps/cuPC-S.h:49  (#define ML 14)

extern "C" {

// ------------------------------------------------------------------------------------------
// Standard normal quantile.  The reference calls boost::math::quantile(normal(0,1), p)
// (src/cuPC_call_prep.cpp:7-11; Boost is an un-vendored system dependency, version not
// pinned by the reference).  Restated here with Wichura's algorithm AS 241 (PPND16,
// Appl. Statist. 37 (1988) 477-484), relative accuracy ~1e-16.
// ------------------------------------------------------------------------------------------
double orc_qnorm(double p)
{
    if (!(p > 0.0 && p < 1.0)) {
        if (p == 0.0) return -INFINITY;
        if (p == 1.0) return INFINITY;
        return NAN;
    }
    const double q = p - 0.5;
    double r, val;
    if (std::fabs(q) <= 0.425) {
        r = 0.180625 - q * q;
        val = q *
              (((((((r * 2509.0809287301226727 + 33430.575583588128105) * r +
                    67265.770927008700853) * r + 45921.953931549871457) * r +
                  13731.693765509461125) * r + 1971.5909503065514427) * r +
                133.14166789178437745) * r + 3.387132872796366608) /
              (((((((r * 5226.495278852545925 + 28729.085735721942674) * r +
                    39307.89580009271061) * r + 21213.794301586595867) * r +
                  5394.1960214247511077) * r + 687.1870074920579083) * r +
                42.313330701600911252) * r + 1.0);
        return val;
    }
    r = (q < 0) ? p : 1.0 - p;
    r = std::sqrt(-std::log(r));
    if (r <= 5.0) {
        r -= 1.6;
        val = (((((((r * 7.7454501427834140764e-4 + 0.0227238449892691845833) * r +
                    0.24178072517745061177) * r + 1.27045825245236838258) * r +
                  3.64784832476320460504) * r + 5.7694972214606914055) * r +
                4.6303378461565452959) * r + 1.42343711074968357734) /
              (((((((r * 1.05075007164441684324e-9 + 5.475938084995344946e-4) * r +
                    0.0151986665636164571966) * r + 0.14810397642748007459) * r +
                  0.68976733498510000455) * r + 1.6763848301838038494) * r +
                2.05319162663775882187) * r + 1.0);
    } else {
        r -= 5.0;
        val = (((((((r * 2.01033439929228813265e-7 + 2.71155556874348757815e-5) * r +
                    0.0012426609473880784386) * r + 0.026532189526576123093) * r +
                  0.29656057182850489123) * r + 1.7848265399172913358) * r +
                5.4637849111641143699) * r + 6.6579046435011037772) /
              (((((((r * 2.04426310338993978564e-15 + 1.4215117583164458887e-7) * r +
                    1.8463183175100546818e-5) * r + 7.868691311456132591e-4) * r +
                  0.0148753612908506148525) * r + 0.13692988092273580531) * r +
                0.59983220655588793769) * r + 1.0);
    }
    return (q < 0.0) ? -val : val;
}

// src/cuPC_call_prep.cpp:13-23: Th[l] = |qnorm(alpha/2)| / sqrt(n - l - 3), l = 0..14.
// Type flow restated: alpha and 0.5f*alpha are float; the quantile is evaluated in double
// and returned as float; the division by sqrt((double)(n-l-3)) is in double and the
// quotient is stored as float.
void orc_threshold_array(int n, float alpha, float *out15)
{
    const float half = 0.5f;
    const float p = half * alpha;
    const float q = (float)orc_qnorm((double)p);
    for (int i = 0; i < ORC_ML + 1; i++) {
        // n - i - 3 is size_t arithmetic in the reference (wraps if n < i + 3)
        size_t d = (size_t)n - (size_t)i - (size_t)3;
        out15[i] = (float)((double)std::fabs(q) / std::sqrt((double)d));
    }
}

// src/cuPC_call_prep.cpp:25-28
float orc_hetcor_threshold(float alpha)
{
    const float p = 0.5f * alpha;
    return std::fabs((float)orc_qnorm((double)p));
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// Combinatorics.  src/cuPC-S.cu:6453-6506 (BINOM, IthCombination) use 32-bit int; here
// 64-bit with saturation -- identical wherever the reference does not overflow.
// ------------------------------------------------------------------------------------------
static const uint64_t ORC_SAT = (uint64_t)1 << 62;

static uint64_t binom64(int n, int k)
{
    if (k < 0 || k > n) return 0;
    int p = std::min(k, n - k);
    int big = n - p;
    unsigned __int128 r = 1;
    for (int i = 1; i <= p; i++) {
        r = r * (unsigned)(big + i) / (unsigned)i;  // exact at every step
        if (r > ORC_SAT) return ORC_SAT;
    }
    return (uint64_t)r;
}

// 1-based rank L (1..C(N,P)) -> 1-based strictly increasing positions, lexicographic
// (src/cuPC-S.cu:6483-6506).
static void ith_combination(int *out, int N, int P, uint64_t L)
{
    uint64_t k = 0;
    int prev = 0;
    for (int i = 0; i < P - 1; i++) {
        int cur = prev;
        uint64_t R = 0;
        while (k < L) {
            cur++;
            R = binom64(N - cur, P - (i + 1));
            k += R;
        }
        k -= R;
        out[i] = cur;
        prev = cur;
    }
    out[P - 1] = prev + (int)(L - k);
}

// lexicographic successor of 0-based positions pos[0..P) over [0,N); false when exhausted
static bool next_combination(int *pos, int N, int P)
{
    int i = P - 1;
    while (i >= 0 && pos[i] == N - P + i) i--;
    if (i < 0) return false;
    pos[i]++;
    for (int j = i + 1; j < P; j++) pos[j] = pos[j - 1] + 1;
    return true;
}

// ------------------------------------------------------------------------------------------
// Small-matrix inverses (row-major, size m <= 14).
// ------------------------------------------------------------------------------------------

// Courrieu's full-rank-Cholesky Moore-Penrose inverse, as used for l = 2, 3
// (src/cuPC-S.cu:3084-3461): A = G'G, L = rank-revealing Cholesky factor of A with
// tolerance min(diag(A)>0) * 1e-20, M = inv(L'L), pinv = L M M L' G'.
static void pinv_courrieu(const double *G, double *out, int m)
{
    double A[ORC_ML * ORC_ML], L[ORC_ML * ORC_ML];
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < m; k++) s += G[i * m + k] * G[k * m + j];
            A[i * m + j] = s;
            L[i * m + j] = 0;
        }
    double tol = 999.99;
    for (int i = 0; i < m; i++)
        if (tol > A[i * m + i] && A[i * m + i] > 0) tol = A[i * m + i];
    tol *= 1e-20;
    int r = 0;
    for (int k = 0; k < m; k++) {
        for (int i = k; i < m; i++) {
            double t = 0;
            for (int q = 0; q < r; q++) t += L[i * m + q] * L[k * m + q];
            L[i * m + r] = A[i * m + k] - t;
        }
        if (L[k * m + r] > tol) {
            L[k * m + r] = std::sqrt(L[k * m + r]);
            for (int i = k + 1; i < m; i++) L[i * m + r] /= L[k * m + r];
            r++;
        }  // else: column r is overwritten by the next k (rank deficient)
    }
    // wipe the unused trailing column that a rejected pivot may have left behind
    for (int i = 0; i < m; i++)
        for (int j = r; j < m; j++) L[i * m + j] = 0;
    // M = inv(L' L)  (r x r, SPD) by Gauss-Jordan
    double M[ORC_ML * ORC_ML], W[ORC_ML * ORC_ML];
    for (int i = 0; i < r; i++)
        for (int j = 0; j < r; j++) {
            double s = 0;
            for (int k = 0; k < m; k++) s += L[k * m + i] * L[k * m + j];
            W[i * r + j] = s;
            M[i * r + j] = (i == j) ? 1.0 : 0.0;
        }
    for (int c = 0; c < r; c++) {
        int piv = c;
        for (int i = c + 1; i < r; i++)
            if (std::fabs(W[i * r + c]) > std::fabs(W[piv * r + c])) piv = i;
        if (piv != c)
            for (int j = 0; j < r; j++) {
                std::swap(W[c * r + j], W[piv * r + j]);
                std::swap(M[c * r + j], M[piv * r + j]);
            }
        double d = 1.0 / W[c * r + c];
        for (int j = 0; j < r; j++) {
            W[c * r + j] *= d;
            M[c * r + j] *= d;
        }
        for (int i = 0; i < r; i++) {
            if (i == c) continue;
            double f = W[i * r + c];
            if (f == 0) continue;
            for (int j = 0; j < r; j++) {
                W[i * r + j] -= f * W[c * r + j];
                M[i * r + j] -= f * M[c * r + j];
            }
        }
    }
    // out = L M M L' G'
    double T1[ORC_ML * ORC_ML], T2[ORC_ML * ORC_ML], T3[ORC_ML * ORC_ML];
    for (int i = 0; i < m; i++)
        for (int j = 0; j < r; j++) {
            double s = 0;
            for (int k = 0; k < r; k++) s += L[i * m + k] * M[k * r + j];
            T1[i * r + j] = s;  // L M
        }
    for (int i = 0; i < r; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < m; k++) s += L[k * m + i] * G[k * m + j];
            T2[i * m + j] = s;  // L' G   (the reference multiplies by M2, symmetric = G')
        }
    for (int i = 0; i < r; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < r; k++) s += M[i * r + k] * T2[k * m + j];
            T3[i * m + j] = s;  // M L' G
        }
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < r; k++) s += T1[i * r + k] * T3[k * m + j];
            out[i * m + j] = s;
        }
}

// l >= 4: the reference inverts through an SVD (Numerical-Recipes svdcmp, Golub-Reinsch,
// <=30 QR sweeps) and forms V diag(1/w) U' WITHOUT any singular-value cutoff
// (src/cuPC-S.cu:3463-3724 and the l5..l14 copies).  Restated with a one-sided Jacobi
// (Hestenes) SVD in double, which yields the same decomposition up to rounding; the inverse
// formula (no cutoff) is kept.
static void inv_svd_nocutoff(const double *Ain, double *out, int m)
{
    double U[ORC_ML * ORC_ML], V[ORC_ML * ORC_ML], w[ORC_ML];
    for (int i = 0; i < m * m; i++) U[i] = Ain[i];
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) V[i * m + j] = (i == j);
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < m - 1; p++)
            for (int q = p + 1; q < m; q++) {
                double a = 0, b = 0, c = 0;
                for (int k = 0; k < m; k++) {
                    a += U[k * m + p] * U[k * m + p];
                    b += U[k * m + q] * U[k * m + q];
                    c += U[k * m + p] * U[k * m + q];
                }
                if (c == 0) continue;
                off = std::max(off, std::fabs(c) / std::sqrt(a * b));
                double zeta = (b - a) / (2.0 * c);
                double t = ((zeta >= 0) ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
                for (int k = 0; k < m; k++) {
                    double up = U[k * m + p], uq = U[k * m + q];
                    U[k * m + p] = cs * up - sn * uq;
                    U[k * m + q] = sn * up + cs * uq;
                    double vp = V[k * m + p], vq = V[k * m + q];
                    V[k * m + p] = cs * vp - sn * vq;
                    V[k * m + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
    for (int j = 0; j < m; j++) {
        double s = 0;
        for (int k = 0; k < m; k++) s += U[k * m + j] * U[k * m + j];
        w[j] = std::sqrt(s);
        for (int k = 0; k < m; k++) U[k * m + j] /= w[j];  // no cutoff: w ~ 0 -> inf/nan
    }
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < m; k++) s += (V[i * m + k] / w[k]) * U[j * m + k];
            out[i * m + j] = s;
        }
}

// ------------------------------------------------------------------------------------------
// Partial correlation of (X,Y) given S, as the reference's level kernels compute it
// (src/cuPC-S.cu:561-566 for l = 1, :669-701 pattern for l >= 2):
//   M2 = C[S,S] with the diagonal taken as exactly 1, M1 = [C[X,S]; C[Y,S]],
//   H = [[1,C_xy],[.,1]] - M1 M2inv M1',  rho = H01 / sqrt|H00 H11|,
//   Z = |0.5 ln|(1+rho)/(1-rho)||.
// l = 2,3 use the Courrieu pseudo-inverse, l >= 4 the SVD inverse.  ONE documented deviation:
// for l >= 4 a singular value below 1e-10 * max is dropped (pseudo-inverse) instead of being
// divided by; the reference's float32 result is undefined (inf/nan) there.  *singular is
// incremented when that happens so callers can exclude such inputs from parity claims.
// ------------------------------------------------------------------------------------------
struct M2Inverse {
    double inv[ORC_ML * ORC_ML];
};

static void invert_m2(const float *C, int n, const int *S, int l, M2Inverse &out, long long *singular)
{
    if (l == 1) {
        out.inv[0] = 1.0;
        return;
    }
    double M2[ORC_ML * ORC_ML];
    for (int i = 0; i < l; i++)
        for (int j = 0; j < l; j++)
            M2[i * l + j] = (i == j) ? 1.0 : (double)C[(size_t)S[i] * n + S[j]];
    if (l <= 3) {
        pinv_courrieu(M2, out.inv, l);
        return;
    }
    // SVD inverse with the degenerate-case guard described above
    double U[ORC_ML * ORC_ML], V[ORC_ML * ORC_ML], w[ORC_ML];
    const int m = l;
    for (int i = 0; i < m * m; i++) U[i] = M2[i];
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) V[i * m + j] = (i == j);
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < m - 1; p++)
            for (int q = p + 1; q < m; q++) {
                double a = 0, b = 0, c = 0;
                for (int k = 0; k < m; k++) {
                    a += U[k * m + p] * U[k * m + p];
                    b += U[k * m + q] * U[k * m + q];
                    c += U[k * m + p] * U[k * m + q];
                }
                if (c == 0 || a == 0 || b == 0) continue;
                off = std::max(off, std::fabs(c) / std::sqrt(a * b));
                double zeta = (b - a) / (2.0 * c);
                double t = ((zeta >= 0) ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                double cs = 1.0 / std::sqrt(1.0 + t * t), sn = cs * t;
                for (int k = 0; k < m; k++) {
                    double up = U[k * m + p], uq = U[k * m + q];
                    U[k * m + p] = cs * up - sn * uq;
                    U[k * m + q] = sn * up + cs * uq;
                    double vp = V[k * m + p], vq = V[k * m + q];
                    V[k * m + p] = cs * vp - sn * vq;
                    V[k * m + q] = sn * vp + cs * vq;
                }
            }
        if (off < 1e-15) break;
    }
    double wmax = 0;
    for (int j = 0; j < m; j++) {
        double s = 0;
        for (int k = 0; k < m; k++) s += U[k * m + j] * U[k * m + j];
        w[j] = std::sqrt(s);
        wmax = std::max(wmax, w[j]);
    }
    bool dropped = false;
    for (int j = 0; j < m; j++) {
        if (w[j] <= 1e-10 * wmax) {
            dropped = true;
            for (int k = 0; k < m; k++) U[k * m + j] = 0;
            w[j] = 1.0;
        } else {
            for (int k = 0; k < m; k++) U[k * m + j] /= w[j];
        }
    }
    if (dropped && singular) {
#pragma omp atomic
        (*singular)++;
    }
    for (int i = 0; i < m; i++)
        for (int j = 0; j < m; j++) {
            double s = 0;
            for (int k = 0; k < m; k++) s += (V[i * m + k] / w[k]) * U[j * m + k];
            out.inv[i * m + j] = s;
        }
}

// Z statistic for (X,Y | S) given the inverse of M2 and v_x = M2inv * C[S,X] precomputed per S.
static inline double fisher_z(double rho)
{
    return std::fabs(0.5 * std::log(std::fabs((1.0 + rho) / (1.0 - rho))));
}

static double z_given_s(const float *C, int n, int X, int Y, const int *S, int l, const M2Inverse &inv,
                        const double *m1x)
{
    const double cxy = (double)C[(size_t)X * n + Y];
    if (l == 1) {
        // src/cuPC-S.cu:557-566
        const double a = m1x[0], b = (double)C[(size_t)Y * n + S[0]];
        const double h00 = 1 - a * a, h01 = cxy - a * b, h11 = 1 - b * b;
        const double rho = h01 / (std::sqrt(std::fabs(h00)) * std::sqrt(std::fabs(h11)));
        return std::fabs(0.5 * (std::log(std::fabs(1 + rho)) - std::log(std::fabs(1 - rho))));
    }
    double m1y[ORC_ML];
    for (int i = 0; i < l; i++) m1y[i] = (double)C[(size_t)Y * n + S[i]];
    // M1 * M2inv (2 x l), then H
    double ax[ORC_ML], ay[ORC_ML];
    for (int c = 0; c < l; c++) {
        double sx = 0, sy = 0;
        for (int k = 0; k < l; k++) {
            sx += m1x[k] * inv.inv[k * l + c];
            sy += m1y[k] * inv.inv[k * l + c];
        }
        ax[c] = sx;
        ay[c] = sy;
    }
    double h00 = 1, h01 = cxy, h11 = 1;
    for (int c = 0; c < l; c++) {
        h00 -= ax[c] * m1x[c];
        h01 -= ax[c] * m1y[c];
        h11 -= ay[c] * m1y[c];
    }
    const double rho = h01 / std::sqrt(std::fabs(h00 * h11));
    return fisher_z(rho);
}

// src/hetcor-cuPC-S.cu:3068-3089 (mean_ess): every unordered pair of var_ixs contributes
// (int)N[later*n + earlier]; isnan() is applied to the int so nothing is ever skipped; a NaN
// converts to 0 on the device (cvt.rzi.s32.f32), values saturate at the int32 range.
static inline long long trunc_ess(float v)
{
    if (std::isnan(v)) return 0;
    if (v >= 2147483648.0f) return 2147483647LL;
    if (v <= -2147483648.0f) return -2147483648LL;
    return (long long)(int)v;
}

static double mean_ess(const float *N, int n, const int *var_ixs, int cnt)
{
    long long s = 0;
    int num = 0;
    for (int i = 0; i < cnt; i++)
        for (int j = 0; j < i; j++) {
            s += trunc_ess(N[(size_t)var_ixs[i] * n + var_ixs[j]]);
            num++;
        }
    return (double)s / (double)num;
}

// src/cuPC-S.cu:6355-6432 (scan_compact): sorted neighbour ids of every row; returns max degree.
static int compact(const int *G, int n, std::vector<std::vector<int>> &adj)
{
    int nprime = 0;
    adj.assign(n, {});
    for (int r = 0; r < n; r++) {
        for (int c = 0; c < n; c++)
            if (G[(size_t)r * n + c] == 1) adj[r].push_back(c);
        nprime = std::max(nprime, (int)adj[r].size());
    }
    return nprime;
}

struct NearRec {
    int x, y, level;
    double z, th;
};

// One level l >= 1 of either engine, PC-stable, exhaustive and schedule independent:
// for every ordered pair (X,Y), Y in adj(X), the conditioning sets S (|S| = l) are drawn from
// adj(X)\{Y} in lexicographic rank order and the first one with Z < th is recorded.
// Canonical (deterministic) reading of the reference's racy bookkeeping
// (src/cuPC-S.cu:568-577): slot [X,Y] holds the lowest-rank separating set of X's row, or
// stays untouched when X's row has none.
// het != 0: hetcor engine (src/hetcor-cuPC-S.cu:379-630 pattern): time filter + per-test
// threshold, nothing recorded.
static void run_level(const float *C, int n, int *G, int l, double th_level, bool het, const float *N,
                      double th_het, const int *time_index, float *pMax, int *SepSet,
                      const std::vector<std::vector<int>> &adj, long long *tests_canonical,
                      long long *tests_executed, long long *singular, double near_band,
                      std::vector<NearRec> *near, int row_rem = 0, int row_mod = 1)
{
    long long t_canon = 0, t_exec = 0;
    std::vector<unsigned char> removed((size_t)n * n, 0);  // directed: found from row X
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : t_canon, t_exec)
    for (int X = 0; X < n; X++) {
        if (row_mod > 1 && (X % row_mod) != row_rem) continue;  // sharded emulation (orc_hetcor_level)
        const std::vector<int> &a = adj[X];
        const int d = (int)a.size();
        if (d <= l) {
            if (!(l == 1 && d == 1)) continue;  // level 1 has no explicit guard but does nothing
            continue;
        }
        const uint64_t ncomb = binom64(d, l);
        t_canon += (long long)std::min<uint64_t>(ncomb, (uint64_t)1 << 40) * (d - l);
        std::vector<char> found(d, 0);
        int n_alive = d;
        int pos[ORC_ML], S[ORC_ML];
        for (int i = 0; i < l; i++) pos[i] = i;
        M2Inverse inv;
        double m1x[ORC_ML];
        do {
            for (int i = 0; i < l; i++) S[i] = a[pos[i]];
            invert_m2(C, n, S, l, inv, singular);
            for (int i = 0; i < l; i++) m1x[i] = (double)C[(size_t)X * n + S[i]];
            int pi = 0;
            for (int yp = 0; yp < d; yp++) {
                if (pi < l && pos[pi] == yp) {
                    pi++;
                    continue;
                }
                if (found[yp]) continue;
                const int Y = a[yp];
                double th = th_level;
                if (het) {
                    // src/hetcor-cuPC-S.cu:3055-3065
                    const int maxtime = std::max(time_index[X], time_index[Y]);
                    bool ok = true;
                    for (int i = 0; i < l; i++)
                        if (time_index[S[i]] > maxtime) ok = false;
                    if (!ok) continue;
                    int var_ixs[ORC_ML + 2];
                    var_ixs[0] = X;
                    var_ixs[1] = Y;
                    for (int i = 0; i < l; i++) var_ixs[2 + i] = S[i];
                    th = th_het / std::sqrt(mean_ess(N, n, var_ixs, l + 2) - (double)l - 3.0);
                }
                t_exec++;
                const double Z = z_given_s(C, n, X, Y, S, l, inv, m1x);
                if (near && std::fabs(Z - th) <= near_band * std::fabs(th)) {
#pragma omp critical
                    if (near->size() < 100000) near->push_back({X, Y, l, Z, th});
                }
                if (Z < th) {
                    found[yp] = 1;
                    n_alive--;
                    removed[(size_t)X * n + Y] = 1;
                    if (!het) {
                        pMax[(size_t)X * n + Y] = (float)Z;
                        int *slot = SepSet + ((size_t)X * n + Y) * ORC_ML;
                        for (int i = 0; i < l; i++) slot[i] = S[i];
                    }
                }
            }
        } while (n_alive > 0 && next_combination(pos, d, l));
    }
    for (int X = 0; X < n; X++)
        for (int Y = 0; Y < n; Y++)
            if (removed[(size_t)X * n + Y]) {
                G[(size_t)X * n + Y] = 0;
                G[(size_t)Y * n + X] = 0;
            }
    *tests_canonical = t_canon;
    *tests_executed = t_exec;
}

extern "C" {

// Canonical full-enumeration test count of a level: sum_X C(d_X,l) (d_X - l)   
long long orc_canonical_tests(const int *deg, int n, int l)
{
    long long t = 0;
    for (int i = 0; i < n; i++)
        if (deg[i] > l) {
            uint64_t c = binom64(deg[i], l);
            if (c > ((uint64_t)1 << 40)) c = (uint64_t)1 << 40;
            t += (long long)c * (deg[i] - l);
        }
    return t;
}

// Restatement of Skeleton() (src/cuPC-S.cu:61-450).
//   C[n*n] in; G[n*n] out (input ignored: level 0 sets 0/1, src/cuPC-S.cu:458-484);
//   Th[15]; *l_out = last level run (src/cuPC-S.cu:99,154-159); pMax[n*n], SepSet[n*n*14] out.
// pMax: the reference leaves never-written device entries undefined; here they start at 0.
//   Post-processing as src/cuPC-S.cu:424-442.
// stats (may be NULL): [0..14] canonical tests per level, [15..29] executed, [30] singular M2
// near (may be NULL): up to near_cap records {x,y,level,z,th} (5 doubles each) of every
//   evaluated test with |Z-th| <= near_band*th; *near_n = count.
void orc_skeleton(const float *C, int n, int *G, const float *Th, int *l_out, int maxlevel, float *pMax,
                  int *SepSet, long long *stats, double near_band, double *near, int near_cap,
                  int *near_n)
{
    std::vector<NearRec> nr;
    std::vector<NearRec> *nrp = near ? &nr : nullptr;
    long long singular = 0;
    if (stats) std::memset(stats, 0, 31 * sizeof(long long));
    for (size_t i = 0; i < (size_t)n * n; i++) pMax[i] = 0.f;
    for (size_t i = 0; i < (size_t)n * n * ORC_ML; i++) SepSet[i] = -1;
    int l;
    bool finish = false;
    std::vector<std::vector<int>> adj;
    for (l = 0; l <= ORC_ML && !finish && l <= maxlevel; l++) {
        if (l == 0) {
            for (int r = 0; r < n; r++)
                for (int c = r + 1; c < n; c++) {
                    const double res = fisher_z((double)C[(size_t)r * n + c]);
                    if (nrp && std::fabs(res - (double)Th[0]) <= near_band * Th[0] && nr.size() < 100000)
                        nr.push_back({r, c, 0, res, (double)Th[0]});
                    if (res < (double)Th[0]) {
                        pMax[(size_t)r * n + c] = pMax[(size_t)c * n + r] = (float)res;
                        G[(size_t)r * n + c] = G[(size_t)c * n + r] = 0;
                    } else {
                        G[(size_t)r * n + c] = G[(size_t)c * n + r] = 1;
                    }
                }
            for (int r = 0; r < n; r++) G[(size_t)r * n + r] = 0;
            if (stats) stats[0] = stats[15] = (long long)n * (n - 1) / 2;
        } else {
            const int nprime = compact(G, n, adj);
            if (nprime - 1 < l) {
                l = l - 1;
                finish = true;
                break;
            }
            long long tc = 0, te = 0;
            run_level(C, n, G, l, (double)Th[l], false, nullptr, 0, nullptr, pMax, SepSet, adj, &tc, &te,
                      &singular, near_band, nrp);
            if (stats) {
                stats[l] = tc;
                stats[15 + l] = te;
            }
        }
    }
    *l_out = l;
    // src/cuPC-S.cu:424-442
    for (int i = 0; i < n; i++) {
        pMax[(size_t)i * n + i] = 1;
        for (int j = i + 1; j < n; j++) {
            if (G[(size_t)i * n + j] == 0) {
                const float t = std::fmax(pMax[(size_t)j * n + i], pMax[(size_t)i * n + j]);
                pMax[(size_t)j * n + i] = pMax[(size_t)i * n + j] = t;
            } else {
                pMax[(size_t)j * n + i] = pMax[(size_t)i * n + j] = -100000;
            }
        }
    }
    if (stats) stats[30] = singular;
    if (near) {
        int k = std::min<int>((int)nr.size(), near_cap);
        for (int i = 0; i < k; i++) {
            near[5 * i + 0] = nr[i].x;
            near[5 * i + 1] = nr[i].y;
            near[5 * i + 2] = nr[i].level;
            near[5 * i + 3] = nr[i].z;
            near[5 * i + 4] = nr[i].th;
        }
        *near_n = k;
    }
}

// Restatement of hetcor_skeleton() (src/hetcor-cuPC-S.cu:75-341).
//   G[n*n] IN/OUT (1 = candidate edge; level 0 only clears, :343-377); N[n*n] effective sample
//   sizes; th = |qnorm(alpha/2)|; time_index[n].
// Deviation, documented in DESIGN.md: the reference's block-level NoEdgeFlag early exit
// (src/hetcor-cuPC-S.cu:550-589) can stop a row before all S were enumerated when
// all remaining candidates of one 64-rank chunk are filtered; that is schedule dependent and
// is NOT restated -- every (X,Y,S) allowed by the time filter is evaluated.
void orc_hetcor_skeleton(const float *C, int n, int *G, const float *N, float th, int *l_out, int maxlevel,
                         const int *time_index, long long *stats, double near_band, double *near,
                         int near_cap, int *near_n)
{
    std::vector<NearRec> nr;
    std::vector<NearRec> *nrp = near ? &nr : nullptr;
    long long singular = 0;
    if (stats) std::memset(stats, 0, 31 * sizeof(long long));
    int l;
    bool finish = false;
    std::vector<std::vector<int>> adj;
    for (l = 0; l <= ORC_ML && !finish && l <= maxlevel; l++) {
        if (l == 0) {
            for (int r = 0; r < n; r++)
                for (int c = r + 1; c < n; c++) {
                    const double res = fisher_z((double)C[(size_t)r * n + c]);
                    const double loc_th = (double)th / std::sqrt((double)N[(size_t)r * n + c] - 3.0);
                    if (nrp && std::fabs(res - loc_th) <= near_band * loc_th && nr.size() < 100000)
                        nr.push_back({r, c, 0, res, loc_th});
                    if (res < loc_th) G[(size_t)r * n + c] = G[(size_t)c * n + r] = 0;
                }
            for (int r = 0; r < n; r++) G[(size_t)r * n + r] = 0;
            if (stats) stats[0] = stats[15] = (long long)n * (n - 1) / 2;
        } else {
            const int nprime = compact(G, n, adj);
            if (nprime - 1 < l) {
                l = l - 1;
                finish = true;
                break;
            }
            long long tc = 0, te = 0;
            run_level(C, n, G, l, 0.0, true, N, (double)th, time_index, nullptr, nullptr, adj, &tc, &te,
                      &singular, near_band, nrp);
            if (stats) {
                stats[l] = tc;
                stats[15 + l] = te;
            }
        }
    }
    *l_out = l;
    if (stats) stats[30] = singular;
    if (near) {
        int k = std::min<int>((int)nr.size(), near_cap);
        for (int i = 0; i < k; i++) {
            near[5 * i + 0] = nr[i].x;
            near[5 * i + 1] = nr[i].y;
            near[5 * i + 2] = nr[i].level;
            near[5 * i + 3] = nr[i].z;
            near[5 * i + 4] = nr[i].th;
        }
        *near_n = k;
    }
}

// One level l >= 1 of the hetcor engine restricted to the rows X with X % row_mod == row_rem, on the
// adjacency G as it stands at the start of the level (compacted here).  Test scaffolding for the
// edge-sharded multi-GPU mode: ranks run disjoint row sets, AND their G's, and must reach the
// single-rank result (PC-stable).  Returns 0 when the level does not run (max degree - 1 < l,
// src/hetcor-cuPC-S.cu:160-166), else 1.
int orc_hetcor_level(const float *C, int n, int *G, const float *N, float th, int l, const int *time_index,
                     int row_rem, int row_mod)
{
    std::vector<std::vector<int>> adj;
    const int nprime = compact(G, n, adj);
    if (nprime - 1 < l) return 0;
    long long tc = 0, te = 0, singular = 0;
    run_level(C, n, G, l, 0.0, true, N, (double)th, time_index, nullptr, nullptr, adj, &tc, &te, &singular, 0.0,
              nullptr, row_rem, row_mod);
    return 1;
}

// Exposed for tests: the two inverse restatements and unranking.
void orc_pinv_courrieu(const double *A, double *out, int m) { pinv_courrieu(A, out, m); }
void orc_inv_svd(const double *A, double *out, int m) { inv_svd_nocutoff(A, out, m); }
void orc_ith_combination(int *out, int N, int P, long long L) { ith_combination(out, N, P, (uint64_t)L); }
long long orc_binom(int n, int k) { return (long long)binom64(n, k); }

// ------------------------------------------------------------------------------------------
// PLINK BED decode (include/mps/bed_lut_gpu.h:3,1030):
//   2-bit code 00 -> value 2, 01 -> missing (value 0, valid 0), 10 -> 1, 11 -> 0; sample j of a
//   column sits in bits 2(j&3).. of byte j>>2.
// ------------------------------------------------------------------------------------------
static inline void bed_code(unsigned char byte, int j, int *val, int *valid)
{
    const int code = (byte >> (2 * j)) & 3;
    static const int V[4] = {2, 0, 1, 0};
    static const int OK[4] = {1, 0, 1, 1};
    *val = V[code];
    *valid = OK[code];
}

// src/prep.cpp:15-77 (compute_bed_col_stats_no_impute): float32 finish, sequential float
// accumulation of squares in sample order.  float32 is part of the on-disk contract (the
// values are written as text with 6 significant digits, src/io.cpp:337-362), so the oracle
// keeps float here.
void orc_bed_col_stats(const unsigned char *bed, size_t num_markers, size_t num_individuals, float *means,
                       float *stds, int *modes, long long *counts4 /* [m*4]: n0,n1,n2,nmiss or NULL */)
{
    const size_t col_bytes = (num_individuals + 3) / 4;
#pragma omp parallel for schedule(static)
    for (long long m = 0; m < (long long)num_markers; m++) {
        const unsigned char *col = bed + (size_t)m * col_bytes;
        int gt_counts[3] = {0, 0, 0};
        size_t gt_sum = 0, nan_count = 0;
        for (size_t i = 0; i < col_bytes; i++)
            for (int j = 0; j < 4 && i * 4 + j < num_individuals; j++) {
                int v, ok;
                bed_code(col[i], j, &v, &ok);
                if (!ok) nan_count++;
                else {
                    gt_counts[v]++;
                    gt_sum += v;
                }
            }
        int mode = 0;
        for (int i = 1; i < 3; i++)
            if (gt_counts[i] > gt_counts[mode]) mode = i;
        const float mean = gt_sum / (float)(num_individuals - nan_count);
        float sum_squares = 0.0f;
        for (size_t i = 0; i < col_bytes; i++)
            for (int j = 0; j < 4 && i * 4 + j < num_individuals; j++) {
                int v, ok;
                bed_code(col[i], j, &v, &ok);
                if (ok) sum_squares += ((float)v - mean) * ((float)v - mean);
            }
        means[m] = mean;
        stds[m] = std::sqrt(sum_squares / (float)(num_individuals - nan_count));
        modes[m] = mode;
        if (counts4) {
            counts4[4 * m + 0] = gt_counts[0];
            counts4[4 * m + 1] = gt_counts[1];
            counts4[4 * m + 2] = gt_counts[2];
            counts4[4 * m + 3] = (long long)nan_count;
        }
    }
}

// 3x3 contingency table of a marker pair, s[3a+b], valid-valid samples only
// (src/corr_kernels.cu:497-509).  Exact integers.
void orc_bed_pair_table(const unsigned char *bed, size_t num_individuals, size_t a, size_t b, long long *s9)
{
    const size_t col_bytes = (num_individuals + 3) / 4;
    for (int i = 0; i < 9; i++) s9[i] = 0;
    const unsigned char *ca = bed + a * col_bytes, *cb = bed + b * col_bytes;
    for (size_t i = 0; i < col_bytes; i++)
        for (int j = 0; j < 4 && i * 4 + j < num_individuals; j++) {
            int va, oa, vb, ob;
            bed_code(ca[i], j, &va, &oa);
            bed_code(cb[i], j, &vb, &ob);
            if (oa && ob) s9[3 * va + vb]++;
        }
}

// src/corr_kernels.cu:548-563: Kendall tau_b from the table, then sin(pi/2 tau).  FP64.
double orc_npn_from_table(const long long *t)
{
    double s[9];
    for (int i = 0; i < 9; i++) s[i] = (double)t[i];
    const double p = (s[0] * (s[4] + s[5] + s[7] + s[8])) + (s[1] * (s[5] + s[8])) + (s[3] * (s[7] + s[8])) +
                     (s[4] * s[8]);
    const double q = (s[1] * (s[3] + s[6])) + (s[2] * (s[3] + s[4] + s[6] + s[7])) + (s[4] * s[6]) +
                     (s[5] * (s[6] + s[7]));
    const double tt = (s[0] * (s[1] + s[2])) + (s[1] * s[2]) + (s[3] * (s[4] + s[5])) + (s[4] * s[5]) +
                      (s[6] * (s[7] + s[8])) + (s[7] * s[8]);
    const double u = (s[0] * (s[3] + s[6])) + (s[1] * (s[4] + s[7])) + (s[2] * (s[5] + s[8])) + (s[3] * s[6]) +
                     (s[4] * s[7]) + (s[5] * s[8]);
    const double kendall = (p - q) / std::sqrt((p + q + tt) * (p + q + u));
    return std::sin(M_PI / 2 * kendall);
}

// cu_corr_pearson_npn (src/corr_host.cu:1094-1197): mxm upper-tri row-major (npn), mxp
// [m*p] (src/corr_kernels.cu:157-238), pxp upper-tri (src/corr_kernels.cu:285-343).
// Outputs in double (truth) -- callers compare after casting as they need.
void orc_corr_npn(const unsigned char *bed, const float *phen, size_t m, size_t n, size_t p, const float *mean,
                  const float *stdv, double *mxm, double *mxp, double *pxp)
{
    const size_t col_bytes = (n + 3) / 4;
    if (mxm) {
#pragma omp parallel for schedule(dynamic, 1)
        for (long long a = 0; a < (long long)m; a++) {
            size_t lin = (size_t)a * (2 * m - a - 1) / 2;
            for (size_t b = a + 1; b < m; b++, lin++) {
                long long t[9];
                orc_bed_pair_table(bed, n, a, b, t);
                mxm[lin] = orc_npn_from_table(t);
            }
        }
    }
    if (mxp) {
#pragma omp parallel for schedule(static)
        for (long long mv = 0; mv < (long long)m; mv++) {
            const unsigned char *col = bed + (size_t)mv * col_bytes;
            for (size_t t = 0; t < p; t++) {
                const float *y = phen + t * n;
                double s_gy = 0, s_y = 0, s_v = 0;
                for (size_t i = 0; i < n; i++) {
                    int v, ok;
                    bed_code(col[i >> 2], (int)(i & 3), &v, &ok);
                    const float yv = y[i];
                    if (!std::isnan(yv)) {
                        s_gy += (double)ok * v * (double)yv;
                        s_y += (double)ok * (double)yv;
                        s_v += ok;
                    }
                }
                mxp[mv * p + t] = (s_gy - (double)mean[mv] * s_y) / (s_v * (double)stdv[mv]);
            }
        }
    }
    if (pxp) {
        size_t lin = 0;
        for (size_t a = 0; a < p; a++)
            for (size_t b = a + 1; b < p; b++, lin++) {
                double s = 0, cnt = 0;
                for (size_t i = 0; i < n; i++) {
                    const float va = phen[a * n + i], vb = phen[b * n + i];
                    if (!(std::isnan(va) || std::isnan(vb))) {
                        s += (double)va * (double)vb;
                        cnt += 1;
                    }
                }
                pxp[lin] = s / cnt;
            }
    }
}

// ---- `mps block` (src/cli.cpp:362-411): banded marker correlations, row sums, LD blocking ----

// cal_mcorrk_banded (src/corr_host.cu:64-110) through cu_marker_corr_pearson_npn_batched_sparse
// (:1199-1319) and bed_marker_corr_pearson_npn_sparse_scan (src/corr_kernels.cu:920-1010):
// corrs[row * W + k] = npn correlation of markers (row, row + 1 + k) for row + 1 + k < m, else 0.
// (The device-memory batching of the reference does not change any value.)  Table exact, formula
// in FP64, stored as float like the reference's result array.
void orc_mcorrk_banded(const unsigned char *bed, size_t m, size_t n, size_t W, float *corrs)
{
#pragma omp parallel for schedule(dynamic, 4)
    for (long long row = 0; row < (long long)m; row++)
        for (size_t k = 0; k < W; k++) {
            const size_t col = (size_t)row + 1 + k;
            float r = 0.0f;
            if (col < m) {
                long long t[9];
                orc_bed_pair_table(bed, n, (size_t)row, col, t);
                r = (float)orc_npn_from_table(t);
            }
            corrs[(size_t)row * W + k] = r;
        }
}

// marker_corr_banded_mat_row_abs_sums (src/corr_host.cu:112-128): float accumulator, columns in order.
void orc_banded_row_abs_sums(const float *corrs, size_t m, size_t W, float *sums)
{
    for (size_t row = 0; row < m; row++) {
        float s = 0.0f;
        for (size_t col = 0; col < W; col++) s += std::fabs(corrs[row * W + col]);
        sums[row] = s;
    }
}

// hann (src/blocking.cpp:8-11): the cosine is taken in single precision (cosf of the double argument
// rounded to float); the rest is double.  window_size 1 gives 0/0 -> NaN weights, as in the reference.
static double orc_hann(int i, int window_size)
{
    return 0.5 - 0.5 * cosf(2.0 * M_PI * (double)i / ((double)window_size - 1.0));
}

// hanning_smoothing (src/blocking.cpp:13-36): res[c] = sum_i w[i] v[c - margin + i] for margin <= c < nv - margin,
// 0 outside.  (The reference compares `int center` against the unsigned nv - margin; every window block_chr
// tries has margin <= nv / 4, where the two readings agree.)
void orc_hanning_smoothing(const float *v, size_t nv, int window_size, double *res)
{
    std::vector<double> w(window_size);
    for (int i = 0; i < window_size; i++) w[i] = orc_hann(i, window_size);
    const long long margin = window_size / 2;
    for (size_t i = 0; i < nv; i++) res[i] = 0.0;
    for (long long c = margin; c < (long long)nv - margin; c++)
        for (int i = 0; i < window_size; i++) res[c] += w[i] * (double)v[c - margin + i];
}

// local_minima (src/blocking.cpp:38-55): i is a minimum when the running maximum since the last
// minimum exceeds v[i] and v[i] < v[i+1]; the running maximum restarts at 0 after each minimum.
static std::vector<size_t> orc_local_minima(const std::vector<double> &v)
{
    std::vector<size_t> res;
    double left = 0.0;
    for (size_t i = 1; i + 1 < v.size(); i++) {
        if (left > v[i] && v[i] < v[i + 1]) {
            res.push_back(i);
            left = 0.0;
        } else if (v[i] > left)
            left = v[i];
    }
    return res;
}

// block_chr_with_window_size (src/blocking.cpp:87-95) + blocks_from_minima (:57-70): blocks
// [prev, minimum], next block starts after the minimum; the last block ends at nv - 1.
static long long orc_blocks_with_window(const float *v, size_t nv, int window_size, std::vector<long long> *first,
                                        std::vector<long long> *last)
{
    std::vector<double> smooth(nv);
    orc_hanning_smoothing(v, nv, window_size, smooth.data());
    const std::vector<size_t> minima = orc_local_minima(smooth);
    first->clear();
    last->clear();
    long long prev = 0, largest = 0;
    for (size_t pos : minima) {
        first->push_back(prev);
        last->push_back((long long)pos);
        prev = (long long)pos + 1;
    }
    first->push_back(prev);
    last->push_back((long long)nv - 1);
    for (size_t i = 0; i < first->size(); i++) largest = std::max(largest, (*last)[i] - (*first)[i] + 1);
    return largest;  // largest_block_size (:72-85)
}

// block_chr (src/blocking.cpp:97-136): bisection on the (odd) window size between 3 and nv until the largest
// block is within MAX_BLOCK_SIZE_TOL = 100 below max_block_size, or the window stops changing.
// first/last need room for nv entries; returns the number of blocks, *final_window the window used.
long long orc_block_chr(const float *v, size_t nv, int max_block_size, long long *first, long long *last,
                        int *final_window)
{
    auto make_odd = [](int x) { return x % 2 == 0 ? x - 1 : x; };
    int too_large = (int)nv, too_small = 3;
    int window = make_odd((too_large + too_small) / 2);
    std::vector<long long> f, l;
    int lbs = (int)orc_blocks_with_window(v, nv, window, &f, &l);
    while (std::abs(lbs - max_block_size) > 100 || lbs > max_block_size) {
        if (lbs > max_block_size) too_large = std::min(too_large, window);
        else too_small = std::max(too_small, window);
        const int next = make_odd((too_large + too_small) / 2);
        if (next == window) break;
        window = next;
        lbs = (int)orc_blocks_with_window(v, nv, window, &f, &l);
    }
    for (size_t i = 0; i < f.size(); i++) {
        first[i] = f[i];
        last[i] = l[i];
    }
    if (final_window) *final_window = window;
    return (long long)f.size();
}

// src/parent_set.cpp:8-53 (subset_variables): all traits + markers within max_depth hops of
// a trait, hop 1 = trait->marker, later hops marker->marker.  out_sel[num_var] 0/1 flags.
void orc_subset_variables(const int *G, int num_var, int num_markers, int max_depth, int *out_sel)
{
    for (int i = 0; i < num_var; i++) out_sel[i] = (i >= num_markers);
    for (int start = num_markers; start < num_var; start++) {
        std::vector<char> in(num_var, 0);
        for (int i = num_markers; i < num_var; i++) in[i] = 1;
        std::vector<int> q{start}, nq;
        for (int depth = 0; depth < max_depth; depth++) {
            nq.clear();
            for (int node : q)
                for (int c = 0; c < num_markers; c++)
                    if (G[(size_t)node * num_var + c] == 1 && !in[c]) {
                        in[c] = 1;
                        nq.push_back(c);
                    }
            q = nq;
        }
        for (int i = 0; i < num_var; i++)
            if (in[i]) out_sel[i] = 1;
    }
}

// bench.py's CPU legs ask for all host cores: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
// otherwise time the CPU baseline on one thread
// ---- reference-emulation counters (SURVEY 7 step 1 / 8c "--ref-emulation") -----------------------
// What the REFERENCE's level-l kernels would do differently from the restatement above on the graph G at the start of
// level l, counted per row, so that a difference against the reference's GPU output can be explained:
//   F10  int32 ranks: BINOM / IthCombination work in int (src/cuPC-S.cu:6453-6506); a row whose C(d, l) does not fit
//        an int enumerates garbage.  out[0] = rows with C(d, l) >= 2^31.
//   F9   NoEdgeFlag (src/hetcor-cuPC-S.cu:550-589, the same in every cal_Indepl*_ess): a row X is served by
//        NumOfBlockForEachNodeL = 2 CTAs of ParGivenL = 64 threads; in pass d1 thread tx of CTA bx takes the rank
//        tx + 64 bx + 128 d1; every thread with a rank sets the CTA's flag to 1 before its set and clears it when it
//        meets a neighbour Y outside S whose edge is still there and whose time allows S (valid_time_conditioning,
//        :3055-3065); a CTA whose flag is still 1 at the top of the next pass returns.  Statically (edges as at level
//        start, which is the best case for the reference: removals during the level only add early exits) a pass leaves
//        the flag at 1 exactly when none of its 64 sets has such a Y.  out[1] = rows where some CTA returns while
//        later passes of that CTA still hold a set with a valid (Y, S): rows the reference cuts short; out[2] = the
//        conditioning sets it never evaluates there; out[3] = rows not analysed (more than `cap` sets).
// The float32 / --use_fast_math arithmetic of the reference is NOT emulated (its GPU intrinsics have no CPU
// equivalent); oracle/_ref (the reference's own kernels built for sm_100) is the evidence for that side.
void orc_ref_quirks(const int *G, int n, const int *time_index, int l, long long cap, long long *out4)
{
    std::vector<std::vector<int>> adj;
    compact(G, n, adj);
    long long rows_f10 = 0, rows_cut = 0, sets_lost = 0, rows_skipped = 0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : rows_f10, rows_cut, sets_lost, rows_skipped)
    for (int X = 0; X < n; X++) {
        const std::vector<int> &a = adj[X];
        const int d = (int)a.size();
        if (d <= l) continue;  // the kernels return at SizeOfArr <= level
        const uint64_t nc = binom64(d, l);
        if (nc >= ((uint64_t)1 << 31)) rows_f10++;
        if ((long long)nc > cap) {
            rows_skipped++;
            continue;
        }
        const int tX = time_index[X];
        // per pass and CTA: does any of the 64 sets have a neighbour Y outside S with max(tX, tY) >= max time of S?
        std::vector<int> pos(l);
        for (int i = 0; i < l; i++) pos[i] = i;
        // largest and second largest max(tX, tY) over the neighbours, to answer "some Y outside S" in O(l)
        const uint64_t npass = (nc + 127) / 128;
        std::vector<unsigned char> resets(2 * npass, 0), has_valid(2 * npass, 0);
        std::vector<uint64_t> valid_sets(2 * npass, 0);
        uint64_t rank = 0;
        bool more = true;
        while (more) {
            int tS = -2147483647 - 1;
            for (int i = 0; i < l; i++) tS = std::max(tS, time_index[a[pos[i]]]);
            bool any = false;
            int pi = 0;
            for (int y = 0; y < d && !any; y++) {
                while (pi < l && pos[pi] < y) pi++;
                if (pi < l && pos[pi] == y) continue;
                if (std::max(tX, time_index[a[y]]) >= tS) any = true;
            }
            const uint64_t pass = rank / 128, cta = (rank % 128) / 64;
            if (any) {
                resets[2 * pass + cta] = 1;
                valid_sets[2 * pass + cta]++;
            }
            rank++;
            more = next_combination(pos.data(), d, l);
        }
        bool cut = false;
        for (int cta = 0; cta < 2; cta++) {
            bool returned = false;
            for (uint64_t ps = 0; ps < npass; ps++) {
                if (returned) {
                    if (valid_sets[2 * ps + cta]) {
                        cut = true;
                        sets_lost += (long long)valid_sets[2 * ps + cta];
                    }
                    continue;
                }
                // the pass runs; it leaves the flag at 1 when it had ranks but no thread cleared it
                const uint64_t first = ps * 128 + (uint64_t)cta * 64;
                if (first < nc && !resets[2 * ps + cta]) returned = true;  // seen at the top of pass ps + 1
            }
        }
        if (cut) rows_cut++;
    }
    out4[0] = rows_f10;
    out4[1] = rows_cut;
    out4[2] = sets_lost;
    out4[3] = rows_skipped;
}

void orc_set_num_threads(int t)
{
#ifdef _OPENMP
    if (t > 0) omp_set_num_threads(t);
#else
    (void)t;
#endif
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
