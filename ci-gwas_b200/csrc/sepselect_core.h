// sepselect_core.h -- the greedy separating-set search of ONE outer pair (i, j)
// (cusk_postprocessing/sepselect.py:262-304, find_maximal_and_min_pcorr_sepsets_incr, with partial_correlation
// :162-164, fisher_z :22-23, independent :30-31).  Device code of sepselect_pairs_kernel (sepselect.cu); it is
// plain C++ so that the tests can also compile it for the host and check it against the reference's outputs
// without a GPU (tests/sepselect_host_core.cpp) -- the product only ever calls it from the kernel.
//
// The reference inverts corr[{i, j} + S, {i, j} + S] with LAPACK for every candidate S and uses entries (0,0),
// (0,1), (1,1) of the inverse.  Those are the inverse of the 2x2 Schur complement
//     H = C[ij,ij] - C[ij,S] C[S,S]^-1 C[S,ij],     -m01 / sqrt|m00 m11| = +-h01 / sqrt|h00 h11|
// (the sign drops out of fisher_z).  S grows by one trait per round, so C[S,S] = L D L^T is extended by one row
// per committed trait and every candidate of a round costs one forward substitution: O(P^2) per round instead
// of O(P^4).
#pragma once
#include <math.h>
#include <stdint.h>

#ifdef __CUDACC__
#define CUSK_SS_HD __host__ __device__ __forceinline__
#else
#define CUSK_SS_HD inline
#endif

#include "../../include/cusk_b200.h"  // cusk_sepselect_result, CUSK_SS_MAXP (trait neighbours are a 32-bit mask)

CUSK_SS_HD int cusk_ss_popc(uint32_t m)
{
#ifdef __CUDA_ARCH__
    return __popc(m);
#else
    return __builtin_popcount(m);
#endif
}
CUSK_SS_HD int cusk_ss_lowest(uint32_t m)  // index of the lowest set bit, m != 0
{
#ifdef __CUDA_ARCH__
    return __ffs((int)m) - 1;
#else
    return __builtin_ctz(m);
#endif
}

// |atanh(v)| as the reference writes it (sepselect.py:22-23)
CUSK_SS_HD double cusk_ss_fisher_z(double v) { return fabs(0.5 * log(fabs((1.0 + v) / (1.0 - v)))); }

// Ct: rows 0..P (the traits and the first marker, see is_marker sepselect.py:542-543) of the symmetric correlation
// matrix, row-major with leading dimension ld; entry (a, b) with a <= P is Ct[a * ld + b].
// thr[l] = qnorm(1 - alpha/2) / sqrt(num_samples - l - 3), l = 0..P.
// Ties between candidates go to the larger trait index (the reference's `<=` over its set iteration).
CUSK_SS_HD void cusk_sepselect_pair(const double *Ct, int64_t ld, int P, int i, int j, uint32_t trait_neighbours_of_i,
                                    const double *thr, cusk_sepselect_result *out)
{
    double L[CUSK_SS_MAXP * (CUSK_SS_MAXP - 1) / 2];  // unit lower factor of C[S,S], packed: row k starts at k(k-1)/2
    double D[CUSK_SS_MAXP], ti[CUSK_SS_MAXP], tj[CUSK_SS_MAXP];  // pivots, L^-1 C[S,i], L^-1 C[S,j]
    double row[CUSK_SS_MAXP], best_row[CUSK_SS_MAXP];
    int cand[CUSK_SS_MAXP];
    // C[i,j]: at least one of the two is <= P (every outer pair has a trait end)
    const double cij = i <= P ? Ct[(int64_t)i * ld + j] : Ct[(int64_t)j * ld + i];
    double h00 = 1.0, h11 = 1.0, h01 = cij;  // Schur complement for the committed set
    uint32_t rem = trait_neighbours_of_i;
    const int max_size = cusk_ss_popc(rem);
    int len = 0;
    bool found_sepset = cusk_ss_fisher_z(h01 / sqrt(fabs(h00 * h11))) < thr[0];
    bool found_minimum = false;
    double last_ref = INFINITY;
    for (int size = 1; size <= max_size; size++) {
        int add = -1;
        double ref = INFINITY, b_d = 0, b_ti = 0, b_tj = 0;
        for (uint32_t m = rem; m; m &= m - 1) {
            const int nb = cusk_ss_lowest(m);
            const double *cn = Ct + (int64_t)nb * ld;  // nb < P: a trait row
            // extend the factorisation by nb: row = L-row of nb, d = its pivot
            double d = 1.0, xi = cn[i], xj = cn[j];
            for (int k = 0; k < len; k++) {
                double s = cn[cand[k]];
                const double *Lk = L + k * (k - 1) / 2;
                for (int q = 0; q < k; q++) s -= row[q] * D[q] * Lk[q];
                s /= D[k];
                row[k] = s;
                d -= s * s * D[k];
                xi -= s * ti[k];
                xj -= s * tj[k];
            }
            const double g00 = h00 - xi * xi / d, g11 = h11 - xj * xj / d, g01 = h01 - xi * xj / d;
            const double pc = cusk_ss_fisher_z(g01 / sqrt(fabs(g00 * g11)));
            if (pc <= ref) {  // NaN never wins
                ref = pc;
                add = nb;
                b_d = d;
                b_ti = xi;
                b_tj = xj;
                for (int k = 0; k < len; k++) best_row[k] = row[k];
            }
        }
        if (add < 0) break;  // every candidate singular (the reference stops with a LinAlgError there)
        if (ref > last_ref && found_sepset && !found_minimum) found_minimum = true;
        const bool loc_idp = ref < thr[size];
        if (!loc_idp && found_sepset) break;
        if (loc_idp) found_sepset = true;
        last_ref = ref;
        for (int k = 0; k < len; k++) L[len * (len - 1) / 2 + k] = best_row[k];
        D[len] = b_d;
        ti[len] = b_ti;
        tj[len] = b_tj;
        h00 -= b_ti * b_ti / b_d;
        h11 -= b_tj * b_tj / b_d;
        h01 -= b_ti * b_tj / b_d;
        cand[len++] = add;
        rem &= ~(1u << add);
    }
    out->len = len;
    out->found_minimum = found_minimum ? 1 : 0;
    for (int k = 0; k < CUSK_SS_MAXP; k++) out->set[k] = k < len ? cand[k] : -1;
}
