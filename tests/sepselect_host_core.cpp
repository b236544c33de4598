// TEST INFRASTRUCTURE, never part of the product: compiles the device routine of the separating-set search
// (ci-gwas_b200/csrc/sepselect_core.h) for the HOST and exports it under the name of the library's GPU entry point.
// tests/test_sepselect_cpu.py preloads this object (LD_PRELOAD) in a child process so that the host side of
// `sepselect` (file parsing, triples, orientation, output) and the arithmetic of the device routine can be checked
// against the reference's golden outputs in a container without a GPU.  The product library itself has no CPU path:
// cusk_sepselect_pairs in libcusk_b200.so launches sepselect_pairs_kernel.
#include "../ci-gwas_b200/csrc/sepselect_core.h"

extern "C" int cusk_sepselect_pairs(cusk_ctx *, const double *Ct, int64_t ld, int P, int, const int *pairs, int64_t npairs,
                                    const uint32_t *tmask, const double *thr, cusk_sepselect_result *out)
{
    for (int64_t p = 0; p < npairs; p++)
        cusk_sepselect_pair(Ct, ld, P, pairs[2 * p], pairs[2 * p + 1], tmask[pairs[2 * p]], thr, &out[p]);
    return 0;
}
