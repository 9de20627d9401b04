/* csim_norm2.h — TEST INFRASTRUCTURE (oracle). Not product code.
 *
 * The one place where the reference path does arithmetic inside a third-party
 * library that is absent from /root/reference: Eigen's VectorXd::norm()
 * (call sites: src/dcanalysis.cpp:276, src/tanalisis.cpp:366,
 * include/solver.hpp:187).  Eigen's summation order is SIMD/version dependent
 * and cannot be checked on this box ("parity unpinned" at this call), so the
 * order is DEFINED here, once, and used by the Eigen stand-in (and therefore by
 * the unmodified reference build in oracle/_ref), by the C restatement and —
 * re-implemented, not included — by the CUDA kernels:
 *
 *   acc[l] = sum over m (ascending) of v[l + 32*m]^2 ,  l = 0..31   (+0 padded)
 *   for s in 16,8,4,2,1:  acc[i] += acc[i+s]  for i < s              (halving tree)
 *   norm = sqrt(acc[0])
 *
 * For n <= 32 this is a plain 32-leaf pairwise tree, i.e. exactly the
 * xor-butterfly a warp does with shuffles.
 */
#ifndef CSIM_NORM2_H
#define CSIM_NORM2_H
#include <math.h>
static inline double csim_norm2(const double* v, long n) {
    double acc[32];
    for (int l = 0; l < 32; ++l) acc[l] = 0.0;
    for (long i = 0; i < n; ++i) {
        double sq = v[i] * v[i];
        acc[i & 31] = acc[i & 31] + sq;
    }
    for (int s = 16; s >= 1; s >>= 1)
        for (int i = 0; i < s; ++i) acc[i] = acc[i] + acc[i + s];
    return sqrt(acc[0]);
}
#endif
