/*
 * oracle/cblas_shim/cblas.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Header-only stand-in for the three CBLAS level-1 routines the reference's
 * CPU path calls, so that /root/reference/seq.cpp and seq_test.cpp compile
 * unmodified in an image with no CBLAS/BLAS installed:
 *     cblas_saxpy   seq.cpp:406, seq_test.cpp (rbf_kernel), svmTrain.cu:705
 *     cblas_snrm2   seq.cpp:408, seq_test.cpp (rbf_kernel)
 *     cblas_sdot    svmTrain.cu:707
 *
 * The reference links an unversioned netlib CBLAS+BLAS tarball (README.md:17-18,
 * Makefile:45-46) that is NOT vendored under /root/reference, and the reference
 * holds no test that pins results at that boundary ("parity unpinned" there).
 * This shim restates the published netlib reference-BLAS algorithms
 * (LAPACK <= 3.9 era, the versions contemporary with the reference):
 *     SAXPY: y[i] += a*x[i]; element-wise, order irrelevant.
 *     SDOT : single fp32 accumulator, elements added in index order (netlib's
 *            unroll-by-5 evaluates left to right, i.e. the same sequence).
 *     SNRM2: the scale / sum-of-squares recurrence (scale=0, ssq=1), fp32.
 * Every golden vector generated through this shim says so in its provenance.
 */
#ifndef DPSVM_ORACLE_CBLAS_SHIM_H
#define DPSVM_ORACLE_CBLAS_SHIM_H

#include <math.h>

#ifdef __cplusplus
extern "C" {
#endif

static inline void cblas_saxpy(const int n, const float a, const float *x, const int incx,
                               float *y, const int incy)
{
    if (n <= 0 || a == 0.0f) return;
    int ix = incx < 0 ? (1 - n) * incx : 0;
    int iy = incy < 0 ? (1 - n) * incy : 0;
    for (int i = 0; i < n; ++i, ix += incx, iy += incy) y[iy] = y[iy] + a * x[ix];
}

static inline float cblas_sdot(const int n, const float *x, const int incx,
                               const float *y, const int incy)
{
    float acc = 0.0f;
    if (n <= 0) return acc;
    int ix = incx < 0 ? (1 - n) * incx : 0;
    int iy = incy < 0 ? (1 - n) * incy : 0;
    for (int i = 0; i < n; ++i, ix += incx, iy += incy) acc = acc + x[ix] * y[iy];
    return acc;
}

static inline float cblas_snrm2(const int n, const float *x, const int incx)
{
    if (n < 1 || incx < 1) return 0.0f;
    if (n == 1) return fabsf(x[0]);
    float scale = 0.0f, ssq = 1.0f;
    for (int i = 0, ix = 0; i < n; ++i, ix += incx) {
        if (x[ix] != 0.0f) {
            const float a = fabsf(x[ix]);
            if (scale < a) {
                const float r = scale / a;
                ssq = 1.0f + ssq * r * r;
                scale = a;
            } else {
                const float r = a / scale;
                ssq = ssq + r * r;
            }
        }
    }
    return scale * sqrtf(ssq);
}

#ifdef __cplusplus
}
#endif
#endif
