// oracle/qt_shim/Accelerate/Accelerate.h — TEST INFRASTRUCTURE ONLY.
// Linux stand-in for the handful of Apple vDSP names pebblelib/decimator.{h,cpp} and
// fftaccelerate.h mention, as plain scalar loops following Apple's documented semantics (the
// pseudo-code the reference itself quotes at decimator.cpp:637-647). Lets decimator.cpp compile
// unmodified; FFTAccelerate is never instantiated (USE_FFTW).
#pragma once
#include <cstddef>
typedef unsigned long vDSP_Length;
typedef long vDSP_Stride;
typedef struct { double real, imag; } DSPDoubleComplex;
typedef struct { double *realp, *imagp; } DSPDoubleSplitComplex;
typedef struct OpaqueFFTSetupD *FFTSetupD;
typedef int FFTRadix;
typedef int FFTDirection;
enum { kFFTRadix2 = 0, kFFTDirection_Forward = 1, kFFTDirection_Inverse = -1 };

// interleaved -> split. Strides are in units of double for C (so 2 = consecutive complex).
static inline void vDSP_ctozD(const DSPDoubleComplex *C, vDSP_Stride IC, const DSPDoubleSplitComplex *Z,
                              vDSP_Stride IZ, vDSP_Length N) {
    const double *c = (const double *)C;
    for (vDSP_Length n = 0; n < N; ++n) {
        Z->realp[n * IZ] = c[n * IC];
        Z->imagp[n * IZ] = c[n * IC + 1];
    }
}
// split -> interleaved
static inline void vDSP_ztocD(const DSPDoubleSplitComplex *Z, vDSP_Stride IZ, DSPDoubleComplex *C,
                              vDSP_Stride IC, vDSP_Length N) {
    double *c = (double *)C;
    for (vDSP_Length n = 0; n < N; ++n) {
        c[n * IC] = Z->realp[n * IZ];
        c[n * IC + 1] = Z->imagp[n * IZ];
    }
}
static inline void vDSP_zvmovD(const DSPDoubleSplitComplex *A, vDSP_Stride IA, const DSPDoubleSplitComplex *C,
                               vDSP_Stride IC, vDSP_Length N) {
    for (vDSP_Length n = 0; n < N; ++n) {
        C->realp[n * IC] = A->realp[n * IA];
        C->imagp[n * IC] = A->imagp[n * IA];
    }
}
// C[n] = sum_p A[n*DF + p] * F[p]   (complex A, real F)
static inline void vDSP_zrdesampD(const DSPDoubleSplitComplex *A, vDSP_Stride DF, const double *F,
                                  const DSPDoubleSplitComplex *C, vDSP_Length N, vDSP_Length P) {
    for (vDSP_Length n = 0; n < N; ++n) {
        double sr = 0, si = 0;
        for (vDSP_Length p = 0; p < P; ++p) {
            sr += A->realp[n * DF + p] * F[p];
            si += A->imagp[n * DF + p] * F[p];
        }
        C->realp[n] = sr;
        C->imagp[n] = si;
    }
}
// C[0] = sum_n A[n*IA] * B[n*IB]   (complex A, real B)
static inline void vDSP_zrdotprD(const DSPDoubleSplitComplex *A, vDSP_Stride IA, const double *B, vDSP_Stride IB,
                                 const DSPDoubleSplitComplex *C, vDSP_Length N) {
    double sr = 0, si = 0;
    for (vDSP_Length n = 0; n < N; ++n) {
        sr += A->realp[n * IA] * B[n * IB];
        si += A->imagp[n * IA] * B[n * IB];
    }
    C->realp[0] = sr;
    C->imagp[0] = si;
}
