// rtp_math.h - FP32 arithmetic with the evaluation order the reference gets from Eigen 3.3.7.
//
// Every hit/miss, tie and shading decision of the reference (src/flyscene.cpp:444-614) is a
// chain of IEEE single-precision +,-,*,/,sqrt with no FMA (the reference builds for baseline
// x86-64).  On the device those chains are written with __fmul_rn/__fadd_rn/... which ptxas
// never contracts into FFMA, so results are bit-identical to the CPU reference regardless of
// -fmad.  The same header compiles as host code (g++ -ffp-contract=off) for the flattener.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RTP_HD __host__ __device__ __forceinline__
#else
#define RTP_HD inline
#endif

namespace rtp {

#if defined(__CUDA_ARCH__)
RTP_HD float fmul(float a, float b) { return __fmul_rn(a, b); }
RTP_HD float fadd(float a, float b) { return __fadd_rn(a, b); }
RTP_HD float fsub(float a, float b) { return __fsub_rn(a, b); }
RTP_HD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
RTP_HD float fsqrt(float a) { return __fsqrt_rn(a); }
#else
RTP_HD float fmul(float a, float b) { return a * b; }
RTP_HD float fadd(float a, float b) { return a + b; }
RTP_HD float fsub(float a, float b) { return a - b; }
RTP_HD float fdiv(float a, float b) { return a / b; }
RTP_HD float fsqrt(float a) { return sqrtf(a); }
#endif

struct v3 {
  float x, y, z;
};
RTP_HD v3 mk3(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
RTP_HD v3 vadd(v3 a, v3 b) { return mk3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
RTP_HD v3 vsub(v3 a, v3 b) { return mk3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
RTP_HD v3 vmuls(v3 a, float s) { return mk3(fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)); }
RTP_HD v3 vdivs(v3 a, float s) { return mk3(fdiv(a.x, s), fdiv(a.y, s), fdiv(a.z, s)); }
RTP_HD v3 vneg(v3 a) { return mk3(-a.x, -a.y, -a.z); }
// 3-term reduction a0 + (a1 + a2): Eigen/src/Core/Redux.h:91-105
RTP_HD float vdot(v3 a, v3 b) { return fadd(fmul(a.x, b.x), fadd(fmul(a.y, b.y), fmul(a.z, b.z))); }
// Eigen/src/Geometry/OrthoMethods.h:43-47
RTP_HD v3 vcross(v3 a, v3 b) {
  return mk3(fsub(fmul(a.y, b.z), fmul(a.z, b.y)), fsub(fmul(a.z, b.x), fmul(a.x, b.z)),
             fsub(fmul(a.x, b.y), fmul(a.y, b.x)));
}
RTP_HD float vnorm(v3 a) { return fsqrt(vdot(a, a)); }
// Eigen/src/Core/Dot.h:124-134
RTP_HD v3 vnormalized(v3 a) {
  float z = vdot(a, a);
  if (z > 0.f) return vdivs(a, fsqrt(z));
  return a;
}
// normalized() of a dynamic-size expression (vertices[i].head(3) - ...): sequential redux
RTP_HD v3 vnormalized_dyn(v3 a) {
  float z = fadd(fadd(fmul(a.x, a.x), fmul(a.y, a.y)), fmul(a.z, a.z));
  if (z > 0.f) return vdivs(a, fsqrt(z));
  return a;
}
// Affine3f * Vector3f: ((col0*x + col1*y) + col2*z) + col3*1 (Eigen/src/Geometry/Transform.h:
// 1372-1392 through the SSE 4x4 * 4x1 product). m: 3x4 row-major.
RTP_HD v3 affine_mul(const float *m, v3 p) {
  return mk3(fadd(fadd(fadd(fmul(m[0], p.x), fmul(m[1], p.y)), fmul(m[2], p.z)), fmul(m[3], 1.0f)),
             fadd(fadd(fadd(fmul(m[4], p.x), fmul(m[5], p.y)), fmul(m[6], p.z)), fmul(m[7], 1.0f)),
             fadd(fadd(fadd(fmul(m[8], p.x), fmul(m[9], p.y)), fmul(m[10], p.z)), fmul(m[11], 1.0f)));
}
// Matrix3f * Vector3f (coefficient-wise lazy product, 3-term redux); m: 3x4 row-major, 3x3 block
RTP_HD v3 linear_mul(const float *m, v3 p) {
  return mk3(fadd(fmul(m[0], p.x), fadd(fmul(m[1], p.y), fmul(m[2], p.z))),
             fadd(fmul(m[4], p.x), fadd(fmul(m[5], p.y), fmul(m[6], p.z))),
             fadd(fmul(m[8], p.x), fadd(fmul(m[9], p.y), fmul(m[10], p.z))));
}
RTP_HD float std_min(float a, float b) { return (b < a) ? b : a; }  // std::min keeps a NaN 1st arg
RTP_HD float std_max(float a, float b) { return (a < b) ? b : a; }  // std::max keeps a NaN 1st arg
RTP_HD float clamp01(float c) { return std_max(std_min(c, 1.f), 0.f); }  // flyscene.cpp:367-369,613

RTP_HD int f2i(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_int(f);
#else
  union { float f; int i; } u; u.f = f; return u.i;
#endif
}
RTP_HD float i2f(int i) {
#if defined(__CUDA_ARCH__)
  return __int_as_float(i);
#else
  union { float f; int i; } u; u.i = i; return u.f;
#endif
}

// Host-only helpers (flatten time): Eigen's cofactor 3x3 inverse and Affine3f::inverse()
// (Eigen/src/LU/InverseImpl.h:120-170, Eigen/src/Geometry/Transform.h:1215-1226).
inline float cof3(const float *a, int lda, int i, int j) {
  int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return a[i1 * lda + j1] * a[i2 * lda + j2] - a[i1 * lda + j2] * a[i2 * lda + j1];
}
inline void inverse3(const float *a, int lda, float *inv, int ldi) {
  float c0 = cof3(a, lda, 0, 0), c1 = cof3(a, lda, 1, 0), c2 = cof3(a, lda, 2, 0);
  float det = c0 * a[0] + (c1 * a[lda] + c2 * a[2 * lda]);
  float invdet = 1.0f / det;
  inv[0] = c0 * invdet; inv[1] = c1 * invdet; inv[2] = c2 * invdet;
  inv[ldi + 0] = cof3(a, lda, 0, 1) * invdet;
  inv[ldi + 1] = cof3(a, lda, 1, 1) * invdet;
  inv[ldi + 2] = cof3(a, lda, 2, 1) * invdet;
  inv[2 * ldi + 0] = cof3(a, lda, 0, 2) * invdet;
  inv[2 * ldi + 1] = cof3(a, lda, 1, 2) * invdet;
  inv[2 * ldi + 2] = cof3(a, lda, 2, 2) * invdet;
}
inline void affine_inverse(const float *m, float *inv) {
  inverse3(m, 4, inv, 4);
  float n[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) n[r * 3 + c] = -inv[r * 4 + c];
  for (int r = 0; r < 3; ++r)
    inv[r * 4 + 3] = n[r * 3] * m[3] + (n[r * 3 + 1] * m[7] + n[r * 3 + 2] * m[11]);
}

}  // namespace rtp
