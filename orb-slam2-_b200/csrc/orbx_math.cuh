// Bit-exact float helpers for the orientation / descriptor stage.
// Every operation is an explicit round-to-nearest intrinsic so that neither nvcc's default
// -fmad=true nor any re-association can change a result.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbx {

// cv::fastAtan2 (OpenCV core, atan_f32): degrees in [0,360); UNFUSED float polynomial. Constants
// are 0.9997878412794807f*(float)(180/pi) etc. evaluated in float.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  const float p1 = 0x1.ca44dep+5f, p3 = -0x1.2aaddcp+4f, p5 = 0x1.1d3f7ep+3f, p7 = -0x1.4515b2p+1f;
  const float eps = 0x1p-52f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a;
  if (ax >= ay) {
    const float c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    const float c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    const float c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    const float c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f,
                  __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0.f) a = __fsub_rn(180.f, a);
  if (y < 0.f) a = __fsub_rn(360.f, a);
  return a;
}

// glibc 2.39 sincosf as the x86-64 ifunc resolves it on FMA hardware (__sincosf_fma; the ARM
// optimized-routines algorithm with every a+b*c fused): double polynomials, quadrant reduction by
// a 2^24-prescaled 2/pi. Branches for |y| >= 120 are unreachable from the ORB path (angles are in
// [0, 2*pi)) and fall back to the small-angle formula's domain clamp below.
__device__ __forceinline__ void sincosf_glibc(float y, float* sinp, float* cosp) {
  const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5,
               C3 = -0x1.6c087e89a359dp-10, C4 = 0x1.99343027bf8c3p-16;
  const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
  const double HPI_INV = 0x1.45F306DC9C883p+23, HPI = 0x1.921FB54442D18p0;
  const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
  double x = (double)y;
  int n = 0;
  double sgn = 1.0, flip = 1.0;  // flip = -1 selects the second table (all cosine coefficients negated)
  if (top < 0x3f4) {             // |y| < pi/4
    if (top < 0x398) {           // |y| < 2^-12
      *sinp = y;
      *cosp = 1.0f;
      return;
    }
  } else {
    const double r = __dmul_rn(x, HPI_INV);
    n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, HPI, x);
    sgn = ((n + 1) & 2) ? -1.0 : 1.0;  // sign table {1,-1,-1,1}[n&3]
    flip = (n & 2) ? -1.0 : 1.0;
  }
  const double x2 = __dmul_rn(x, x);  // (x*s)^2 is never formed: glibc passes x*x
  const double xs = __dmul_rn(x, sgn);
  const double x4 = __dmul_rn(x2, x2), x3 = __dmul_rn(x2, xs);
  const double c2 = __fma_rn(x2, flip * C4, flip * C3);
  const double s1 = __fma_rn(x2, S3, S2);
  const double c1 = __fma_rn(x2, flip * C1, flip * C0);
  const double x5 = __dmul_rn(x3, x2), x6 = __dmul_rn(x4, x2);
  const double s = __fma_rn(x3, S1, xs);
  const double c = __fma_rn(x4, flip * C2, c1);
  const float sv = __double2float_rn(__fma_rn(x5, s1, s));
  const float cv = __double2float_rn(__fma_rn(x6, c2, c));
  if (n & 1) { *sinp = cv; *cosp = sv; }
  else       { *sinp = sv; *cosp = cv; }
}

}  // namespace orbx
