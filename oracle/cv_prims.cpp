// TEST INFRASTRUCTURE (oracle) -- see cv_prims.h. Build with -ffp-contract=off: every float
// operation below is meant as written (OpenCV's baseline build has no FMA contraction).
#include "cv_prims.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace orb_oracle {

// ---------------------------------------------------------------------------------------------
// cvRound: OpenCV uses _mm_cvtsd_si32 / _mm_cvtss_si32 -> round to nearest, ties to even.
// ---------------------------------------------------------------------------------------------
int cv_round(double v) { return (int)std::nearbyint(v); }
int cv_round_f(float v) { return (int)std::nearbyintf(v); }

static inline int cv_floor_f(float v) {
  int i = (int)v;
  return i - (i > v);
}

static inline short saturate_short(int v) {
  return (short)std::min(std::max(v, -32768), 32767);
}

// ---------------------------------------------------------------------------------------------
// cv::resize INTER_LINEAR, CV_8UC1 (imgproc/src/resize.cpp: resizeGeneric_ with
// HResizeLinear<uchar,int,short,2048> and VResizeLinear<uchar,int,short,FixedPtCast<22>>).
//  - coefficient tables: 11-bit fixed point, computed in float from a double scale
//  - horizontal pass keeps int32 (8+11 bits), vertical pass drops 4 bits before the multiply
// ---------------------------------------------------------------------------------------------
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                      uint8_t* dst, int dw, int dh, int dstride) {
  const int COEF = 2048;
  const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
  const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;

  std::vector<int> xofs(dw), yofs(dh);
  std::vector<short> ialpha(2 * dw), ibeta(2 * dh);
  int xmax = dw;
  for (int dx = 0; dx < dw; dx++) {
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = cv_floor_f(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx + 1 >= sw) {
      xmax = std::min(xmax, dx);
      if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    }
    xofs[dx] = sx;
    ialpha[2 * dx] = saturate_short(cv_round_f((1.f - fx) * COEF));
    ialpha[2 * dx + 1] = saturate_short(cv_round_f(fx * COEF));
  }
  for (int dy = 0; dy < dh; dy++) {
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = cv_floor_f(fy);
    fy -= sy;
    yofs[dy] = sy;
    ibeta[2 * dy] = saturate_short(cv_round_f((1.f - fy) * COEF));
    ibeta[2 * dy + 1] = saturate_short(cv_round_f(fy * COEF));
  }

  std::vector<int> row0(dw), row1(dw);
  auto hresize = [&](int sy, std::vector<int>& out) {
    const uint8_t* S = src + (size_t)sy * sstride;
    int dx = 0;
    for (; dx < xmax; dx++) {
      int sx = xofs[dx];
      out[dx] = S[sx] * ialpha[2 * dx] + S[sx + 1] * ialpha[2 * dx + 1];
    }
    for (; dx < dw; dx++) out[dx] = S[xofs[dx]] * COEF;
  };
  auto clip = [](int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; };

  for (int dy = 0; dy < dh; dy++) {
    int sy0 = clip(yofs[dy], 0, sh), sy1 = clip(yofs[dy] + 1, 0, sh);
    hresize(sy0, row0);
    hresize(sy1, row1);
    const int b0 = ibeta[2 * dy], b1 = ibeta[2 * dy + 1];
    uint8_t* D = dst + (size_t)dy * dstride;
    for (int x = 0; x < dw; x++) {
      int v = (((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2;
      D[x] = (uint8_t)std::min(std::max(v, 0), 255);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// cv::GaussianBlur 7x7 sigma 2, CV_8U: OpenCV >= 3.4/4.x takes the fixed-point path
// (imgproc/src/smooth.dispatch.cpp, ufixedpoint16 kernel): the 8.8 fixed-point kernel for
// (ksize 7, sigma 2) is {18,34,48,56,48,34,18}/256; horizontal pass exact in 16 bits,
// vertical pass 16.16 with a single round-to-nearest at the end. BORDER_REFLECT_101.
// ---------------------------------------------------------------------------------------------
static inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
  static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
  std::vector<uint16_t> hbuf((size_t)w * h);
  for (int y = 0; y < h; y++) {
    const uint8_t* S = src + (size_t)y * sstride;
    for (int x = 0; x < w; x++) {
      int acc = 0;
      for (int k = 0; k < 7; k++) acc += K[k] * S[reflect101(x + k - 3, w)];
      hbuf[(size_t)y * w + x] = (uint16_t)acc;  // <= 255*256
    }
  }
  for (int y = 0; y < h; y++) {
    uint8_t* D = dst + (size_t)y * dstride;
    for (int x = 0; x < w; x++) {
      uint32_t acc = 0;
      for (int k = 0; k < 7; k++) acc += (uint32_t)K[k] * hbuf[(size_t)reflect101(y + k - 3, h) * w + x];
      D[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// cv::FAST TYPE_9_16 (features2d/src/fast.cpp FAST_t<16> + fast_score.cpp cornerScore<16>).
// Literal restatement: threshold table test, 25-entry circular run count, score as the largest
// threshold that keeps the pixel a corner, 3-row rolling buffers for the 3x3 strict-> NMS.
// ---------------------------------------------------------------------------------------------
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

static int corner_score16(const uint8_t* ptr, const int* pixel, int threshold) {
  const int K = 8, N = K * 3 + 1;
  int v = ptr[0];
  short d[N];
  for (int k = 0; k < N; k++) d[k] = (short)(v - ptr[pixel[k]]);

  int a0 = threshold;
  for (int k = 0; k < 16; k += 2) {
    int a = std::min((int)d[k + 1], (int)d[k + 2]);
    a = std::min(a, (int)d[k + 3]);
    if (a <= a0) continue;
    a = std::min(a, (int)d[k + 4]);
    a = std::min(a, (int)d[k + 5]);
    a = std::min(a, (int)d[k + 6]);
    a = std::min(a, (int)d[k + 7]);
    a = std::min(a, (int)d[k + 8]);
    a0 = std::max(a0, std::min(a, (int)d[k]));
    a0 = std::max(a0, std::min(a, (int)d[k + 9]));
  }
  int b0 = -a0;
  for (int k = 0; k < 16; k += 2) {
    int b = std::max((int)d[k + 1], (int)d[k + 2]);
    b = std::max(b, (int)d[k + 3]);
    b = std::max(b, (int)d[k + 4]);
    b = std::max(b, (int)d[k + 5]);
    if (b >= b0) continue;
    b = std::max(b, (int)d[k + 6]);
    b = std::max(b, (int)d[k + 7]);
    b = std::max(b, (int)d[k + 8]);
    b0 = std::min(b0, std::max(b, (int)d[k]));
    b0 = std::min(b0, std::max(b, (int)d[k + 9]));
  }
  return -b0 - 1;
}

void fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, bool nms,
              std::vector<FastPoint>& out) {
  out.clear();
  const int K = 8, N = 16 + K + 1;
  int pixel[25];
  for (int k = 0; k < 16; k++) pixel[k] = kRingDx[k] + kRingDy[k] * stride;
  for (int k = 16; k < 25; k++) pixel[k] = pixel[k - 16];
  threshold = std::min(std::max(threshold, 0), 255);

  uint8_t tab[512];
  for (int i = -255; i <= 255; i++) tab[i + 255] = (uint8_t)(i < -threshold ? 1 : i > threshold ? 2 : 0);

  if (w < 7 || h < 7) return;
  std::vector<uint8_t> bufmem((size_t)w * 3, 0);
  std::vector<int> cpmem((size_t)(w + 1) * 3, 0);
  uint8_t* buf[3] = {bufmem.data(), bufmem.data() + w, bufmem.data() + 2 * w};
  int* cpbuf[3] = {cpmem.data() + 1, cpmem.data() + (w + 1) + 1, cpmem.data() + 2 * (w + 1) + 1};

  for (int i = 3; i < h - 2; i++) {
    const uint8_t* ptr = img + (size_t)i * stride + 3;
    uint8_t* curr = buf[(i - 3) % 3];
    int* cornerpos = cpbuf[(i - 3) % 3];
    std::memset(curr, 0, w);
    int ncorners = 0;

    if (i < h - 3) {
      for (int j = 3; j < w - 3; j++, ptr++) {
        int v = ptr[0];
        const uint8_t* t = &tab[0] - v + 255;
        int d = t[ptr[pixel[0]]] | t[ptr[pixel[8]]];
        if (d == 0) continue;
        d &= t[ptr[pixel[2]]] | t[ptr[pixel[10]]];
        d &= t[ptr[pixel[4]]] | t[ptr[pixel[12]]];
        d &= t[ptr[pixel[6]]] | t[ptr[pixel[14]]];
        if (d == 0) continue;
        d &= t[ptr[pixel[1]]] | t[ptr[pixel[9]]];
        d &= t[ptr[pixel[3]]] | t[ptr[pixel[11]]];
        d &= t[ptr[pixel[5]]] | t[ptr[pixel[13]]];
        d &= t[ptr[pixel[7]]] | t[ptr[pixel[15]]];

        if (d & 1) {
          int vt = v - threshold, count = 0;
          for (int k = 0; k < N; k++) {
            int x = ptr[pixel[k]];
            if (x < vt) {
              if (++count > K) {
                cornerpos[ncorners++] = j;
                if (nms) curr[j] = (uint8_t)corner_score16(ptr, pixel, threshold);
                break;
              }
            } else {
              count = 0;
            }
          }
        }
        if (d & 2) {
          int vt = v + threshold, count = 0;
          for (int k = 0; k < N; k++) {
            int x = ptr[pixel[k]];
            if (x > vt) {
              if (++count > K) {
                cornerpos[ncorners++] = j;
                if (nms) curr[j] = (uint8_t)corner_score16(ptr, pixel, threshold);
                break;
              }
            } else {
              count = 0;
            }
          }
        }
      }
    }
    cornerpos[-1] = ncorners;
    if (i == 3) continue;

    const uint8_t* prev = buf[(i - 4 + 3) % 3];
    const uint8_t* pprev = buf[(i - 5 + 3) % 3];
    cornerpos = cpbuf[(i - 4 + 3) % 3];
    ncorners = cornerpos[-1];
    for (int k = 0; k < ncorners; k++) {
      int j = cornerpos[k];
      int score = prev[j];
      if (!nms || (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] &&
                   score > pprev[j] && score > pprev[j + 1] && score > curr[j - 1] &&
                   score > curr[j] && score > curr[j + 1])) {
        out.push_back({j, i - 1, score});
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// cv::fastAtan2 (core/src/mathfuncs_core.simd.hpp atan_f32): 7th-order odd polynomial in degrees.
// ---------------------------------------------------------------------------------------------
float fast_atan2(float y, float x) {
  static const float p1 = 0.9997878412794807f * (float)(180 / M_PI);
  static const float p3 = -0.3258083974640975f * (float)(180 / M_PI);
  static const float p5 = 0.1555786518463281f * (float)(180 / M_PI);
  static const float p7 = -0.04432655554792128f * (float)(180 / M_PI);
  float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + (float)DBL_EPSILON);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + (float)DBL_EPSILON);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ---------------------------------------------------------------------------------------------
// glibc 2.39 sincosf (sysdeps/ieee754/flt-32/s_sincosf.c, the ARM optimized-routines algorithm),
// as the x86_64 ifunc selects it on FMA hardware (__sincosf_fma): double-precision polynomials
// with every "a + b*c" fused. Only the two branches reachable for |y| < 120 are restated; the
// ORB path feeds angles in [0, 2*pi).
// ---------------------------------------------------------------------------------------------
namespace {
struct SinCosTab {
  double sign[4];
  double hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3;
};
const SinCosTab kTab[2] = {
    {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1p0,
     -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10, 0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13},
    {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, -0x1p0,
     0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10, -0x1.99343027bf8c3p-16,
     -0x1.555545995a603p-3, 0x1.1107605230bc4p-7, -0x1.994eb3774cf24p-13}};

inline void sincos_poly(double x, double x2, const SinCosTab& p, int n, float* sinp, float* cosp) {
  double x4 = x2 * x2, x3 = x2 * x;
  double c2 = std::fma(x2, p.c4, p.c3);
  double s1 = std::fma(x2, p.s3, p.s2);
  float* tmp = (n & 1) ? cosp : sinp;
  cosp = (n & 1) ? sinp : cosp;
  sinp = tmp;
  double c1 = std::fma(x2, p.c1, p.c0);
  double x5 = x3 * x2, x6 = x4 * x2;
  double s = std::fma(x3, p.s1, x);
  double c = std::fma(x4, p.c2, c1);
  *sinp = (float)std::fma(x5, s1, s);
  *cosp = (float)std::fma(x6, c2, c);
}
inline uint32_t abstop12(float x) {
  uint32_t u;
  std::memcpy(&u, &x, 4);
  return (u >> 20) & 0x7ff;
}
}  // namespace

void sincosf_port(float y, float* sinp, float* cosp) {
  double x = y;
  const SinCosTab* p = &kTab[0];
  if (abstop12(y) < abstop12(0x1.921FB6p-1f /* pi/4 */)) {
    double x2 = x * x;
    if (abstop12(y) < abstop12(0x1p-12f)) {
      *sinp = y;
      *cosp = 1.0f;
      return;
    }
    sincos_poly(x, x2, *p, 0, sinp, cosp);
  } else if (abstop12(y) < abstop12(120.0f)) {
    double r = x * p->hpi_inv;
    int n = ((int32_t)r + 0x800000) >> 24;
    x = std::fma(-(double)n, p->hpi, x);
    double s = p->sign[n & 3];
    if (n & 2) p = &kTab[1];
    sincos_poly(x * s, x * x, *p, n, sinp, cosp);
  } else {
    // Large-argument reduction is unreachable from the ORB path; defer to libm.
    ::sincosf(y, sinp, cosp);
  }
}

}  // namespace orb_oracle

// ---- cv::undistortPoints(src, dst, K, distCoeffs, noArray(), K) -- OpenCV 4.13 imgproc, cvUndistortPointsInternal ----
// Called by Frame::UndistortKeyPoints (R/src/Frame.cpp:470-504) and ComputeImageBounds (:506-540). Double arithmetic,
// x = (u - cx) * (1/fx), five iterations (default TermCriteria(COUNT, 5, 0.01): no EPS test), identity tilt and
// rectification, projection with P = K, one rounding to float. This file is compiled with -ffp-contract=off.
// PINNED: tests/test_oracle_prims.py against cv2 4.13.0 live and against tests/golden/undistort_cv2.npz.
namespace cvp {
void undistort_points(const float* xy, long long n, const float* K4, const float* dist, int ndist, float* out) {
  const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
  const double ifx = 1. / fx, ify = 1. / fy;
  double k[12];
  for (int i = 0; i < 12; i++) k[i] = i < ndist ? (double)dist[i] : 0.0;
  for (long long i = 0; i < n; i++) {
    double x = ((double)xy[2 * i] - cx) * ifx, y = ((double)xy[2 * i + 1] - cy) * ify;
    const double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      const double r2 = x * x + y * y;
      const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
      if (icdist < 0) { x = x0; y = y0; break; }
      const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
      const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    const double xx = fx * x + 0.0 * y + cx, yy = 0.0 * x + fy * y + cy, ww = 1. / (0.0 * x + 0.0 * y + 1.0);
    out[2 * i] = (float)(xx * ww);
    out[2 * i + 1] = (float)(yy * ww);
  }
}
}  // namespace cvp

extern "C" void orb_oracle_undistort_points(const float* xy, long long n, const float* K4, const float* dist, int ndist,
                                            float* out) {
  cvp::undistort_points(xy, n, K4, dist, ndist, out);
}
