// Frame epilogue on the device: Frame::UndistortKeyPoints (R/src/Frame.cpp:470-504), i.e.
// cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) on the keypoint coordinates.
//
// OpenCV 4.13 (imgproc undistort, cvUndistortPointsInternal) works in double: x = (u-cx)/fx through the reciprocal,
// five fixed-point iterations of the radial + tangential model (the default TermCriteria(COUNT, 5, 0.01) has no EPS
// part), then the projection with P = K, and one rounding to float at the end. Restated with explicit round-to-nearest
// operations and no contraction; bit-exact against cv2 4.13 (tests/test_oracle_prims.py pins the oracle's restatement,
// tests/test_gpu_frame.py this kernel against the oracle). 1000 keypoints x ~150 FP64 operations: the kernel is
// launch-latency, the point is that the keypoints stay in HBM between extraction and matching.
#include "orbx_common.cuh"

namespace orbx {

struct UndistortParams {
  double fx, fy, cx, cy, ifx, ify;
  double k[12];  // k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4 (OpenCV order); the reference passes 4 or 5 coefficients
};

__device__ __forceinline__ void undistort_point(float u, float v, const UndistortParams& p, float* ox, float* oy) {
  const double* k = p.k;
  double x = __dmul_rn(__dsub_rn((double)u, p.cx), p.ifx);
  double y = __dmul_rn(__dsub_rn((double)v, p.cy), p.ify);
  const double x0 = x, y0 = y;
#pragma unroll 1
  for (int j = 0; j < 5; j++) {
    const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
    const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[7], r2), k[6]), r2), k[5]), r2));
    const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[4], r2), k[1]), r2), k[0]), r2));
    const double icdist = __ddiv_rn(num, den);
    if (icdist < 0) { x = x0; y = y0; break; }   // "(x - cx)*ifx" again in OpenCV: the same value
    // deltaX = 2*k[2]*x*y + k[3]*(r2 + 2*x*x) + k[8]*r2 + k[9]*r2*r2, left to right
    const double dX = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, k[2]), x), y),
                                                    __dmul_rn(k[3], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x)))),
                                          __dmul_rn(k[8], r2)),
                                __dmul_rn(__dmul_rn(k[9], r2), r2));
    // deltaY = k[2]*(r2 + 2*y*y) + 2*k[3]*x*y + k[10]*r2 + k[11]*r2*r2
    const double dY = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(k[2], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                                    __dmul_rn(__dmul_rn(__dmul_rn(2.0, k[3]), x), y)),
                                          __dmul_rn(k[10], r2)),
                                __dmul_rn(__dmul_rn(k[11], r2), r2));
    x = __dmul_rn(__dsub_rn(x0, dX), icdist);
    y = __dmul_rn(__dsub_rn(y0, dY), icdist);
  }
  // RR = P * I: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1/(0*x + 0*y + 1)
  const double xx = __dadd_rn(__dadd_rn(__dmul_rn(p.fx, x), __dmul_rn(0.0, y)), p.cx);
  const double yy = __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(p.fy, y)), p.cy);
  const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(__dmul_rn(0.0, x), __dmul_rn(0.0, y)), 1.0));
  *ox = (float)__dmul_rn(xx, ww);
  *oy = (float)__dmul_rn(yy, ww);
}

// mvKeysUn[i] = mvKeys[i] with pt replaced (R/src/Frame.cpp:497-503)
__global__ void undistort_keys_kernel(const orbx_keypoint* __restrict__ in, orbx_keypoint* __restrict__ out, int n,
                                      const __grid_constant__ UndistortParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  orbx_keypoint kp = in[i];
  undistort_point(kp.x, kp.y, p, &kp.x, &kp.y);
  out[i] = kp;
}

__global__ void undistort_points_kernel(const float2* __restrict__ in, float2* __restrict__ out, long long n,
                                        const __grid_constant__ UndistortParams p) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 q = in[i];
  float2 r;
  undistort_point(q.x, q.y, p, &r.x, &r.y);
  out[i] = r;
}

static UndistortParams make_params(const float* K4, const float* dist, int ndist) {
  UndistortParams p;
  p.fx = (double)K4[0]; p.fy = (double)K4[1]; p.cx = (double)K4[2]; p.cy = (double)K4[3];
  p.ifx = 1.0 / p.fx; p.ify = 1.0 / p.fy;
  for (int i = 0; i < 12; i++) p.k[i] = i < ndist ? (double)dist[i] : 0.0;
  return p;
}

void launch_undistort_keys(const orbx_keypoint* in, orbx_keypoint* out, int n, const float* K4, const float* dist, int ndist,
                           cudaStream_t s) {
  if (n <= 0) return;
  undistort_keys_kernel<<<(n + 127) / 128, 128, 0, s>>>(in, out, n, make_params(K4, dist, ndist));
}

void launch_undistort_points(const float* in, float* out, long long n, const float* K4, const float* dist, int ndist,
                             cudaStream_t s) {
  if (n <= 0) return;
  undistort_points_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(reinterpret_cast<const float2*>(in),
                                                                      reinterpret_cast<float2*>(out), n,
                                                                      make_params(K4, dist, ndist));
}

}  // namespace orbx
