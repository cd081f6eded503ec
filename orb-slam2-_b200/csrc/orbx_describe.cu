// Orientation + rBRIEF descriptors + final keypoint records (SURVEY.md 8a rows E7, E9, E1;
// R/src/ORBextractor.cpp IC_Angle :79-108, computeOrbDescriptor :113-155, operator() :1153-1186).
//
// One warp per kept keypoint.
//  * IC_Angle: the 31 rows of the circular patch are walked with the lanes across u (one coalesced
//    31-byte row segment per step) on the UNBLURRED level; m10/m01 are exact int32, reduced by
//    shuffles; angle = cv::fastAtan2 (orbx_math.cuh).
//  * Descriptor: lane b produces byte b (16 samples of the BLURRED level at the rotated pattern
//    positions); the rotation uses glibc-exact sincosf and the fused forms the reference's
//    -O3 -march=native build emits: row = fma(x,b,y*a), col = fma(x,a,-(y*b)), cvRound = rint.
//  * Output row = (rows of lower levels) + list position: level-major like the reference's concat.
#include "orbx_common.cuh"
#include "orbx_math.cuh"

namespace orbx {

__constant__ int8_t cPattern[1024] = {
#include "orb_pattern.inc"
};
// umax of the 15-radius patch (R/src/ORBextractor.cpp:480-500), a function of HALF_PATCH_SIZE only
__constant__ int cUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__device__ __forceinline__ float ic_angle_warp(const uint8_t* __restrict__ center, int pitch, int lane) {
  // lane -> u = lane-15 (lane 31 idle)
  const int u = lane - kHalfPatch;
  int m10 = 0, m01 = 0;
  if (lane < 31) {
#pragma unroll 1
    for (int v = -kHalfPatch; v <= kHalfPatch; v++) {
      const int d = cUmax[v < 0 ? -v : v];
      if (u >= -d && u <= d) {
        const int val = center[(long long)v * pitch + u];
        m10 += u * val;
        m01 += v * val;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m10 += __shfl_xor_sync(0xffffffffu, m10, o);
    m01 += __shfl_xor_sync(0xffffffffu, m01, o);
  }
  return fast_atan2_deg((float)m01, (float)m10);
}

__device__ __forceinline__ uint8_t brief_byte(const uint8_t* __restrict__ center, int pitch, float a, float b,
                                              int byteIdx) {
  const int8_t* pat = &cPattern[byteIdx * 32];
  int val = 0;
#pragma unroll
  for (int bit = 0; bit < 8; bit++) {
    int t[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const float px = (float)pat[4 * bit + 2 * k], py = (float)pat[4 * bit + 2 * k + 1];
      const int r = __float2int_rn(__fmaf_rn(px, b, __fmul_rn(py, a)));
      const int c = __float2int_rn(__fmaf_rn(px, a, -__fmul_rn(py, b)));
      t[k] = center[(long long)r * pitch + c];
    }
    val |= (t[0] < t[1]) << bit;
  }
  return (uint8_t)val;
}

__global__ void __launch_bounds__(256)
describe_kernel(PyramidPtrs pyr, const uint8_t* __restrict__ blurArena, const FrameSetGeom* __restrict__ g,
                const uint32_t* __restrict__ kept, const int* __restrict__ keptCount,
                orbx_keypoint* __restrict__ outKps, uint8_t* __restrict__ outDesc, int cap,
                int* __restrict__ outN, int keptPerFrameMax) {
  const int f = blockIdx.y;
  const int nl = g->nlevels;
  const int lane = threadIdx.x & 31;
  const int warpInFrame = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  // locate (level, position) of this warp's keypoint among the frame's kept lists
  int level = 0, base = 0, total = 0;
  int myLevel = -1, myPos = 0, myBase = 0;
  for (level = 0; level < nl; level++) {
    const int c = max(0, min(keptCount[f * nl + level], g->lv[level].keptCap));
    if (myLevel < 0 && warpInFrame < total + c) { myLevel = level; myPos = warpInFrame - total; myBase = total; }
    total += c;
  }
  (void)base; (void)myBase;
  if (blockIdx.x == 0 && threadIdx.x == 0) outN[f] = total;
  if (myLevel < 0) return;
  const int row = warpInFrame;  // level-major concatenation
  if (row >= cap) return;

  const LevelGeom& L = g->lv[myLevel];
  const uint32_t k = kept[L.keptOff + (long long)f * L.keptCap + myPos];
  const int x = cand_x(k) + kMinBorder, y = cand_y(k) + kMinBorder;  // level coordinates (:909-910)
  const int pitch = pyr.pitch[myLevel];
  const uint8_t* plane = pyr.lvl[myLevel] + (long long)f * pyr.frameStride[myLevel];
  const float angle = ic_angle_warp(plane + (long long)y * pitch + x, pitch, lane);

  const float factorPI = 0x1.1df46ap-6f;  // (float)(CV_PI/180.f)
  float sn, cs;
  sincosf_glibc(__fmul_rn(angle, factorPI), &sn, &cs);
  const uint8_t* bplane = blurArena + L.planeOff + (long long)f * L.planeBytes;
  const uint8_t d = brief_byte(bplane + (long long)y * L.pitch + x, L.pitch, cs, sn, lane);
  outDesc[((long long)f * cap + row) * 32 + lane] = d;

  if (lane == 0) {
    orbx_keypoint kp;
    const float fx = (float)x, fy = (float)y;
    kp.x = myLevel ? __fmul_rn(fx, L.scale) : fx;  // pt *= scale for level != 0 (:1178-1184)
    kp.y = myLevel ? __fmul_rn(fy, L.scale) : fy;
    kp.size = L.patchSize;
    kp.angle = angle;
    kp.response = (float)cand_score(k);
    kp.octave = myLevel;
    kp.class_id = -1;
    outKps[(long long)f * cap + row] = kp;
  }
}

void launch_describe(const PyramidPtrs& pyr, const uint8_t* blurArena, const FrameSetGeom* dGeom,
                     const FrameSetGeom& hGeom, const uint32_t* kept, const int* keptCount,
                     orbx_keypoint* outKps, uint8_t* outDesc, int cap, int* outN, int nframes,
                     cudaStream_t s) {
  const int warpsPerBlock = 8;
  const int maxKp = hGeom.keptPerFrame;
  dim3 grid((maxKp + warpsPerBlock - 1) / warpsPerBlock, nframes);
  describe_kernel<<<grid, warpsPerBlock * 32, 0, s>>>(pyr, blurArena, dGeom, kept, keptCount, outKps, outDesc, cap,
                                                      outN, maxKp);
}

// ---- array versions of the float helpers, for the unit parity tests ----------------------------
__global__ void atan2_kernel(const float* y, const float* x, float* out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = fast_atan2_deg(y[i], x[i]);
}
__global__ void sincosf_kernel(const float* a, float* s, float* c, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sincosf_glibc(a[i], &s[i], &c[i]);
}
__global__ void sincosf_sweep_kernel(uint32_t lo, long long n, float* s, float* c) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) sincosf_glibc(__uint_as_float(lo + (uint32_t)i), &s[i], &c[i]);
}
void launch_atan2(const float* y, const float* x, float* out, long long n, cudaStream_t s) {
  if (n > 0) atan2_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(y, x, out, n);
}
void launch_sincosf(const float* a, float* s_out, float* c_out, long long n, cudaStream_t s) {
  if (n > 0) sincosf_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(a, s_out, c_out, n);
}
void launch_sincosf_sweep(uint32_t lo, uint32_t hi, float* s_out, float* c_out, cudaStream_t s) {
  const long long n = (long long)hi - lo;
  if (n > 0) sincosf_sweep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(lo, n, s_out, c_out);
}

}  // namespace orbx
