// Orientation + rBRIEF descriptors + final keypoint records (SURVEY.md 8a rows E7, E9, E1;
// R/src/ORBextractor.cpp IC_Angle :79-108, computeOrbDescriptor :113-155, operator() :1153-1186).
//
// One warp per kept keypoint, several keypoints per warp (v2).
//  * IC_Angle on the UNBLURRED level: the 31x31 patch is cut into 31 rows x 8 chunks of 4 pixels;
//    lane = chunk + 8*(row mod 4). A chunk is fetched as two aligned words realigned by a funnel
//    shift, and the moments come from two IDP.4A per chunk against per-(row,chunk) weight words
//    (u inside the disc, else 0 / 1 inside the disc, else 0) kept in shared memory:
//    m10 += dot(px, u), m01 += v * dot(px, 1). Exact int32, reduced by shuffles;
//    angle = cv::fastAtan2 (orbx_math.cuh).
//  * Descriptor on the BLURRED level: the 39x39 neighbourhood the rotated pattern can reach (|x|,|y| <= 13 =>
//    radius <= 18.4) is first staged into shared memory with coalesced 8-byte loads -- 5 rows x 6 vectors per
//    warp instruction, 8 instructions per keypoint -- and the 512 samples are gathered from there. Gathering them
//    from global memory made the kernel L1-bound (ncu r1_x: l1tex 86 %, every warp-wide byte gather touches up
//    to 32 sectors); from shared memory a gather costs its bank conflicts only.
//    Lane b produces byte b (16 samples). The pattern sits in
//    shared memory as float4 (x0,y0,x1,y1) per (bit, lane) -- conflict-free LDS.128 instead of
//    the 32-way divergent constant loads of v1. Rotation uses glibc-exact sincosf and the fused forms
//    the reference's -O3 -march=native build emits: row = fma(x,b,y*a), col = fma(x,a,-(y*b));
//    cvRound = round-half-even done as (t + 1.5*2^23) - 1.5*2^23 on the FADD pipe instead of F2I.
//  * Output row = (rows of lower levels) + list position: level-major like the reference's concat.
#include "orbx_common.cuh"
#include "orbx_math.cuh"

namespace orbx {

__constant__ int8_t cPattern[1024] = {
#include "orb_pattern.inc"
};
// umax of the 15-radius patch (R/src/ORBextractor.cpp:480-500), a function of HALF_PATCH_SIZE only
__constant__ int cUmax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

constexpr int kDescWarps = 8;
constexpr int kDescKpPerWarp = 4;      // batches: the CTA prologue (pattern + weights into shared memory) is shared by 32 keypoints
constexpr int kDescKpPerWarpSmall = 1; // a few frames: four times the CTAs, each warp done after one keypoint (latency)
constexpr int kPatchR = 19;                      // EDGE_THRESHOLD: reach of the rotated pattern (R/src/ORBextractor.cpp:76)
constexpr int kPatchRows = 2 * kPatchR + 1;      // 39
constexpr int kPatchVecs = 6;                    // 8-byte vectors per staged row (39 bytes + up to 7 of misalignment)
constexpr int kPatchPitch = kPatchVecs * 8;      // 48 bytes
constexpr int kPatchRowsPerPass = 32 / kPatchVecs;                                        // 5
constexpr int kPatchPasses = (kPatchRows + kPatchRowsPerPass - 1) / kPatchRowsPerPass;    // 8

__device__ __forceinline__ int round_half_even_small(float t) {  // |t| < 2^22
  return __float_as_int(__fadd_rn(t, 12582912.f)) - 0x4B400000;
}

__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {
  int r;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}

template <int KPW>
__global__ void __launch_bounds__(kDescWarps * 32)
describe_kernel(const __grid_constant__ PyramidPtrs pyr, const uint8_t* __restrict__ blurArena,
                const __grid_constant__ FrameSetGeom g, const uint32_t* __restrict__ kept,
                const int* __restrict__ keptCount, const int* __restrict__ candCount,
                orbx_keypoint* __restrict__ outKps, uint8_t* __restrict__ outDesc, int cap, int* __restrict__ outN) {
  pdl_wait();
  __shared__ float4 sPat[8 * 32];        // [bit][lane] = (x0, y0, x1, y1) of pattern pair lane*8+bit
  __shared__ uint32_t sWU[32 * 8];       // [row v+15][chunk]: u (s8) inside the disc, else 0
  __shared__ uint32_t sWO[32 * 8];       // [row][chunk]: 1 inside the disc, else 0
  __shared__ int sStart[kMaxLevels + 1]; // first output row of each level in this frame
  __shared__ uint2 sPatch[kDescWarps][kPatchRows * kPatchVecs];   // per warp: the blurred 39x39 neighbourhood

  const int f = blockIdx.y;
  const int nl = g.nlevels;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  {  // pattern pair p = lane*8 + bit -> sPat[bit*32 + lane]; thread tid handles pair tid
    const int8_t* pp = &cPattern[tid * 4];
    sPat[(tid & 7) * 32 + (tid >> 3)] = make_float4((float)pp[0], (float)pp[1], (float)pp[2], (float)pp[3]);
    const int row = tid >> 3, chunk = tid & 7;  // 32 x 8 weight words (row 31 is unused padding)
    uint32_t wu = 0, wo = 0;
    if (row < 31) {
      const int v = row - kHalfPatch;
      const int d = cUmax[v < 0 ? -v : v];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int u = -16 + 4 * chunk + i;
        if (u >= -d && u <= d) {
          wu |= ((uint32_t)u & 0xffu) << (8 * i);
          wo |= 1u << (8 * i);
        }
      }
    }
    sWU[tid] = wu;
    sWO[tid] = wo;
    if (tid == 0) {
      int total = 0;
      // a level whose FAST candidates did not fit candCap (the dropped subset depends on atomics) or whose octree
      // list hit its guard makes the frame's result meaningless: n[f] = -1 tells the device-resident caller
      bool truncated = false;
      for (int l = 0; l < nl; l++) {
        sStart[l] = total;
        const int kc = keptCount[f * nl + l];
        truncated = truncated || kc < 0 || candCount[f * nl + l] > g.lv[l].candCap;
        total += max(0, min(kc, g.lv[l].keptCap));
      }
      sStart[nl] = total;
      if (blockIdx.x == 0) outN[f] = truncated ? -1 : total;
    }
  }
  __syncthreads();
  const int total = sStart[nl];
  const float factorPI = 0x1.1df46ap-6f;  // (float)(CV_PI/180.f)

#pragma unroll 1
  for (int it = 0; it < KPW; it++) {
    const int row = blockIdx.x * (kDescWarps * KPW) + it * kDescWarps + warp;  // level-major concatenation
    if (row >= total || row >= cap) break;
    int level = 0;
    while (level + 1 < nl && row >= sStart[level + 1]) level++;
    const LevelGeom& L = g.lv[level];
    const uint32_t k = kept[L.keptOff + (long long)f * L.keptCap + (row - sStart[level])];
    const int x = cand_x(k) + kMinBorder, y = cand_y(k) + kMinBorder;  // level coordinates (:909-910)

    // stage rows y-19 .. y+19, bytes from the 8-byte boundary at or below x-19 (planes are 16-byte pitched)
    const int bp = L.pitch;
    const int x0 = (x - kPatchR) & ~7, xo = (x - kPatchR) - x0;
    const int prow = lane / kPatchVecs, pvec = lane - prow * kPatchVecs;
    const bool pload = lane < kPatchRowsPerPass * kPatchVecs && pvec * 8 < xo + kPatchRows;
    // this lane's first vector; every pass moves on by kPatchRowsPerPass rows (one 64-bit add instead of a fresh
    // row * pitch product per pass)
    const uint8_t* bsrc = blurArena + L.planeOff + (long long)f * L.planeBytes + (long long)(y - kPatchR + prow) * bp + x0 + pvec * 8;
    const long long bstep = (long long)kPatchRowsPerPass * bp;
    uint2* bdst = &sPatch[warp][prow * kPatchVecs + pvec];
    __syncwarp();  // the previous keypoint's gathers are done
#pragma unroll
    for (int i = 0; i < kPatchPasses; i++) {
      if (pload && i * kPatchRowsPerPass + prow < kPatchRows)
        bdst[i * kPatchRowsPerPass * kPatchVecs] = *reinterpret_cast<const uint2*>(bsrc);
      bsrc += bstep;
    }

    // ---- IC_Angle ----
    const int pitch = pyr.pitch[level];
    const int chunk = lane & 7, sub = lane >> 3;
    // rows sub, sub+4, ... of the 31-row disc: one pointer, advanced by four rows per step
    const uint8_t* ip = pyr.lvl[level] + (long long)f * pyr.frameStride[level] + (long long)(y + sub - kHalfPatch) * pitch + x +
                        (4 * chunk - 16);
    const long long istep = 4LL * pitch;
    int m10 = 0, m01 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int r = 4 * i + sub;  // row index 0..31, v = r - 15
      if (r < 31) {
        const uintptr_t addr = reinterpret_cast<uintptr_t>(ip);
        const uint32_t* ap = reinterpret_cast<const uint32_t*>(addr & ~(uintptr_t)3);
        const uint32_t px = __funnelshift_r(ap[0], ap[1], (uint32_t)(addr & 3) * 8);
        m10 = dp4a_us(px, sWU[r * 8 + chunk], m10);
        m01 += (r - kHalfPatch) * (int)__dp4a(px, sWO[r * 8 + chunk], 0u);
      }
      ip += istep;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rBRIEF byte `lane` ----
    float sn, cs;
    sincosf_glibc(__fmul_rn(angle, factorPI), &sn, &cs);
    const float a = cs, b = sn;
    __syncwarp();  // the staged rows are visible
    const uint8_t* bc = reinterpret_cast<const uint8_t*>(sPatch[warp]) + kPatchR * kPatchPitch + kPatchR + xo;
    int val = 0;
#pragma unroll
    for (int bit = 0; bit < 8; bit++) {
      const float4 q = sPat[bit * 32 + lane];
      const int r0 = round_half_even_small(__fmaf_rn(q.x, b, __fmul_rn(q.y, a)));
      const int c0 = round_half_even_small(__fmaf_rn(q.x, a, -__fmul_rn(q.y, b)));
      const int r1 = round_half_even_small(__fmaf_rn(q.z, b, __fmul_rn(q.w, a)));
      const int c1 = round_half_even_small(__fmaf_rn(q.z, a, -__fmul_rn(q.w, b)));
      const int t0 = bc[r0 * kPatchPitch + c0], t1 = bc[r1 * kPatchPitch + c1];
      val |= (t0 < t1) << bit;
    }
    outDesc[((long long)f * cap + row) * 32 + lane] = (uint8_t)val;

    if (lane < 7) {  // the 28-byte record as 7 words, one per lane
      const float fx = (float)x, fy = (float)y;
      uint32_t wv;
      switch (lane) {
        case 0: wv = __float_as_uint(level ? __fmul_rn(fx, L.scale) : fx); break;  // pt *= scale (:1178-1184)
        case 1: wv = __float_as_uint(level ? __fmul_rn(fy, L.scale) : fy); break;
        case 2: wv = __float_as_uint(L.patchSize); break;
        case 3: wv = __float_as_uint(angle); break;
        case 4: wv = __float_as_uint((float)cand_score(k)); break;
        case 5: wv = (uint32_t)level; break;
        default: wv = 0xffffffffu; break;  // class_id = -1
      }
      reinterpret_cast<uint32_t*>(outKps + (long long)f * cap + row)[lane] = wv;
    }
  }
}

void launch_describe(const PyramidPtrs& pyr, const uint8_t* blurArena, const FrameSetGeom& hGeom,
                     const uint32_t* kept, const int* keptCount, const int* candCount, orbx_keypoint* outKps,
                     uint8_t* outDesc, int cap, int* outN, int nframes, cudaStream_t s) {
  const int maxKp = hGeom.keptPerFrame;
  if (nframes <= 4) {
    constexpr int perCta = kDescWarps * kDescKpPerWarpSmall;
    launch_pdl(describe_kernel<kDescKpPerWarpSmall>, dim3((maxKp + perCta - 1) / perCta, nframes), dim3(kDescWarps * 32), 0, s, pyr,
               blurArena, hGeom, kept, keptCount, candCount, outKps, outDesc, cap, outN);
  } else {
    constexpr int perCta = kDescWarps * kDescKpPerWarp;
    launch_pdl(describe_kernel<kDescKpPerWarp>, dim3((maxKp + perCta - 1) / perCta, nframes), dim3(kDescWarps * 32), 0, s, pyr,
               blurArena, hGeom, kept, keptCount, candCount, outKps, outDesc, cap, outN);
  }
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
