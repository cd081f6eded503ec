// Issue-rate probe for the packed-integer / half2 instructions the FAST, blur and Hamming kernels are built
// from: which of them share an issue pipe on sm_100a. Every kernel runs 8 independent dependency chains per
// thread, 1024 threads per CTA, one CTA per SM, and reports warp-instructions per clock per SM sub-partition
// (SMSP) from clock64(). A mix "A+B" that runs at the sum of the two rates means A and B are on different
// pipes; a mix that runs at the harmonic combination means they share one.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/pipe_probe tools/pipe_probe.cu && /tmp/pipe_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

// ptxas fuses the two chained 2-input ops into one VIMNMX3 (this is what __vimax3_u16x2 expands to)
#define OP_VIMNMX3(x, y, z) asm volatile("{.reg .b32 t; max.u16x2 t, %0, %1; max.u16x2 %0, t, %2;}" : "+r"(x) : "r"(y), "r"(z))
#define OP_VIMNMX(x, y) asm volatile("min.u16x2 %0, %0, %1;" : "+r"(x) : "r"(y))
#define OP_HMNMX2(x, y) asm volatile("max.f16x2 %0, %0, %1;" : "+r"(x) : "r"(y))
#define OP_FMNMX(x, y) asm volatile("max.f32 %0, %0, %1;" : "+f"(x) : "f"(y))
#define OP_FMNMX3(x, y, z) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x) : "f"(y), "f"(z))
#define OP_HFMA2(x, y) asm volatile("fma.rn.f16x2 %0, %0, %1, %1;" : "+r"(x) : "r"(y))
#define OP_IMAD(x, y) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x) : "r"(y))
#define OP_PRMT(x, y) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(x) : "r"(y))
#define OP_LOP3(x, y, z) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(y), "r"(z))
#define OP_DP4A(x, y) asm volatile("dp4a.u32.u32 %0, %1, %1, %0;" : "+r"(x) : "r"(y))
#define OP_POPC(x, y) asm volatile("{.reg .u32 t; popc.b32 t, %0; xor.b32 %0, t, %1;}" : "+r"(x) : "r"(y))
#define OP_FFMA(x, y) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x) : "f"(y))
#define OP_VABSDIFF4(x, y) asm volatile("vabsdiff4.u32.u32.u32.add %0, %0, %1, %0;" : "+r"(x) : "r"(y))

constexpr int kIters = 2048;

template <int MODE>
__global__ void __launch_bounds__(1024) probe(unsigned* out, long long* cycles, unsigned seed) {
  unsigned a[8], b[8];
  float fa[8], fb[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    a[i] = seed * (threadIdx.x + 1) + i * 0x10203u;
    b[i] = seed ^ (0x01010101u * i);
    fa[i] = (float)(a[i] & 0xffff);
    fb[i] = (float)(b[i] & 0xffff) * 1e-3f;
  }
  const unsigned y = seed | 0x00030005u, z = seed ^ 0x00110022u;
  const float fy = (float)seed, fz = 0.5f;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) OP_VIMNMX3(a[i], y, z);
      if (MODE == 1) OP_VIMNMX(a[i], y);
      if (MODE == 2) OP_HMNMX2(a[i], y);
      if (MODE == 3) OP_FMNMX(fa[i], fy);
      if (MODE == 4) OP_FMNMX3(fa[i], fy, fz);
      if (MODE == 5) OP_HFMA2(a[i], y);
      if (MODE == 6) OP_IMAD(a[i], y);
      if (MODE == 7) OP_PRMT(a[i], y);
      if (MODE == 8) OP_LOP3(a[i], y, z);
      if (MODE == 9) OP_DP4A(a[i], y);
      if (MODE == 10) OP_POPC(a[i], y);
      if (MODE == 11) OP_FFMA(fa[i], fy);
      if (MODE == 12) OP_VABSDIFF4(a[i], y);
      // mixes: one op of each kind per chain slot, on separate chains
      if (MODE == 20) { OP_VIMNMX3(a[i], y, z); OP_HMNMX2(b[i], y); }
      if (MODE == 21) { OP_VIMNMX3(a[i], y, z); OP_IMAD(b[i], y); }
      if (MODE == 22) { OP_VIMNMX3(a[i], y, z); OP_HFMA2(b[i], y); }
      if (MODE == 23) { OP_VIMNMX3(a[i], y, z); OP_PRMT(b[i], y); }
      if (MODE == 24) { OP_VIMNMX3(a[i], y, z); OP_FMNMX3(fb[i], fy, fz); }
      if (MODE == 25) { OP_VIMNMX3(a[i], y, z); OP_DP4A(b[i], y); }
      if (MODE == 26) { OP_POPC(a[i], y); OP_IMAD(b[i], y); }
      if (MODE == 27) { OP_POPC(a[i], y); OP_LOP3(b[i], y, z); }
      if (MODE == 28) { OP_DP4A(a[i], y); OP_PRMT(b[i], y); }
      if (MODE == 29) { OP_DP4A(a[i], y); OP_IMAD(b[i], y); }
      if (MODE == 30) { OP_HMNMX2(a[i], y); OP_IMAD(b[i], y); }
      if (MODE == 31) { OP_HMNMX2(a[i], y); OP_PRMT(b[i], y); }
      if (MODE == 32) { OP_VIMNMX3(a[i], y, z); OP_FFMA(fb[i], fy); }
      if (MODE == 33) { OP_VIMNMX3(a[i], y, z); OP_VABSDIFF4(b[i], y); }
      if (MODE == 34) { OP_VABSDIFF4(a[i], y); OP_IMAD(b[i], y); }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  unsigned acc = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) acc ^= a[i] ^ b[i] ^ __float_as_uint(fa[i]) ^ __float_as_uint(fb[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

struct Case { int mode; const char* name; int opsPerSlot; };

template <int MODE>
static void run(const char* name, int opsPerSlot, unsigned* out, long long* cyc, int nsm) {
  probe<MODE><<<nsm, 1024>>>(out, cyc, 12345u);
  probe<MODE><<<nsm, 1024>>>(out, cyc, 12345u);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("%s: launch failed\n", name); exit(1); }
  long long h[256];
  cudaMemcpy(h, cyc, sizeof(long long) * nsm, cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < nsm; i++) mean += (double)h[i];
  mean /= nsm;
  // 32 warps per CTA = 8 warps per SMSP; each issues kIters * 8 * opsPerSlot instructions
  const double warpInstr = 8.0 * kIters * 8.0 * opsPerSlot;
  printf("%-28s %8.3f warp-instr/clk/SMSP   (%.0f cycles)\n", name, warpInstr / mean, mean);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount;
  printf("%s, %d SMs\n", p.name, nsm);
  unsigned* out; long long* cyc;
  cudaMalloc(&out, sizeof(unsigned) * nsm * 1024);
  cudaMalloc(&cyc, sizeof(long long) * nsm);
  run<0>("VIMNMX3.U16x2", 1, out, cyc, nsm);
  run<1>("VIMNMX.U16x2", 1, out, cyc, nsm);
  run<2>("HMNMX2", 1, out, cyc, nsm);
  run<3>("FMNMX", 1, out, cyc, nsm);
  run<4>("FMNMX3", 1, out, cyc, nsm);
  run<5>("HFMA2", 1, out, cyc, nsm);
  run<6>("IMAD", 1, out, cyc, nsm);
  run<7>("PRMT", 1, out, cyc, nsm);
  run<8>("LOP3", 1, out, cyc, nsm);
  run<9>("IDP.4A", 1, out, cyc, nsm);
  run<10>("POPC+LOP3", 2, out, cyc, nsm);
  run<11>("FFMA", 1, out, cyc, nsm);
  run<12>("VABSDIFF4", 1, out, cyc, nsm);
  run<20>("VIMNMX3 + HMNMX2", 2, out, cyc, nsm);
  run<21>("VIMNMX3 + IMAD", 2, out, cyc, nsm);
  run<22>("VIMNMX3 + HFMA2", 2, out, cyc, nsm);
  run<23>("VIMNMX3 + PRMT", 2, out, cyc, nsm);
  run<24>("VIMNMX3 + FMNMX3", 2, out, cyc, nsm);
  run<25>("VIMNMX3 + IDP.4A", 2, out, cyc, nsm);
  run<26>("POPC+LOP3 + IMAD", 3, out, cyc, nsm);
  run<27>("POPC+LOP3 + LOP3", 3, out, cyc, nsm);
  run<28>("IDP.4A + PRMT", 2, out, cyc, nsm);
  run<29>("IDP.4A + IMAD", 2, out, cyc, nsm);
  run<30>("HMNMX2 + IMAD", 2, out, cyc, nsm);
  run<31>("HMNMX2 + PRMT", 2, out, cyc, nsm);
  run<32>("VIMNMX3 + FFMA", 2, out, cyc, nsm);
  run<33>("VIMNMX3 + VABSDIFF4", 2, out, cyc, nsm);
  run<34>("VABSDIFF4 + IMAD", 2, out, cyc, nsm);
  return 0;
}
