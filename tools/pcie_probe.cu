// Host<->device transfer probe for the host-buffer path (orbx_extract_batch): what limits the H2D leg on
// this box? Compares copy-engine transfers (1..4 streams, plain / write-combined pinned memory) with SM
// loads from mapped pinned memory (zero copy) and SM stores to it.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/pcie_probe.bin tools/pcie_probe.cu && tools/pcie_probe.bin
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void pull_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

static void fill(void* p, size_t bytes) {
  unsigned* u = (unsigned*)p;
  for (size_t i = 0; i < bytes / 4; i += 1024) u[i] = (unsigned)i;
}

int main() {
  const size_t bytes = 512ull << 20;
  void *hPlain, *hWC, *hMapped, *dev;
  CK(cudaSetDeviceFlags(cudaDeviceMapHost));
  CK(cudaMalloc(&dev, bytes));
  CK(cudaHostAlloc(&hPlain, bytes, cudaHostAllocDefault));
  CK(cudaHostAlloc(&hWC, bytes, cudaHostAllocWriteCombined));
  CK(cudaHostAlloc(&hMapped, bytes, cudaHostAllocMapped));
  memset(hPlain, 1, bytes); memset(hWC, 2, bytes); memset(hMapped, 3, bytes);
  cudaStream_t st[8];
  for (int i = 0; i < 8; i++) CK(cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking));

  auto copy_test = [&](const char* name, void* host, bool h2d, int nstreams, size_t piece) {
    for (int rep = 0; rep < 2; rep++) {
      CK(cudaDeviceSynchronize());
      const double t0 = now();
      size_t off = 0; int k = 0;
      while (off < bytes) {
        const size_t n = bytes - off < piece ? bytes - off : piece;
        if (h2d) CK(cudaMemcpyAsync((char*)dev + off, (char*)host + off, n, cudaMemcpyHostToDevice, st[k % nstreams]));
        else CK(cudaMemcpyAsync((char*)host + off, (char*)dev + off, n, cudaMemcpyDeviceToHost, st[k % nstreams]));
        off += n; k++;
      }
      CK(cudaDeviceSynchronize());
      const double dt = now() - t0;
      if (rep) printf("%-44s %2d stream(s) piece %5zu KiB: %6.1f GB/s\n", name, nstreams, piece >> 10, bytes / dt / 1e9);
    }
  };
  for (int ns : {1, 2, 4}) {
    copy_test("H2D copy engine, pinned", hPlain, true, ns, 9600ull << 10);
    copy_test("H2D copy engine, pinned write-combined", hWC, true, ns, 9600ull << 10);
  }
  copy_test("H2D copy engine, pinned", hPlain, true, 1, bytes);
  copy_test("H2D copy engine, pinned", hPlain, true, 4, 300ull << 10);
  for (int ns : {1, 2, 4}) copy_test("D2H copy engine, pinned", hPlain, false, ns, 2048ull << 10);

  void* dMapped;
  CK(cudaHostGetDevicePointer(&dMapped, hMapped, 0));
  for (int blocks : {148, 296, 592, 1184}) {
    for (int rep = 0; rep < 2; rep++) {
      CK(cudaDeviceSynchronize());
      const double t0 = now();
      pull_kernel<<<blocks, 256, 0, st[0]>>>((const uint4*)dMapped, (uint4*)dev, bytes / 16);
      CK(cudaDeviceSynchronize());
      const double dt = now() - t0;
      if (rep) printf("%-44s %4d CTAs: %6.1f GB/s\n", "H2D SM loads from mapped pinned memory", blocks, bytes / dt / 1e9);
    }
  }
  for (int blocks : {148, 592}) {
    for (int rep = 0; rep < 2; rep++) {
      CK(cudaDeviceSynchronize());
      const double t0 = now();
      pull_kernel<<<blocks, 256, 0, st[0]>>>((const uint4*)dev, (uint4*)dMapped, bytes / 16);
      CK(cudaDeviceSynchronize());
      const double dt = now() - t0;
      if (rep) printf("%-44s %4d CTAs: %6.1f GB/s\n", "D2H SM stores to mapped pinned memory", blocks, bytes / dt / 1e9);
    }
  }
  // both directions at once (the pipeline overlaps chunk c's D2H with chunk c+1's H2D)
  for (int rep = 0; rep < 2; rep++) {
    CK(cudaDeviceSynchronize());
    const double t0 = now();
    CK(cudaMemcpyAsync(dev, hPlain, bytes / 2, cudaMemcpyHostToDevice, st[0]));
    CK(cudaMemcpyAsync((char*)hWC, (char*)dev + bytes / 2, bytes / 2, cudaMemcpyDeviceToHost, st[1]));
    CK(cudaDeviceSynchronize());
    const double dt = now() - t0;
    if (rep) printf("%-44s: %6.1f GB/s per direction\n", "H2D + D2H concurrently", bytes / 2 / dt / 1e9);
  }
  // pageable source for reference
  void* hPage = malloc(bytes);
  fill(hPage, bytes);
  memset(hPage, 5, bytes);
  {
    const double t0 = now();
    CK(cudaMemcpy(dev, hPage, bytes, cudaMemcpyHostToDevice));
    printf("%-44s: %6.1f GB/s\n", "H2D pageable", bytes / (now() - t0) / 1e9);
  }
  return 0;
}
