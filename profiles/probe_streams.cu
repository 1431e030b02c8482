// Probe (not part of the product): what four concurrent streams -- two read, two written, 16 bytes per element, the packed
// GAE's traffic mix -- reach on this GPU with the simplest possible kernel (c = a + b, d = a - b), against one read +
// one write stream (a copy).  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o probe_streams probe_streams.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) two_in_two_out(const float4* __restrict__ a, const float4* __restrict__ b,
                                                      float4* __restrict__ c, float4* __restrict__ d, long long n4) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n4; i += gridDim.x * 256ll) {
    const float4 x = __ldg(a + i), y = __ldg(b + i);
    c[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
    d[i] = make_float4(x.x - y.x, x.y - y.y, x.z - y.z, x.w - y.w);
  }
}
__global__ void __launch_bounds__(256) copy2(const float4* __restrict__ a, float4* __restrict__ c, long long n4) {
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n4; i += gridDim.x * 256ll) c[i] = __ldg(a + i);
}
// the same four streams in 2 KB pieces per CTA (512 tokens), one CTA per piece, like the scan kernel's tiles
__global__ void __launch_bounds__(64) tiles_64(const float4* __restrict__ a, const float4* __restrict__ b,
                                               float4* __restrict__ c, float4* __restrict__ d, long long n4) {
  const long long i0 = blockIdx.x * 128ll + threadIdx.x * 2;
  if (i0 + 1 < n4) {
    const float4 x0 = __ldg(a + i0), x1 = __ldg(a + i0 + 1), y0 = __ldg(b + i0), y1 = __ldg(b + i0 + 1);
    c[i0] = make_float4(x0.x + y0.x, x0.y + y0.y, x0.z + y0.z, x0.w + y0.w);
    c[i0 + 1] = make_float4(x1.x + y1.x, x1.y + y1.y, x1.z + y1.z, x1.w + y1.w);
    d[i0] = make_float4(x0.x - y0.x, x0.y - y0.y, x0.z - y0.z, x0.w - y0.w);
    d[i0 + 1] = make_float4(x1.x - y1.x, x1.y - y1.y, x1.z - y1.z, x1.w - y1.w);
  }
}

// ... with a DEPENDENT second load (b's address comes out of a's data, like values after the tile descriptor), and the
// same with every CTA asking L2 for the lines the CTA `dist` blocks later will load
template <bool kPrefetch>
__global__ void __launch_bounds__(64) chain_64(const float4* __restrict__ a, const float4* __restrict__ b,
                                               float4* __restrict__ c, float4* __restrict__ d, long long n4, long long dist) {
  const long long i0 = blockIdx.x * 128ll + threadIdx.x * 2;
  if (kPrefetch) {
    const long long j = (blockIdx.x + dist) * 128ll;
    if (j + 128 <= n4 && threadIdx.x < 32) {
      const char* q = reinterpret_cast<const char*>((threadIdx.x < 16 ? a : b) + j) + (threadIdx.x & 15) * 128;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(q));
    }
  }
  if (i0 + 1 < n4) {
    const float4 x0 = __ldg(a + i0), x1 = __ldg(a + i0 + 1);
    const long long off = static_cast<long long>(fminf(fabsf(x0.x), 0.f));        // 0, but only known once a has arrived
    const float4 y0 = __ldg(b + i0 + off), y1 = __ldg(b + i0 + 1 + off);
    c[i0] = make_float4(x0.x + y0.x, x0.y + y0.y, x0.z + y0.z, x0.w + y0.w);
    c[i0 + 1] = make_float4(x1.x + y1.x, x1.y + y1.y, x1.z + y1.z, x1.w + y1.w);
    d[i0] = make_float4(x0.x - y0.x, x0.y - y0.y, x0.z - y0.z, x0.w - y0.w);
    d[i0 + 1] = make_float4(x1.x - y1.x, x1.y - y1.y, x1.z - y1.z, x1.w - y1.w);
  }
}

int main() {
  const long long n = 1ll << 27, n4 = n / 4;
  float4 *a, *b, *c, *d;
  cudaMalloc(&a, n * 4); cudaMalloc(&b, n * 4); cudaMalloc(&c, n * 4); cudaMalloc(&d, n * 4);
  cudaMemset(a, 1, n * 4); cudaMemset(b, 2, n * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float ms;
  for (int per_sm = 2; per_sm <= 8; per_sm *= 2) {
    const int grid = 148 * per_sm;
    for (int i = 0; i < 3; ++i) two_in_two_out<<<grid, 256>>>(a, b, c, d, n4);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) two_in_two_out<<<grid, 256>>>(a, b, c, d, n4);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("2 in + 2 out, %d CTAs/SM persistent: %.1f us, %.0f GB/s\n", per_sm, ms * 100, 16.0 * n / (ms / 10 * 1e-3) / 1e9);
  }
  for (int i = 0; i < 3; ++i) tiles_64<<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4);
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) tiles_64<<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4);
  cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  printf("2 in + 2 out, one 64-thread CTA per 512 elements: %.1f us, %.0f GB/s\n", ms * 100, 16.0 * n / (ms / 10 * 1e-3) / 1e9);
  for (int i = 0; i < 3; ++i) chain_64<false><<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4, 0);
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) chain_64<false><<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4, 0);
  cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  printf("the same with a dependent second load: %.1f us, %.0f GB/s\n", ms * 100, 16.0 * n / (ms / 10 * 1e-3) / 1e9);
  for (long long dist = 2048; dist <= 16384; dist *= 2) {
    for (int i = 0; i < 3; ++i) chain_64<true><<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4, dist);
    cudaEventRecord(e0);
    for (int i = 0; i < 10; ++i) chain_64<true><<<(unsigned)(n4 / 128), 64>>>(a, b, c, d, n4, dist);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("  + prefetch.global.L2 %lld CTAs ahead: %.1f us, %.0f GB/s\n", dist, ms * 100, 16.0 * n / (ms / 10 * 1e-3) / 1e9);
  }
  // is prefetch.global.L2 effective at all?  The same chain at ONE resident CTA per SM (200 KB of dynamic shared memory):
  // latency-bound, so turning the next CTAs' DRAM misses into L2 hits must show if the prefetches are honoured
  cudaFuncSetAttribute(chain_64<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  cudaFuncSetAttribute(chain_64<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  {
    const long long m4 = n4 / 16;
    cudaEventRecord(e0);
    for (int i = 0; i < 3; ++i) chain_64<false><<<(unsigned)(m4 / 128), 64, 200 * 1024>>>(a, b, c, d, m4, 0);
    cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
    printf("1 CTA/SM, dependent chain, no prefetch: %.1f us for 2^23 elements, %.0f GB/s\n", ms / 3 * 1e3, 16.0 * n / 16 / (ms / 3 * 1e-3) / 1e9);
    for (long long dist = 148; dist <= 148 * 16; dist *= 4) {
      cudaEventRecord(e0);
      for (int i = 0; i < 3; ++i) chain_64<true><<<(unsigned)(m4 / 128), 64, 200 * 1024>>>(a, b, c, d, m4, dist);
      cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
      printf("  + prefetch.global.L2 %lld CTAs ahead: %.1f us, %.0f GB/s\n", dist, ms / 3 * 1e3, 16.0 * n / 16 / (ms / 3 * 1e-3) / 1e9);
    }
  }
  for (int i = 0; i < 3; ++i) copy2<<<148 * 8, 256>>>(a, c, n4);
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) copy2<<<148 * 8, 256>>>(a, c, n4);
  cudaEventRecord(e1); cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
  printf("1 in + 1 out (copy of 2^27 floats): %.1f us, %.0f GB/s\n", ms * 100, 8.0 * n / (ms / 10 * 1e-3) / 1e9);
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
