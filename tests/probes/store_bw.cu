// Probe: store throughput per SM for the epilogue's pattern (each warp instruction writes one 128-byte line, consecutive
// instructions 512 bytes apart) against 512-byte-per-instruction stores, 16 warps per SM, one CTA per SM.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512, 1) k(float* out, int mode, int iters, long long* clk) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3, cq = warp >> 2;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float* tile = out + ((size_t)blockIdx.x * iters + it) * (640 * 128);
    if (mode == 0) {          // [f][128 g]: warp (q, cq) writes rows f = cq*32 .. +31 (of 5 heads), 128 B each
      for (int hs = 0; hs < 5; ++hs) {
        float* zp = tile + ((hs * 128 + cq * 32) * 128) + q * 32 + lane;
#pragma unroll
        for (int j = 0; j < 32; ++j) zp[j * 128] = (float)j;
      }
    } else if (mode == 1) {   // same bytes, float4 per lane: 512 B contiguous per instruction
      for (int hs = 0; hs < 5; ++hs) {
        float4* zp = reinterpret_cast<float4*>(tile + (hs * 128 + cq * 32) * 128 + q * 1024) + lane;
#pragma unroll
        for (int j = 0; j < 8; ++j) zp[j * 32] = make_float4(1.f, 2.f, 3.f, 4.f);
      }
    } else {                  // mode 2: 128 B per instruction but consecutive instructions contiguous (row = 4 KB per warp)
      for (int hs = 0; hs < 5; ++hs) {
        float* zp = tile + (hs * 128 + cq * 32) * 128 + q * 1024 + lane;
#pragma unroll
        for (int j = 0; j < 32; ++j) zp[j * 32] = (float)j;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) clk[blockIdx.x] = clock64() - t0;
}
int main() {
  const int iters = 50;
  float* out; long long* clk;
  cudaMalloc(&out, (size_t)148 * iters * 640 * 128 * 4);
  cudaMallocManaged(&clk, 148 * 8);
  for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0);
      k<<<148, 512>>>(out, mode, iters, clk);
      cudaEventRecord(e1);
      cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double bytes = (double)148 * iters * 640 * 128 * 4;
      if (rep) printf("mode %d: %.3f ms, %.0f GB/s, %.1f B/clk/SM (block 0: %lld clk)\n", mode, ms, bytes / ms / 1e6,
                      (double)iters * 640 * 128 * 4 / clk[0], clk[0]);
    }
  }
  return 0;
}
