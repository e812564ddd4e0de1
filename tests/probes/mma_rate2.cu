// Microbenchmark 2: throughput of tcgen05.mma.kind::tf32 for M in {64,128}, N in {64,128,256}, K=8, issued by one
// elected thread in an unrolled loop (8 MMAs per iteration, descriptors precomputed), accumulators rotating over
// `nacc` TMEM regions so that consecutive MMAs are independent. Answers: is the cost proportional to M and N?
#include <cstdio>
#include <cuda_runtime.h>
#include "../../e-d3dgs_b200/csrc/tc_gemm.cuh"
using namespace ed3dgs;

__global__ void __launch_bounds__(128, 1) probe(int M, int N, int iters, int nacc, long long* out) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.001f * (i & 1023);
  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  if (warp == 1) {
    const uint32_t idesc = umma_idesc_tf32(M, N);
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    uint64_t da[8], db[8];
    for (int i = 0; i < 8; ++i) {
      da[i] = umma_desc(a0 + (uint32_t)i * 4096u, 128 * 16, 128);
      db[i] = umma_desc(b0 + (uint32_t)i * 8192u, N * 16, 128);
    }
    long long t0 = clock64(), t1 = t0;
    if (elect_one()) {
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) umma_tf32(tmem + (uint32_t)((i % nacc) * N), da[i], db[i], idesc, it > 0);
      }
      umma_commit(&bar);
      t1 = clock64();
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 32 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 16);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int M : {64, 128})
    for (int N : {64, 128, 256})
      for (int nacc : {1, 2}) {
        if (nacc * N > 512) continue;
        const int iters = 1024;
        probe<<<148, 128, 200 * 1024>>>(M, N, iters, nacc, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("M=%3d N=%3d nacc=%d : issue %.1f clk/mma, complete %.1f clk/mma\n", M, N, nacc,
               (double)out[0] / (iters * 8), (double)out[1] / (iters * 8));
      }
  return 0;
}
