// Microbenchmark: cycles per tcgen05.mma.kind::tf32 (M=128, N in {64,128,256}, K=8) with the
// no-swizzle K-major canonical layout used by the deformation GEMMs. One CTA per SM.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../e-d3dgs_b200/csrc/tc_gemm.cuh"
using namespace ed3dgs;

__global__ void __launch_bounds__(128, 1) probe(int N, int iters, int distinct, long long* out) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 48 * 1024; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.001f * (i & 1023);
  if (warp == 0) tmem_alloc(&s_tmem, 256);
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  if (warp == 1) {
    const uint32_t idesc = umma_idesc_tf32(128, N);
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (elect_one()) {
        const uint32_t off = distinct ? (uint32_t)(i & 7) * 4096u : 0u;
        const uint64_t da = umma_desc(a0 + off, 128 * 16, 128);
        const uint64_t db = umma_desc(b0 + off * (N / 128 > 0 ? N / 128 : 1), N * 16, 128);
        umma_tf32(tmem, da, db, idesc, i > 0);
      }
      __syncwarp();
    }
    if (elect_one()) umma_commit(&bar);
    __syncwarp();
    long long t1 = clock64();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    if (threadIdx.x == 32 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 16);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int distinct = 0; distinct < 2; ++distinct)
    for (int N : {64, 128, 256})
      for (int grid : {1, 148}) {
        const int iters = 4096;
        probe<<<grid, 128, 200 * 1024>>>(N, iters, distinct, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("N=%3d grid=%3d distinct=%d : issue %.1f clk/mma, complete %.1f clk/mma\n", N, grid, distinct,
               (double)out[0] / iters, (double)out[1] / iters);
      }
  return 0;
}
