// Probe: tcgen05.mma kind::tf32 with the A operand in TENSOR MEMORY (written by tcgen05.st), B in shared memory.
// (1) correctness: D[128 x N] = A[128 x 32] * B[N x 32]^T with A[m][k] at TMEM lane m, column a0 + k;
// (2) rate: clocks per 128 x N x 8 MMA for N in {48, 64, 128, 256} (TS) next to the SS form;
// (3) the in-place chain step: accumulator -> registers -> same TMEM columns -> A operand of the next MMA.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../e-d3dgs_b200/csrc/tc_gemm.cuh"
using namespace ed3dgs;

__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
      "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
      "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
      "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
      "r"(__float_as_uint(v[15])), "r"(__float_as_uint(v[16])), "r"(__float_as_uint(v[17])), "r"(__float_as_uint(v[18])),
      "r"(__float_as_uint(v[19])), "r"(__float_as_uint(v[20])), "r"(__float_as_uint(v[21])), "r"(__float_as_uint(v[22])),
      "r"(__float_as_uint(v[23])), "r"(__float_as_uint(v[24])), "r"(__float_as_uint(v[25])), "r"(__float_as_uint(v[26])),
      "r"(__float_as_uint(v[27])), "r"(__float_as_uint(v[28])), "r"(__float_as_uint(v[29])), "r"(__float_as_uint(v[30])),
      "r"(__float_as_uint(v[31]))
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// B[n][k] in the K-major no-swizzle image of tc_prepare_b: float4 index (k/4)*NT + n.
__device__ __forceinline__ int b_off(int n, int k, int NT) { return ((k >> 2) * NT + n) * 4 + (k & 3); }

__global__ void __launch_bounds__(128, 1) correctness(int N, float* out, float* out2) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  float* B = reinterpret_cast<float*>(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < N * 32; i += 128) {
    const int n = i % N, k = i / N;
    B[b_off(n, k, N)] = (float)(((n * 3 + k * 5) % 11) - 5) * 0.25f;
  }
  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const int m = threadIdx.x;   // TMEM lane = row of A and D
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  float a[32];
#pragma unroll
  for (int k = 0; k < 32; ++k) a[k] = (float)(((m * 7 + k * 3) % 13) - 6) * 0.5f;
  tmem_st32(lane_base + 0u, a);          // A at columns 0..31
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_tf32(128, N);
      for (int ks = 0; ks < 4; ++ks) {
        const uint64_t bd = umma_desc(smem_u32(B) + (uint32_t)ks * 2u * (uint32_t)N * 16u, (uint32_t)N * 16u, 128u);
        umma_tf32_ts(tmem + 256u, tmem + (uint32_t)ks * 8u, bd, idesc, ks > 0);
      }
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 32) {
    float v[32];
    tmem_ld32(lane_base + 256u + (uint32_t)c0, v);
    for (int j = 0; j < 32 && c0 + j < N; ++j) out[m * N + c0 + j] = v[j];
  }
  // chain step: relu(D[:, 0:32]) written back over columns 256..287, used as A (K = 32) of a second product
  // D2 = relu(D[:, 0:32]) * B^T into columns 0.. (the first A is dead)
  {
    float v[32];
    tmem_ld32(lane_base + 256u, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
    tmem_st32(lane_base + 256u, v);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_tf32(128, N);
      for (int ks = 0; ks < 4; ++ks) {
        const uint64_t bd = umma_desc(smem_u32(B) + (uint32_t)ks * 2u * (uint32_t)N * 16u, (uint32_t)N * 16u, 128u);
        umma_tf32_ts(tmem + 0u, tmem + 256u + (uint32_t)ks * 8u, bd, idesc, ks > 0);
      }
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 1);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 32) {
    float v[32];
    tmem_ld32(lane_base + (uint32_t)c0, v);
    for (int j = 0; j < 32 && c0 + j < N; ++j) out2[m * N + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// rate: mode 0 = SS (A and B from shared memory), 1 = TS (A from TMEM); nacc accumulators in rotation
__global__ void __launch_bounds__(128, 1) rate(int mode, int N, int iters, int nacc, long long* out) {
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
    const uint32_t idesc = umma_idesc_tf32(128, N);
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    uint64_t da[8], db[8];
    for (int i = 0; i < 8; ++i) {
      da[i] = umma_desc(a0 + (uint32_t)i * 4096u, 128 * 16, 128);
      db[i] = umma_desc(b0 + (uint32_t)i * 8192u, N * 16, 128);
    }
    long long t0 = clock64(), t1 = t0;
    if (elect_one()) {
      // accumulators at columns 0 .. nacc*N; the TMEM A operand at columns 448..511
      for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t acc = tmem + (uint32_t)((i % nacc) * N);
          if (mode == 0) umma_tf32(acc, da[i], db[i], idesc, it > 0);
          else umma_tf32_ts(acc, tmem + 448u + (uint32_t)(i * 8), db[i], idesc, it > 0);
        }
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
  int bad = 0;
  for (int N : {48, 128, 256}) {
    float *out, *out2;
    cudaMallocManaged(&out, 128 * N * 4);
    cudaMallocManaged(&out2, 128 * N * 4);
    cudaFuncSetAttribute(correctness, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    correctness<<<1, 128, 64 * 1024>>>(N, out, out2);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("correctness N=%d: error %s\n", N, cudaGetErrorString(e)); return 1; }
    double maxerr = 0, maxerr2 = 0;
    for (int m = 0; m < 128; ++m) {
      double d1[256];
      for (int n = 0; n < N; ++n) {
        double s = 0;
        for (int k = 0; k < 32; ++k)
          s += (double)((((m * 7 + k * 3) % 13) - 6) * 0.5f) * (double)((((n * 3 + k * 5) % 11) - 5) * 0.25f);
        d1[n] = s;
        const double err = fabs(s - out[m * N + n]);
        if (err > maxerr) maxerr = err;
      }
      for (int n = 0; n < N; ++n) {
        double s = 0;
        for (int k = 0; k < 32; ++k) s += (d1[k] > 0 ? d1[k] : 0) * (double)((((n * 3 + k * 5) % 11) - 5) * 0.25f);
        const double err = fabs(s - out2[m * N + n]) / (1.0 + fabs(s));
        if (err > maxerr2) maxerr2 = err;
      }
    }
    printf("TS correctness N=%3d: max abs err %.3g ; chained (relu in place) max rel err %.3g\n", N, maxerr, maxerr2);
    if (maxerr > 1e-3 || maxerr2 > 2e-3) bad = 1;   // second product rounds relu(D) to tf32
    cudaFree(out); cudaFree(out2);
  }
  long long* t;
  cudaMallocManaged(&t, 16);
  cudaFuncSetAttribute(rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int mode : {0, 1})
    for (int N : {48, 64, 128, 256})
      for (int nacc : {1, 2}) {
        if (nacc * N > 448) continue;
        const int iters = 1024;
        rate<<<148, 128, 200 * 1024>>>(mode, N, iters, nacc, t);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("rate: error %s\n", cudaGetErrorString(e)); return 1; }
        printf("%s M=128 N=%3d nacc=%d : issue %.1f clk/mma, complete %.1f clk/mma\n", mode ? "TS" : "SS", N, nacc,
               (double)t[0] / (iters * 8), (double)t[1] / (iters * 8));
      }
  printf(bad ? "TS PROBE FAILED\n" : "TS PROBE OK\n");
  return bad;
}
