// Probe: does tcgen05.mma kind::tf32 accept MN-major SWIZZLE_128B operands on this GPU, and with
// which LBO/SBO meaning? One CTA, D[128 x 128] = A[128 x 8] * B[128 x 8]^T, operands written by the
// threads in the MN-major SW128 layout: tile of 32 rows(m) = 8 k-rows of 128 B, 16-byte chunk c of
// k-row k stored at chunk c ^ (k & 7).
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include "../../e-d3dgs_b200/csrc/tc_gemm.cuh"
using namespace ed3dgs;

struct Variant { uint32_t idesc_extra, lbo, sbo, layout, block_stride; };

__global__ void __launch_bounds__(128, 1) probe(Variant v, float* out) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t s_tmem;
  unsigned char* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  float* A = reinterpret_cast<float*>(sm);
  float* B = reinterpret_cast<float*>(sm + 16384);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 8192; i += 128) { A[i] = 0.f; B[i] = 0.f; }
  __syncthreads();
  // element (m, k): block m/32 at block_stride bytes, k-row k (128 B), chunk (m%32)/4 ^ (k&7), lane m%4
  for (int i = threadIdx.x; i < 128 * 8; i += 128) {
    const int m = i & 127, k = i >> 7;
    int off;
    if (v.layout == 1)   // 128B_BASE32B: 32-byte chunks XORed with (k & 3)
      off = (m >> 5) * (v.block_stride / 4) + k * 32 + ((((m & 31) >> 3) ^ (k & 3)) << 3) + (m & 7);
    else
      off = (m >> 5) * (v.block_stride / 4) + k * 32 + ((((m & 31) >> 2) ^ (k & 7)) << 2) + (m & 3);
    A[off] = (float)((m % 7) - 3 + k);
    B[off] = (float)(((m % 5) - 2) * (k + 1));
  }
  if (warp == 0) tmem_alloc(&s_tmem, 128);
  if (threadIdx.x == 32) { mbar_init(&bar, 1); fence_mbar_init(); }
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_tf32(128, 128) | v.idesc_extra;
      auto desc = [&](uint32_t addr) {
        uint64_t d = 0;
        d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
        d |= (uint64_t)((v.lbo >> 4) & 0x3FFFu) << 16;
        d |= (uint64_t)((v.sbo >> 4) & 0x3FFFu) << 32;
        d |= (uint64_t)1 << 46;
        d |= (uint64_t)v.layout << 61;
        return d;
      };
      umma_tf32(tmem, desc(smem_u32(A)), desc(smem_u32(B)), idesc, 0u);
      umma_commit(&bar);
    }
    __syncwarp();
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 128; c0 += 32) {
    float r[32];
    tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
    for (int j = 0; j < 32; ++j) out[(warp * 32 + (threadIdx.x & 31)) * 128 + c0 + j] = r[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 128);
}

int main() {
  float* out;
  cudaMallocManaged(&out, 128 * 128 * 4);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  const Variant vs[] = {
      {(1u << 15) | (1u << 16), 2048, 512, 1, 2048},
      {(1u << 15) | (1u << 16), 512, 2048, 1, 2048},
      {(1u << 15) | (1u << 16), 2048, 1024, 1, 2048},
      {(1u << 15) | (1u << 16), 1024, 2048, 1, 2048},
      {(1u << 15) | (1u << 16), 2048, 256, 1, 2048},
      {(1u << 15) | (1u << 16), 2048, 128, 1, 2048},
      {(1u << 15) | (1u << 16), 1024, 512, 1, 1024},
      {(1u << 15) | (1u << 16), 512, 1024, 1, 1024},
  };
  for (const Variant& v : vs) {
    memset(out, 0, 128 * 128 * 4);
    probe<<<1, 128, 48 * 1024>>>(v, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    double maxerr = 0, maxref = 0; int nz = 0;
    for (int m = 0; m < 128; ++m)
      for (int n = 0; n < 128; ++n) {
        double ref = 0;
        for (int k = 0; k < 8; ++k) ref += (double)((m % 7) - 3 + k) * (double)(((n % 5) - 2) * (k + 1));
        const double d = out[m * 128 + n];
        if (d != 0) ++nz;
        if (fabs(d - ref) > maxerr) maxerr = fabs(d - ref);
        if (fabs(ref) > maxref) maxref = fabs(ref);
      }
    printf("idesc_extra=%08x lbo=%4u sbo=%4u layout=%u block_stride=%u : max_err=%g (max |ref| %g), nonzero=%d  D[1][1]=%g D[5][3]=%g\n",
           v.idesc_extra, v.lbo, v.sbo, v.layout, v.block_stride, maxerr, maxref, nz, out[129], out[5 * 128 + 3]);
  }
  return 0;
}
