// Shared pieces of the 3xTF32 tensor-core kernels of the deformation network (sm_100a, tcgen05 + TMEM):
// argument structs, PTX wrappers (mbarrier, tcgen05, cp.async, bulk copies), UMMA descriptors and the
// weight preparation kernel. The kernels themselves:
//   tc_rows.cuh   D^T[f][g] = sum_k W[f][k] X[g][k]   forward layers and input gradients (256 Gaussians / tile)
//   tc_wgrad.cuh  D[m][n]   = sum_g P[g][m] Q[g][n]   weight gradients (contraction over Gaussians)
//
// Every product is evaluated as  A_lo*B_hi + A_hi*B_lo + A_hi*B_hi  with hi = x rounded to tf32 and
// lo = x - hi, so a result carries ~21 mantissa bits per factor and fp32 accumulation: fp32-grade,
// which is what the reference computes (nn.Linear on cuBLAS SGEMM, TF32 off).
// No transposed or split copy of an activation matrix is ever written to HBM.
#pragma once
#include <cuda.h>           // CUtensorMap (type only; the encoder is fetched through the runtime)
#include <cuda_runtime.h>
#include <stdint.h>

namespace ed3dgs {

// Layout of an operand in GLOBAL memory. In shared memory both operands are always K-major.
enum TcMajor : int {
  TC_K_MAJOR = 0,   // element (row, k) at src[row * ld + k]
  TC_MN_MAJOR = 1   // element (row, k) at src[k * ld + row]
};

enum TcEpilogue : int {
  TC_EPI_BIAS_ACT = 0,   // C = act(acc + bias)                       (forward layers)
  TC_EPI_MASK = 1,       // C = aux > 0 ? acc : 0                     (activation backward)
  TC_EPI_ACCUM = 2,      // C = (accumulate ? C : 0) + acc            (input gradient)
  TC_EPI_HEAD = 3,       // C = aux + (acc + bias) * scale            (deformation heads)
  TC_EPI_ATOMIC = 4,     // test hook only: weight-gradient form, C[m][n] += acc
  TC_EPI_ATOMIC_T = 5,   // test hook only: weight-gradient form, C[n][m] += acc
};

constexpr int TC_MAX_BATCH = 5;
constexpr int TC_M = 128;        // rows per CTA = TMEM lanes
constexpr int TC_KC = 16;       // K elements per weight slab: two UMMA k-steps
constexpr int TC_PC = TC_KC / 4; // 16-byte pieces per row of a slab
// Where the generic->async proxy fence for stager-written operand stages is executed: by the MMA
// warp after it acquired the stage (true) or by every stager thread before it arrives (false).
#ifndef ED3DGS_FENCE_AT_CONSUMER
#define ED3DGS_FENCE_AT_CONSUMER 0
#endif
constexpr bool FENCE_AT_CONSUMER = ED3DGS_FENCE_AT_CONSUMER != 0;

struct TcBatch {
  const float* A;
  const float* B;
  float* C;
  const float* bias;     // BIAS_ACT / HEAD: [n]
  const float* aux;      // MASK: mask matrix; HEAD: base matrix
  float* bias_grad;      // MASK: += column sums of the stored values ([n], bias gradient of the layer below); may be null
  long ldc, ldaux;
  int n_valid;           // columns of D that exist (rows of B)
  int K;                 // contraction length of this batch entry (0: TcArgs::K)
  int kb_valid;          // MN-major B: rows of B along k that exist (0: same as K)
  float scale;           // HEAD
  const float4* Bprep;   // set by tc_launch: per-chunk UMMA images of B (weights), see tc_prepare_b
  // ReLU masks as bits: word [Gaussian / 32][mask_ld] at column mask_col + feature, bit = Gaussian % 32
  uint32_t* mask_out;        // BIAS_ACT + relu: written; may be null
  const uint32_t* mask_in;   // MASK: read instead of aux when not null
  long mask_ld;
  int mask_col;
};

// Per-tile programme of the persistent rows kernel (tc_rows.cuh), built on the host by tc_launch:
// a list of 16-wide K slabs; each one is an activation slab to stage (hi/lo), a prepared weight
// slab to copy and 3 or 6 UMMAs.
constexpr int TC_MAX_STEPS = 48;
enum TcStepFlags : int { TC_STEP_FIRST = 1, TC_STEP_LAST = 2 };

struct TcArgs {
  TcBatch b[TC_MAX_BATCH];
  long lda, ldb;
  int M;          // K-major A: number of valid rows (Gaussians). MN-major A: valid rows of D (features)
  int K;          // contraction length (MN-major A: number of Gaussians)
  int NT;         // rows of the prepared weight images (set by tc_launch: 128)
  int k_per_cta;  // unused (test hook compatibility)
  int nbatch;     // batch entries (heads) processed per tile, in order
  int epilogue;
  int relu;       // BIAS_ACT
  int accumulate; // ACCUM
  int ones_col;   // unused (test hook compatibility)
  int single;     // precision "tf32": one tf32 MMA per product (hi*hi only) instead of the 3xTF32 triple
  int dbg;        // timing knobs of builds with -DED3DGS_TIMING_KNOBS (16: skip the epilogue, 64: wrap addresses); else unused
  // rows kernel programme (K-major A only)
  int nsteps, ntiles;
  uint32_t steps[TC_MAX_STEPS];   // batch | slab << 8 | npieces << 16 | nk << 20 | flags << 24
  // one 2-D tensor map per batch entry: dims {K_b floats, M rows}, box {16, 256}, SWIZZLE_64B (tc_rows_launch)
  alignas(64) CUtensorMap xmap[TC_MAX_BATCH];
};

// ---- raw PTX wrappers ------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Bounded wait: a kernel that would hang traps instead (a hang on the GPU box is a strike).
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
#pragma unroll 1
  for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// One lane of a converged warp (the issuing lane of tcgen05.mma / bulk copies).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bulk copy global -> shared through the TMA unit (no tensor map), completion on an mbarrier.
__device__ __forceinline__ void bulk_copy_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// Bulk prefetch of a contiguous global range into L2 (size: multiple of 16 bytes).
__device__ __forceinline__ void prefetch_l2_bulk(const void* gmem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem_src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// 16-byte async copy global -> shared; src_bytes = 0 zero-fills the destination (predication).
__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gmem_src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit_group() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Shared-memory matrix descriptor, no swizzle (cute/arch/mma_sm100_desc.hpp: SmemDescriptor).
//  bits [0,14) start address >> 4, [16,30) leading byte offset >> 4, [32,46) stride byte offset >> 4,
//  [46,48) version = 1 on sm_100, [61,64) layout type = 0 (SWIZZLE_NONE / interleave)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}

// Instruction descriptor for kind::tf32, fp32 accumulate, both operands K-major.
__device__ __forceinline__ uint32_t umma_idesc_tf32(int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4;                       // D format F32
  d |= 2u << 7;                       // A format TF32
  d |= 2u << 10;                      // B format TF32
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

__device__ __forceinline__ void split_store(float4 x, float4* hi, float4* lo) {
  float4 h, l;
  h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
  l.x = tf32_rna(x.x - h.x); l.y = tf32_rna(x.y - h.y); l.z = tf32_rna(x.z - h.z); l.w = tf32_rna(x.w - h.w);
  *hi = h;
  *lo = l;
}

// ---- weights: split once per launch into per-chunk UMMA images ---------------------------------
// image of chunk c (NT*8 float4 = NT*128 B): [hi: (k/4)*NT + row][lo: same], so the GEMM CTAs copy
// it into shared memory with plain 16-byte cp.async and never touch it with the ALU.
template <int MAJ>
__global__ void __launch_bounds__(256) tc_prepare_b(const TcArgs p, float4* __restrict__ dst, long batch_stride_f4) {
  const TcBatch bt = p.b[blockIdx.y];
  const int NT = p.NT;
  const int Ktot = bt.K > 0 ? bt.K : p.K;
  const int Kb = bt.kb_valid > 0 ? min(bt.kb_valid, Ktot) : Ktot;
  const int c = blockIdx.x;
  const int k0 = c * TC_KC;
  if (k0 >= Ktot) return;
  float4* img = dst + (long)blockIdx.y * batch_stride_f4 + (long)c * NT * (2 * TC_PC);
  for (int q = threadIdx.x; q < NT * TC_PC; q += 256) {
    const int c4 = q / NT, row = q - c4 * NT;
    const int k = k0 + c4 * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < bt.n_valid) {
      if (MAJ == TC_K_MAJOR) {
        if (k < Kb) x = *reinterpret_cast<const float4*>(bt.B + (long)row * p.ldb + k);
      } else {
        float* xs = &x.x;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (k + i < Kb) xs[i] = bt.B[(long)(k + i) * p.ldb + row];
      }
    }
    split_store(x, img + q, img + NT * TC_PC + q);
  }
}

}  // namespace ed3dgs

