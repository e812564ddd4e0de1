// 3xTF32 tensor-core GEMM building block for the deformation network (sm_100a, tcgen05 + TMEM).
//
//   D[128 x NT] (+)= A[128 x K] * B[NT x K]^T        per CTA, fp32 accumulate in tensor memory
//
// Every product is evaluated as  A_lo*B_hi + A_hi*B_lo + A_hi*B_hi  with hi = rna_tf32(x) and
// lo = rna_tf32(x - hi), so the result carries ~22 mantissa bits per factor: fp32-grade, which
// is what the reference computes (nn.Linear on cuBLAS SGEMM, TF32 off).
//
// Pipeline of one CTA (256 threads, 2 CTAs per SM):
//   global --cp.async 16 B, 2-3 chunks ahead--> raw fp32 ring in shared memory
//   raw ring --all threads: split hi/lo (+ transpose when the contraction index is the slow one
//             in global memory)--> two UMMA operand stages (canonical no-swizzle K-major layout)
//   one elected thread: 6 x tcgen05.mma.kind::tf32 per 16-wide K chunk, tcgen05.commit -> mbarrier
//   epilogue: 8 warps read the accumulator from TMEM (tcgen05.ld 32x32b.x16), fused
//             bias/ReLU | ReLU-mask | x += delta*coef | atomic weight-gradient (+ ones column = bias)
// No transposed or split copy of any operand is ever written to HBM.
#pragma once
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
  TC_EPI_ATOMIC = 4,     // C[m][n] += acc      (weight gradient, split over Gaussians)
  TC_EPI_ATOMIC_T = 5,   // C[n][m] += acc      (weight gradient, transposed)
};

constexpr int TC_MAX_BATCH = 5;
constexpr int TC_M = 128;        // rows per CTA = TMEM lanes
#ifndef ED3DGS_TC_KC
#define ED3DGS_TC_KC 16
#endif
constexpr int TC_KC = ED3DGS_TC_KC;   // K elements per pipeline chunk (8 or 16: one or two UMMA k-steps)
constexpr int TC_PC = TC_KC / 4;      // 16-byte pieces per row of a chunk
// K-major raw tile: 16-byte pieces XOR-swizzled inside a row so the conversion pass (8 consecutive
// rows per quarter-warp) reads conflict-free.
__device__ __forceinline__ int tc_raw_kmajor(int row, int c4) {
  return row * TC_KC + ((c4 ^ ((row >> (TC_PC == 4 ? 1 : 2)) & (TC_PC - 1))) << 2);
}
constexpr int TC_THREADS = 256;

struct TcBatch {
  const float* A;
  const float* B;
  float* C;
  const float* bias;     // BIAS_ACT / HEAD: [n]
  const float* aux;      // MASK: mask matrix; HEAD: base matrix
  float* bias_grad;      // ATOMIC*: receives the ones-column sums ([m]), may be null
  long ldc, ldaux;
  int n_valid;           // columns of D that exist (rows of B)
  int K;                 // contraction length of this batch entry (0: TcArgs::K)
  int kb_valid;          // MN-major B: rows of B along k that exist (0: same as K)
  float scale;           // HEAD
  const float4* Bprep;   // set by tc_launch: per-chunk UMMA images of B (weights), see tc_prepare_b
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
  int NT;         // tile width, multiple of 16, <= 256
  int k_per_cta;  // MN-major A only: contraction slice per CTA
  int nbatch;     // batch entries; blockIdx.x = tile * nbatch + batch (entries of a tile run adjacently -> L2 reuse of A)
  int epilogue;
  int relu;       // BIAS_ACT
  int accumulate; // ACCUM
  int ones_col;   // MN-major B: column of D that sums A over the contraction (-1: none)
  int dbg;        // test hook: 2 = single pass (hi*hi only)
  // rows kernel programme (K-major A only)
  int nsteps, ntiles;
  uint32_t steps[TC_MAX_STEPS];   // batch | slab << 8 | npieces << 16 | nk << 20 | flags << 24
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

// ---- stage 1: global -> raw ring (cp.async) ----------------------------------------------------
// K-major source, tile [rows x 16k]: raw float index = row*16 + ((c4 ^ ((row>>1)&3)) << 2)
//   (16-byte pieces XOR-swizzled inside the 64-byte row so the conversion pass reads conflict-free)
// MN-major source, tile [16k x rows]: raw float index = kl*rows + row  (straight lines)
// ROWS_T > 0 makes the tile height a compile-time constant (no integer divisions, loops unrolled).
template <int MAJ, int ROWS_T>
__device__ __forceinline__ void raw_issue(float* raw, const float* __restrict__ src, long ld, int rows_rt,
                                          int rows_valid, int k0, int kend) {
  const int rows = ROWS_T ? ROWS_T : rows_rt;
  if (MAJ == TC_K_MAJOR) {
    const int total = rows * TC_PC;
#pragma unroll
    for (int q = threadIdx.x; q < total; q += TC_THREADS) {
      const int row = q / TC_PC, c4 = q % TC_PC;
      const int k = k0 + c4 * 4;
      const bool ok = row < rows_valid && k < kend;
      const float* s = ok ? (src + (long)row * ld + k) : src;
      cp_async16_zfill(raw + tc_raw_kmajor(row, c4), s, ok ? 16 : 0);
    }
  } else {
    const int groups = rows >> 2;
    const int total = TC_KC * groups;
#pragma unroll
    for (int q = threadIdx.x; q < total; q += TC_THREADS) {
      const int kl = q / groups, g = q - kl * groups;
      const int k = k0 + kl;
      const bool ok = k < kend && g * 4 < rows_valid;
      const float* s = ok ? (src + (long)k * ld + g * 4) : src;
      cp_async16_zfill(raw + kl * rows + g * 4, s, ok ? 16 : 0);
    }
  }
}

// ---- stage 2: raw ring -> UMMA operand stage (hi / lo), K-major tile ----------------------------
// UMMA tile [rows x 16k]: float4 index = (k/4) * rows + row -> LBO (next K chunk) = rows*16 B, SBO = 128 B.
template <int MAJ, int ROWS_T>
__device__ __forceinline__ void raw_convert(const float* raw, int rows_rt, int k0, int kend, int ones_row, float4* hi,
                                            float4* lo) {
  const int rows = ROWS_T ? ROWS_T : rows_rt;
  const int total = rows * TC_PC;
#pragma unroll
  for (int q = threadIdx.x; q < total; q += TC_THREADS) {
    const int c4 = q / rows, row = q - c4 * rows;
    float4 x;
    if (MAJ == TC_K_MAJOR) {
      x = *reinterpret_cast<const float4*>(raw + tc_raw_kmajor(row, c4));
    } else {
      const float* r0 = raw + (c4 * 4) * rows + row;
      x = make_float4(r0[0], r0[rows], r0[2 * rows], r0[3 * rows]);
      if (row == ones_row) {
        const int k = k0 + c4 * 4;
        x = make_float4(k < kend ? 1.f : 0.f, k + 1 < kend ? 1.f : 0.f, k + 2 < kend ? 1.f : 0.f, k + 3 < kend ? 1.f : 0.f);
      }
    }
    split_store(x, hi + q, lo + q);
  }
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

inline size_t tc_raw_stages(int NT) { return NT <= 128 ? 3 : 2; }
// Shared memory of one CTA. bprep: B arrives as prepared images (ring of 3), A through a raw ring of 2.
inline size_t tc_gemm_smem_bytes(int NT, bool bprep) {
  if (bprep) {
    return (size_t)2 * TC_M * TC_KC * sizeof(float)                 // raw A ring
           + (size_t)2 * 2 * TC_M * (TC_KC / 4) * sizeof(float4)     // UMMA A stages (hi, lo)
           + (size_t)3 * NT * (2 * TC_PC) * sizeof(float4);          // prepared-B ring
  }
  const size_t raw = tc_raw_stages(NT) * (size_t)(TC_M + NT) * TC_KC * sizeof(float);
  const size_t umma = (size_t)2 * 2 * (TC_M + NT) * (TC_KC / 4) * sizeof(float4);
  return raw + umma;
}

}  // namespace ed3dgs

#include "tc_gemm_kernel.cuh"
