// 3xTF32 tensor-core GEMM building block for the deformation network (sm_100a, tcgen05 + TMEM).
//
//   D[128 x NT] (+)= A[128 x K] * B[NT x K]^T        per CTA, fp32 accumulate in tensor memory
//
// Every product is evaluated as  A_lo*B_hi + A_hi*B_lo + A_hi*B_hi  with hi = rna_tf32(x) and
// lo = rna_tf32(x - hi), so the result carries ~22 mantissa bits per factor: fp32-grade, which
// is what the reference computes (nn.Linear on cuBLAS SGEMM, TF32 off). Operands are staged by
// the CTA's threads into the canonical no-swizzle K-major UMMA shared-memory layout, one elected
// thread issues tcgen05.mma.kind::tf32, tcgen05.commit signals an mbarrier per stage, and the 128
// threads read their own accumulator row back with tcgen05.ld for the fused epilogue.
// Operands whose contraction index is the slow one in global memory (weight gradients, W^T) are
// transposed while staging, so no transposed copy is ever written to HBM.
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
constexpr int TC_KC = 16;        // K elements staged per pipeline stage (two UMMA k-steps)
constexpr int TC_THREADS = 128;

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
};

struct TcArgs {
  TcBatch b[TC_MAX_BATCH];
  long lda, ldb;
  int M;          // K-major A: number of valid rows (Gaussians). MN-major A: valid rows of D (features)
  int K;          // contraction length (MN-major A: number of Gaussians)
  int NT;         // tile width, multiple of 16, <= 256
  int k_per_cta;  // MN-major A only: contraction slice per CTA (split over blockIdx.x)
  int epilogue;
  int relu;       // BIAS_ACT
  int accumulate; // ACCUM
  int ones_col;   // MN-major B: column of D that sums A over the contraction (-1: none)
  int dbg;        // test hook: 2 = single pass (hi*hi only)
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

__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

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

// Instruction descriptor for kind::tf32, fp32 accumulate, both operands K-major
// (InstrDescriptor in the same header).
__device__ __forceinline__ uint32_t umma_idesc_tf32(int M, int N) {
  uint32_t d = 0;
  d |= 1u << 4;                       // D format F32
  d |= 2u << 7;                       // A format TF32
  d |= 2u << 10;                      // B format TF32
  d |= (uint32_t)(N >> 3) << 17;
  d |= (uint32_t)(M >> 4) << 24;
  return d;
}

// ---- operand staging ---------------------------------------------------------------------------
// K-major tile [ROWS x 16k] in shared memory: float index = ((k/4) * ROWS + row) * 4 + k%4
//   -> core matrix (8 rows x 16 B) contiguous, LBO (next K chunk) = ROWS*16 B, SBO (next 8 rows) = 128 B.
__device__ __forceinline__ void split_store(float4 x, float4* hi, float4* lo) {
  float4 h, l;
  h.x = tf32_rna(x.x); h.y = tf32_rna(x.y); h.z = tf32_rna(x.z); h.w = tf32_rna(x.w);
  l.x = tf32_rna(x.x - h.x); l.y = tf32_rna(x.y - h.y); l.z = tf32_rna(x.z - h.z); l.w = tf32_rna(x.w - h.w);
  *hi = h;
  *lo = l;
}

// Source K-major: element (row, k) = src[row * ld + k]. rows = tile rows (128 for A, NT for B).
__device__ __forceinline__ void stage_k_major(const float* __restrict__ src, long ld, int rows, int rows_valid,
                                              int k0, int K, float4* hi, float4* lo) {
  const int total = rows * (TC_KC / 4);
  for (int q = threadIdx.x; q < total; q += TC_THREADS) {
    const int c4 = q / rows, row = q - c4 * rows;
    const int k = k0 + c4 * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < rows_valid && k < K) x = *reinterpret_cast<const float4*>(src + (long)row * ld + k);
    split_store(x, hi + q, lo + q);
  }
}

// Source MN-major: element (row, k) = src[k * ld + row] (rows contiguous). The tensor core only
// accepted K-major tf32 operands without swizzle in our probes (MN-major descriptors produced
// zeros), so this stager transposes on the fly into the same K-major tile: a float4 of 4
// consecutive rows is scattered as 4 scalars. Lanes are mapped (k%4, row-group) and each lane
// rotates its store order so that a warp's 32 scalar stores hit 32 distinct banks.
__device__ __forceinline__ void stage_transposed(const float* __restrict__ src, long ld, int rows, int rows_valid,
                                                 int k0, int K, int ones_row, float* hi, float* lo) {
  const int groups = rows / 4;
  const int total = TC_KC * groups;
  for (int q = threadIdx.x; q < total; q += TC_THREADS) {
    const int kq = q & 3;
    const int g = (q >> 2) % groups;
    const int k4 = (q >> 2) / groups;
    const int k = k0 + k4 * 4 + kq;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K) {
      if (g * 4 < rows_valid) x = *reinterpret_cast<const float4*>(src + (long)k * ld + g * 4);
      if (ones_row >= 0 && (ones_row >> 2) == g) (&x.x)[ones_row & 3] = 1.0f;
    }
    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = (j + (g >> 1)) & 3;
      const float v = xs[i];
      const float h = tf32_rna(v);
      const int idx = (k4 * rows + g * 4 + i) * 4 + kq;
      hi[idx] = h;
      lo[idx] = tf32_rna(v - h);
    }
  }
}

template <int AMAJ, int BMAJ>
__global__ void __launch_bounds__(TC_THREADS) tc_gemm_kernel(const TcArgs p) {
  extern __shared__ __align__(128) unsigned char tc_smem[];
  __shared__ __align__(8) uint64_t s_bar[2];
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5;
  const TcBatch bt = p.b[blockIdx.y];
  const int NT = p.NT;
  const int Ktot = bt.K > 0 ? bt.K : p.K;
  const int Kb = bt.kb_valid > 0 ? min(bt.kb_valid, Ktot) : Ktot;
  const uint32_t a_f4 = TC_M * (TC_KC / 4);          // float4s per A stage half (hi or lo)
  const uint32_t b_f4 = (uint32_t)NT * (TC_KC / 4);
  float4* sA_hi = reinterpret_cast<float4*>(tc_smem);            // [2][a_f4]
  float4* sA_lo = sA_hi + 2 * a_f4;
  float4* sB_hi = sA_lo + 2 * a_f4;                              // [2][b_f4]
  float4* sB_lo = sB_hi + 2 * b_f4;

  uint32_t ncols = 32;
  while (ncols < (uint32_t)NT) ncols <<= 1;
  if (warp == 0) tmem_alloc(&s_tmem, ncols);
  if (tid == 0) {
    mbar_init(&s_bar[0], 1);
    mbar_init(&s_bar[1], 1);
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  // Tile origin and contraction range.
  int m0 = 0, kbeg = 0, kend = Ktot;
  const float* Abase = bt.A;
  if (AMAJ == TC_K_MAJOR) {
    m0 = blockIdx.x * TC_M;
    Abase += (long)m0 * p.lda;
  } else {
    kbeg = blockIdx.x * p.k_per_cta;
    kend = min(Ktot, kbeg + p.k_per_cta);
  }
  const int a_rows_valid = (AMAJ == TC_K_MAJOR) ? max(0, min(TC_M, p.M - m0)) : min(TC_M, p.M);
  const int nchunks = (kend - kbeg + TC_KC - 1) / TC_KC;
  const int kend_b = min(kend, Kb);

  const uint32_t idesc = umma_idesc_tf32(TC_M, NT);
  const uint32_t a_lbo = TC_M * 16u, b_lbo = (uint32_t)NT * 16u;
  const uint32_t sbo = 128u;
  // bytes to advance the descriptor start address per UMMA k-step (8 tf32 = two 16-byte K chunks)
  const uint32_t a_kstep = 2u * TC_M * 16u;
  const uint32_t b_kstep = 2u * (uint32_t)NT * 16u;

  for (int c = 0; c < nchunks; ++c) {
    const int s = c & 1;
    if (c >= 2) mbar_wait(&s_bar[s], (uint32_t)(((c >> 1) - 1) & 1));  // MMAs of chunk c-2 done with stage s
    const int k0 = kbeg + c * TC_KC;
    if (AMAJ == TC_K_MAJOR)
      stage_k_major(Abase, p.lda, TC_M, a_rows_valid, k0, kend, sA_hi + s * a_f4, sA_lo + s * a_f4);
    else
      stage_transposed(Abase, p.lda, TC_M, a_rows_valid, k0, kend, -1, reinterpret_cast<float*>(sA_hi + s * a_f4),
                       reinterpret_cast<float*>(sA_lo + s * a_f4));
    if (BMAJ == TC_K_MAJOR)
      stage_k_major(bt.B, p.ldb, NT, bt.n_valid, k0, kend_b, sB_hi + s * b_f4, sB_lo + s * b_f4);
    else
      stage_transposed(bt.B, p.ldb, NT, bt.n_valid, k0, kend_b, p.ones_col, reinterpret_cast<float*>(sB_hi + s * b_f4),
                       reinterpret_cast<float*>(sB_lo + s * b_f4));
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t ah = smem_u32(sA_hi + s * a_f4), al = smem_u32(sA_lo + s * a_f4);
      const uint32_t bh = smem_u32(sB_hi + s * b_f4), bl = smem_u32(sB_lo + s * b_f4);
#pragma unroll
      for (int ks = 0; ks < TC_KC / 8; ++ks) {
        const uint64_t dah = umma_desc(ah + ks * a_kstep, a_lbo, sbo);
        const uint64_t dal = umma_desc(al + ks * a_kstep, a_lbo, sbo);
        const uint64_t dbh = umma_desc(bh + ks * b_kstep, b_lbo, sbo);
        const uint64_t dbl = umma_desc(bl + ks * b_kstep, b_lbo, sbo);
        const uint32_t first = (c == 0 && ks == 0) ? 0u : 1u;
        if (p.dbg & 2) {
          umma_tf32(tmem, dah, dbh, idesc, first);
        } else {
          umma_tf32(tmem, dal, dbh, idesc, first);  // small terms first
          umma_tf32(tmem, dah, dbl, idesc, 1u);
          umma_tf32(tmem, dah, dbh, idesc, 1u);
        }
      }
      umma_commit(&s_bar[s]);
    }
  }
  if (nchunks > 0) {
    const int last = nchunks - 1;
    mbar_wait(&s_bar[last & 1], (uint32_t)((last >> 1) & 1));
  }
  tc_fence_after();

  // ---- epilogue: thread t owns accumulator row t (TMEM lane t) ----
  const int row = tid;
  const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);
  const long grow = (long)m0 + row;  // global row for K-major A
  const bool row_ok = row < a_rows_valid;
  for (int n0 = 0; n0 < NT; n0 += 16) {
    float v[16];
    if (nchunks > 0) {
      tmem_ld16(lane_addr + (uint32_t)n0, v);
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = 0.f;
    }
    if (!row_ok) continue;
    if (AMAJ == TC_K_MAJOR) {
      const bool full = (n0 + 16 <= bt.n_valid);
      switch (p.epilogue) {
        case TC_EPI_BIAS_ACT: {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int n = n0 + i;
            if (n < bt.n_valid) {
              float x = v[i] + (bt.bias ? bt.bias[n] : 0.f);
              if (p.relu) x = fmaxf(x, 0.f);
              v[i] = x;
            }
          }
          if (full && (bt.ldc & 3) == 0) {
            float4* dst = reinterpret_cast<float4*>(bt.C + grow * bt.ldc + n0);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          } else {
            for (int i = 0; i < 16; ++i)
              if (n0 + i < bt.n_valid) bt.C[grow * bt.ldc + n0 + i] = v[i];
          }
        } break;
        case TC_EPI_MASK: {
          if (full && (bt.ldc & 3) == 0 && (bt.ldaux & 3) == 0) {
            const float4* msk = reinterpret_cast<const float4*>(bt.aux + grow * bt.ldaux + n0);
            float4* dst = reinterpret_cast<float4*>(bt.C + grow * bt.ldc + n0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 m = msk[i];
              dst[i] = make_float4(m.x > 0.f ? v[4 * i] : 0.f, m.y > 0.f ? v[4 * i + 1] : 0.f,
                                   m.z > 0.f ? v[4 * i + 2] : 0.f, m.w > 0.f ? v[4 * i + 3] : 0.f);
            }
          } else {
            for (int i = 0; i < 16; ++i) {
              const int n = n0 + i;
              if (n < bt.n_valid) bt.C[grow * bt.ldc + n] = (bt.aux[grow * bt.ldaux + n] > 0.f) ? v[i] : 0.f;
            }
          }
        } break;
        case TC_EPI_ACCUM: {
          for (int i = 0; i < 16; ++i) {
            const int n = n0 + i;
            if (n < bt.n_valid) {
              float* dst = bt.C + grow * bt.ldc + n;
              *dst = (p.accumulate ? *dst : 0.f) + v[i];
            }
          }
        } break;
        case TC_EPI_HEAD: {
          for (int i = 0; i < 16; ++i) {
            const int n = n0 + i;
            if (n < bt.n_valid)
              bt.C[grow * bt.ldc + n] = bt.aux[grow * bt.ldaux + n] + (v[i] + bt.bias[n]) * bt.scale;
          }
        } break;
        default: break;
      }
    } else {
      // weight gradients: D row = feature m, contraction was over Gaussians
      for (int i = 0; i < 16; ++i) {
        const int n = n0 + i;
        if (n < bt.n_valid) {
          float* dst = (p.epilogue == TC_EPI_ATOMIC_T) ? (bt.C + (long)n * bt.ldc + row) : (bt.C + (long)row * bt.ldc + n);
          atomicAdd(dst, v[i]);
        } else if (n == p.ones_col && bt.bias_grad != nullptr) {
          atomicAdd(bt.bias_grad + row, v[i]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, ncols);
}

inline size_t tc_gemm_smem_bytes(int NT) { return (size_t)2 * 2 * (TC_M + NT) * (TC_KC / 4) * sizeof(float4); }

}  // namespace ed3dgs
