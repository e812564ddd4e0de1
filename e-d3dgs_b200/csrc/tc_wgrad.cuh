// Persistent, warp-specialised 3xTF32 weight-gradient GEMM (sm_100a, tcgen05 + TMEM): contraction
// over Gaussians.
//
//   D[m][n] = sum_g P[g][m] * Q[g][n]        m < 128 features of P (TMEM lanes), n < 256 features of Q
//
// P and Q are activation matrices stored [Gaussian][feature]: the contraction index is the slow one,
// i.e. both operands are MN-major for the UMMA. With the 128-byte / 32-byte-atom swizzle that layout is exactly
// what a TMA tile of 32 features x 16 Gaussians produces (rows of 128 B, one per Gaussian, atoms of 4 rows),
// so nothing is transposed and no thread touches the raw data except to derive the lo part:
//
//   loader    global --TMA 2-D tiles {32 features, 16 Gaussians}, SWIZZLE_128B_ATOM_32B--> hi ring (raw fp32; the
//             tensor core reads it as tf32 by ignoring the low 13 bits)
//   stagers   hi slab --lo = x - trunc_tf32(x), same offsets--> lo ring
//   MMA       per 16-Gaussian slab: 2 k-steps x (P_lo*Q_hi + P_hi*Q_lo + P_hi*Q_hi), N = 256 (two heads
//             side by side keep the tensor pipe at full rate, see tc_rows.cuh)
//   epilogue  TMEM (2 accumulators, ping-pong) -> red.global.add.f32 into the weight gradients, one
//             flush per work item
//
// Work item = (slice of WG_SLICE Gaussians, group); a group is one (P, Q) pair. One CTA per SM walks
// items blockIdx.x, blockIdx.x + gridDim.x, ...; groups of a slice are adjacent so they share P through L2.
// (The first version transposed in the stagers -- cp.async rows, ld.shared.b32 columns -- because
// MN-major tf32 descriptors of every other layout type return zeros; it was shared-memory and barrier bound.)
#pragma once

namespace ed3dgs {

constexpr int WG_SLICE = 2048;                    // default Gaussians per work item (tensor-core accumulation chain)
constexpr int WG_MAX_GROUPS = 12;                 // one stage's dW1 heads (or pairs), dW2 heads and dW0 in a single launch
constexpr int WG_EPI_WARPS = 8;                   // warps 0..7 (TMEM lane quarter = warp % 4)
constexpr int WG_MMA_WARP = 8;
constexpr int WG_LOAD_WARP = 9;
constexpr int WG_CONV_WARP0 = 10;                 // 6 warps: 16 warps in all = 128 registers
constexpr int WG_CONV_THREADS = 192;
constexpr int WG_THREADS = WG_CONV_WARP0 * 32 + WG_CONV_THREADS;
constexpr int WG_BLOCK_BYTES = 16 * 128;          // one TMA tile: 16 Gaussians x 32 features
constexpr int WG_P_BYTES = 4 * WG_BLOCK_BYTES;    // 128 features
constexpr int WG_Q_BYTES = 8 * WG_BLOCK_BYTES;    // 256 features
constexpr int WG_SLAB_BYTES = WG_P_BYTES + WG_Q_BYTES;
constexpr int WG_HI_SLOTS = 5;
constexpr int WG_LO_SLOTS = 4;
constexpr size_t WG_SMEM_BYTES = (size_t)(WG_HI_SLOTS + WG_LO_SLOTS) * WG_SLAB_BYTES + 1024;   // + slack to align to 1024

// Where a block of D goes: C[(n - n0) * ldn + (m - m0) * ldm] += D[m][n] for n0 <= n < n0 + ncnt,
// m0 <= m < m0 + mcnt.
struct WgSegment {
  float* C;
  long ldn, ldm;
  int n0, ncnt, m0, mcnt;
};
struct WgGroup {
  const float* P;   // [N][ldp], columns 0..pcols-1 are the M side (pcols <= 128, multiple of 4)
  const float* Q;   // [N][ldq], columns 0..qcols-1 are the N side (qcols <= 256, multiple of 4)
  long ldp, ldq;
  int pcols, qcols;
  int nseg;
  WgSegment seg[2];
  // Feature-major operands (written by the stage-resident kernels, fused_mlp.cuh): P / Q point at row 0 of a
  // [tile][rows_total][128 Gaussians] matrix; pk / qk select that layout per operand (K-major for the UMMA).
  // A feature-major Q of qcols <= 128 rows is loaded and multiplied at its own width (qbox = qcols rounded up to 16, set
  // by tc_wgrad_launch): the second-layer weight gradients have 16 or 48 columns, not 128.
  int pk, qk, qbox;
  long prow_total, qrow_total;
};
// Row-major (MN-major) Q: blocks of 32 features, as many as its columns need (1..4, or 8 for two heads side by side).
__device__ __forceinline__ int wg_q_blocks(const WgGroup& g) { return g.qcols > 128 ? 8 : (g.qcols + 31) >> 5; }
__device__ __forceinline__ int wg_q_bytes(const WgGroup& g) { return g.qk ? g.qbox * 64 : wg_q_blocks(g) * WG_BLOCK_BYTES; }
__device__ __forceinline__ int wg_mma_n(const WgGroup& g) { return g.qk ? g.qbox : 32 * wg_q_blocks(g); }
struct WgArgs {
  WgGroup g[WG_MAX_GROUPS];
  int ngroups;
  int N;            // Gaussians
  int slice;        // Gaussians per work item (multiple of 128)
  int single;       // precision "tf32": one tf32 MMA per product (the lo ring is not filled)
  // tensor maps of P and Q of every group: dims {cols, N}, box {32, 16}, SWIZZLE_128B_ATOM_32B (tc_wgrad_launch)
  alignas(64) CUtensorMap pmap[WG_MAX_GROUPS];
  alignas(64) CUtensorMap qmap[WG_MAX_GROUPS];
};

struct WgBars {
  uint64_t hi_full[WG_HI_SLOTS], hi_free[WG_HI_SLOTS];
  uint64_t x_full[WG_LO_SLOTS], lo_free[WG_LO_SLOTS];
  uint64_t acc_full[2], acc_free[2];
};

// UMMA instruction descriptor, kind::tf32, fp32 accumulate; bit 15 / 16: A / B operand MN-major.
__device__ __forceinline__ uint32_t umma_idesc_tf32_mn(int M, int N, bool a_mn = true, bool b_mn = true) {
  return umma_idesc_tf32(M, N) | (a_mn ? (1u << 15) : 0u) | (b_mn ? (1u << 16) : 0u);
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const void* tmap, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
      : "memory");
}

__global__ void __launch_bounds__(WG_THREADS, 1) tc_wgrad_kernel(const __grid_constant__ WgArgs p) {
  extern __shared__ __align__(128) unsigned char wg_smem_raw[];
  __shared__ WgBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // 128-byte-swizzle atoms are 1024 bytes: the tiles must sit on that grid
  unsigned char* sHi = wg_smem_raw + ((1024u - (smem_u32(wg_smem_raw) & 1023u)) & 1023u);
  unsigned char* sLo = sHi + (size_t)WG_HI_SLOTS * WG_SLAB_BYTES;

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < WG_HI_SLOTS; ++i) { mbar_init(&bars.hi_full[i], 1); mbar_init(&bars.hi_free[i], 1); }
    for (int i = 0; i < WG_LO_SLOTS; ++i) { mbar_init(&bars.x_full[i], WG_CONV_THREADS); mbar_init(&bars.lo_free[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], WG_EPI_WARPS * 32); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int nslices = (p.N + p.slice - 1) / p.slice;
  const int nitems = nslices * p.ngroups;
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  // item k of this CTA -> group, first Gaussian, number of 16-Gaussian slabs
  auto item_setup = [&](int k, int& group, long& row0, int& nslab) {
    const int item = (int)blockIdx.x + k * (int)gridDim.x;
    const int slice = item / p.ngroups;
    group = item - slice * p.ngroups;
    row0 = (long)slice * p.slice;
    const int rows = (int)min((long)p.slice, (long)p.N - row0);
    nslab = (rows + 15) >> 4;
  };

  if (warp >= WG_CONV_WARP0) {
    // ------------------------------------------------------------------ stagers (192 threads)
    // lo = x - trunc_tf32(x), elementwise: same offsets in the lo slot, so the swizzle never shows
    const int ct = tid - WG_CONV_WARP0 * 32;
    constexpr int NIT = (WG_SLAB_BYTES / 16 + WG_CONV_THREADS - 1) / WG_CONV_THREADS;
    RingPos<WG_HI_SLOTS> hr;
    RingPos<WG_LO_SLOTS> lr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const int nchunks = (WG_P_BYTES + wg_q_bytes(p.g[group])) / 16;
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.hi_full[hr.idx], hr.phase);       // TMA bytes of this slab have landed
        const float4* src = reinterpret_cast<const float4*>(sHi + (size_t)hr.idx * WG_SLAB_BYTES);
        float4 r[NIT];
        if (!p.single) {
#pragma unroll
          for (int i = 0; i < NIT; ++i) {
            const int q = ct + i * WG_CONV_THREADS;
            if (q < nchunks) r[i] = src[q];
          }
        }
        mbar_wait(&bars.lo_free[lr.idx], lr.phase ^ 1u);  // the UMMAs that read this lo slot are done
        float4* lo = reinterpret_cast<float4*>(sLo + (size_t)lr.idx * WG_SLAB_BYTES);
        if (!p.single) {
#pragma unroll
          for (int i = 0; i < NIT; ++i) {
            const int q = ct + i * WG_CONV_THREADS;
            if (q < nchunks) lo[q] = split_lo(r[i]);
          }
        }
        fence_async_smem();
        mbar_arrive(&bars.x_full[lr.idx]);
        hr.advance();
        lr.advance();
      }
    }
  } else if (warp == WG_LOAD_WARP) {
    // ------------------------------------------------------------------ loader (converged warp, elected lane issues)
    RingPos<WG_HI_SLOTS> hr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const int nqb = wg_q_blocks(p.g[group]);
      const uint32_t bytes = (uint32_t)(WG_P_BYTES + wg_q_bytes(p.g[group]));
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.hi_free[hr.idx], hr.phase ^ 1u);
        if (elect_one()) {
          uint64_t* bar = &bars.hi_full[hr.idx];
          mbar_arrive_expect_tx(bar, bytes);
          unsigned char* dst = sHi + (size_t)hr.idx * WG_SLAB_BYTES;
          const int row = (int)row0 + s * 16;
          // feature-major operand: one box {16 Gaussians, 128 or 256 feature rows} of tile row / 128
          if (p.g[group].pk) tma_load_3d(dst, &p.pmap[group], row & 127, 0, row >> 7, bar);
          else {
#pragma unroll
            for (int b = 0; b < 4; ++b) tma_load_2d(dst + b * WG_BLOCK_BYTES, &p.pmap[group], 32 * b, row, bar);
          }
          if (p.g[group].qk) tma_load_3d(dst + WG_P_BYTES, &p.qmap[group], row & 127, 0, row >> 7, bar);
          else
            for (int b = 0; b < nqb; ++b) tma_load_2d(dst + WG_P_BYTES + b * WG_BLOCK_BYTES, &p.qmap[group], 32 * b, row, bar);
        }
        __syncwarp();
        hr.advance();
      }
    }
  } else if (warp == WG_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer
    // MN-major operands of 32-bit elements: only layout type 1 (SWIZZLE_128B_BASE32B: 32-byte chunks of a
    // 128-byte row XORed with the row index & 3) produces non-zero products on B200; every other layout
    // type returns zeros with the MN-major bits set (tests/probes/mn_probe.cu). A tile of 32 features is
    // 16 rows of 128 B: SBO = 512 B to the next 4 Gaussians, LBO = 2048 B to the next 32 features,
    // k-step (8 Gaussians) = +1024 B.
    const uint32_t desc_hi = (512u >> 4) | (1u << 14) | (1u << 29);
    const uint32_t lbo = (2048u >> 4) << 16;
    const uint32_t hi0 = ((smem_u32(sHi) & 0x3FFFFu) >> 4) | lbo;
    const uint32_t lo0 = ((smem_u32(sLo) & 0x3FFFFu) >> 4) | lbo;
    constexpr uint32_t SLAB = WG_SLAB_BYTES >> 4, QOFF = WG_P_BYTES >> 4, KS = 1024u >> 4;
    RingPos<WG_HI_SLOTS> hr;
    RingPos<WG_LO_SLOTS> lr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const bool pk = p.g[group].pk != 0, qk = p.g[group].qk != 0;
      const uint32_t idesc = umma_idesc_tf32_mn(128, wg_mma_n(p.g[group]), !pk, !qk);
      // feature-major operand = K-major, SWIZZLE_64B rows of 16 Gaussians (as the activation slabs of tc_rows.cuh):
      // SBO = 512 B between 8-row atoms, a k-step (8 Gaussians) advances the start by 32 B
      const uint32_t kdesc_hi = (512u >> 4) | (1u << 14) | (4u << 29);
      const uint32_t p_hi = pk ? kdesc_hi : desc_hi, q_hi = qk ? kdesc_hi : desc_hi;
      const uint32_t p_fix = pk ? ((1u << 16) - lbo) : 0u, q_fix = qk ? ((1u << 16) - lbo) : 0u;   // LBO field
      const uint32_t p_ks = pk ? (32u >> 4) : KS, q_ks = qk ? (32u >> 4) : KS;
      const int buf = k & 1;
      mbar_wait(&bars.acc_free[buf], (uint32_t)(((k >> 1) & 1) ^ 1));
      const uint32_t acc = tmem + (uint32_t)buf * 256u;
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.x_full[lr.idx], lr.phase);        // lo written (and the stagers saw the hi slab land)
        tc_fence_after();
        if (elect_one()) {
          const uint32_t ph = hi0 + (uint32_t)hr.idx * SLAB + p_fix, pl = lo0 + (uint32_t)lr.idx * SLAB + p_fix;
          const uint32_t qh = hi0 + (uint32_t)hr.idx * SLAB + QOFF + q_fix, ql = lo0 + (uint32_t)lr.idx * SLAB + QOFF + q_fix;
          const uint32_t first = s == 0 ? 0u : 1u;
          if (!p.single) {
            umma_tf32(acc, desc64(p_hi, pl), desc64(q_hi, qh), idesc, first);  // small terms first
            umma_tf32(acc, desc64(p_hi, ph), desc64(q_hi, ql), idesc, 1u);
            umma_tf32(acc, desc64(p_hi, ph), desc64(q_hi, qh), idesc, 1u);
            umma_tf32(acc, desc64(p_hi, pl + p_ks), desc64(q_hi, qh + q_ks), idesc, 1u);
            umma_tf32(acc, desc64(p_hi, ph + p_ks), desc64(q_hi, ql + q_ks), idesc, 1u);
            umma_tf32(acc, desc64(p_hi, ph + p_ks), desc64(q_hi, qh + q_ks), idesc, 1u);
          } else {
            umma_tf32(acc, desc64(p_hi, ph), desc64(q_hi, qh), idesc, first);
            umma_tf32(acc, desc64(p_hi, ph + p_ks), desc64(q_hi, qh + q_ks), idesc, 1u);
          }
          umma_commit(&bars.lo_free[lr.idx]);
          umma_commit(&bars.hi_free[hr.idx]);
          if (s == nslab - 1) umma_commit(&bars.acc_full[buf]);
        }
        __syncwarp();
        hr.advance();
        lr.advance();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 0..7)
    const int q = warp & 3, par = warp >> 2;
    const int m = q * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const WgGroup& G = p.g[group];
      const int buf = k & 1;
      mbar_wait(&bars.acc_full[buf], (uint32_t)((k >> 1) & 1));
      tc_fence_after();
      if (q * 32 < G.pcols) {
        const int nblk32 = (G.qcols + 31) >> 5;
        for (int blk = par; blk < nblk32; blk += 2) {
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(buf * 256 + blk * 32), v);
          for (int sg = 0; sg < G.nseg; ++sg) {
            const WgSegment& S = G.seg[sg];
            if (m < S.m0 || m >= S.m0 + S.mcnt) continue;
            float* base = S.C + (long)(m - S.m0) * S.ldm;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int n = blk * 32 + j - S.n0;
              if (n >= 0 && n < S.ncnt) atomicAdd(base + (long)n * S.ldn, v[j]);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&bars.acc_free[buf]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace ed3dgs
