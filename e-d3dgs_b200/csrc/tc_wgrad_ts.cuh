// Weight gradients of the stage-resident path with the M-side operand in TENSOR MEMORY (sm_100a, tcgen05 TS form).
//
//   D[m][n] = sum_g P[m][g] * Q[n][g]        P feature-major (written by fused_mlp.cuh), m < 128 = TMEM lanes
//
// tc_wgrad.cuh keeps both operands in shared memory; ncu (profiles/r02_v1_ncu.md) shows that kernel bound by the
// shared-memory data pipe (LSU 38 % + tensor-core operand fetch 35 % of its peak, the two clients colliding): per
// 16-Gaussian slab of a 128 x 256 product it moves 144 KB through shared memory (TMA fill 24, lo split 24 + 24, UMMA
// operand fetch 72) for 768 clocks of UMMA. P is feature-major here, i.e. element (m, g) of a tile already sits where a
// TMEM-resident A operand wants it (lane = m, column = g), so:
//
//   loader    TMA: P box {16 Gaussians, 128 features} and Q box of the group -> hi ring (raw fp32)
//   fill      4 warps, thread = feature m: its 64-byte row of the P box -> tcgen05.st of hi (raw) and lo (x - trunc) into
//             an 8-slab ring of TMEM columns; the UMMA reads A from there: no lo copy of P in shared memory, no operand
//             fetch of A through the shared-memory pipe
//   stagers   lo = x - trunc_tf32(x) of the Q box only
//   MMA       per slab 2 k-steps x (P_lo*Q_hi + P_hi*Q_lo + P_hi*Q_hi), A from TMEM, N = columns of Q (16 .. 128)
//   flush     4 warps: TMEM (2 accumulators of 128 columns, ping-pong per work item) -> red.global.add.f32
//
// One head per group (N <= 128) so that two accumulators and the A ring fit the 512 TMEM columns:
//   [0,128) ACC0   [128,256) ACC1   [256,512) A ring: slot i = 32 columns, hi at +0, lo at +16
#pragma once
#include "tc_wgrad.cuh"

namespace ed3dgs {

constexpr int WT_FLUSH_WARP0 = 0;                 // warps 0..3 (TMEM lane quarter = warp % 4)
constexpr int WT_FILL_WARP0 = 4;                  // warps 4..7
constexpr int WT_MMA_WARP = 8;
constexpr int WT_LOAD_WARP = 9;
constexpr int WT_CONV_WARP0 = 10;                 // 6 warps
constexpr int WT_CONV_THREADS = 192;
constexpr int WT_THREADS = WT_CONV_WARP0 * 32 + WT_CONV_THREADS;
constexpr int WT_P_BYTES = 16 * 128 * 4;          // P box: 128 features x 16 Gaussians
constexpr int WT_Q_BYTES = 16 * 128 * 4;          // Q box: up to 128 features
constexpr int WT_HI_SLAB = WT_P_BYTES + WT_Q_BYTES;
constexpr int WT_HI_SLOTS = 8;                    // = slots of the TMEM A ring (same index, same phase)
constexpr int WT_LO_SLOTS = WT_HI_SLOTS;           // lo slot i belongs to hi slot i: one ring position, one pair of barriers
constexpr uint32_t WT_A_BASE = 256;
constexpr size_t WT_SMEM_BYTES = (size_t)WT_HI_SLOTS * WT_HI_SLAB + (size_t)WT_LO_SLOTS * WT_Q_BYTES + 1024;

// The MMA warp's issue loop is the critical path of this kernel (one thread: waits, fence, descriptors, six UMMAs, commits
// per 16 Gaussians), so a slab costs it ONE wait and ONE commit: `ready` collects the fill warps and the stagers, `free`
// releases the shared-memory slot, its lo slot and the TMEM A slot together.
struct WtBars {
  uint64_t landed[WT_HI_SLOTS];   // TMA bytes of the slab (waited for by fill warps and stagers)
  uint64_t ready[WT_HI_SLOTS];    // A slot in tensor memory written (128 threads) + lo of Q written (192 threads)
  uint64_t free_[WT_HI_SLOTS];    // the UMMAs of the slab have completed
  uint64_t acc_full[2], acc_free[2];
};

__global__ void __launch_bounds__(WT_THREADS, 1) tc_wgrad_ts_kernel(const __grid_constant__ WgArgs p) {
  extern __shared__ __align__(128) unsigned char wt_smem_raw[];
  __shared__ WtBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sHi = wt_smem_raw + ((1024u - (smem_u32(wt_smem_raw) & 1023u)) & 1023u);
  unsigned char* sLo = sHi + (size_t)WT_HI_SLOTS * WT_HI_SLAB;

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < WT_HI_SLOTS; ++i) {
      mbar_init(&bars.landed[i], 1);
      mbar_init(&bars.ready[i], 128 + WT_CONV_THREADS);
      mbar_init(&bars.free_[i], 1);
    }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], 128); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int nslices = (p.N + p.slice - 1) / p.slice;
  const int nitems = nslices * p.ngroups;
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  auto item_setup = [&](int k, int& group, long& row0, int& nslab) {
    const int item = (int)blockIdx.x + k * (int)gridDim.x;
    const int slice = item / p.ngroups;
    group = item - slice * p.ngroups;
    row0 = (long)slice * p.slice;
    const int rows = (int)min((long)p.slice, (long)p.N - row0);
    nslab = (rows + 15) >> 4;
  };

  if (warp >= WT_CONV_WARP0) {
    // ------------------------------------------------------------------ stagers: lo part of the Q box
    const int ct = tid - WT_CONV_WARP0 * 32;
    constexpr int NIT = (WT_Q_BYTES / 16 + WT_CONV_THREADS - 1) / WT_CONV_THREADS;
    RingPos<WT_HI_SLOTS> hr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const int nchunks = wg_q_bytes(p.g[group]) / 16;
      for (int s = 0; s < nslab; ++s) {
        // a landed box also says that the previous round's UMMAs on this slot (and its lo slot) are done: the loader
        // refilled it only after `free_`
        mbar_wait(&bars.landed[hr.idx], hr.phase);
        const float4* src = reinterpret_cast<const float4*>(sHi + (size_t)hr.idx * WT_HI_SLAB + WT_P_BYTES);
        float4 r[NIT];
        if (!p.single) {
#pragma unroll
          for (int i = 0; i < NIT; ++i) {
            const int q = ct + i * WT_CONV_THREADS;
            if (q < nchunks) r[i] = src[q];
          }
        }
        float4* lo = reinterpret_cast<float4*>(sLo + (size_t)hr.idx * WT_Q_BYTES);
        if (!p.single) {
#pragma unroll
          for (int i = 0; i < NIT; ++i) {
            const int q = ct + i * WT_CONV_THREADS;
            if (q < nchunks) lo[q] = split_lo(r[i]);
          }
        }
        fence_async_smem();
        mbar_arrive(&bars.ready[hr.idx]);
        hr.advance();
      }
    }
  } else if (warp == WT_LOAD_WARP) {
    // ------------------------------------------------------------------ loader
    RingPos<WT_HI_SLOTS> hr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const int nqb = wg_q_blocks(p.g[group]);
      const uint32_t bytes = (uint32_t)(WT_P_BYTES + wg_q_bytes(p.g[group]));
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.free_[hr.idx], hr.phase ^ 1u);
        if (elect_one()) {
          uint64_t* bar = &bars.landed[hr.idx];
          mbar_arrive_expect_tx(bar, bytes);
          unsigned char* dst = sHi + (size_t)hr.idx * WT_HI_SLAB;
          const int row = (int)row0 + s * 16;
          tma_load_3d(dst, &p.pmap[group], row & 127, 0, row >> 7, bar);
          if (p.g[group].qk) tma_load_3d(dst + WT_P_BYTES, &p.qmap[group], row & 127, 0, row >> 7, bar);
          else
            for (int b = 0; b < nqb; ++b) tma_load_2d(dst + WT_P_BYTES + b * WG_BLOCK_BYTES, &p.qmap[group], 32 * b, row, bar);
        }
        __syncwarp();
        hr.advance();
      }
    }
  } else if (warp == WT_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer (A from tensor memory)
    const uint32_t desc_hi = (512u >> 4) | (1u << 14) | (1u << 29);          // MN-major Q, SWIZZLE_128B_BASE32B
    const uint32_t kdesc_hi = (512u >> 4) | (1u << 14) | (4u << 29);         // K-major Q, SWIZZLE_64B
    const uint32_t lbo = (2048u >> 4) << 16;
    const uint32_t hi0 = (((smem_u32(sHi) + WT_P_BYTES) & 0x3FFFFu) >> 4) | lbo;
    const uint32_t lo0 = ((smem_u32(sLo) & 0x3FFFFu) >> 4) | lbo;
    constexpr uint32_t HSLAB = WT_HI_SLAB >> 4, LSLAB = WT_Q_BYTES >> 4, KS = 1024u >> 4;
    RingPos<WT_HI_SLOTS> hr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const bool qk = p.g[group].qk != 0;
      const uint32_t idesc = umma_idesc_tf32_mn(128, wg_mma_n(p.g[group]), false, !qk);
      const uint32_t q_hi = qk ? kdesc_hi : desc_hi;
      const uint32_t q_fix = qk ? ((1u << 16) - lbo) : 0u;
      const uint32_t q_ks = qk ? (32u >> 4) : KS;
      const int buf = k & 1;
      mbar_wait(&bars.acc_free[buf], (uint32_t)(((k >> 1) & 1) ^ 1));
      tc_fence_after();
      const uint32_t acc = tmem + (uint32_t)buf * 128u;
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.ready[hr.idx], hr.phase);         // hi / lo of P in tensor memory, lo of Q written (its hi landed)
        tc_fence_after();
        if (elect_one()) {
          const uint32_t a_hi = tmem + WT_A_BASE + (uint32_t)hr.idx * 32u, a_lo = a_hi + 16u;
          const uint32_t qh = hi0 + (uint32_t)hr.idx * HSLAB + q_fix, ql = lo0 + (uint32_t)hr.idx * LSLAB + q_fix;
          const uint32_t first = s == 0 ? 0u : 1u;
          if (!p.single) {
            umma_tf32_ts(acc, a_lo, desc64(q_hi, qh), idesc, first);  // small terms first
            umma_tf32_ts(acc, a_hi, desc64(q_hi, ql), idesc, 1u);
            umma_tf32_ts(acc, a_hi, desc64(q_hi, qh), idesc, 1u);
            umma_tf32_ts(acc, a_lo + 8u, desc64(q_hi, qh + q_ks), idesc, 1u);
            umma_tf32_ts(acc, a_hi + 8u, desc64(q_hi, ql + q_ks), idesc, 1u);
            umma_tf32_ts(acc, a_hi + 8u, desc64(q_hi, qh + q_ks), idesc, 1u);
          } else {
            umma_tf32_ts(acc, a_hi, desc64(q_hi, qh), idesc, first);
            umma_tf32_ts(acc, a_hi + 8u, desc64(q_hi, qh + q_ks), idesc, 1u);
          }
          umma_commit(&bars.free_[hr.idx]);               // shared-memory slot, its lo slot and the TMEM A slot of this index
          if (s == nslab - 1) umma_commit(&bars.acc_full[buf]);
        }
        __syncwarp();
        hr.advance();
      }
    }
  } else if (warp >= WT_FILL_WARP0) {
    // ------------------------------------------------------------------ fill: P box -> TMEM A ring (thread = feature)
    const int q = warp & 3;
    const int m = q * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16) + WT_A_BASE;
    // K-major SWIZZLE_64B box: row m at byte 64 m, its 16-byte chunk c (Gaussians 4c..4c+3) at position c ^ ((m >> 1) & 3)
    const int sw = (m >> 1) & 3;
    RingPos<WT_HI_SLOTS> hr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      for (int s = 0; s < nslab; ++s) {
        // the loader refills slot i only after the UMMAs that read A slot i have completed (free_), so a landed box
        // also says that the TMEM columns of this slot are free
        mbar_wait(&bars.landed[hr.idx], hr.phase);
        const float4* rowp = reinterpret_cast<const float4*>(sHi + (size_t)hr.idx * WT_HI_SLAB + m * 64);
        float v[16];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const float4 x = rowp[c ^ sw];
          v[4 * c] = x.x; v[4 * c + 1] = x.y; v[4 * c + 2] = x.z; v[4 * c + 3] = x.w;
        }
        tc_fence_after();
        const uint32_t a = lane_base + (uint32_t)hr.idx * 32u;
        tmem_st16(a, v);
        if (!p.single) {
          float lo[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) lo[j] = v[j] - trunc_tf32(v[j]);
          tmem_st16(a + 16u, lo);
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars.ready[hr.idx]);
        hr.advance();
      }
    }
  } else {
    // ------------------------------------------------------------------ flush (warps 0..3)
    const int q = warp & 3;
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
        for (int blk = 0; blk < nblk32; ++blk) {
          float v[32];
          tmem_ld32(lane_base + (uint32_t)(buf * 128 + blk * 32), v);
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
