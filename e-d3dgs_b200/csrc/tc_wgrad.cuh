// Persistent, warp-specialised 3xTF32 weight-gradient GEMM (sm_100a, tcgen05 + TMEM): contraction
// over Gaussians.
//
//   D[m][n] = sum_g P[g][m] * Q[g][n]        m < 128 features of P (TMEM lanes), n < 256 features of Q
//
// P and Q are activation matrices stored [Gaussian][feature], i.e. the contraction index is the slow
// one, and the UMMA wants it contiguous (K-major; MN-major tf32 descriptors without swizzle return
// zeros on B200). The stagers therefore transpose on the way:
//
//   global --cp.async 16 B, whole feature rows (512 B - 1 KB contiguous per Gaussian)--> raw ring [16 g][384]
//   raw ring --4 x ld.shared.b32 down a column = one float4 of 4 Gaussians--> split hi/lo -->
//            UMMA K-major stages P [128 x 16 g], Q [256 x 16 g]          (all accesses conflict-free)
//
// Work item = (slice of WG_SLICE Gaussians, group); a group is one (P, Q) pair, e.g. the trunk
// activations against the pre-activation gradients of two heads side by side (N = 256 keeps the
// tensor pipe at full rate, see tc_rows.cuh). One CTA per SM walks items blockIdx.x,
// blockIdx.x + gridDim.x, ...; groups of a slice are adjacent so they share P through L2.
//
//   warps 9-16  stagers     as above, raw ring of 3 slabs, per-thread cp.async groups
//   warp 8      MMA issuer  per 16-Gaussian slab: 2 k-steps x (P_lo*Q_hi + P_hi*Q_lo + P_hi*Q_hi)
//   warps 0-7   epilogue    TMEM (2 accumulators, ping-pong) -> red.global.add.f32 into the weight
//                           gradients; one flush per item, so atomics are ~1/128 of the old per-slice scheme
#pragma once

namespace ed3dgs {

// hi = x rounded to tf32 (round half away from zero on the bit pattern: two integer instructions,
// where cvt.rna.tf32 costs four because of its NaN/Inf path), lo = x - hi (exact). The tensor core
// ignores the low 13 bits of lo, which perturbs a product by < 2^-21 relative.
__device__ __forceinline__ float tf32_round_bits(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void split_store_fast(float4 x, float4* hi, float4* lo) {
  float4 h;
  h.x = tf32_round_bits(x.x); h.y = tf32_round_bits(x.y); h.z = tf32_round_bits(x.z); h.w = tf32_round_bits(x.w);
  *hi = h;
  *lo = make_float4(x.x - h.x, x.y - h.y, x.z - h.z, x.w - h.w);
}

constexpr int WG_SLICE = 2048;                    // Gaussians per work item (tensor-core accumulation chain)
constexpr int WG_MAX_GROUPS = 3;
constexpr int WG_EPI_WARPS = 8;
constexpr int WG_MMA_WARP = 8;
constexpr int WG_CONV_WARP0 = 9;
constexpr int WG_CONV_THREADS = 256;
constexpr int WG_THREADS = (WG_CONV_WARP0 + 8) * 32;
constexpr int WG_STAGES = 3;
constexpr int WG_P_STAGE_BYTES = 128 * 16 * 4 * 2;  // hi + lo of [128 features x 16 g]
constexpr int WG_Q_STAGE_BYTES = 256 * 16 * 4 * 2;  // hi + lo of [256 features x 16 g]
constexpr int WG_RAW_SLOTS = 3;
constexpr int WG_RAW_LD = 384;                      // floats per Gaussian row of a raw slab: P 128 | Q 256
constexpr int WG_RAW_SLOT_BYTES = 16 * WG_RAW_LD * 4;
constexpr size_t WG_SMEM_BYTES =
    (size_t)WG_STAGES * (WG_P_STAGE_BYTES + WG_Q_STAGE_BYTES) + (size_t)WG_RAW_SLOTS * WG_RAW_SLOT_BYTES;

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
};
struct WgArgs {
  WgGroup g[WG_MAX_GROUPS];
  int ngroups;
  int N;            // Gaussians
  int dbg;          // 2 = single pass (hi*hi only)
};

struct WgBars {
  uint64_t x_full[WG_STAGES], x_free[WG_STAGES];
  uint64_t acc_full[2], acc_free[2];
};

__global__ void __launch_bounds__(WG_THREADS, 1) tc_wgrad_kernel(const __grid_constant__ WgArgs p) {
  extern __shared__ __align__(128) unsigned char wg_smem[];
  __shared__ WgBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sP = wg_smem;
  unsigned char* sQ = sP + (size_t)WG_STAGES * WG_P_STAGE_BYTES;
  unsigned char* sRaw = sQ + (size_t)WG_STAGES * WG_Q_STAGE_BYTES;

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < WG_STAGES; ++i) { mbar_init(&bars.x_full[i], WG_CONV_THREADS); mbar_init(&bars.x_free[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], WG_EPI_WARPS * 32); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int nslices = (p.N + WG_SLICE - 1) / WG_SLICE;
  const int nitems = nslices * p.ngroups;
  const int my_items = (nitems - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  // item k of this CTA -> group, first Gaussian, number of 16-Gaussian slabs
  auto item_setup = [&](int k, int& group, long& row0, int& nslab) {
    const int item = (int)blockIdx.x + k * (int)gridDim.x;
    const int slice = item / p.ngroups;
    group = item - slice * p.ngroups;
    row0 = (long)slice * WG_SLICE;
    const int rows = (int)min((long)WG_SLICE, (long)p.N - row0);
    nslab = (rows + 15) >> 4;
  };

  if (warp >= WG_CONV_WARP0) {
    // ------------------------------------------------------------------ stagers (256 threads)
    const int ct = tid - WG_CONV_WARP0 * 32;
    // copy mapping: Gaussian row ct / 16 of the slab, 16-byte chunks (ct % 16) + 16 i, i < 6:
    //   chunks 0..31 = P columns, 32..95 = Q columns -> 256 contiguous bytes per half-warp
    const int cg = ct >> 4, cw = ct & 15;
    // convert mapping: feature ct % 128 of a 128-feature block, Gaussian quads ct / 128 and ct / 128 + 2
    const int f = ct & 127, gq0 = ct >> 7;
    float* const raw = reinterpret_cast<float*>(sRaw);

    int ik = 0, is_ = 0, igroup = 0, inslab = 0, islot = 0;
    long irow0 = 0;
    if (my_items > 0) item_setup(0, igroup, irow0, inslab);
    auto issue = [&]() {
      const WgGroup& G = p.g[igroup];
      const long grow = irow0 + is_ * 16 + cg;
      const bool rok = grow < (long)p.N;
      float* dst = raw + islot * (WG_RAW_SLOT_BYTES / 4) + cg * WG_RAW_LD;
      const float* prow = G.P + grow * G.ldp;
      const float* qrow = G.Q + grow * G.ldq;
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        const int w = cw + 16 * i;
        const int col = (i < 2) ? 4 * w : 4 * (w - 32);
        const bool ok = rok && col < ((i < 2) ? G.pcols : G.qcols);
        const float* src = (i < 2) ? prow + col : qrow + col;
        cp_async16_zfill(dst + 4 * w, ok ? src : G.P, ok ? 16 : 0);
      }
      if (++islot == WG_RAW_SLOTS) islot = 0;
      if (++is_ == inslab) {
        is_ = 0;
        if (++ik < my_items) item_setup(ik, igroup, irow0, inslab);
      }
    };
    for (int u = 0; u < WG_RAW_SLOTS - 1; ++u) {
      if (ik < my_items) issue();
      cp_async_commit_group();
    }
    RingPos<WG_STAGES> xr;
    int rslot = 0;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const int nblk = 1 + ((p.g[group].qcols + 127) >> 7);
      for (int s = 0; s < nslab; ++s) {
        cp_async_wait_group<WG_RAW_SLOTS - 2>();        // this thread's pieces of the current slab landed
        // The raw slab is read across threads (a column holds pieces copied by 4 different threads), so
        // all stagers meet here: everybody's copies of this slab are visible, and everybody is done
        // reading the previous slab, whose slot the next copy refills.
        asm volatile("bar.sync 1, %0;" ::"n"(WG_CONV_THREADS) : "memory");
        if (ik < my_items) issue();
        cp_async_commit_group();                        // always commit: keeps the group count uniform
        const float* src = raw + rslot * (WG_RAW_SLOT_BYTES / 4) + f;
        mbar_wait(&bars.x_free[xr.idx], xr.phase ^ 1u);
        tc_fence_after();
        float4* pst = reinterpret_cast<float4*>(sP + (size_t)xr.idx * WG_P_STAGE_BYTES);
        float4* qst = reinterpret_cast<float4*>(sQ + (size_t)xr.idx * WG_Q_STAGE_BYTES);
        for (int b = 0; b < nblk; ++b) {
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int gq = gq0 + 2 * j;
            const float* c0 = src + (4 * gq) * WG_RAW_LD + b * 128;
            const float4 x = make_float4(c0[0], c0[WG_RAW_LD], c0[2 * WG_RAW_LD], c0[3 * WG_RAW_LD]);
            if (b == 0) split_store_fast(x, pst + gq * 128 + f, pst + 4 * 128 + gq * 128 + f);
            else split_store_fast(x, qst + gq * 256 + (b - 1) * 128 + f, qst + 4 * 256 + gq * 256 + (b - 1) * 128 + f);
          }
        }
        if (!FENCE_AT_CONSUMER) fence_async_smem();
        mbar_arrive(&bars.x_full[xr.idx]);
        xr.advance();
        if (++rslot == WG_RAW_SLOTS) rslot = 0;
      }
    }
    cp_async_wait_group<0>();
  } else if (warp == WG_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);
    const uint32_t p_lo0 = ((smem_u32(sP) & 0x3FFFFu) >> 4) | ((128u * 16u >> 4) << 16);
    const uint32_t q_lo0 = ((smem_u32(sQ) & 0x3FFFFu) >> 4) | ((256u * 16u >> 4) << 16);
    constexpr uint32_t P_STAGE = WG_P_STAGE_BYTES >> 4, P_LO = P_STAGE / 2, P_KS = (2u * 128u * 16u) >> 4;
    constexpr uint32_t Q_STAGE = WG_Q_STAGE_BYTES >> 4, Q_LO = Q_STAGE / 2, Q_KS = (2u * 256u * 16u) >> 4;
    const bool single = (p.dbg & 2) != 0;
    RingPos<WG_STAGES> xr;
    for (int k = 0; k < my_items; ++k) {
      int group, nslab;
      long row0;
      item_setup(k, group, row0, nslab);
      const uint32_t idesc = umma_idesc_tf32(128, p.g[group].qcols > 128 ? 256 : 128);
      const int buf = k & 1;
      mbar_wait(&bars.acc_free[buf], (uint32_t)(((k >> 1) & 1) ^ 1));
      const uint32_t acc = tmem + (uint32_t)buf * 256u;
      for (int s = 0; s < nslab; ++s) {
        mbar_wait(&bars.x_full[xr.idx], xr.phase);
        tc_fence_after();
        if (FENCE_AT_CONSUMER) fence_async_smem();   // generic-proxy stage writes (acquired above) -> async proxy
        if (elect_one()) {
          const uint32_t ph = p_lo0 + (uint32_t)xr.idx * P_STAGE, qh = q_lo0 + (uint32_t)xr.idx * Q_STAGE;
          const uint32_t first = s == 0 ? 0u : 1u;
          if (single) {
            umma_tf32(acc, desc64(desc_hi, ph), desc64(desc_hi, qh), idesc, first);
            umma_tf32(acc, desc64(desc_hi, ph + P_KS), desc64(desc_hi, qh + Q_KS), idesc, 1u);
          } else {
            umma_tf32(acc, desc64(desc_hi, ph + P_LO), desc64(desc_hi, qh), idesc, first);  // small terms first
            umma_tf32(acc, desc64(desc_hi, ph), desc64(desc_hi, qh + Q_LO), idesc, 1u);
            umma_tf32(acc, desc64(desc_hi, ph), desc64(desc_hi, qh), idesc, 1u);
            umma_tf32(acc, desc64(desc_hi, ph + P_KS + P_LO), desc64(desc_hi, qh + Q_KS), idesc, 1u);
            umma_tf32(acc, desc64(desc_hi, ph + P_KS), desc64(desc_hi, qh + Q_KS + Q_LO), idesc, 1u);
            umma_tf32(acc, desc64(desc_hi, ph + P_KS), desc64(desc_hi, qh + Q_KS), idesc, 1u);
          }
          umma_commit(&bars.x_free[xr.idx]);
          if (s == nslab - 1) umma_commit(&bars.acc_full[buf]);
        }
        __syncwarp();
        xr.advance();
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
