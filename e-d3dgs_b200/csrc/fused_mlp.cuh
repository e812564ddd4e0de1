// Stage-resident deformation MLP (sm_100a, tcgen05 + TMEM): one persistent kernel per stage and direction.
//
// Replaces scene/deformation.py:72-106 of the reference (query_time -> feature_out -> the five heads) for the
// shape every dataset config uses (net_width 128, defor_depth <= 1, 32-wide Gaussian embedding).
//
// Orientation: the 128 Gaussians of a tile sit on the UMMA M side = the 128 TMEM lanes, output features on
// the N side. An accumulator D[g][f] is then exactly the layout the tensor core accepts as the A operand of
// the NEXT layer (A from tensor memory, K = f along the columns; tests/probes/ts_probe.cu: bit-exact, and a
// 128 x 128 x 8 tf32 MMA takes 73.5 clocks against 104.5 with both operands in shared memory), so the chain
//      E --L0--> h --W1_i--> z_i --W2_i--> delta_i          never leaves the SM:
// the epilogue warps read an accumulator (tcgen05.ld), add bias, ReLU, split hi/lo and write the operand back
// (tcgen05.st); only the weights stream through shared memory (prepared hi/lo slab images, cp.async.bulk).
// 3xTF32 as everywhere (tc_gemm.cuh): A_lo*B_hi + A_hi*B_lo + A_hi*B_hi with A_hi the raw fp32 word.
//
// TMEM (512 columns):  [0,128) ACC0  [128,256) ACC1  [256,384) H_hi / OUT  [384,512) H_lo / Z4_lo
#pragma once
#include "tc_gemm.cuh"
#include "tc_rows.cuh"

namespace ed3dgs {

constexpr int FM_TG = 128;                       // Gaussians per tile = UMMA M
constexpr int FM_EPI_WARPS = 16;                 // warp w: TMEM lanes 32*(w&3).., column quarter w>>2
constexpr int FM_MMA_WARP = 16;
constexpr int FM_WLOAD_WARP = 17;
constexpr int FM_ELOAD_WARP = 18;
constexpr int FM_THREADS = 19 * 32;
constexpr int FM_W_SLOTS = 8;
constexpr int FMF_W_SLOTS = 7;                   // forward kernel: one slot less, the space holds the head-output partial sums
constexpr int FM_W_SLOT_BYTES = 128 * 16 * 4 * 2;   // hi + lo of [128 rows x 16 k], K-major no-swizzle image
constexpr int FM_E_SLAB_BYTES = 128 * 16 * 4;       // [128 Gaussians x 16 k], 64-byte rows, SWIZZLE_64B (TMA)
constexpr int FM_E_BYTES = 2 * FM_E_SLAB_BYTES;     // K = 32
constexpr uint32_t FM_ACC0 = 0, FM_H_HI = 256, FM_H_LO = 384;
constexpr int FM_MAX_SLABS = 64;

// ---- weight images -------------------------------------------------------------------------------------
// One slab = 16 consecutive k of a weight matrix for NT output rows: [hi: (k/4)*NT + row][lo: same] float4,
// NT*128 bytes, the layout the UMMA reads as a K-major no-swizzle operand (LBO = NT*16, SBO = 128).
struct PrepSlab {
  const float* src;
  long ld;
  int maj;        // TC_K_MAJOR: element (row, k) at src[row*ld + k];  TC_MN_MAJOR: src[k*ld + row]
  int n_valid;    // rows that exist (others are zero)
  int NT;         // rows of the image
  int k0, kvalid; // first k of the slab, number of k that exist in the matrix
  uint32_t dst_f4;
};
struct PrepArgs {
  PrepSlab s[FM_MAX_SLABS];
  float4* dst;
};

__global__ void __launch_bounds__(128) fused_prepare(const __grid_constant__ PrepArgs p) {
  const PrepSlab& sl = p.s[blockIdx.x];
  const int NT = sl.NT;
  float4* img = p.dst + sl.dst_f4;
  for (int q = threadIdx.x; q < NT * 4; q += 128) {
    const int c4 = q / NT, row = q - c4 * NT;
    const int k = sl.k0 + c4 * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < sl.n_valid) {
      float* xs = &x.x;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (k + i < sl.kvalid)
          xs[i] = sl.maj == TC_K_MAJOR ? sl.src[(long)row * sl.ld + k + i] : sl.src[(long)(k + i) * sl.ld + row];
    }
    split_store(x, img + q, img + NT * 4 + q);
  }
}

// ---- PTX wrappers the layered kernels do not need -----------------------------------------------------------
__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate)
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
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
      "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
      "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
      "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
      "r"(__float_as_uint(v[15]))
      : "memory");
}
// lo part (x - trunc_tf32(x)) of 32 values to 32 TMEM columns, 16 at a time (keeps the register peak at 48)
__device__ __forceinline__ void tmem_st_lo32(uint32_t taddr, const float* v) {
  float lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) lo[j] = v[j] - __uint_as_float(__float_as_uint(v[j]) & 0xffffe000u);
  tmem_st16(taddr, lo);
#pragma unroll
  for (int j = 0; j < 16; ++j) lo[j] = v[16 + j] - __uint_as_float(__float_as_uint(v[16 + j]) & 0xffffe000u);
  tmem_st16(taddr + 16u, lo);
}
__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 512;" ::: "memory"); }
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// =============================================================================================================
// Forward of one stage.
struct FusedFwdArgs {
  const float4* wimg;            // slab images: [0,2) layer 0, [2 + 8*s, +8) hidden layer of slot s, then 8 slabs of the
  uint32_t slab_off[2 + 8 * 6];  //               rgb output layer (NT = 48); float4 offsets
  int na, has_rgb;               // active heads in slot order; the rgb head, when active, is the last slot
  int N;
  int single;                    // precision "tf32": one tf32 MMA per product instead of the 3xTF32 triple
  long npad;                     // N rounded up to whole tiles
  const float* c1;               // [128] layer-0 bias with the temporal half of the input folded in
  const float* b1[5];            // per slot
  const float* w2[5];            // [k][128] fp32 (small heads are evaluated in the epilogue)
  const float* b2[5];
  int hdim[5];
  const float* base[5];
  float* dst[5];
  float coef[5];
  float* zT;                     // [tile][na*128][128]  post-ReLU hidden activations of the heads, feature-major per tile
  float* hT;                     // [tile][128][128]     post-ReLU trunk activations
                                 // (columns of Gaussians >= N hold finite junk: every consumer multiplies them by a zero gradient)
  uint32_t* Mg;                  // [npad][na][4]   ReLU masks of the heads, bit = feature % 32
  uint32_t* MgH;                 // [npad][4]       ReLU mask of the trunk
  alignas(64) CUtensorMap emap;  // embedding [N][32]: box {16, 128}, SWIZZLE_64B
};

struct FwdBars {
  uint64_t w_full[FMF_W_SLOTS], w_free[FMF_W_SLOTS];
  uint64_t part[3][4];           // partial sums of a small head parked by the four column-quarter warps of a lane quarter
  uint64_t e_tma[2], e_full[2], e_free[2];
  uint64_t acc_full[2], acc_free[2];
  uint64_t h_ready, z4_ready, out_full;
};

constexpr size_t FM_FWD_SMEM = (size_t)FMF_W_SLOTS * FM_W_SLOT_BYTES + 4 * FM_E_BYTES + 5 * 128 * 16 +
                               (128 + 5 * 128 + 5 * 48) * 4 + 3 * 4 * 128 * 16 + 1024;

__global__ void __launch_bounds__(FM_THREADS, 1) fused_fwd_kernel(const __grid_constant__ FusedFwdArgs p) {
  extern __shared__ __align__(128) unsigned char fm_smem_raw[];
  __shared__ FwdBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sW = fm_smem_raw + ((1024u - (smem_u32(fm_smem_raw) & 1023u)) & 1023u);
  unsigned char* sEh = sW + (size_t)FMF_W_SLOTS * FM_W_SLOT_BYTES;
  unsigned char* sEl = sEh + 2 * FM_E_BYTES;
  float4* s_w2 = reinterpret_cast<float4*>(sEl + 2 * FM_E_BYTES);   // [slot][feature] -> (w2[0..3][feature])
  float* s_c1 = reinterpret_cast<float*>(s_w2 + 5 * 128);
  float* s_b1 = s_c1 + 128;                                         // [slot][128]
  float* s_b2 = s_b1 + 5 * 128;                                     // [slot][48]
  float4* s_part = reinterpret_cast<float4*>(s_b2 + 5 * 48);        // [3][4 column quarters][128 rows]: small-head partial outputs

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < FMF_W_SLOTS; ++i) { mbar_init(&bars.w_full[i], 1); mbar_init(&bars.w_free[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars.e_tma[i], 1); mbar_init(&bars.e_full[i], 32); mbar_init(&bars.e_free[i], 1);
      mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], FM_EPI_WARPS * 32);
    }
    mbar_init(&bars.h_ready, FM_EPI_WARPS * 32);
    mbar_init(&bars.z4_ready, FM_EPI_WARPS * 32);
    mbar_init(&bars.out_full, 1);
    for (int i = 0; i < 12; ++i) mbar_init(&bars.part[i / 4][i & 3], 128);
    fence_mbar_init();
  }
  for (int i = tid; i < 5 * 128; i += FM_THREADS) {
    const int s = i >> 7, f = i & 127;
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (s < p.na && !(p.has_rgb && s == p.na - 1)) {
      const int k = p.hdim[s];
      w.x = p.w2[s][f];
      if (k > 1) w.y = p.w2[s][128 + f];
      if (k > 2) w.z = p.w2[s][256 + f];
      if (k > 3) w.w = p.w2[s][384 + f];
    }
    s_w2[i] = w;
    s_b1[i] = s < p.na ? p.b1[s][f] : 0.f;
  }
  for (int i = tid; i < 128; i += FM_THREADS) s_c1[i] = p.c1[i];
  for (int i = tid; i < 5 * 48; i += FM_THREADS) {
    const int s = i / 48, o = i - s * 48;
    s_b2[i] = (s < p.na && o < p.hdim[s]) ? p.b2[s][o] : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int ntiles = (p.N + FM_TG - 1) / FM_TG;
  const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int na = p.na;
  const bool has_rgb = p.has_rgb != 0;

  if (warp == FM_ELOAD_WARP) {
    // ------------------------------------------------------------------ embedding tiles: TMA (hi = raw) + lo split
    for (int tl = 0; tl < my_tiles; ++tl) {
      const int eb = tl & 1;
      const uint32_t ph = (uint32_t)((tl >> 1) & 1);
      const long row = ((long)blockIdx.x + (long)tl * gridDim.x) * FM_TG;
      mbar_wait(&bars.e_free[eb], ph ^ 1u);
      if (elect_one()) {
        mbar_arrive_expect_tx(&bars.e_tma[eb], FM_E_BYTES);
        tma_load_2d(sEh + (size_t)eb * FM_E_BYTES, &p.emap, 0, (int)row, &bars.e_tma[eb]);
        tma_load_2d(sEh + (size_t)eb * FM_E_BYTES + FM_E_SLAB_BYTES, &p.emap, 16, (int)row, &bars.e_tma[eb]);
      }
      __syncwarp();
      mbar_wait(&bars.e_tma[eb], ph);
      const float4* hi = reinterpret_cast<const float4*>(sEh + (size_t)eb * FM_E_BYTES);
      float4* lo = reinterpret_cast<float4*>(sEl + (size_t)eb * FM_E_BYTES);
      if (!p.single) {
#pragma unroll 8
        for (int i = 0; i < FM_E_BYTES / 16 / 32; ++i) lo[lane + 32 * i] = split_lo(hi[lane + 32 * i]);
      }
      fence_async_smem();
      mbar_arrive(&bars.e_full[eb]);
    }
  } else if (warp == FM_WLOAD_WARP) {
    // ------------------------------------------------------------------ weight slabs, in the order the MMA warp uses them
    RingPos<FMF_W_SLOTS> wr;
    auto push = [&](int slab, uint32_t bytes) {
      mbar_wait(&bars.w_free[wr.idx], wr.phase ^ 1u);
      if (elect_one()) {
        mbar_arrive_expect_tx(&bars.w_full[wr.idx], bytes);
        bulk_copy_g2s(sW + (size_t)wr.idx * FM_W_SLOT_BYTES, p.wimg + p.slab_off[slab], bytes, &bars.w_full[wr.idx]);
      }
      __syncwarp();
      wr.advance();
    };
    if (my_tiles > 0) { push(0, FM_W_SLOT_BYTES); push(1, FM_W_SLOT_BYTES); }
    for (int tl = 0; tl < my_tiles; ++tl) {
      for (int s = 0; s < na * 8; ++s) push(2 + s, FM_W_SLOT_BYTES);
      if (tl + 1 < my_tiles) { push(0, FM_W_SLOT_BYTES); push(1, FM_W_SLOT_BYTES); }
      if (has_rgb)
        for (int s = 0; s < 8; ++s) push(2 + na * 8 + s, 48 * 128);
    }
  } else if (warp == FM_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer (converged warp, elected lane issues)
    const uint32_t idesc128 = umma_idesc_tf32(128, 128), idesc48 = umma_idesc_tf32(128, 48);
    const uint32_t w_desc_hi = (128u >> 4) | (1u << 14);                 // SBO = 128 B between 8-row groups
    const uint32_t e_desc_hi = (512u >> 4) | (1u << 14) | (4u << 29);    // SWIZZLE_64B, 8-row atoms of 512 B
    const uint32_t w128_lo0 = ((smem_u32(sW) & 0x3FFFFu) >> 4) | ((128u * 16u >> 4) << 16);
    const uint32_t w48_lo0 = ((smem_u32(sW) & 0x3FFFFu) >> 4) | ((48u * 16u >> 4) << 16);
    const uint32_t eh_lo0 = ((smem_u32(sEh) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t el_lo0 = ((smem_u32(sEl) & 0x3FFFFu) >> 4) | (1u << 16);
    constexpr uint32_t W_SLOT = FM_W_SLOT_BYTES >> 4;
    constexpr uint32_t W128_LO = (128u * 64u) >> 4, W128_KS = (2u * 128u * 16u) >> 4;
    constexpr uint32_t W48_LO = (48u * 64u) >> 4, W48_KS = (2u * 48u * 16u) >> 4;
    constexpr uint32_t E_SLAB = FM_E_SLAB_BYTES >> 4, E_KS = 32u >> 4;
    RingPos<FMF_W_SLOTS> wr;
    int go = 0;
    const bool single = p.single != 0;

    auto layer0 = [&](int tl) {      // SS: ACC[g][f] = sum_k E[g][k] W0e[f][k], K = 32
      const int eb = tl & 1, buf = go & 1;
      mbar_wait(&bars.acc_free[buf], (uint32_t)(((go >> 1) & 1) ^ 1));
      mbar_wait(&bars.e_full[eb], (uint32_t)((tl >> 1) & 1));
      const uint32_t acc = tmem + FM_ACC0 + (uint32_t)buf * 128u;
      for (int s = 0; s < 2; ++s) {
        mbar_wait(&bars.w_full[wr.idx], wr.phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t wh = w128_lo0 + (uint32_t)wr.idx * W_SLOT;
          const uint32_t eh = eh_lo0 + (uint32_t)(eb * 2 + s) * E_SLAB, el = el_lo0 + (uint32_t)(eb * 2 + s) * E_SLAB;
#pragma unroll
          for (uint32_t ks = 0; ks < 2; ++ks) {
            const uint32_t first = (s | ks) ? 1u : 0u;
            if (!single) {
              umma_tf32(acc, desc64(e_desc_hi, el + ks * E_KS), desc64(w_desc_hi, wh + ks * W128_KS), idesc128, first);
              umma_tf32(acc, desc64(e_desc_hi, eh + ks * E_KS), desc64(w_desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
            }
            umma_tf32(acc, desc64(e_desc_hi, eh + ks * E_KS), desc64(w_desc_hi, wh + ks * W128_KS), idesc128, single ? first : 1u);
          }
          umma_commit(&bars.w_free[wr.idx]);
          if (s == 1) { umma_commit(&bars.e_free[eb]); umma_commit(&bars.acc_full[buf]); }
        }
        __syncwarp();
        wr.advance();
      }
      ++go;
    };

    if (my_tiles > 0) layer0(0);
    for (int tl = 0; tl < my_tiles; ++tl) {
      mbar_wait(&bars.h_ready, (uint32_t)(tl & 1));
      tc_fence_after();
      int rgb_buf = 0;
      for (int hs = 0; hs < na; ++hs) {   // TS: ACC[g][f] = sum_k H[g][k] W1[f][k], K = 128
        const int buf = go & 1;
        rgb_buf = buf;
        mbar_wait(&bars.acc_free[buf], (uint32_t)(((go >> 1) & 1) ^ 1));
        tc_fence_after();
        const uint32_t acc = tmem + FM_ACC0 + (uint32_t)buf * 128u;
        for (int s = 0; s < 8; ++s) {
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w128_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
            for (uint32_t ks = 0; ks < 2; ++ks) {
              const uint32_t col = (uint32_t)s * 16u + ks * 8u;
              const uint32_t first = (s | (int)ks) ? 1u : 0u;
              if (!single) {
                umma_tf32_ts(acc, tmem + FM_H_LO + col, desc64(w_desc_hi, wh + ks * W128_KS), idesc128, first);
                umma_tf32_ts(acc, tmem + FM_H_HI + col, desc64(w_desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
              }
              umma_tf32_ts(acc, tmem + FM_H_HI + col, desc64(w_desc_hi, wh + ks * W128_KS), idesc128, single ? first : 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == 7) umma_commit(&bars.acc_full[buf]);
          }
          __syncwarp();
          wr.advance();
        }
        ++go;
      }
      if (tl + 1 < my_tiles) layer0(tl + 1);   // keeps the tensor pipe busy while the epilogue converts the rgb hidden layer
      if (has_rgb) {                           // TS: OUT[g][o] = sum_f Z4[g][f] W2[o][f], N = 48
        mbar_wait(&bars.z4_ready, (uint32_t)(tl & 1));
        tc_fence_after();
        const uint32_t a_hi = tmem + FM_ACC0 + (uint32_t)rgb_buf * 128u, a_lo = tmem + FM_H_LO, out = tmem + FM_H_HI;
        for (int s = 0; s < 8; ++s) {
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w48_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
            for (uint32_t ks = 0; ks < 2; ++ks) {
              const uint32_t col = (uint32_t)s * 16u + ks * 8u;
              const uint32_t first = (s | (int)ks) ? 1u : 0u;
              if (!single) {
                umma_tf32_ts(out, a_lo + col, desc64(w_desc_hi, wh + ks * W48_KS), idesc48, first);
                umma_tf32_ts(out, a_hi + col, desc64(w_desc_hi, wh + W48_LO + ks * W48_KS), idesc48, 1u);
              }
              umma_tf32_ts(out, a_hi + col, desc64(w_desc_hi, wh + ks * W48_KS), idesc48, single ? first : 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == 7) umma_commit(&bars.out_full);
          }
          __syncwarp();
          wr.advance();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 0..15)
    // warp w: Gaussians (TMEM lanes) 32*(w&3).., features (columns) 32*(w>>2)..: one 32 x 32 block of every accumulator
    const int q = warp & 3, cq = warp >> 2;
    const int col = cq * 32;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    int go = 0, np = 0;
    for (int tl = 0; tl < my_tiles; ++tl) {
      const long tile = (long)blockIdx.x + (long)tl * gridDim.x;
      const int gl = q * 32 + lane;
      const long g = tile * FM_TG + gl;
      const bool valid = g < (long)p.N;
      if (tl + 1 < my_tiles) {   // next tile's input rows: into L2 now
        const long gn = g + (long)gridDim.x * FM_TG;
        if (gn < (long)p.N) {
          if (cq < na && !(has_rgb && cq == na - 1)) prefetch_l2(p.base[cq] + gn * p.hdim[cq]);
          if (has_rgb && cq < 2) prefetch_l2(p.base[na - 1] + gn * 48 + cq * 32);
        }
      }
      // the previous tile's rgb output MMAs read H_lo and write H_hi: they must be done before H is rewritten
      if (has_rgb && tl > 0) { mbar_wait(&bars.out_full, (uint32_t)((tl - 1) & 1)); tc_fence_after(); }
      {  // ---- layer 0: h = relu(acc + c1) -> H_hi / H_lo (A operand of the five hidden layers), hT, trunk mask
        const int buf = go & 1;
        mbar_wait(&bars.acc_full[buf], (uint32_t)((go >> 1) & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + FM_ACC0 + (uint32_t)(buf * 128 + col), v);
        const float4* c4 = reinterpret_cast<const float4*>(s_c1 + col);
        float* hp = p.hT + (tile * 128 + col) * 128 + gl;
        uint32_t word = 0;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c = c4[j4];
          const float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int j = j4 * 4 + i;
            const float h = fmaxf(v[j] + cc[i], 0.f);
            v[j] = h;
            word |= (h > 0.f ? 1u : 0u) << j;
          }
        }
        tmem_st32(lane_base + FM_H_HI + (uint32_t)col, v);
        if (!p.single) tmem_st_lo32(lane_base + FM_H_LO + (uint32_t)col, v);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars.h_ready);
        mbar_arrive(&bars.acc_free[buf]);
        ++go;
#pragma unroll
        for (int j = 0; j < 32; ++j) hp[j * 128] = v[j];   // off the critical path: the hidden-layer UMMAs are running
        if (valid) p.MgH[g * 4 + cq] = word;
      }
      // ---- hidden layer of every head: z = relu(acc + b1) -> zT, ReLU mask; the small heads' output layer
      //      (k <= 4) is evaluated right here on the FP32 pipe, the rgb head's z goes back to TMEM for its UMMA
#pragma unroll 1
      for (int hs = 0; hs < na; ++hs) {
        const bool is_rgb = has_rgb && hs == na - 1;
        const int buf = go & 1;
        mbar_wait(&bars.acc_full[buf], (uint32_t)((go >> 1) & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + FM_ACC0 + (uint32_t)(buf * 128 + col), v);
        const float4* b4 = reinterpret_cast<const float4*>(s_b1 + hs * 128 + col);
        uint32_t word = 0;
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
          const float4 c = b4[j4];
          const float cc[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int j = j4 * 4 + i;
            const float z = fmaxf(v[j] + cc[i], 0.f);
            v[j] = z;
            word |= (z > 0.f ? 1u : 0u) << j;
          }
        }
        if (is_rgb) {
          tmem_st32(lane_base + FM_ACC0 + (uint32_t)(buf * 128 + col), v);   // Z4_hi in place
          if (!p.single) tmem_st_lo32(lane_base + FM_H_LO + (uint32_t)col, v);   // the trunk operand is dead by now
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(&bars.z4_ready);
        }
        tc_fence_before();
        mbar_arrive(&bars.acc_free[buf]);
        ++go;
        float* zp = p.zT + ((tile * na + hs) * 128 + col) * 128 + gl;
#pragma unroll
        for (int j = 0; j < 32; ++j) zp[j * 128] = v[j];
        if (valid) p.Mg[(g * na + hs) * 4 + cq] = word;
        if (!is_rgb) {
          const int k = p.hdim[hs];
          const bool fin = cq == (hs & 3);      // this warp finishes the head: its base row is requested before the FMAs
          float bsv[4] = {0.f, 0.f, 0.f, 0.f};
          if (fin && valid) {
            const float* bs = p.base[hs] + g * k;
#pragma unroll
            for (int o = 0; o < 4; ++o)
              if (o < k) bsv[o] = bs[o];
          }
          const float4* w2 = s_w2 + hs * 128 + col;
          float o4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float4 w = w2[j];
            o4[0] = fmaf(v[j], w.x, o4[0]);
            o4[1] = fmaf(v[j], w.y, o4[1]);
            o4[2] = fmaf(v[j], w.z, o4[2]);
            o4[3] = fmaf(v[j], w.w, o4[3]);
          }
          // the four column quarters of a Gaussian meet in shared memory; the quarter (hs & 3) warp finishes the head:
          // x = base + (delta + b2) * coef. Three buffers: a buffer is rewritten three heads later, by which time its
          // reader has gone through two more accumulator hand-overs that every warp takes part in.
          const int pb = np % 3;
          const uint32_t pph = (uint32_t)((np / 3) & 1);
          ++np;
          s_part[(pb * 4 + cq) * 128 + gl] = make_float4(o4[0], o4[1], o4[2], o4[3]);
          mbar_arrive(&bars.part[pb][q]);
          if (fin) {
            mbar_wait(&bars.part[pb][q], pph);
            float t[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const float4 o2 = s_part[(pb * 4 + c) * 128 + gl];
              t[0] += o2.x; t[1] += o2.y; t[2] += o2.z; t[3] += o2.w;
            }
            if (valid) {
              const float cf = p.coef[hs];
              float* ds = p.dst[hs] + g * k;
#pragma unroll
              for (int o = 0; o < 4; ++o)
                if (o < k) ds[o] = bsv[o] + (t[o] + s_b2[hs * 48 + o]) * cf;
            }
          }
        }
      }
      if (has_rgb && cq < 2) {
        // rgb accumulator: columns [0,32) by the quarter-0 warps, [32,48) by the quarter-1 warps -- the warps that
        // rewrite those H_hi columns in the next tile's layer-0 epilogue
        const int s = na - 1;
        float4 bsr[8];                     // the base row is requested before the wait for the output UMMAs
#pragma unroll
        for (int o = 0; o < 8; ++o) bsr[o] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid) {
          const float4* bs = reinterpret_cast<const float4*>(p.base[s] + g * 48 + cq * 32);
#pragma unroll
          for (int o = 0; o < 8; ++o)
            if (cq == 0 || o < 4) bsr[o] = bs[o];
        }
        mbar_wait(&bars.out_full, (uint32_t)(tl & 1));
        tc_fence_after();
        float v[32];
        if (cq == 0) tmem_ld32(lane_base + FM_H_HI, v);
        else tmem_ld16(lane_base + FM_H_HI + 32u, v);
        if (valid) {
          float4* ds = reinterpret_cast<float4*>(p.dst[s] + g * 48 + cq * 32);
          const float cf = p.coef[s];
          const float4* b2 = reinterpret_cast<const float4*>(s_b2 + s * 48 + cq * 32);
#pragma unroll
          for (int o = 0; o < 8; ++o) {
            if (cq == 1 && o >= 4) break;
            float4 b = bsr[o];
            const float4 c = b2[o];
            b.x += (v[4 * o + 0] + c.x) * cf;
            b.y += (v[4 * o + 1] + c.y) * cf;
            b.z += (v[4 * o + 2] + c.z) * cf;
            b.w += (v[4 * o + 3] + c.w) * cf;
            ds[o] = b;
          }
        }
        tc_fence_before();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

// =============================================================================================================
// Backward of one stage, the per-Gaussian chain:
//      upstream gradients --W2_i--> dz_i (masked) --W1_i--> dh (masked) --W0e--> dE
// Same orientation as the forward (Gaussians on the TMEM lanes). The gradient of a small head's hidden layer is
// k <= 4 FMAs per element and is formed in registers; the rgb head's (k = 48) is a UMMA on the staged upstream
// gradients. Each half head (64 features) is written to TMEM as hi/lo and consumed as the A operand of
// dh += dz_i W1_i, the sum over all heads stays in one accumulator, its masked value is the A operand of dE = dh W0e.
// Side products: the feature-major matrices dzT, dhT, GT the weight-gradient GEMMs contract over Gaussians, and the
// bias gradients (column sums, transposed-reduced across the warp and collected in shared memory).
//
// TMEM: [0,128) DH acc / dh_hi   [128,256) DZ[0] = hi 64 + lo 64 (DE acc reuses its first 32 columns)
//       [256,384) DZ[1]          [384,512) dh_lo
constexpr uint32_t FB_DH = 0, FB_DZ = 128, FB_DH_LO = 384;
constexpr int FB_G_BYTES = 12 * 128 * 16;   // rgb upstream gradients [128 Gaussians][48], K-major no-swizzle image
constexpr int FB_GT_ROWS = 64;              // rows of a GT tile (59 used)
constexpr int FB_THREADS = 18 * 32;         // 16 epilogue warps + MMA + weight loader (register allocation rounds to 20 warps: 96 per thread)

struct FusedBwdArgs {
  const float4* wimg;
  uint32_t seq_off[64];      // weight slabs of one tile in the order the MMA warp consumes them (float4 offsets)
  uint16_t seq_bytes[64];
  int nseq;
  int na, has_rgb, N;
  int use_sub, accumulate_emb, store_dz;
  int single;                // precision "tf32": one tf32 MMA per product
  long long* trace;          // ED3DGS_TRACE builds: (event, clock) pairs of block 0
  const float* w2[5];        // by slot, [k][128]
  int hdim[5];
  int goff[5];               // first row of the slot in a GT tile
  float coef[5];
  const float* d_out[5];     // may be null
  const float* d_sub[5];     // may be null
  const uint32_t* Mg;
  const uint32_t* MgH;
  float* dzT;                // [tile][na*128][128]
  float* dhT;                // [tile][128][128]
  float* GT;                 // [tile][64][128]
  float* d_emb;              // [N][32]
  float* db1[5];             // += column sums of dz_i (may be null)
  float* db2[5];             // += column sums of the scaled upstream gradients (may be null)
  float* dc1;                // += column sums of dh
};

struct BwdBars {
  uint64_t w_full[FM_W_SLOTS], w_free[FM_W_SLOTS];
  uint64_t g_ready;
  uint64_t dzacc_full[2], dz_ready[2];
  uint64_t dh_full, dh_ready, de_full;
};

constexpr int FB_GS_BYTES = 8 * 128 * 16;   // small-head upstream gradients [128 Gaussians][4 slots x 8 k], K-major image
constexpr size_t FM_BWD_SMEM = (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES + 2 * FB_G_BYTES + 2 * FB_GS_BYTES +
                               4 * (5 * 128 + 128 + 64) * 4 + 1024;

#ifdef ED3DGS_TRACE
#define FM_TRACE(ev) do { if (blockIdx.x == 0 && lane == 0 && p.trace && tcount < 4000) { p.trace[tbase + 2 * tcount] = (ev); p.trace[tbase + 2 * tcount + 1] = clock64(); ++tcount; } } while (0)
#else
#define FM_TRACE(ev) do {} while (0)
#endif

// Column sums of a 32 x 32 block held one row per lane: lane l returns sum over lanes of v[l]. 31 shuffles.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float keep = up ? v[i + off] : v[i];
      const float send = up ? v[i] : v[i + off];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

__global__ void __launch_bounds__(FB_THREADS, 1) fused_bwd_kernel(const __grid_constant__ FusedBwdArgs p) {
  extern __shared__ __align__(128) unsigned char fm_smem_raw[];
  __shared__ BwdBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sW = fm_smem_raw + ((1024u - (smem_u32(fm_smem_raw) & 1023u)) & 1023u);
  unsigned char* sGh = sW + (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES;   // rgb gradients, hi / lo
  unsigned char* sGl = sGh + FB_G_BYTES;
  unsigned char* sSh = sGl + FB_G_BYTES;                            // small-head gradients, hi / lo
  unsigned char* sSl = sSh + FB_GS_BYTES;
  // bias-gradient partial sums, one private copy per lane quarter (a warp only ever touches its own addresses:
  // shared-memory float atomics are a compare-and-swap loop on this architecture)
  float* s_db1 = reinterpret_cast<float*>(sSl + FB_GS_BYTES);      // [4][slot][128]
  float* s_dc1 = s_db1 + 4 * 5 * 128;                              // [4][128]
  float* s_db2 = s_dc1 + 4 * 128;                                  // [4][64] rows of GT

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < FM_W_SLOTS; ++i) { mbar_init(&bars.w_full[i], 1); mbar_init(&bars.w_free[i], 1); }
    mbar_init(&bars.g_ready, FM_EPI_WARPS * 32);
    for (int i = 0; i < 2; ++i) { mbar_init(&bars.dzacc_full[i], 1); mbar_init(&bars.dz_ready[i], FM_EPI_WARPS / 2 * 32); }
    mbar_init(&bars.dh_full, 1); mbar_init(&bars.dh_ready, FM_EPI_WARPS * 32); mbar_init(&bars.de_full, 1);
    fence_mbar_init();
  }
  // the padding k-pieces of the small-gradient image stay zero for the whole kernel
  for (int i = tid; i < 2 * FB_GS_BYTES / 16; i += FB_THREADS) reinterpret_cast<float4*>(sSh)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int i = tid; i < 4 * (5 * 128 + 128 + 64); i += FB_THREADS) s_db1[i] = 0.f;   // the three arrays are adjacent
  fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int ntiles = (p.N + FM_TG - 1) / FM_TG;
  const int my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int na = p.na;
  const bool has_rgb = p.has_rgb != 0;
  const int nunits = 2 * na;

  if (warp == FM_WLOAD_WARP) {
    RingPos<FM_W_SLOTS> wr;
    for (int tl = 0; tl < my_tiles; ++tl)
      for (int i = 0; i < p.nseq; ++i) {
        mbar_wait(&bars.w_free[wr.idx], wr.phase ^ 1u);
        if (elect_one()) {
          const uint32_t bytes = (uint32_t)p.seq_bytes[i];
          mbar_arrive_expect_tx(&bars.w_full[wr.idx], bytes);
          bulk_copy_g2s(sW + (size_t)wr.idx * FM_W_SLOT_BYTES, p.wimg + p.seq_off[i], bytes, &bars.w_full[wr.idx]);
        }
        __syncwarp();
        wr.advance();
      }
  } else if (warp == FM_MMA_WARP) {
    const uint32_t idesc128 = umma_idesc_tf32(128, 128), idesc64 = umma_idesc_tf32(128, 64), idesc32 = umma_idesc_tf32(128, 32);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);   // K-major no-swizzle: SBO = 128 B between 8-row groups
    const uint32_t sw = (smem_u32(sW) & 0x3FFFFu) >> 4;
    const uint32_t w128_lo0 = sw | ((128u * 16u >> 4) << 16), w64_lo0 = sw | ((64u * 16u >> 4) << 16),
                   w32_lo0 = sw | ((32u * 16u >> 4) << 16);
    const uint32_t lbo128 = (128u * 16u >> 4) << 16;
    const uint32_t gh_lo0 = ((smem_u32(sGh) & 0x3FFFFu) >> 4) | lbo128, gl_lo0 = ((smem_u32(sGl) & 0x3FFFFu) >> 4) | lbo128;
    const uint32_t sh_lo0 = ((smem_u32(sSh) & 0x3FFFFu) >> 4) | lbo128, sl_lo0 = ((smem_u32(sSl) & 0x3FFFFu) >> 4) | lbo128;
    constexpr uint32_t W_SLOT = FM_W_SLOT_BYTES >> 4;
    constexpr uint32_t W128_LO = (128u * 64u) >> 4, W128_KS = (2u * 128u * 16u) >> 4;
    constexpr uint32_t W64_LO = (64u * 64u) >> 4, W64_KS = (2u * 64u * 16u) >> 4;
    constexpr uint32_t W32_LO = (32u * 64u) >> 4, W32_KS = (2u * 32u * 16u) >> 4;
    constexpr uint32_t G_PIECE = (128u * 16u) >> 4;
    RingPos<FM_W_SLOTS> wr;
    int u = 0;
    const bool single = p.single != 0;
    int tcount = 0; const int tbase = 0; (void)tcount; (void)tbase;
    for (int tl = 0; tl < my_tiles; ++tl) {
      FM_TRACE(100 + tl);
      // SS: DZ[hf].hi[g][f] = sum_o G[g][o] W2[o][f] for the 64 features of a half head; K = 8 (small heads, one slab of
      // which one k-step is used) or 48 (rgb, three slabs)
      auto ss_dz = [&](int unit) {
        const int hs = unit >> 1, hf = unit & 1;
        const bool is_rgb = has_rgb && hs == na - 1;
        const uint32_t acc = tmem + FB_DZ + (uint32_t)hf * 128u;
        const int nsl = is_rgb ? 3 : 1;
        for (int s = 0; s < nsl; ++s) {
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w64_lo0 + (uint32_t)wr.idx * W_SLOT;
            if (is_rgb) {
#pragma unroll
              for (uint32_t ks = 0; ks < 2; ++ks) {
                const uint32_t go = ((uint32_t)s * 4u + ks * 2u) * G_PIECE;
                const uint32_t first = (s | (int)ks) ? 1u : 0u;
                if (!single) {
                  umma_tf32(acc, desc64(desc_hi, gl_lo0 + go), desc64(desc_hi, wh + ks * W64_KS), idesc64, first);
                  umma_tf32(acc, desc64(desc_hi, gh_lo0 + go), desc64(desc_hi, wh + W64_LO + ks * W64_KS), idesc64, 1u);
                }
                umma_tf32(acc, desc64(desc_hi, gh_lo0 + go), desc64(desc_hi, wh + ks * W64_KS), idesc64, single ? first : 1u);
              }
            } else {
              const uint32_t go = (uint32_t)(2 * hs) * G_PIECE;
              if (!single) {
                umma_tf32(acc, desc64(desc_hi, sl_lo0 + go), desc64(desc_hi, wh), idesc64, 0u);
                umma_tf32(acc, desc64(desc_hi, sh_lo0 + go), desc64(desc_hi, wh + W64_LO), idesc64, 1u);
              }
              umma_tf32(acc, desc64(desc_hi, sh_lo0 + go), desc64(desc_hi, wh), idesc64, single ? 0u : 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == nsl - 1) umma_commit(&bars.dzacc_full[hf]);
          }
          __syncwarp();
          wr.advance();
        }
      };
      mbar_wait(&bars.g_ready, (uint32_t)(tl & 1));
      tc_fence_after();
      ss_dz(0);
      for (int uu = 0; uu < nunits; ++uu, ++u) {
        const int b = u & 1;
        if (uu + 1 < nunits) ss_dz(uu + 1);   // its buffer was last read by the UMMAs of unit uu - 1, already in the queue
        FM_TRACE(1);
        mbar_wait(&bars.dz_ready[b], (uint32_t)((u >> 1) & 1));
        tc_fence_after();
        FM_TRACE(2);
        const uint32_t a_hi = tmem + FB_DZ + (uint32_t)b * 128u, a_lo = a_hi + 64u;
        for (int s = 0; s < 4; ++s) {   // TS: DH[g][k] += sum_f DZ[g][f] W1[f][k], 64 f of this unit
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w128_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
            for (uint32_t ks = 0; ks < 2; ++ks) {
              const uint32_t col = (uint32_t)s * 16u + ks * 8u;
              const uint32_t first = (uu | s | (int)ks) ? 1u : 0u;
              if (!single) {
                umma_tf32_ts(tmem + FB_DH, a_lo + col, desc64(desc_hi, wh + ks * W128_KS), idesc128, first);
                umma_tf32_ts(tmem + FB_DH, a_hi + col, desc64(desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
              }
              umma_tf32_ts(tmem + FB_DH, a_hi + col, desc64(desc_hi, wh + ks * W128_KS), idesc128, single ? first : 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == 3 && uu == nunits - 1) umma_commit(&bars.dh_full);
          }
          __syncwarp();
          wr.advance();
        }
      }
      FM_TRACE(3);
      mbar_wait(&bars.dh_ready, (uint32_t)(tl & 1));
      tc_fence_after();
      FM_TRACE(4);
      for (int s = 0; s < 8; ++s) {     // TS: DE[g][j] = sum_k dh[g][k] W0e[k][j], N = 32
        mbar_wait(&bars.w_full[wr.idx], wr.phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t wh = w32_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
          for (uint32_t ks = 0; ks < 2; ++ks) {
            const uint32_t col = (uint32_t)s * 16u + ks * 8u;
            const uint32_t first = (s | (int)ks) ? 1u : 0u;
            if (!single) {
              umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH_LO + col, desc64(desc_hi, wh + ks * W32_KS), idesc32, first);
              umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH + col, desc64(desc_hi, wh + W32_LO + ks * W32_KS), idesc32, 1u);
            }
            umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH + col, desc64(desc_hi, wh + ks * W32_KS), idesc32, single ? first : 1u);
          }
          umma_commit(&bars.w_free[wr.idx]);
          if (s == 7) umma_commit(&bars.de_full);
        }
        __syncwarp();
        wr.advance();
      }
    }
  } else if (warp < FM_EPI_WARPS) {
    // warp w: Gaussians 32*(w&3).., column quarter cq = w>>2. A head's features cq*32.. are this warp's; the half
    // head hf = cq>>1 (buffer DZ[hf]) is its unit, so the two warp groups alternate units.
    const int q = warp & 3, cq = warp >> 2, hf = cq >> 1, ch = cq & 1;
    const int col = cq * 32;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t dz_hi = lane_base + FB_DZ + (uint32_t)(hf * 128 + ch * 32), dz_lo = dz_hi + 64u;
    int n = 0;   // units this group has produced
    // Per-tile inputs of this thread (ReLU mask words, upstream gradient rows). They are requested one tile ahead, right
    // after the trunk-gradient hand-over of the previous tile, so that their latency hides behind that tile's stores
    // (a trace showed the tensor pipe idle a third of the time behind these loads when they were issued at the tile top).
    uint32_t mword[5] = {0u, 0u, 0u, 0u, 0u};
    uint32_t hword = 0u;
    float tsm[4] = {0.f, 0.f, 0.f, 0.f};                 // small slot cq of this Gaussian (this warp stages it)
    float4 xr[3];                                       // rgb outputs 12*cq .. 12*cq+11
    const bool my_small = cq < na && !(has_rgb && cq == na - 1);
    auto request_tile = [&](int tl) {
      const long g = ((long)blockIdx.x + (long)tl * gridDim.x) * FM_TG + q * 32 + lane;
      const bool valid = g < (long)p.N;
#pragma unroll
      for (int s = 0; s < 5; ++s) mword[s] = 0u;
      hword = 0u;
#pragma unroll
      for (int o = 0; o < 4; ++o) tsm[o] = 0.f;
      if (valid) {
#pragma unroll
        for (int s = 0; s < 5; ++s)
          if (s < na) mword[s] = p.Mg[(g * na + s) * 4 + cq];
        hword = p.MgH[g * 4 + cq];
      }
      if (my_small && valid) {
        const int k = p.hdim[cq];
        const float* a = p.d_out[cq] ? p.d_out[cq] + g * k : nullptr;
        const float* bsub = (p.use_sub && p.d_sub[cq]) ? p.d_sub[cq] + g * k : nullptr;
#pragma unroll
        for (int o = 0; o < 4; ++o)
          if (o < k) tsm[o] = (a ? a[o] : 0.f) + (bsub ? bsub[o] : 0.f);
      }
#pragma unroll
      for (int i = 0; i < 3; ++i) xr[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (has_rgb && valid) {
        const int s = na - 1;
        if (p.d_out[s]) {
          const float4* a = reinterpret_cast<const float4*>(p.d_out[s] + g * 48) + cq * 3;
#pragma unroll
          for (int i = 0; i < 3; ++i) xr[i] = a[i];
        }
        if (p.use_sub && p.d_sub[s]) {
          const float4* bsub = reinterpret_cast<const float4*>(p.d_sub[s] + g * 48) + cq * 3;
#pragma unroll
          for (int i = 0; i < 3; ++i) { const float4 y = bsub[i]; xr[i].x += y.x; xr[i].y += y.y; xr[i].z += y.z; xr[i].w += y.w; }
        }
      }
    };
    if (my_tiles > 0) request_tile(0);
    int tcount = 0; const int tbase = 8000 * (1 + (warp >> 2)); (void)tcount; (void)tbase;
    for (int tl = 0; tl < my_tiles; ++tl) {
      if (q == 0) FM_TRACE(100 + tl);
      const long tile = (long)blockIdx.x + (long)tl * gridDim.x;
      const int gl = q * 32 + lane;
      const long g = tile * FM_TG + gl;
      const bool valid = g < (long)p.N;
      // ---- upstream gradients scaled by the annealing coefficients: GT rows, bias-gradient sums, and the K-major A
      //      operands of the dz UMMAs (column quarter cq stages small slot cq and rgb outputs 12*cq .. 12*cq+11)
      if (q == 0) FM_TRACE(20);
      float red[32];   // the gradient columns whose sums (second-layer bias gradients) this warp owns
#pragma unroll
      for (int i = 0; i < 32; ++i) red[i] = 0.f;
      float* gt = p.GT + tile * (FB_GT_ROWS * 128) + gl;
      if (my_small) {
        const float cf = p.coef[cq];
        const int k = p.hdim[cq];
#pragma unroll
        for (int o = 0; o < 4; ++o) tsm[o] *= cf;
        float* gr = gt + p.goff[cq] * 128;
        gr[0] = tsm[0];
        if (k > 1) gr[128] = tsm[1];
        if (k > 2) gr[256] = tsm[2];
        if (k > 3) gr[384] = tsm[3];
        red[0] = tsm[0]; red[1] = tsm[1]; red[2] = tsm[2]; red[3] = tsm[3];
        const float4 x = make_float4(tsm[0], tsm[1], tsm[2], tsm[3]);
        reinterpret_cast<float4*>(sSh)[(2 * cq) * 128 + gl] = x;
        reinterpret_cast<float4*>(sSl)[(2 * cq) * 128 + gl] = split_lo(x);
      }
      if (has_rgb) {
        const int s = na - 1;
        const float cf = p.coef[s];
        float4* gh = reinterpret_cast<float4*>(sGh) + gl;
        float4* glo = reinterpret_cast<float4*>(sGl) + gl;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int pc = cq * 3 + i;
          float4 x = xr[i];
          x.x *= cf; x.y *= cf; x.z *= cf; x.w *= cf;
          gh[pc * 128] = x;
          glo[pc * 128] = split_lo(x);
          float* gr = gt + (p.goff[s] + 4 * pc) * 128;
          gr[0] = x.x; gr[128] = x.y; gr[256] = x.z; gr[384] = x.w;
          red[16 + 4 * i] = x.x; red[17 + 4 * i] = x.y; red[18 + 4 * i] = x.z; red[19 + 4 * i] = x.w;
        }
      }
      if (q == 0) FM_TRACE(21);
      fence_async_smem();
      mbar_arrive(&bars.g_ready);
      if (q == 0) FM_TRACE(22);
      {
        const float cs = warp_colsum32(red, lane);
        // columns 0..3: small slot cq; 16..27: rgb outputs 12*cq + (lane - 16)
        int row = -1;
        if (lane < 4) { if (my_small && lane < p.hdim[cq]) row = p.goff[cq] + lane; }
        else if (has_rgb && lane >= 16 && lane < 28) row = p.goff[na - 1] + 12 * cq + lane - 16;
        if (row >= 0) s_db2[q * 64 + row] += cs;
      }
      if (q == 0) FM_TRACE(23);
      // ---- this group's half of every head: accumulator -> ReLU mask -> hi / lo A operand of dh += dz W1
#pragma unroll 1
      for (int hs = 0; hs < na; ++hs, ++n) {
        uint32_t word = mword[0];
        if (hs == 1) word = mword[1];
        if (hs == 2) word = mword[2];
        if (hs == 3) word = mword[3];
        if (hs == 4) word = mword[4];
        if (q == 0) FM_TRACE(10);
        mbar_wait(&bars.dzacc_full[hf], (uint32_t)(n & 1));
        tc_fence_after();
        if (q == 0) FM_TRACE(11);
        float v[32];
        tmem_ld32(dz_hi, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = ((word >> j) & 1u) ? v[j] : 0.f;
        tmem_st32(dz_hi, v);
        if (!p.single) tmem_st_lo32(dz_lo, v);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars.dz_ready[hf]);
        if (q == 0) FM_TRACE(12);
        if (p.store_dz) {
          float* zp = p.dzT + ((tile * na + hs) * 128 + col) * 128 + gl;
#pragma unroll
          for (int j = 0; j < 32; ++j) zp[j * 128] = v[j];
        }
        const float cs = warp_colsum32(v, lane);
        s_db1[q * 640 + hs * 128 + col + lane] += cs;
      }
      // ---- dh = (sum over heads) masked by the trunk ReLU -> dhT, bias gradient, A operand of dE
      {
        if (q == 0) FM_TRACE(13);
        mbar_wait(&bars.dh_full, (uint32_t)(tl & 1));
        tc_fence_after();
        if (q == 0) FM_TRACE(14);
        float v[32];
        tmem_ld32(lane_base + FB_DH + (uint32_t)col, v);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = ((hword >> j) & 1u) ? v[j] : 0.f;
        tmem_st32(lane_base + FB_DH + (uint32_t)col, v);
        if (!p.single) tmem_st_lo32(lane_base + FB_DH_LO + (uint32_t)col, v);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars.dh_ready);
        if (q == 0) FM_TRACE(15);
        if (tl + 1 < my_tiles) request_tile(tl + 1);   // latency hides behind the stores and the column sums below
        float* hp = p.dhT + (tile * 128 + col) * 128 + gl;
#pragma unroll
        for (int j = 0; j < 32; ++j) hp[j * 128] = v[j];
        if (q == 0) FM_TRACE(16);
        const float cs = warp_colsum32(v, lane);
        s_dc1[q * 128 + col + lane] += cs;
        if (q == 0) FM_TRACE(17);
      }
      // ---- dE row of this Gaussian (the first 32 DZ[0] columns are the quarter-0 warps' own)
      if (cq == 0) {
        float4* de = reinterpret_cast<float4*>(p.d_emb + g * 32);
        float4 old[8];
#pragma unroll
        for (int o = 0; o < 8; ++o) old[o] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid && p.accumulate_emb) {
#pragma unroll
          for (int o = 0; o < 8; ++o) old[o] = de[o];
        }
        mbar_wait(&bars.de_full, (uint32_t)(tl & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + FB_DZ, v);
        if (valid) {
#pragma unroll
          for (int o = 0; o < 8; ++o)
            de[o] = make_float4(old[o].x + v[4 * o], old[o].y + v[4 * o + 1], old[o].z + v[4 * o + 2], old[o].w + v[4 * o + 3]);
        }
        tc_fence_before();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  // bias gradients: one atomic per column and CTA
  for (int i = tid; i < 5 * 128; i += FB_THREADS) {
    const int s = i >> 7;
    const float x = s_db1[i] + s_db1[640 + i] + s_db1[1280 + i] + s_db1[1920 + i];
    if (s < na && p.db1[s] && x != 0.f) atomicAdd(p.db1[s] + (i & 127), x);
  }
  for (int i = tid; i < 128; i += FB_THREADS)
    if (p.dc1) atomicAdd(p.dc1 + i, s_dc1[i] + s_dc1[128 + i] + s_dc1[256 + i] + s_dc1[384 + i]);
  for (int i = tid; i < 5 * 48; i += FB_THREADS) {
    const int s = i / 48, o = i - s * 48;
    if (s < na && o < p.hdim[s] && p.db2[s]) {
      const int r = p.goff[s] + o;
      const float x = s_db2[r] + s_db2[64 + r] + s_db2[128 + r] + s_db2[192 + r];
      if (x != 0.f) atomicAdd(p.db2[s] + o, x);
    }
  }
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace ed3dgs
