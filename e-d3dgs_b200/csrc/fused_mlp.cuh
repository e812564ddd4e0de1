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
constexpr int FM_EPI_WARPS = 8;                  // warp w: TMEM lanes 32*(w&3).., column half w>>2
constexpr int FM_MMA_WARP = 8;
constexpr int FM_WLOAD_WARP = 9;
constexpr int FM_ELOAD_WARP = 10;
constexpr int FM_THREADS = 11 * 32;
constexpr int FM_W_SLOTS = 8;
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
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }

// =============================================================================================================
// Forward of one stage.
struct FusedFwdArgs {
  const float4* wimg;            // slab images: [0,2) layer 0, [2 + 8*s, +8) hidden layer of slot s, then 8 slabs of the
  uint32_t slab_off[2 + 8 * 6];  //               rgb output layer (NT = 48); float4 offsets
  int na, has_rgb;               // active heads in slot order; the rgb head, when active, is the last slot
  int N;
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
  uint64_t w_full[FM_W_SLOTS], w_free[FM_W_SLOTS];
  uint64_t e_tma[2], e_full[2], e_free[2];
  uint64_t acc_full[2], acc_free[2];
  uint64_t h_ready, z4_ready, out_full;
};

constexpr size_t FM_FWD_SMEM = (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES + 4 * FM_E_BYTES + 5 * 128 * 16 +
                               (128 + 5 * 128 + 5 * 48) * 4 + 2 * 128 * 16 * 4 + 1024;

__global__ void __launch_bounds__(FM_THREADS, 1) fused_fwd_kernel(const __grid_constant__ FusedFwdArgs p) {
  extern __shared__ __align__(128) unsigned char fm_smem_raw[];
  __shared__ FwdBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sW = fm_smem_raw + ((1024u - (smem_u32(fm_smem_raw) & 1023u)) & 1023u);
  unsigned char* sEh = sW + (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES;
  unsigned char* sEl = sEh + 2 * FM_E_BYTES;
  float4* s_w2 = reinterpret_cast<float4*>(sEl + 2 * FM_E_BYTES);   // [slot][feature] -> (w2[0..3][feature])
  float* s_c1 = reinterpret_cast<float*>(s_w2 + 5 * 128);
  float* s_b1 = s_c1 + 128;                                         // [slot][128]
  float* s_b2 = s_b1 + 5 * 128;                                     // [slot][48]
  float* s_part = s_b2 + 5 * 48;                                    // [2][128 rows][16]

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < FM_W_SLOTS; ++i) { mbar_init(&bars.w_full[i], 1); mbar_init(&bars.w_free[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars.e_tma[i], 1); mbar_init(&bars.e_full[i], 32); mbar_init(&bars.e_free[i], 1);
      mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], FM_EPI_WARPS * 32);
    }
    mbar_init(&bars.h_ready, FM_EPI_WARPS * 32);
    mbar_init(&bars.z4_ready, FM_EPI_WARPS * 32);
    mbar_init(&bars.out_full, 1);
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
#pragma unroll 8
      for (int i = 0; i < FM_E_BYTES / 16 / 32; ++i) lo[lane + 32 * i] = split_lo(hi[lane + 32 * i]);
      fence_async_smem();
      mbar_arrive(&bars.e_full[eb]);
    }
  } else if (warp == FM_WLOAD_WARP) {
    // ------------------------------------------------------------------ weight slabs, in the order the MMA warp uses them
    RingPos<FM_W_SLOTS> wr;
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
    RingPos<FM_W_SLOTS> wr;
    int go = 0;

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
            umma_tf32(acc, desc64(e_desc_hi, el + ks * E_KS), desc64(w_desc_hi, wh + ks * W128_KS), idesc128,
                      (s | ks) ? 1u : 0u);
            umma_tf32(acc, desc64(e_desc_hi, eh + ks * E_KS), desc64(w_desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
            umma_tf32(acc, desc64(e_desc_hi, eh + ks * E_KS), desc64(w_desc_hi, wh + ks * W128_KS), idesc128, 1u);
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
              umma_tf32_ts(acc, tmem + FM_H_LO + col, desc64(w_desc_hi, wh + ks * W128_KS), idesc128, (s | (int)ks) ? 1u : 0u);
              umma_tf32_ts(acc, tmem + FM_H_HI + col, desc64(w_desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
              umma_tf32_ts(acc, tmem + FM_H_HI + col, desc64(w_desc_hi, wh + ks * W128_KS), idesc128, 1u);
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
              umma_tf32_ts(out, a_lo + col, desc64(w_desc_hi, wh + ks * W48_KS), idesc48, (s | (int)ks) ? 1u : 0u);
              umma_tf32_ts(out, a_hi + col, desc64(w_desc_hi, wh + W48_LO + ks * W48_KS), idesc48, 1u);
              umma_tf32_ts(out, a_hi + col, desc64(w_desc_hi, wh + ks * W48_KS), idesc48, 1u);
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
    // ------------------------------------------------------------------ epilogue (warps 0..7)
    const int q = warp & 3, ch = warp >> 2;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    int go = 0, npart = 0;
    for (int tl = 0; tl < my_tiles; ++tl) {
      const long tile = (long)blockIdx.x + (long)tl * gridDim.x;
      const int gl = q * 32 + lane;
      const long g = tile * FM_TG + gl;
      const bool valid = g < (long)p.N;
      // the previous tile's rgb output MMAs read H_lo and write H_hi: they must be done before H is rewritten
      if (has_rgb && tl > 0 && ch == 1) { mbar_wait(&bars.out_full, (uint32_t)((tl - 1) & 1)); tc_fence_after(); }
      {  // ---- layer 0: h = relu(acc + c1) -> H_hi / H_lo (A operand of the five hidden layers), hT, trunk mask
        const int buf = go & 1;
        mbar_wait(&bars.acc_full[buf], (uint32_t)((go >> 1) & 1));
        tc_fence_after();
        uint32_t hw0 = 0, hw1 = 0;
#pragma unroll 1
        for (int blk = 0; blk < 2; ++blk) {
          const int col = ch * 64 + blk * 32;
          float v[32], lo[32];
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
              lo[j] = h - trunc_tf32(h);
              word |= (h > 0.f ? 1u : 0u) << j;
              hp[j * 128] = h;
            }
          }
          tmem_st32(lane_base + FM_H_HI + (uint32_t)col, v);
          tmem_st32(lane_base + FM_H_LO + (uint32_t)col, lo);
          if (blk == 0) hw0 = word; else hw1 = word;
        }
        tmem_st_wait();
        if (valid) *reinterpret_cast<uint2*>(p.MgH + g * 4 + ch * 2) = make_uint2(hw0, hw1);
        tc_fence_before();
        mbar_arrive(&bars.h_ready);
        mbar_arrive(&bars.acc_free[buf]);
        ++go;
      }
      // ---- hidden layer of every head: z = relu(acc + b1) -> zT, both mask layouts; the small heads' output layer
      //      (k <= 4) is evaluated right here on the FP32 pipe, the rgb head's z goes back to TMEM for its UMMA
#pragma unroll 1
      for (int hs = 0; hs < na; ++hs) {
        const bool is_rgb = has_rgb && hs == na - 1;
        const int buf = go & 1;
        mbar_wait(&bars.acc_full[buf], (uint32_t)((go >> 1) & 1));
        tc_fence_after();
        uint32_t mw0 = 0, mw1 = 0;
        float o4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
        for (int blk = 0; blk < 2; ++blk) {
          const int col = ch * 64 + blk * 32;
          float v[32];
          tmem_ld32(lane_base + FM_ACC0 + (uint32_t)(buf * 128 + col), v);
          const float4* b4 = reinterpret_cast<const float4*>(s_b1 + hs * 128 + col);
          float* zp = p.zT + ((tile * na + hs) * 128 + col) * 128 + gl;
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
              const bool pos = z > 0.f;
              word |= (pos ? 1u : 0u) << j;
              zp[j * 128] = z;
            }
          }
          if (blk == 0) mw0 = word; else mw1 = word;
          if (!is_rgb) {
            const float4* w2 = s_w2 + hs * 128 + col;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float4 w = w2[j];
              o4[0] = fmaf(v[j], w.x, o4[0]);
              o4[1] = fmaf(v[j], w.y, o4[1]);
              o4[2] = fmaf(v[j], w.z, o4[2]);
              o4[3] = fmaf(v[j], w.w, o4[3]);
            }
          } else {
            float lo[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) lo[j] = v[j] - trunc_tf32(v[j]);
            tmem_st32(lane_base + FM_ACC0 + (uint32_t)(buf * 128 + col), v);   // Z4_hi in place
            tmem_st32(lane_base + FM_H_LO + (uint32_t)col, lo);               // the trunk operand is dead by now
          }
        }
        if (valid) *reinterpret_cast<uint2*>(p.Mg + (g * na + hs) * 4 + ch * 2) = make_uint2(mw0, mw1);
        if (is_rgb) {
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(&bars.z4_ready);
        }
        tc_fence_before();
        mbar_arrive(&bars.acc_free[buf]);
        ++go;
        if (!is_rgb) {
          // x = base + (delta + b2) * coef: the column-half 0 warps park their partial sums, the column-half 1 warps finish
          float4* part = reinterpret_cast<float4*>(s_part) + (size_t)(npart & 1) * 128 + gl;
          ++npart;
          if (ch == 0) *part = make_float4(o4[0], o4[1], o4[2], o4[3]);
          epi_bar_sync();
          if (ch == 1 && valid) {
            const float4 o2 = *part;
            const float t[4] = {o4[0] + o2.x, o4[1] + o2.y, o4[2] + o2.z, o4[3] + o2.w};
            const int k = p.hdim[hs];
            const float cf = p.coef[hs];
            const float* bs = p.base[hs] + g * k;
            float* ds = p.dst[hs] + g * k;
#pragma unroll
            for (int o = 0; o < 4; ++o)
              if (o < k) ds[o] = bs[o] + (t[o] + s_b2[hs * 48 + o]) * cf;
          }
        }
      }
      if (has_rgb && ch == 0) {
        // the rgb accumulator sits in the columns these warps rewrite in the next tile's layer-0 epilogue
        const int s = na - 1;
        mbar_wait(&bars.out_full, (uint32_t)(tl & 1));
        tc_fence_after();
        float v[48];
        tmem_ld32(lane_base + FM_H_HI, v);
        tmem_ld16(lane_base + FM_H_HI + 32u, v + 32);
        if (valid) {
          const float4* bs = reinterpret_cast<const float4*>(p.base[s] + g * 48);
          float4* ds = reinterpret_cast<float4*>(p.dst[s] + g * 48);
          const float cf = p.coef[s];
          const float4* b2 = reinterpret_cast<const float4*>(s_b2 + s * 48);
#pragma unroll
          for (int o = 0; o < 12; ++o) {
            float4 b = bs[o];
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

struct FusedBwdArgs {
  const float4* wimg;
  uint32_t seq_off[64];      // weight slabs of one tile in the order the MMA warp consumes them (float4 offsets)
  uint16_t seq_bytes[64];
  int nseq;
  int na, has_rgb, N;
  int use_sub, accumulate_emb, store_dz;
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
  uint64_t dzacc_full[2], dz_ready[2], dz_free[2];
  uint64_t dh_full, dh_ready, de_full;
};

constexpr size_t FM_BWD_SMEM = (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES + 2 * FB_G_BYTES + 5 * 128 * 16 +
                               (5 * 128 + 128 + 64) * 4 + 1024;

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

__global__ void __launch_bounds__(FM_THREADS, 1) fused_bwd_kernel(const __grid_constant__ FusedBwdArgs p) {
  extern __shared__ __align__(128) unsigned char fm_smem_raw[];
  __shared__ BwdBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* sW = fm_smem_raw + ((1024u - (smem_u32(fm_smem_raw) & 1023u)) & 1023u);
  unsigned char* sGh = sW + (size_t)FM_W_SLOTS * FM_W_SLOT_BYTES;
  unsigned char* sGl = sGh + FB_G_BYTES;
  float4* s_w2 = reinterpret_cast<float4*>(sGl + FB_G_BYTES);
  float* s_db1 = reinterpret_cast<float*>(s_w2 + 5 * 128);   // [slot][128]
  float* s_dc1 = s_db1 + 5 * 128;                            // [128]
  float* s_db2 = s_dc1 + 128;                                // [64] rows of GT

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < FM_W_SLOTS; ++i) { mbar_init(&bars.w_full[i], 1); mbar_init(&bars.w_free[i], 1); }
    mbar_init(&bars.g_ready, FM_EPI_WARPS * 32);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars.dzacc_full[i], 1); mbar_init(&bars.dz_ready[i], FM_EPI_WARPS * 32); mbar_init(&bars.dz_free[i], 1);
    }
    mbar_init(&bars.dh_full, 1); mbar_init(&bars.dh_ready, FM_EPI_WARPS * 32); mbar_init(&bars.de_full, 1);
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
    s_db1[i] = 0.f;
  }
  for (int i = tid; i < 128 + 64; i += FM_THREADS) s_dc1[i] = 0.f;   // s_dc1 and s_db2 are adjacent
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
    const uint32_t gh_lo0 = ((smem_u32(sGh) & 0x3FFFFu) >> 4) | ((128u * 16u >> 4) << 16);
    const uint32_t gl_lo0 = ((smem_u32(sGl) & 0x3FFFFu) >> 4) | ((128u * 16u >> 4) << 16);
    constexpr uint32_t W_SLOT = FM_W_SLOT_BYTES >> 4;
    constexpr uint32_t W128_LO = (128u * 64u) >> 4, W128_KS = (2u * 128u * 16u) >> 4;
    constexpr uint32_t W64_LO = (64u * 64u) >> 4, W64_KS = (2u * 64u * 16u) >> 4;
    constexpr uint32_t W32_LO = (32u * 64u) >> 4, W32_KS = (2u * 32u * 16u) >> 4;
    constexpr uint32_t G_PIECE = (128u * 16u) >> 4;
    RingPos<FM_W_SLOTS> wr;
    int u = 0;
    for (int tl = 0; tl < my_tiles; ++tl) {
      auto ss_dz = [&](int hf) {   // SS: DZ[hf].hi[g][f] = sum_o G[g][o] W2[o][f], f in the half, K = 48
        if (hf == 0) { mbar_wait(&bars.g_ready, (uint32_t)(tl & 1)); tc_fence_after(); }
        const uint32_t acc = tmem + FB_DZ + (uint32_t)hf * 128u;
        for (int s = 0; s < 3; ++s) {
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w64_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
            for (uint32_t ks = 0; ks < 2; ++ks) {
              const uint32_t go = ((uint32_t)s * 4u + ks * 2u) * G_PIECE;
              umma_tf32(acc, desc64(desc_hi, gl_lo0 + go), desc64(desc_hi, wh + ks * W64_KS), idesc64, (s | (int)ks) ? 1u : 0u);
              umma_tf32(acc, desc64(desc_hi, gh_lo0 + go), desc64(desc_hi, wh + W64_LO + ks * W64_KS), idesc64, 1u);
              umma_tf32(acc, desc64(desc_hi, gh_lo0 + go), desc64(desc_hi, wh + ks * W64_KS), idesc64, 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == 2) umma_commit(&bars.dzacc_full[hf]);
          }
          __syncwarp();
          wr.advance();
        }
      };
      for (int uu = 0; uu < nunits; ++uu, ++u) {
        const int b = u & 1;
        if (has_rgb && uu + 1 < nunits && ((uu + 1) >> 1) == na - 1) ss_dz((uu + 1) & 1);
        mbar_wait(&bars.dz_ready[b], (uint32_t)((u >> 1) & 1));
        tc_fence_after();
        const uint32_t a_hi = tmem + FB_DZ + (uint32_t)b * 128u, a_lo = a_hi + 64u;
        for (int s = 0; s < 4; ++s) {   // TS: DH[g][k] += sum_f DZ[g][f] W1[f][k], 64 f of this unit
          mbar_wait(&bars.w_full[wr.idx], wr.phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t wh = w128_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
            for (uint32_t ks = 0; ks < 2; ++ks) {
              const uint32_t col = (uint32_t)s * 16u + ks * 8u;
              umma_tf32_ts(tmem + FB_DH, a_lo + col, desc64(desc_hi, wh + ks * W128_KS), idesc128, (uu | s | (int)ks) ? 1u : 0u);
              umma_tf32_ts(tmem + FB_DH, a_hi + col, desc64(desc_hi, wh + W128_LO + ks * W128_KS), idesc128, 1u);
              umma_tf32_ts(tmem + FB_DH, a_hi + col, desc64(desc_hi, wh + ks * W128_KS), idesc128, 1u);
            }
            umma_commit(&bars.w_free[wr.idx]);
            if (s == 3) {
              umma_commit(&bars.dz_free[b]);
              if (uu == nunits - 1) umma_commit(&bars.dh_full);
            }
          }
          __syncwarp();
          wr.advance();
        }
      }
      mbar_wait(&bars.dh_ready, (uint32_t)(tl & 1));
      tc_fence_after();
      for (int s = 0; s < 8; ++s) {     // TS: DE[g][j] = sum_k dh[g][k] W0e[k][j], N = 32
        mbar_wait(&bars.w_full[wr.idx], wr.phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t wh = w32_lo0 + (uint32_t)wr.idx * W_SLOT;
#pragma unroll
          for (uint32_t ks = 0; ks < 2; ++ks) {
            const uint32_t col = (uint32_t)s * 16u + ks * 8u;
            umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH_LO + col, desc64(desc_hi, wh + ks * W32_KS), idesc32, (s | (int)ks) ? 1u : 0u);
            umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH + col, desc64(desc_hi, wh + W32_LO + ks * W32_KS), idesc32, 1u);
            umma_tf32_ts(tmem + FB_DZ, tmem + FB_DH + col, desc64(desc_hi, wh + ks * W32_KS), idesc32, 1u);
          }
          umma_commit(&bars.w_free[wr.idx]);
          if (s == 7) umma_commit(&bars.de_full);
        }
        __syncwarp();
        wr.advance();
      }
    }
  } else if (warp < FM_EPI_WARPS) {
    const int q = warp & 3, ch = warp >> 2;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    int u = 0;
    for (int tl = 0; tl < my_tiles; ++tl) {
      const long tile = (long)blockIdx.x + (long)tl * gridDim.x;
      const int gl = q * 32 + lane;
      const long g = tile * FM_TG + gl;
      const bool valid = g < (long)p.N;
      // ---- upstream gradients of this Gaussian, scaled by the annealing coefficients
      float4 G4[4];
      float red[32];   // the 32 gradient columns whose sums (second-layer bias gradients) this warp owns
#pragma unroll
      for (int i = 0; i < 32; ++i) red[i] = 0.f;
      float* gt = p.GT + tile * (FB_GT_ROWS * 128) + gl;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        float t[4] = {0.f, 0.f, 0.f, 0.f};
        if (s < na && !(has_rgb && s == na - 1)) {
          const int k = p.hdim[s];
          if (valid) {
            const float* a = p.d_out[s] ? p.d_out[s] + g * k : nullptr;
            const float* bsub = (p.use_sub && p.d_sub[s]) ? p.d_sub[s] + g * k : nullptr;
#pragma unroll
            for (int o = 0; o < 4; ++o)
              if (o < k) t[o] = ((a ? a[o] : 0.f) + (bsub ? bsub[o] : 0.f)) * p.coef[s];
          }
          if (ch == 0) {
#pragma unroll
            for (int o = 0; o < 4; ++o)
              if (o < k) gt[(p.goff[s] + o) * 128] = t[o];
          }
        }
        G4[s] = make_float4(t[0], t[1], t[2], t[3]);
        if (ch == 0) { red[4 * s] = t[0]; red[4 * s + 1] = t[1]; red[4 * s + 2] = t[2]; red[4 * s + 3] = t[3]; }
      }
      if (has_rgb) {
        // rgb rows: column-half 0 warps take o = 0..15, column-half 1 warps o = 16..47; staged as the K-major A operand
        const int s = na - 1;
        const float cf = p.coef[s];
        const float4* a = (valid && p.d_out[s]) ? reinterpret_cast<const float4*>(p.d_out[s] + g * 48) : nullptr;
        const float4* bsub = (valid && p.use_sub && p.d_sub[s]) ? reinterpret_cast<const float4*>(p.d_sub[s] + g * 48) : nullptr;
        float4* gh = reinterpret_cast<float4*>(sGh) + gl;
        float4* glo = reinterpret_cast<float4*>(sGl) + gl;
        auto stage_piece = [&](int pc, float* r4) {
          float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
          if (a) x = a[pc];
          if (bsub) { const float4 y = bsub[pc]; x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w; }
          x.x *= cf; x.y *= cf; x.z *= cf; x.w *= cf;
          gh[pc * 128] = x;
          glo[pc * 128] = split_lo(x);
          float* gr = gt + (p.goff[s] + 4 * pc) * 128;
          gr[0] = x.x; gr[128] = x.y; gr[256] = x.z; gr[384] = x.w;
          r4[0] = x.x; r4[1] = x.y; r4[2] = x.z; r4[3] = x.w;
        };
        if (ch == 0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) stage_piece(i, red + 16 + 4 * i);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) stage_piece(4 + i, red + 4 * i);
        }
        fence_async_smem();
        mbar_arrive(&bars.g_ready);
      }
      {
        const float cs = warp_colsum32(red, lane);
        // column-half 0: columns 0..15 are the small slots (4 each), 16..31 rgb outputs 0..15; column-half 1: rgb 16..47
        int row = -1;
        if (ch == 0) {
          if (lane < 16) { const int s = lane >> 2, o = lane & 3; if (s < na && !(has_rgb && s == na - 1) && o < p.hdim[s]) row = p.goff[s] + o; }
          else if (has_rgb) row = p.goff[na - 1] + lane - 16;
        } else if (has_rgb) row = p.goff[na - 1] + 16 + lane;
        if (row >= 0) atomicAdd(&s_db2[row], cs);
      }
      // ---- half heads
#pragma unroll 1
      for (int hs = 0; hs < na; ++hs) {
        const bool is_rgb = has_rgb && hs == na - 1;
        uint4 mw = make_uint4(0u, 0u, 0u, 0u);
        if (valid) mw = *reinterpret_cast<const uint4*>(p.Mg + (g * na + hs) * 4);
        float4 gs = G4[0];
        if (hs == 1) gs = G4[1];
        if (hs == 2) gs = G4[2];
        if (hs == 3) gs = G4[3];
#pragma unroll 1
        for (int hf = 0; hf < 2; ++hf, ++u) {
          const int b = u & 1;
          const int col = hf * 64 + ch * 32;
          const uint32_t word = hf == 0 ? (ch == 0 ? mw.x : mw.y) : (ch == 0 ? mw.z : mw.w);
          mbar_wait(&bars.dz_free[b], (uint32_t)(((u >> 1) & 1) ^ 1));   // the UMMAs that read this buffer two units ago
          tc_fence_after();
          float v[32], lo[32];
          if (!is_rgb) {
            const float4* w2 = s_w2 + hs * 128 + col;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float4 w = w2[j];
              v[j] = fmaf(gs.x, w.x, fmaf(gs.y, w.y, fmaf(gs.z, w.z, gs.w * w.w)));
            }
          } else {
            mbar_wait(&bars.dzacc_full[b], (uint32_t)(tl & 1));
            tc_fence_after();
            tmem_ld32(lane_base + FB_DZ + (uint32_t)(b * 128 + ch * 32), v);
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] = ((word >> j) & 1u) ? v[j] : 0.f;
            lo[j] = v[j] - trunc_tf32(v[j]);
          }
          tmem_st32(lane_base + FB_DZ + (uint32_t)(b * 128 + ch * 32), v);
          tmem_st32(lane_base + FB_DZ + (uint32_t)(b * 128 + 64 + ch * 32), lo);
          tmem_st_wait();
          tc_fence_before();
          mbar_arrive(&bars.dz_ready[b]);
          if (p.store_dz) {
            float* zp = p.dzT + ((tile * na + hs) * 128 + col) * 128 + gl;
#pragma unroll
            for (int j = 0; j < 32; ++j) zp[j * 128] = v[j];
          }
          const float cs = warp_colsum32(v, lane);
          atomicAdd(&s_db1[hs * 128 + col + lane], cs);
        }
      }
      // ---- dh = (sum over heads) masked by the trunk ReLU -> dhT, bias gradient, A operand of dE
      {
        mbar_wait(&bars.dh_full, (uint32_t)(tl & 1));
        tc_fence_after();
        uint2 hw = make_uint2(0u, 0u);
        if (valid) hw = *reinterpret_cast<const uint2*>(p.MgH + g * 4 + ch * 2);
#pragma unroll 1
        for (int blk = 0; blk < 2; ++blk) {
          const int col = ch * 64 + blk * 32;
          const uint32_t word = blk == 0 ? hw.x : hw.y;
          float v[32], lo[32];
          tmem_ld32(lane_base + FB_DH + (uint32_t)col, v);
          float* hp = p.dhT + (tile * 128 + col) * 128 + gl;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            v[j] = ((word >> j) & 1u) ? v[j] : 0.f;
            lo[j] = v[j] - trunc_tf32(v[j]);
            hp[j * 128] = v[j];
          }
          tmem_st32(lane_base + FB_DH + (uint32_t)col, v);
          tmem_st32(lane_base + FB_DH_LO + (uint32_t)col, lo);
          const float cs = warp_colsum32(v, lane);
          atomicAdd(&s_dc1[col + lane], cs);
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars.dh_ready);
      }
      // ---- dE row of this Gaussian (the first 32 DZ[0] columns belong to the column-half 0 warps)
      if (ch == 0) {
        mbar_wait(&bars.de_full, (uint32_t)(tl & 1));
        tc_fence_after();
        float v[32];
        tmem_ld32(lane_base + FB_DZ, v);
        if (valid) {
          float4* de = reinterpret_cast<float4*>(p.d_emb + g * 32);
#pragma unroll
          for (int o = 0; o < 8; ++o) {
            float4 x = make_float4(v[4 * o], v[4 * o + 1], v[4 * o + 2], v[4 * o + 3]);
            if (p.accumulate_emb) { const float4 y = de[o]; x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w; }
            de[o] = x;
          }
        }
        tc_fence_before();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  // bias gradients: one atomic per column and CTA
  for (int i = tid; i < 5 * 128; i += FM_THREADS) {
    const int s = i >> 7;
    if (s < na && p.db1[s] && s_db1[i] != 0.f) atomicAdd(p.db1[s] + (i & 127), s_db1[i]);
  }
  for (int i = tid; i < 128; i += FM_THREADS)
    if (p.dc1 && s_dc1[i] != 0.f) atomicAdd(p.dc1 + i, s_dc1[i]);
  for (int i = tid; i < 5 * 48; i += FM_THREADS) {
    const int s = i / 48, o = i - s * 48;
    if (s < na && o < p.hdim[s] && p.db2[s]) {
      const float x = s_db2[p.goff[s] + o];
      if (x != 0.f) atomicAdd(p.db2[s] + o, x);
    }
  }
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace ed3dgs
