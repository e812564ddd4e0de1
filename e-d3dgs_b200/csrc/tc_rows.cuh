// Persistent, warp-specialised 3xTF32 GEMM over tiles of 256 Gaussians (sm_100a, tcgen05 + TMEM).
//
//   for every tile of 256 Gaussians g, for every batch entry (head):
//       D^T[f][g] = sum_k W[f][k] * X[g][k]          f < 128 output features, then a fused epilogue
//
// The Gaussians sit on the N side of the UMMA (N = 256) and the <= 128 output features on the M
// side: on B200 a 128 x N x 8 tf32 UMMA costs 128 clocks for every N <= 256 (tests/probes/mma_rate.cu),
// so only N = 256 runs the tensor pipe at full rate. It also makes the epilogue naturally
// coalesced: a TMEM lane is a feature, so the 32 lanes of a warp write 128 contiguous bytes of one
// Gaussian's row.
//
// One CTA per SM walks the tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the work of a tile is a
// list of 16-wide K slabs built on the host (TcArgs::steps, tc_launch.cuh). Four roles run that
// list concurrently and meet only through mbarriers:
//
//   warps 10-17 X stagers   global --cp.async 16 B, 4 slabs ahead--> raw ring (5 slabs, thread-private
//                           pieces) --split hi/lo--> UMMA K-major operand stages (ring of 3 x [256 x 16 k])
//   warp 9      W loader    prepared hi/lo weight slabs (tc_prepare_b) --cp.async.bulk--> ring of 3
//   warp 8      MMA issuer  per slab: 2 k-steps x (W_lo*X_hi + W_hi*X_lo + W_hi*X_hi) tcgen05.mma;
//                           tcgen05.commit releases the X stage / the W slot / the accumulator
//   warps 0-7   epilogue    TMEM (2 accumulators of 256 columns, ping-pong) -> registers -> fused
//                           bias/ReLU | ReLU-mask (+ column sums = bias gradient) | accumulate |
//                           x + delta*coef -> global
//
// Measured limits (ncu, tests/gemm_time.py): with global traffic switched off a slab takes ~1260 clocks
// against 768 clocks of UMMA time: the shared-memory pipe is saturated (UMMA operand reads 576
// wavefronts + stager traffic ~650 per slab). The SM is also instruction-issue bound long before it is
// tensor bound, so every role keeps its per-slab instruction count small: descriptors are affine in
// the ring index, the step table is one 32-bit word per slab, stagers keep running offsets, the
// epilogue is specialised on the row stride, waits and issues are warp-uniform with one elected lane.
#pragma once

namespace ed3dgs {

constexpr int ROWS_TG = 256;                             // Gaussians per tile = UMMA N
static_assert(ROWS_TG == 256, "the epilogue assumes four 32-Gaussian blocks per warp");
constexpr int ROWS_EPI_WARPS = 8;                        // warps 0..7 (TMEM lane quarter = warp % 4)
constexpr int ROWS_MMA_WARP = 8;
constexpr int ROWS_LOAD_WARP = 9;
constexpr int ROWS_CONV_WARP0 = 10;                      // 6 warps: 16 warps in all = 4 per SM sub-partition = 128 registers
constexpr int ROWS_CONV_THREADS = 192;
constexpr int ROWS_THREADS = ROWS_CONV_WARP0 * 32 + ROWS_CONV_THREADS;
constexpr int ROWS_XH_SLOTS = 7;                         // raw activation slabs written by TMA = the hi operand
constexpr int ROWS_XL_SLOTS = 4;                         // lo parts written by the stagers
constexpr int ROWS_X_SLAB_BYTES = ROWS_TG * 16 * 4;      // [256 Gaussians][16 k] fp32, 64-byte rows, SWIZZLE_64B
constexpr int ROWS_W_SLOTS = 3;
constexpr int ROWS_W_SLOT_BYTES = 128 * 16 * 4 * 2;      // hi + lo of [128 x 16], no-swizzle K-major
constexpr size_t ROWS_SMEM_BYTES = (size_t)(ROWS_XH_SLOTS + ROWS_XL_SLOTS) * ROWS_X_SLAB_BYTES +
                                   (size_t)ROWS_W_SLOTS * ROWS_W_SLOT_BYTES + 1024;   // + slack to align to 1024

struct RowsBars {
  uint64_t xh_full[ROWS_XH_SLOTS], xh_free[ROWS_XH_SLOTS];
  uint64_t x_full[ROWS_XL_SLOTS], xl_free[ROWS_XL_SLOTS];
  uint64_t w_full[ROWS_W_SLOTS], w_free[ROWS_W_SLOTS];
  uint64_t acc_full[2], acc_free[2];
};

// The tensor core reads a raw fp32 word as tf32 by ignoring its low 13 bits, so the raw slab IS the
// hi part (hi = trunc(x)) and only lo = x - trunc(x) has to be written. lo keeps 11 of its <= 13
// significant bits: a product is perturbed by < 2^-20 relative (dropped lo*lo term included),
// inside the rounding noise of an fp32 GEMM of this length (tests/gemm_check.py: 1.4e-6 at K = 128).
__device__ __forceinline__ float4 split_lo(float4 x) {
  float4 l;
  l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
  l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
  l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xffffe000u);
  l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xffffe000u);
  return l;
}

// Epilogue of one block of 32 Gaussians for this lane's feature: v[j] = D^T[f][g0 + j].
// LD > 0: compile-time row stride of C (and of aux), so every access is base + immediate.
// FULL: all 32 Gaussians of the block exist (no per-row predicate).
// Returns the sum of the values it stored (MASK only: the bias gradient of the layer below).
template <int LD, bool FULL, int EPI>
__device__ __forceinline__ float rows_epilogue_block(const TcArgs& p, const TcBatch& bt, const float* v, int f, long g0,
                                                    int jmax, uint32_t mword) {
  const long ldc = LD ? LD : bt.ldc;
  const long ldx = LD ? LD : bt.ldaux;
  float* dst = bt.C + g0 * ldc + f;
  float csum = 0.f;
  if constexpr (EPI == TC_EPI_BIAS_ACT) {
    const float b = bt.bias != nullptr ? __ldg(bt.bias + f) : 0.f;
    if (p.relu) {
      if (bt.mask_out != nullptr) {
        // sign bits of this feature for the 32 Gaussians of the block, one word per (block, feature): the
        // ReLU mask the backward pass reads instead of the fp32 activations
        uint32_t word = 0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float x = fmaxf(v[j] + b, 0.f);
          word |= (x > 0.f ? 1u : 0u) << j;
          if (FULL || j < jmax) dst[j * ldc] = x;
        }
        bt.mask_out[(g0 >> 5) * bt.mask_ld + bt.mask_col + f] = word;
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (FULL || j < jmax) dst[j * ldc] = fmaxf(v[j] + b, 0.f);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (FULL || j < jmax) dst[j * ldc] = v[j] + b;
    }
  } else if (EPI == TC_EPI_MASK && bt.mask_in != nullptr) {
    const uint32_t word = mword;   // loaded by the caller before it waited for the accumulator
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float r = ((word >> j) & 1u) ? v[j] : 0.f;
      if (FULL || j < jmax) { dst[j * ldc] = r; csum += r; }
    }
  } else {
    // the three other epilogues read one value per output first; a clamped row index keeps the
    // loads unconditional, so they stay batched ahead of the stores
    const float* src = (EPI == TC_EPI_ACCUM) ? dst : (bt.aux + g0 * ldx + f);
    const long lds = (EPI == TC_EPI_ACCUM) ? ldc : ldx;
    float o[32];
    if (EPI == TC_EPI_ACCUM && !p.accumulate) {
#pragma unroll
      for (int j = 0; j < 32; ++j) o[j] = 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) o[j] = src[(FULL ? j : min(j, jmax - 1)) * lds];
    }
    if (EPI == TC_EPI_MASK) {
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float r = o[j] > 0.f ? v[j] : 0.f;
        if (FULL || j < jmax) { dst[j * ldc] = r; csum += r; }
      }
    } else if (EPI == TC_EPI_ACCUM) {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (FULL || j < jmax) dst[j * ldc] = o[j] + v[j];
    } else {  // TC_EPI_HEAD
      const float b = __ldg(bt.bias + f);
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (FULL || j < jmax) dst[j * ldc] = o[j] + (v[j] + b) * bt.scale;
    }
  }
  return csum;
}

template <int LD, int EPI>
__device__ __forceinline__ float rows_epilogue_dispatch(const TcArgs& p, const TcBatch& bt, const float* v, int f, long g0,
                                                        int jmax, uint32_t mword) {
  if (jmax == 32) return rows_epilogue_block<LD, true, EPI>(p, bt, v, f, g0, 32, mword);
  return rows_epilogue_block<LD, false, EPI>(p, bt, v, f, g0, jmax, mword);
}

// Step word (TcArgs::steps): batch | slab << 8 | npieces << 16 | nk << 20 | flags << 24
__device__ __forceinline__ int step_batch(uint32_t w) { return (int)(w & 0xff); }
__device__ __forceinline__ int step_slab(uint32_t w) { return (int)((w >> 8) & 0xff); }
__device__ __forceinline__ int step_npieces(uint32_t w) { return (int)((w >> 16) & 0xf); }
__device__ __forceinline__ int step_nk(uint32_t w) { return (int)((w >> 20) & 0xf); }
__device__ __forceinline__ int step_flags(uint32_t w) { return (int)(w >> 24); }

__device__ __forceinline__ uint64_t desc64(uint32_t hi, uint32_t lo) { return ((uint64_t)hi << 32) | lo; }

// Position in a ring of N slots plus the mbarrier phase of the current round. Consumers wait on
// full[idx] with `phase`; producers wait on free[idx] with `phase ^ 1` (passes at once in round 0).
template <int N>
struct RingPos {
  int idx = 0;
  uint32_t phase = 0;
  __device__ __forceinline__ void advance() {
    if (++idx == N) { idx = 0; phase ^= 1u; }
  }
};

// 2-D TMA tile load (tensor map in kernel parameter space), completion on an mbarrier.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const void* tmap, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

// EPI (TC_EPI_*) is a template parameter: the epilogue is a 32-fold unrolled block per kind, and a run-time switch kept
// the registers of all kinds alive in every launch.
template <int EPI>
__global__ void __launch_bounds__(ROWS_THREADS, 1) tc_rows_kernel(const __grid_constant__ TcArgs p) {
  extern __shared__ __align__(128) unsigned char rows_smem_raw[];
  __shared__ RowsBars bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // SWIZZLE_64B patterns repeat every 512 bytes: the slabs must sit on that grid
  unsigned char* sXh = rows_smem_raw + ((1024u - (smem_u32(rows_smem_raw) & 1023u)) & 1023u);
  unsigned char* sXl = sXh + (size_t)ROWS_XH_SLOTS * ROWS_X_SLAB_BYTES;
  unsigned char* sW = sXl + (size_t)ROWS_XL_SLOTS * ROWS_X_SLAB_BYTES;

  if (warp == 0) tmem_alloc(&s_tmem, 512);
  if (tid == 32) {
    for (int i = 0; i < ROWS_XH_SLOTS; ++i) { mbar_init(&bars.xh_full[i], 1); mbar_init(&bars.xh_free[i], 1); }
    for (int i = 0; i < ROWS_XL_SLOTS; ++i) { mbar_init(&bars.x_full[i], ROWS_CONV_THREADS); mbar_init(&bars.xl_free[i], 1); }
    for (int i = 0; i < ROWS_W_SLOTS; ++i) { mbar_init(&bars.w_full[i], 1); mbar_init(&bars.w_free[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars.acc_full[i], 1); mbar_init(&bars.acc_free[i], ROWS_EPI_WARPS * 32); }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  const int my_tiles = (p.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int nsteps = p.nsteps;
  const int total = my_tiles * nsteps;   // slabs this CTA walks; every role counts them the same way

  if (warp >= ROWS_CONV_WARP0) {
    // ------------------------------------------------------------------ X stagers (256 threads)
    // TMA has dropped slab g (256 Gaussians x 16 k, 64-byte rows, 64B swizzle) into the hi ring; the
    // stagers only add the lo part in the same layout. 16-byte chunk c of row r sits at
    // r*64 + ((c ^ ((r >> 1) & 3)) << 4); 8 consecutive lanes take 8 consecutive rows of one chunk, which
    // the swizzle spreads over all 32 banks.
    const int ct = tid - ROWS_CONV_WARP0 * 32;
    constexpr int NIT = (ROWS_TG * 4 + ROWS_CONV_THREADS - 1) / ROWS_CONV_THREADS;   // 16-byte chunks per thread
    int off[NIT];                                      // float4 index inside a slab (-1: none)
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
      const int q = ct + i * ROWS_CONV_THREADS;        // chunk q >> 8 of row q & 255: consecutive lanes, consecutive rows
      const int r = q & (ROWS_TG - 1), c = q >> 8;
      off[i] = (q < ROWS_TG * 4) ? r * 4 + (c ^ ((r >> 1) & 3)) : -1;
    }
    const float4* const hi0 = reinterpret_cast<const float4*>(sXh);
    float4* const lo0 = reinterpret_cast<float4*>(sXl);
    RingPos<ROWS_XH_SLOTS> hr;
    RingPos<ROWS_XL_SLOTS> lr;
    for (int g = 0; g < total; ++g) {
      mbar_wait(&bars.xh_full[hr.idx], hr.phase);       // TMA bytes of slab g have landed
      const float4* src = hi0 + hr.idx * (ROWS_X_SLAB_BYTES / 16);
      float4 r[NIT];
#pragma unroll
      for (int i = 0; i < NIT; ++i)
        if (off[i] >= 0) r[i] = src[off[i]];
      mbar_wait(&bars.xl_free[lr.idx], lr.phase ^ 1u);  // the UMMAs that read this lo slot are done
      float4* lo = lo0 + lr.idx * (ROWS_X_SLAB_BYTES / 16);
#pragma unroll
      for (int i = 0; i < NIT; ++i)
        if (off[i] >= 0) lo[off[i]] = split_lo(r[i]);
      fence_async_smem();
      mbar_arrive(&bars.x_full[lr.idx]);
      hr.advance();
      lr.advance();
    }
  } else if (warp == ROWS_LOAD_WARP) {
    // ------------------------------------------------------------------ loader (converged warp, elected lane issues)
    // per slab: one bulk copy of the prepared weight slab, one TMA tile of the activations. The
    // activation ring is deeper than the weight ring (weights come from L2, activations from HBM), so
    // the activation loads run XA slabs ahead of the weight loads: slot (g + XA) % 7 and weight slot
    // g % 3 are both released by the UMMAs of slab g - 3.
    constexpr int XA = ROWS_XH_SLOTS - ROWS_W_SLOTS;
    int sx = 0, sw = 0;
    long xtile_row = (long)blockIdx.x * ROWS_TG;
    RingPos<ROWS_W_SLOTS> wr;
    RingPos<ROWS_XH_SLOTS> hr;
    auto load_x = [&]() {
      const uint32_t st = p.steps[sx];
      mbar_wait(&bars.xh_free[hr.idx], hr.phase ^ 1u);
      if (elect_one()) {
        mbar_arrive_expect_tx(&bars.xh_full[hr.idx], ROWS_X_SLAB_BYTES);
        tma_load_2d(sXh + (size_t)hr.idx * ROWS_X_SLAB_BYTES, &p.xmap[step_batch(st)], step_slab(st) * 16, (int)xtile_row,
                    &bars.xh_full[hr.idx]);
      }
      __syncwarp();
      hr.advance();
      if (++sx == nsteps) { sx = 0; xtile_row += (long)gridDim.x * ROWS_TG; }
    };
    for (int g = 0; g < XA && g < total; ++g) load_x();
    for (int g = 0; g < total; ++g) {
      const uint32_t st = p.steps[sw];
      if (++sw == nsteps) sw = 0;
      mbar_wait(&bars.w_free[wr.idx], wr.phase ^ 1u);
      if (elect_one()) {
        mbar_arrive_expect_tx(&bars.w_full[wr.idx], ROWS_W_SLOT_BYTES);
        bulk_copy_g2s(sW + (size_t)wr.idx * ROWS_W_SLOT_BYTES, p.b[step_batch(st)].Bprep + (size_t)step_slab(st) * (128 * 8),
                      ROWS_W_SLOT_BYTES, &bars.w_full[wr.idx]);
      }
      __syncwarp();
      wr.advance();
      if (g + XA < total) load_x();
    }
  } else if (warp == ROWS_MMA_WARP) {
    // ------------------------------------------------------------------ MMA issuer (converged warp, elected lane issues)
    // Descriptor low word = start address >> 4 | LBO >> 4 << 16, high word = SBO >> 4 | version | layout.
    // W: no-swizzle K-major (LBO = 128 rows x 16 B between k-pieces, SBO = 128 B between 8-row groups).
    // X: SWIZZLE_64B K-major (8-row atoms of 512 B -> SBO = 512 B; a k-step advances the start by 32 B).
    const uint32_t idesc = umma_idesc_tf32(128, ROWS_TG);
    const uint32_t w_desc_hi = (128u >> 4) | (1u << 14);
    const uint32_t x_desc_hi = (512u >> 4) | (1u << 14) | (4u << 29);
    const uint32_t w_lo0 = ((smem_u32(sW) & 0x3FFFFu) >> 4) | ((128u * 16u >> 4) << 16);
    const uint32_t xh_lo0 = ((smem_u32(sXh) & 0x3FFFFu) >> 4) | (1u << 16);
    const uint32_t xl_lo0 = ((smem_u32(sXl) & 0x3FFFFu) >> 4) | (1u << 16);
    constexpr uint32_t W_SLOT = ROWS_W_SLOT_BYTES >> 4, W_LO = W_SLOT / 2, W_KS = (2u * 128u * 16u) >> 4;
    constexpr uint32_t X_SLAB = ROWS_X_SLAB_BYTES >> 4, X_KS = 32u >> 4;
    const bool single = p.single != 0;
    int s = 0, go = 0;
    uint32_t acc = tmem;
    RingPos<ROWS_XH_SLOTS> hr;
    RingPos<ROWS_XL_SLOTS> xr;
    RingPos<ROWS_W_SLOTS> wr;
    for (int g = 0; g < total; ++g) {
      const uint32_t st = p.steps[s];
      if (++s == nsteps) s = 0;
      const int fl = step_flags(st);
      if (fl & TC_STEP_FIRST) {
        mbar_wait(&bars.acc_free[go & 1], (uint32_t)(((go >> 1) & 1) ^ 1));
        acc = tmem + (uint32_t)(go & 1) * ROWS_TG;
      }
      mbar_wait(&bars.x_full[xr.idx], xr.phase);        // lo written (and the stagers saw the hi slab land)
      mbar_wait(&bars.w_full[wr.idx], wr.phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t wh = w_lo0 + (uint32_t)wr.idx * W_SLOT;
        const uint32_t xh = xh_lo0 + (uint32_t)hr.idx * X_SLAB, xl = xl_lo0 + (uint32_t)xr.idx * X_SLAB;
        const uint32_t first = (fl & TC_STEP_FIRST) ? 0u : 1u;
        if (single) {
          umma_tf32(acc, desc64(w_desc_hi, wh), desc64(x_desc_hi, xh), idesc, first);
          if (step_nk(st) == 2) umma_tf32(acc, desc64(w_desc_hi, wh + W_KS), desc64(x_desc_hi, xh + X_KS), idesc, 1u);
        } else {
          umma_tf32(acc, desc64(w_desc_hi, wh + W_LO), desc64(x_desc_hi, xh), idesc, first);  // small terms first
          umma_tf32(acc, desc64(w_desc_hi, wh), desc64(x_desc_hi, xl), idesc, 1u);
          umma_tf32(acc, desc64(w_desc_hi, wh), desc64(x_desc_hi, xh), idesc, 1u);
          if (step_nk(st) == 2) {
            umma_tf32(acc, desc64(w_desc_hi, wh + W_KS + W_LO), desc64(x_desc_hi, xh + X_KS), idesc, 1u);
            umma_tf32(acc, desc64(w_desc_hi, wh + W_KS), desc64(x_desc_hi, xl + X_KS), idesc, 1u);
            umma_tf32(acc, desc64(w_desc_hi, wh + W_KS), desc64(x_desc_hi, xh + X_KS), idesc, 1u);
          }
        }
        umma_commit(&bars.w_free[wr.idx]);
        umma_commit(&bars.xl_free[xr.idx]);
        umma_commit(&bars.xh_free[hr.idx]);
        if (fl & TC_STEP_LAST) umma_commit(&bars.acc_full[go & 1]);
      }
      __syncwarp();
      hr.advance();
      xr.advance();
      wr.advance();
      if (fl & TC_STEP_LAST) ++go;
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 0..7)
    // warp e: features 32*(e&3) .. +31 (its TMEM lane quarter); blocks of 32 Gaussians alternate
    // between the two warps that share a quarter.
    const int q = warp & 3, par = warp >> 2;
    const int f = q * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    int go = 0;
    for (int tl = 0; tl < my_tiles; ++tl) {
      const long tile = (long)blockIdx.x + (long)tl * gridDim.x;
      for (int o = 0; o < p.nbatch; ++o, ++go) {
        const TcBatch& bt = p.b[o];
        const int buf = go & 1;
        // ReLU-mask words, loaded before the accumulator wait: a dependent global load per block otherwise, and
        // the dZ launches (K = 16 per head) are epilogue-latency bound
        const bool use_mask = EPI == TC_EPI_MASK && bt.mask_in != nullptr && f < bt.n_valid;
        auto mask_word = [&](int blk) -> uint32_t {
          long g0 = tile * ROWS_TG + blk * 32;
          if (!use_mask || g0 >= (long)p.M) return 0u;
#ifdef ED3DGS_TIMING_KNOBS
          if (p.dbg & 64) g0 &= 16383;
#endif
          return __ldg(bt.mask_in + (g0 >> 5) * bt.mask_ld + bt.mask_col + f);
        };
        // the mask kind has registers to spare: all four words of this warp's blocks before the wait
        uint32_t mnext = 0u, mw1 = 0u, mw2 = 0u, mw3 = 0u;
        if constexpr (EPI == TC_EPI_MASK) {
          mnext = mask_word(par); mw1 = mask_word(par + 2); mw2 = mask_word(par + 4); mw3 = mask_word(par + 6);
        }
        mbar_wait(&bars.acc_full[buf], (uint32_t)((go >> 1) & 1));
        tc_fence_after();
        if (q * 32 < bt.n_valid) {
          const bool same_ld = (EPI == TC_EPI_BIAS_ACT || EPI == TC_EPI_ACCUM || bt.ldaux == bt.ldc);
          const int ld_case = !same_ld ? 0 : (bt.ldc == 640 ? 640 : (bt.ldc == 128 ? 128 : 0));
          float colsum = 0.f;
          for (int bi = 0; bi < ROWS_TG / 64; ++bi) {
            const int blk = par + 2 * bi;
            const uint32_t mword = bi == 0 ? mnext : (bi == 1 ? mw1 : (bi == 2 ? mw2 : mw3));
            long g0 = tile * ROWS_TG + blk * 32;
            const int jmax = (int)min(32L, (long)p.M - g0);
            if (jmax <= 0) break;
#ifdef ED3DGS_TIMING_KNOBS
            if (p.dbg & 64) g0 &= 16383;   // timing knob: outputs stay in an L2-resident window
#endif
            float v[32];
            tmem_ld32(lane_base + (uint32_t)(buf * ROWS_TG + blk * 32), v);
#ifdef ED3DGS_TIMING_KNOBS
            if (f < bt.n_valid && !(p.dbg & 16)) {
#else
            if (f < bt.n_valid) {
#endif
              if (ld_case == 640) colsum += rows_epilogue_dispatch<640, EPI>(p, bt, v, f, g0, jmax, mword);
              else if (ld_case == 128) colsum += rows_epilogue_dispatch<128, EPI>(p, bt, v, f, g0, jmax, mword);
              else colsum += rows_epilogue_dispatch<0, EPI>(p, bt, v, f, g0, jmax, mword);
            }
          }
          // column sums of a masked gradient = bias gradient of the layer that produced the mask
          if (EPI == TC_EPI_MASK && bt.bias_grad != nullptr && f < bt.n_valid) atomicAdd(bt.bias_grad + f, colsum);
        }
        tc_fence_before();
        mbar_arrive(&bars.acc_free[buf]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

}  // namespace ed3dgs
