// Stable LSD radix sort of (u64 key, u32 value) pairs over key bits [0, end_bit), 8 bits per pass, hand-written for
// sm_100a. Replaces cub::DeviceRadixSort::SortPairs at rasterizer_impl.cu:381-386 of the reference. Stability
// matters: equal (tile, depth) keys must stay in emission order (ascending Gaussian index) or the blend order -- and
// with it point_list -- would differ from the reference.
//
// One-sweep organisation: the keys are read ONCE up front for the digit histograms of ALL passes, and every pass is a
// single kernel that reads a 4096-key tile, ranks it, obtains the tile's global digit offsets from its predecessors by
// a decoupled look-back, and scatters:
//   radix_hist_all   : global 256-bin histogram of every pass in one read of the keys -> digit_base[pass][256]
//   radix_digit_scan : exclusive scan of each pass's 256 bins (one CTA per pass)
//   radix_onesweep   : per pass. Tiles are handed out by an atomic ticket, so every predecessor of a running tile is
//                      itself running or done. Thread d owns digit d: it publishes the tile's count of d
//                      (status[tile][d] = count | AGGREGATE), walks back over the predecessors' status words adding
//                      aggregates until it meets an inclusive PREFIX, publishes its own inclusive prefix, and the tile
//                      then writes its keys to digit_base[d] + exclusive prefix + local rank. Count and flag share one
//                      32-bit word, so no fence separates them. Inside the tile: warp match ranking in (warp, round,
//                      lane) order + shared-memory reorder, coalesced writes.
// HBM traffic: 8 B/key once + (12 B read + 12 B write) per pass; the previous three-kernel-per-pass version read the
// keys twice per pass and spent five launches on each.
#include "raster_state.cuh"

namespace ed3dgs {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = SORT_TILE / SORT_THREADS;          // 16 keys per thread
constexpr int SORT_WARP_KEYS = SORT_TILE / SORT_WARPS;        // 512 keys per warp

// Exclusive scan of `count` u32 in place by one CTA of 1024 threads.
__global__ void __launch_bounds__(1024) scan_exclusive_1cta(uint32_t* __restrict__ data, uint32_t count,
                                                           uint32_t* __restrict__ total_out) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  // each thread owns 4 consecutive elements per sweep -> 4096 per sweep
  for (uint32_t base = 0; base < count; base += 4096) {
    const uint32_t i0 = base + threadIdx.x * 4;
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (i0 + k < count) ? data[i0 + k] : 0u;
    const uint32_t tsum = v[0] + v[1] + v[2] + v[3];
    uint32_t inc = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      s_warp[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const uint32_t carry = s_carry;
    uint32_t excl = carry + (warp ? s_warp[warp - 1] : 0u) + (inc - tsum);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (i0 + k < count) data[i0 + k] = excl;
      excl += v[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + s_warp[31];
    __syncthreads();
  }
  if (total_out && threadIdx.x == 0) *total_out = s_carry;
}

struct ScatterSmem {
  uint64_t keys[SORT_TILE];
  uint32_t vals[SORT_TILE];
  uint32_t warp_cnt[SORT_WARPS][256];
  uint32_t local_start[256];
  uint32_t global_base[256];
  uint32_t warp_tot[SORT_WARPS];
  uint32_t tile;
};

constexpr uint32_t OS_FLAG_AGG = 1u << 30, OS_FLAG_PREFIX = 2u << 30, OS_COUNT_MASK = (1u << 30) - 1u;

// Global digit histograms of all passes in one read of the keys. grid-stride over tiles; per-CTA shared histograms.
// The depth digits (key bits below 32) of neighbouring keys are as good as random: plain shared-memory atomics, hardly any
// same-address collisions. The tile digits (bits 32 and up) of keys that follow each other in emission order are mostly
// equal: those are aggregated over the warp first (match.any costs ~2 issue slots per distinct value, which made it the
// bottleneck when it was applied to the depth digits too: ncu showed this kernel at 0.6 TB/s with the SMs 84 % busy).
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_all(const uint64_t* __restrict__ keys, uint32_t n, int npasses, int end_bit, uint32_t* __restrict__ digit_base) {
  __shared__ uint32_t s_hist[SORT_MAX_PASSES][256];
  for (int i = threadIdx.x; i < SORT_MAX_PASSES * 256; i += SORT_THREADS) (&s_hist[0][0])[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const uint32_t nround = (n + SORT_THREADS - 1) / SORT_THREADS;
  for (uint32_t r = blockIdx.x; r < nround; r += gridDim.x) {
    const uint32_t idx = r * SORT_THREADS + threadIdx.x;
    const bool valid = idx < n;
    const uint64_t k = valid ? keys[idx] : 0ull;
    for (int p = 0; p < npasses; ++p) {
      const int nbits = min(8, end_bit - 8 * p);
      const uint32_t d = valid ? ((uint32_t)(k >> (8 * p)) & ((1u << nbits) - 1u)) : (256u + lane);
      if (8 * p + 8 <= 32) {
        if (valid) atomicAdd(&s_hist[p][d], 1u);
      } else {
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (valid && lane == __ffs(peers) - 1) atomicAdd(&s_hist[p][d], (uint32_t)__popc(peers));
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < npasses * 256; i += SORT_THREADS) {
    const uint32_t c = (&s_hist[0][0])[i];
    if (c) atomicAdd(digit_base + i, c);
  }
}

// Exclusive scan of the 256 bins of pass blockIdx.x, in place.
__global__ void __launch_bounds__(256) radix_digit_scan(uint32_t* __restrict__ digit_base) {
  __shared__ uint32_t s_warp[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* h = digit_base + blockIdx.x * 256;
  const uint32_t c = h[threadIdx.x];
  uint32_t inc = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t before = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w)
    if (w < warp) before += s_warp[w];
  h[threadIdx.x] = before + inc - c;
}

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(SORT_THREADS)
radix_onesweep(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
               uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
               int shift, uint32_t mask, const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ status,
               uint32_t* __restrict__ ticket) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  ScatterSmem& s = *reinterpret_cast<ScatterSmem*>(smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;

  if (threadIdx.x == 0) s.tile = atomicAdd(ticket, 1u);
#pragma unroll
  for (int w = 0; w < SORT_WARPS; ++w) s.warp_cnt[w][threadIdx.x] = 0;
  __syncthreads();
  const uint32_t tile = s.tile;

  // Stable order inside the tile = (warp, round, lane); a warp owns 512 consecutive keys.
  const uint32_t wbase = tile * SORT_TILE + warp * SORT_WARP_KEYS;
  uint64_t k[SORT_ITEMS];
  uint32_t v[SORT_ITEMS];
  uint32_t rank[SORT_ITEMS];  // rank of the key among equal digits inside this warp
#pragma unroll
  for (int r = 0; r < SORT_ITEMS; ++r) {
    const uint32_t idx = wbase + r * 32 + lane;
    const bool valid = idx < n;
    k[r] = valid ? keys_in[idx] : ~0ull;
    v[r] = valid ? vals_in[idx] : 0u;
  }
#pragma unroll
  for (int r = 0; r < SORT_ITEMS; ++r) {
    const uint32_t idx = wbase + r * 32 + lane;
    const bool valid = idx < n;
    const uint32_t d = valid ? ((uint32_t)(k[r] >> shift) & mask) : (256u + lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(peers) - 1;
    uint32_t prev = 0;
    if (valid && lane == leader) {
      prev = s.warp_cnt[warp][d];
      s.warp_cnt[warp][d] = prev + __popc(peers);
    }
    prev = __shfl_sync(0xffffffffu, prev, leader);
    rank[r] = prev + __popc(peers & lt_mask);
    __syncwarp();
  }
  __syncthreads();

  // Per digit (thread d): exclusive offsets over warps and the tile's count; publish it, then look back.
  {
    const int d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) {
      const uint32_t t = s.warp_cnt[w][d];
      s.warp_cnt[w][d] = run;
      run += t;
    }
    uint32_t* my = status + (size_t)tile * 256 + d;
    st_relaxed(my, run | (tile == 0 ? OS_FLAG_PREFIX : OS_FLAG_AGG));
    // local exclusive scan over digits while the predecessors publish
    uint32_t inc = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s.warp_tot[warp] = inc;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w)
      if (w < warp) before += s.warp_tot[w];
    s.local_start[d] = before + inc - run;
    // decoupled look-back: sum the aggregates of the predecessors down to the first inclusive prefix
    uint32_t excl = 0;
    if (tile > 0) {
      const uint32_t* st = status + (size_t)(tile - 1) * 256 + d;
      for (;;) {
        const uint32_t x = ld_relaxed(st);
        if ((x >> 30) == 0u) continue;           // predecessor has not published yet
        excl += x & OS_COUNT_MASK;
        if (x & OS_FLAG_PREFIX) break;
        st -= 256;
      }
      st_relaxed(my, (excl + run) | OS_FLAG_PREFIX);
    }
    s.global_base[d] = digit_base[d] + excl;
  }
  __syncthreads();

#pragma unroll
  for (int r = 0; r < SORT_ITEMS; ++r) {
    const uint32_t idx = wbase + r * 32 + lane;
    if (idx < n) {
      const uint32_t d = (uint32_t)(k[r] >> shift) & mask;
      const uint32_t pos = s.local_start[d] + s.warp_cnt[warp][d] + rank[r];
      s.keys[pos] = k[r];
      s.vals[pos] = v[r];
    }
  }
  __syncthreads();

  const uint32_t tile_n = min((uint32_t)SORT_TILE, n - tile * SORT_TILE);
  for (uint32_t i = threadIdx.x; i < tile_n; i += SORT_THREADS) {
    const uint64_t key = s.keys[i];
    const uint32_t d = (uint32_t)(key >> shift) & mask;
    const uint32_t dst = s.global_base[d] + (i - s.local_start[d]);
    keys_out[dst] = key;
    vals_out[dst] = s.vals[i];
  }
}

// Sorts pairs; returns the index (0/1) of the ping-pong half that holds the result, -1 on a CUDA error.
int radix_sort_pairs(BinningState& b, uint32_t n, int end_bit, cudaStream_t stream) {
  if (n == 0) return 0;
  const int npasses = (end_bit + 7) / 8;
  if (npasses > SORT_MAX_PASSES || n > OS_COUNT_MASK) return -1;
  // a per-device attribute, set on every call (cheap; a process-wide flag would miss the second GPU of the process)
  if (cudaFuncSetAttribute(radix_onesweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScatterSmem)) !=
      cudaSuccess)
    return -1;
  const uint32_t nblocks = (uint32_t)sort_blocks(n);
  // aux layout: digit_base[SORT_MAX_PASSES][256] | ticket[SORT_MAX_PASSES] (padded to 256) | status[pass][nblocks][256]
  uint32_t* digit_base = b.hist;
  uint32_t* ticket = b.hist + SORT_MAX_PASSES * 256;
  uint32_t* status = ticket + 256;
  if (cudaMemsetAsync(b.hist, 0, sort_aux_words(n, npasses) * sizeof(uint32_t), stream) != cudaSuccess) return -1;
  const uint32_t hist_grid = min(nblocks * (SORT_TILE / SORT_THREADS), 148u * 8u);
  radix_hist_all<<<hist_grid, SORT_THREADS, 0, stream>>>(b.keys[2], n, npasses, end_bit, digit_base);
  radix_digit_scan<<<npasses, 256, 0, stream>>>(digit_base);
  int cur = 2, dst = 0;  // first pass reads the (preserved) emission buffers
  for (int p = 0; p < npasses; ++p) {
    const int shift = 8 * p;
    const int nbits = min(8, end_bit - shift);
    const uint32_t mask = (1u << nbits) - 1u;
    radix_onesweep<<<nblocks, SORT_THREADS, sizeof(ScatterSmem), stream>>>(
        b.keys[cur], b.vals[cur], b.keys[dst], b.vals[dst], n, shift, mask, digit_base + p * 256,
        status + (size_t)p * nblocks * 256, ticket + p);
    cur = dst;
    dst ^= 1;
  }
  return cudaGetLastError() == cudaSuccess ? cur : -1;
}

}  // namespace ed3dgs
