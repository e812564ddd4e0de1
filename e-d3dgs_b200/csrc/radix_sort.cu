// Stable LSD radix sort of (u64 key, u32 value) pairs over key bits [0, end_bit), 8 bits per
// pass, hand-written for sm_100a. Replaces cub::DeviceRadixSort::SortPairs at
// rasterizer_impl.cu:381-386 of the reference. Stability matters: equal (tile, depth) keys must
// stay in emission order (ascending Gaussian index) or the blend order -- and with it
// point_list -- would differ from the reference.
//
// Per pass, three launches, no inter-block spinning:
//   radix_hist    : 256-bin digit histogram of each 4096-key block -> hist[digit][block]
//   radix_scan    : exclusive scan of the flattened [digit][block] table (one CTA)
//   radix_scatter : block-stable ranking with warp match + shared-memory reorder, then
//                   coalesced writes to base[digit][block] + local rank
// HBM traffic per pass: 8 B/key (hist) + 12 B read + 12 B write.
#include "raster_state.cuh"

namespace ed3dgs {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = SORT_TILE / SORT_THREADS;          // 16 keys per thread
constexpr int SORT_WARP_KEYS = SORT_TILE / SORT_WARPS;        // 512 keys per warp

__global__ void __launch_bounds__(SORT_THREADS)
radix_hist(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t mask,
           uint32_t nblocks, uint32_t* __restrict__ hist) {
  __shared__ uint32_t s_hist[256];
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * SORT_TILE;
#pragma unroll 4
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const uint32_t idx = base + i * SORT_THREADS + threadIdx.x;
    if (idx < n) {
      const uint32_t d = (uint32_t)(keys[idx] >> shift) & mask;
      atomicAdd(&s_hist[d], 1u);
    }
  }
  __syncthreads();
  hist[threadIdx.x * nblocks + blockIdx.x] = s_hist[threadIdx.x];
}

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

// Parallel in-place exclusive scan: 2048-element chunks scanned by 256-thread CTAs, the chunk
// totals by one CTA, then the chunk prefix is added back. `sums` needs ceil(n/2048)+1 words.
constexpr int XS_TILE = 2048;
constexpr int XS_THREADS = 256;
constexpr int XS_ITEMS = XS_TILE / XS_THREADS;

__global__ void __launch_bounds__(XS_THREADS)
scan_chunk_exclusive(uint32_t* __restrict__ data, uint32_t n, uint32_t* __restrict__ sums) {
  __shared__ uint32_t s_warp[XS_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t i0 = blockIdx.x * XS_TILE + threadIdx.x * XS_ITEMS;
  uint32_t v[XS_ITEMS];
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < XS_ITEMS; ++k) {
    v[k] = (i0 + k < n) ? data[i0 + k] : 0u;
    tsum += v[k];
  }
  uint32_t inc = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t before = 0;
#pragma unroll
  for (int w = 0; w < XS_THREADS / 32; ++w)
    if (w < warp) before += s_warp[w];
  uint32_t run = before + inc - tsum;
#pragma unroll
  for (int k = 0; k < XS_ITEMS; ++k) {
    if (i0 + k < n) data[i0 + k] = run;
    run += v[k];
  }
  if (threadIdx.x == XS_THREADS - 1) sums[blockIdx.x] = run;
}

__global__ void __launch_bounds__(XS_THREADS)
scan_chunk_add(uint32_t* __restrict__ data, uint32_t n, const uint32_t* __restrict__ prefix) {
  const uint32_t add = prefix[blockIdx.x];
  if (add == 0) return;
  const uint32_t i0 = blockIdx.x * XS_TILE + threadIdx.x * XS_ITEMS;
#pragma unroll
  for (int k = 0; k < XS_ITEMS; ++k)
    if (i0 + k < n) data[i0 + k] += add;
}

void exclusive_scan_inplace(uint32_t* data, uint32_t n, uint32_t* sums, cudaStream_t stream) {
  const uint32_t nchunks = (n + XS_TILE - 1) / XS_TILE;
  scan_chunk_exclusive<<<nchunks, XS_THREADS, 0, stream>>>(data, n, sums);
  if (nchunks > 1) {
    scan_exclusive_1cta<<<1, 1024, 0, stream>>>(sums, nchunks, nullptr);
    scan_chunk_add<<<nchunks, XS_THREADS, 0, stream>>>(data, n, sums);
  }
}

struct ScatterSmem {
  uint64_t keys[SORT_TILE];
  uint32_t vals[SORT_TILE];
  uint32_t warp_cnt[SORT_WARPS][256];
  uint32_t local_start[256];
  uint32_t global_base[256];
  uint32_t warp_tot[SORT_WARPS];
};

__global__ void __launch_bounds__(SORT_THREADS)
radix_scatter(const uint64_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
              uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, uint32_t n,
              int shift, uint32_t mask, uint32_t nblocks, const uint32_t* __restrict__ hist) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  ScatterSmem& s = *reinterpret_cast<ScatterSmem*>(smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;

#pragma unroll
  for (int w = 0; w < SORT_WARPS; ++w) s.warp_cnt[w][threadIdx.x] = 0;
  s.global_base[threadIdx.x] = hist[threadIdx.x * nblocks + blockIdx.x];
  __syncthreads();

  // Stable order inside the block = (warp, round, lane); a warp owns 512 consecutive keys.
  const uint32_t wbase = blockIdx.x * SORT_TILE + warp * SORT_WARP_KEYS;
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

  // Per digit: exclusive offsets over warps, block total, then exclusive scan over digits.
  {
    const int d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; ++w) {
      const uint32_t t = s.warp_cnt[w][d];
      s.warp_cnt[w][d] = run;
      run += t;
    }
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

  const uint32_t tile_n = min((uint32_t)SORT_TILE, n - blockIdx.x * SORT_TILE);
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
  // a per-device attribute, set on every call (cheap; a process-wide flag would miss the second GPU of the process)
  if (cudaFuncSetAttribute(radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScatterSmem)) !=
      cudaSuccess)
    return -1;
  const uint32_t nblocks = (uint32_t)sort_blocks(n);
  int cur = 2, dst = 0;  // first pass reads the (preserved) emission buffers
  for (int shift = 0; shift < end_bit; shift += 8) {
    const int nbits = min(8, end_bit - shift);
    const uint32_t mask = (1u << nbits) - 1u;
    radix_hist<<<nblocks, SORT_THREADS, 0, stream>>>(b.keys[cur], n, shift, mask, nblocks, b.hist);
    exclusive_scan_inplace(b.hist, 256u * nblocks, b.hist + 256u * nblocks, stream);
    radix_scatter<<<nblocks, SORT_THREADS, sizeof(ScatterSmem), stream>>>(
        b.keys[cur], b.vals[cur], b.keys[dst], b.vals[dst], n, shift, mask, nblocks, b.hist);
    cur = dst;
    dst ^= 1;
  }
  return cur;
}

}  // namespace ed3dgs
