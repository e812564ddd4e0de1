// Deformation network (coarse + fine stage) of E-D3DGS on sm_100a tensor cores.
// Replaces scene/deformation.py:15-148 of the reference (22 nn.Linear calls on cuBLAS SGEMM plus
// ~40 elementwise kernels per render). What changes structurally:
//  * The 256-wide temporal half of the input row is identical for every Gaussian
//    (deformation.py:65 `.repeat`), so layer 0 is  E[N,32] * W0[:,256:]^T + (W0[:,:256] * tau + b0):
//    per-Gaussian K drops from 288 to 32 and the [N,288] concat is never materialised.
//  * All GEMMs (forward, input-gradient, weight-gradient) run through one tcgen05/TMEM building
//    kernel pair (tc_rows.cuh, tc_wgrad.cuh) with bias / ReLU / ReLU-mask / head-update epilogues fused in.
//  * Bias gradients are the column sums of the ReLU-masked gradients, accumulated in that epilogue.
#include <cstdint>
#include <cstdio>
#include "raster_api.h"
#include "tc_gemm.cuh"
#include "profiler.h"
#include "tc_launch.cuh"
#include <vector>
#include <cstdlib>

namespace ed3dgs {

// ---------------------------------------------------------------------------------------------
// ---------------------------------------------------------------------------------------------
// Temporal embedding (deformation.py:53-67): F.interpolate(bilinear, align_corners=True) of
// weight[max_emb, dim] to n_rows rows, then F.grid_sample(bilinear, reflection,
// align_corners=True) at x = column centres, y = (t - 0.5) * 2. One thread per column.
struct TembTap {
  int r0a, r1a, r0b, r1b;   // source rows of resized rows y0 (a) and y1 (b)
  float la, lb;             // row lerp weights of the resize
  int y0, y1;               // resized rows
  float wy0, wy1;           // grid_sample row weights (0 when the row is out of range)
  float dy_dt;              // d iy / d t (0 where clipped)
};

__device__ __forceinline__ float reflect_coord(float in, int twice_low, int twice_high, float& mult) {
  // ATen reflect_coordinates_set_grad (GridSampler.cuh)
  if (twice_low == twice_high) { mult = 0.f; return 0.f; }
  int grad_in_mult;
  const float mn = (float)twice_low / 2;
  const float span = (float)(twice_high - twice_low) / 2;
  in = in - mn;
  if (in < 0.f) { grad_in_mult = -1; in = -in; } else { grad_in_mult = 1; }
  const float extra = fmodf(in, span);
  const int flips = (int)floorf(in / span);
  if (flips % 2 == 0) { mult = (float)grad_in_mult; return extra + mn; }
  mult = (float)(-grad_in_mult);
  return span - extra + mn;
}

__device__ __forceinline__ float grid_coord(float g, int size, float& dmult) {
  // unnormalize (align_corners=True) -> reflect over [0, size-1] -> clip
  float c = ((g + 1.f) / 2) * (size - 1);
  float m_unnorm = (float)(size - 1) / 2;
  float m_ref;
  c = reflect_coord(c, 0, 2 * (size - 1), m_ref);
  float m_clip = 1.f;
  if (c <= 0.f) { c = 0.f; m_clip = 0.f; }
  else if (c >= (float)(size - 1)) { c = (float)(size - 1); m_clip = 0.f; }
  dmult = m_unnorm * m_ref * m_clip;
  return c;
}

__device__ __forceinline__ void resize_row(int r, int n_rows, int max_emb, int& r0, int& r1, float& lam) {
  const float scale = (n_rows > 1) ? (float)(max_emb - 1) / (n_rows - 1) : 0.f;
  const float src = scale * r;
  r0 = (int)src;
  r1 = r0 + ((r0 < max_emb - 1) ? 1 : 0);
  lam = src - r0;
}

__device__ __forceinline__ TembTap temb_taps(float t, int n_rows, int max_emb) {
  TembTap tp;
  float dm;
  const float gy = (t - 0.5f) * 2;
  const float iy = grid_coord(gy, n_rows, dm);
  tp.dy_dt = dm * 2.f;
  tp.y0 = (int)floorf(iy);
  tp.y1 = tp.y0 + 1;
  const float fy = iy - tp.y0;
  tp.wy0 = (tp.y0 >= 0 && tp.y0 < n_rows) ? (1.f - fy) : 0.f;
  tp.wy1 = (tp.y1 >= 0 && tp.y1 < n_rows) ? fy : 0.f;
  const int ya = min(max(tp.y0, 0), n_rows - 1), yb = min(max(tp.y1, 0), n_rows - 1);
  resize_row(ya, n_rows, max_emb, tp.r0a, tp.r1a, tp.la);
  resize_row(yb, n_rows, max_emb, tp.r0b, tp.r1b, tp.lb);
  return tp;
}

__global__ void temb_forward(const float* __restrict__ weight, int max_emb, int dim, int n_rows,
                             const float* __restrict__ t_dev, float* __restrict__ tau) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= dim) return;
  const TembTap tp = temb_taps(*t_dev, n_rows, max_emb);
  float dmx;
  const float gx = ((float)c / (dim - 1) - 0.5f) * 2;
  const float ix = grid_coord(gx, dim, dmx);
  const int x0 = (int)floorf(ix), x1 = x0 + 1;
  const float fx = ix - x0;
  auto resized = [&](int r0, int r1, float lam, int x) -> float {
    if (x < 0 || x >= dim) return 0.f;
    return (1.f - lam) * weight[(long)r0 * dim + x] + lam * weight[(long)r1 * dim + x];
  };
  const float va = (1.f - fx) * resized(tp.r0a, tp.r1a, tp.la, x0) + fx * resized(tp.r0a, tp.r1a, tp.la, x1);
  const float vb = (1.f - fx) * resized(tp.r0b, tp.r1b, tp.lb, x0) + fx * resized(tp.r0b, tp.r1b, tp.lb, x1);
  tau[c] = tp.wy0 * va + tp.wy1 * vb;
}

__global__ void temb_backward(const float* __restrict__ weight, int max_emb, int dim, int n_rows,
                              const float* __restrict__ t_dev, const float* __restrict__ d_tau,
                              float* __restrict__ d_weight, float* __restrict__ d_t) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  float dt_local = 0.f;
  if (c < dim) {
    const TembTap tp = temb_taps(*t_dev, n_rows, max_emb);
    float dmx;
    const float gx = ((float)c / (dim - 1) - 0.5f) * 2;
    const float ix = grid_coord(gx, dim, dmx);
    const int x0 = (int)floorf(ix), x1 = x0 + 1;
    const float fx = ix - x0;
    const float g = d_tau[c];
    auto resized = [&](int r0, int r1, float lam, int x) -> float {
      if (x < 0 || x >= dim) return 0.f;
      return (1.f - lam) * weight[(long)r0 * dim + x] + lam * weight[(long)r1 * dim + x];
    };
    auto scatter = [&](int r0, int r1, float lam, int x, float w) {
      if (x < 0 || x >= dim || w == 0.f) return;
      atomicAdd(d_weight + (long)r0 * dim + x, w * (1.f - lam));
      atomicAdd(d_weight + (long)r1 * dim + x, w * lam);
    };
    scatter(tp.r0a, tp.r1a, tp.la, x0, g * tp.wy0 * (1.f - fx));
    scatter(tp.r0a, tp.r1a, tp.la, x1, g * tp.wy0 * fx);
    scatter(tp.r0b, tp.r1b, tp.lb, x0, g * tp.wy1 * (1.f - fx));
    scatter(tp.r0b, tp.r1b, tp.lb, x1, g * tp.wy1 * fx);
    const float va = (1.f - fx) * resized(tp.r0a, tp.r1a, tp.la, x0) + fx * resized(tp.r0a, tp.r1a, tp.la, x1);
    const float vb = (1.f - fx) * resized(tp.r0b, tp.r1b, tp.lb, x0) + fx * resized(tp.r0b, tp.r1b, tp.lb, x1);
    const float in_a = (tp.y0 >= 0 && tp.y0 < n_rows) ? va : 0.f;
    const float in_b = (tp.y1 >= 0 && tp.y1 < n_rows) ? vb : 0.f;
    dt_local = g * (in_b - in_a) * tp.dy_dt;
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) dt_local += __shfl_xor_sync(0xffffffffu, dt_local, o);
  if ((threadIdx.x & 31) == 0 && dt_local != 0.f) atomicAdd(d_t, dt_local);
}

// ---------------------------------------------------------------------------------------------
// c1[f] = b0[f] + sum_j W0[f][j] * tau[j]   (one warp per feature)
__global__ void layer0_temporal_bias(const float* __restrict__ W0, const float* __restrict__ b0, long ld, int temb,
                                     const float* __restrict__ tau, float* __restrict__ c1, int W) {
  const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (f >= W) return;
  float acc = 0.f;
  for (int j = lane; j < temb; j += 32) acc += W0[(long)f * ld + j] * tau[j];
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) c1[f] = acc + b0[f];
}

// db0 += dc1 ; dW0[f][j] += dc1[f] * tau[j] ; dtau[j] = sum_f W0[f][j] * dc1[f]
// One block per 32 temporal columns j; its 8 warps split the W features, partial sums meet in shared memory.
__global__ void __launch_bounds__(256)
layer0_temporal_backward(const float* __restrict__ W0, long ld, int temb, int W, const float* __restrict__ tau,
                         const float* __restrict__ dc1, float* __restrict__ dW0, float* __restrict__ db0,
                         float* __restrict__ dtau) {
  __shared__ float s_part[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (j < temb) {
    const float tj = tau[j];
    for (int f = warp; f < W; f += 8) {
      const float g = dc1[f];
      acc += W0[(long)f * ld + j] * g;
      if (dW0) dW0[(long)f * ld + j] += g * tj;
    }
  }
  s_part[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && j < temb) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_part[w][lane];
    dtau[j] = t;
  }
  if (db0 && blockIdx.x == 0)
    for (int i = threadIdx.x; i < W; i += blockDim.x) db0[i] += dc1[i];
}

// ---------------------------------------------------------------------------------------------
// Upstream gradients -> G[N][80] (head h at column 16*?): pos 0, scales 8, rotations 16, opacity 24,
// rgb 32..79, each scaled by its coefficient and zero padded; plus column sums (second-layer bias
// gradients).
__constant__ int kHeadDim[5] = {3, 3, 4, 1, 48};
__constant__ int kHeadOff[5] = {0, 8, 16, 24, 32};
static const int hHeadDim[5] = {3, 3, 4, 1, 48};
static const int hHeadOff[5] = {0, 8, 16, 24, 32};
static const int hHeadPad[5] = {8, 8, 8, 8, 48};
constexpr int GW = 80;

struct GradPtrs {
  const float* d_out[5];
  const float* d_sub[5];
  float coef[5];
};

// Thread = (row, 16-byte chunk of the 80-float row): chunks 0..7 hold the four small heads (one head per pair of
// chunks, second chunk all padding), chunks 8..19 the 48 rgb columns as aligned float4 loads. A block covers 16
// rows per round, four rounds in flight per thread; column sums go thread -> shared -> one atomic per column.
constexpr int BHG_THREADS = 320;                 // 16 rows x 20 chunks
constexpr int BHG_UNROLL = 4;

__device__ __forceinline__ float4 head_grad_chunk(const float* __restrict__ src, int h, int c, long n) {
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (src == nullptr) return v;
  if (h == 4) return __ldg(reinterpret_cast<const float4*>(src + n * 48) + (c - 8));
  if (c & 1) return v;                           // columns 4..7 of a small head: padding
  if (h == 2) return __ldg(reinterpret_cast<const float4*>(src) + n);
  if (h == 3) { v.x = __ldg(src + n); return v; }
  v.x = __ldg(src + n * 3); v.y = __ldg(src + n * 3 + 1); v.z = __ldg(src + n * 3 + 2);
  return v;
}

__global__ void __launch_bounds__(BHG_THREADS)
build_head_grads(int N, GradPtrs gp, int use_sub, float* __restrict__ G, float* __restrict__ colsum) {
  __shared__ float s_sum[GW];
  for (int i = threadIdx.x; i < GW; i += blockDim.x) s_sum[i] = 0.f;
  __syncthreads();
  const int c = threadIdx.x % 20, r = threadIdx.x / 20;
  const int h = c >= 8 ? 4 : (c >> 1);
  const float* src = gp.d_out[h];
  const float* src2 = use_sub ? gp.d_sub[h] : nullptr;
  const float cf = gp.coef[h];
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const long S = (long)gridDim.x * 16;
  for (long n0 = (long)blockIdx.x * 16 + r; n0 < N; n0 += S * BHG_UNROLL) {
    float4 v[BHG_UNROLL];
#pragma unroll
    for (int u = 0; u < BHG_UNROLL; ++u) {
      const long n = n0 + u * S;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (n < N) {
        const float4 a = head_grad_chunk(src, h, c, n), b = head_grad_chunk(src2, h, c, n);
        v[u] = make_float4((a.x + b.x) * cf, (a.y + b.y) * cf, (a.z + b.z) * cf, (a.w + b.w) * cf);
      }
    }
#pragma unroll
    for (int u = 0; u < BHG_UNROLL; ++u) {
      const long n = n0 + u * S;
      if (n < N) {
        reinterpret_cast<float4*>(G + n * GW)[c] = v[u];
        acc.x += v[u].x; acc.y += v[u].y; acc.z += v[u].z; acc.w += v[u].w;
      }
    }
  }
  atomicAdd(&s_sum[c * 4 + 0], acc.x);
  atomicAdd(&s_sum[c * 4 + 1], acc.y);
  atomicAdd(&s_sum[c * 4 + 2], acc.z);
  atomicAdd(&s_sum[c * 4 + 3], acc.w);
  __syncthreads();
  for (int i = threadIdx.x; i < GW; i += blockDim.x)
    if (s_sum[i] != 0.f) atomicAdd(colsum + i, s_sum[i]);
}

__global__ void add_vec(float* __restrict__ dst, const float* __restrict__ src, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}

// ---------------------------------------------------------------------------------------------
struct DeformWork {
  float* R[2][ED3DGS_DEFORM_MAX_DEPTH];  // post-ReLU outputs of feature_out layers, [N, W]
  float* Rz[2];                          // post-ReLU head hidden layers, [N, 5W] (active heads packed)
  float* G[2];                           // [N, 80]
  float* dZ;                             // [N, 5W]
  float* dH[2];                          // [N, W]
  float* Wpack;                          // [5W, W]
  float* Wprep;                          // per-launch hi/lo UMMA images of the weight operand
  float* c1[2];                          // [W]
  float* dc1;                            // [W]
  float* colsum[2];                      // [80]
  // ReLU masks as bits: one word per (block of 32 Gaussians, feature)
  uint32_t* Mr[2][ED3DGS_DEFORM_MAX_DEPTH];  // trunk layers: [ceil(N/32)][W]
  uint32_t* Mz[2];                           // head hidden layers: [ceil(N/32)][5W]
  // stage-resident path (fused_mlp.cuh): feature-major activations, both mask layouts, weight slab images
  long npad;
  float* zT[2];                              // [5*128][npad]
  float* hT[2];                              // [128][npad]
  uint32_t* Mg[2];                           // [npad][5][4]
  uint32_t* MgH[2];                          // [npad][4]
  float* wimg[2];                            // forward slab images of a stage
  float* wimg_b;                             // backward slab images (one stage at a time)
  float* dzT;                                // [tile][5*128][128] gradients of the head hidden layers (one stage at a time)
  float* dhT;                                // [tile][128][128]
  float* GT;                                 // [tile][64][128] scaled upstream gradients
};

static inline void carve_f(char*& p, float*& out, size_t count) {
  p = reinterpret_cast<char*>((reinterpret_cast<size_t>(p) + 255) / 256 * 256);
  out = reinterpret_cast<float*>(p);
  p += count * sizeof(float);
}

static DeformWork carve_deform(char*& p, int W, int depth, size_t N) {
  DeformWork w{};
  for (int s = 0; s < 2; ++s) {
    for (int l = 0; l < depth; ++l) carve_f(p, w.R[s][l], N * W);
    carve_f(p, w.Rz[s], N * 5 * W);
    carve_f(p, w.G[s], N * GW);
    carve_f(p, w.c1[s], W);
    carve_f(p, w.colsum[s], GW);
  }
  const size_t nblk32 = (N + 31) / 32;
  for (int s = 0; s < 2; ++s) {
    float* t;
    for (int l = 0; l < depth; ++l) { carve_f(p, t, nblk32 * W); w.Mr[s][l] = reinterpret_cast<uint32_t*>(t); }
    carve_f(p, t, nblk32 * 5 * W);
    w.Mz[s] = reinterpret_cast<uint32_t*>(t);
  }
  carve_f(p, w.dZ, N * 5 * W);
  carve_f(p, w.dH[0], N * W);
  carve_f(p, w.dH[1], N * W);
  carve_f(p, w.Wpack, (size_t)5 * W * W);
  carve_f(p, w.Wprep, tc_prep_scratch_floats());
  carve_f(p, w.dc1, W);
  if (W == 128 && depth == 1) {
    const size_t npad = (N + FM_TG - 1) / FM_TG * FM_TG;
    w.npad = (long)npad;
    for (int s = 0; s < 2; ++s) {
      float* t;
      carve_f(p, w.zT[s], 5 * 128 * npad);
      carve_f(p, w.hT[s], 128 * npad);
      carve_f(p, t, npad * 5 * 4); w.Mg[s] = reinterpret_cast<uint32_t*>(t);
      carve_f(p, t, npad * 4); w.MgH[s] = reinterpret_cast<uint32_t*>(t);
      carve_f(p, w.wimg[s], (size_t)FM_MAX_SLABS * FM_W_SLOT_BYTES / 4);
    }
    carve_f(p, w.wimg_b, (size_t)FM_MAX_SLABS * FM_W_SLOT_BYTES / 4);
    carve_f(p, w.dzT, 5 * 128 * npad);
    carve_f(p, w.dhT, 128 * npad);
    carve_f(p, w.GT, FB_GT_ROWS * npad);
  }
  return w;
}

static int check_weights(const ed3dgs_deform_weights* w) {
  if (!w) return ED3DGS_OK;
  if (w->W < 16 || w->W > 256 || (w->W % 16) != 0) {
    set_error("deform: net_width %d unsupported (multiple of 16, <= 256)", w->W);
    return ED3DGS_ERR_ARG;
  }
  if (w->depth < 1 || w->depth > ED3DGS_DEFORM_MAX_DEPTH || (w->temb_dim % 4) || (w->gemb_dim % 16) || w->gemb_dim > 240) {
    set_error("deform: unsupported depth/temb_dim/gemb_dim (%d, %d, %d)", w->depth, w->temb_dim, w->gemb_dim);
    return ED3DGS_ERR_ARG;
  }
  if (!w->head_w1[0]) { set_error("deform: the position head cannot be disabled"); return ED3DGS_ERR_ARG; }
  return ED3DGS_OK;
}

#define TC_TRY(expr)                                                                         \
  do {                                                                                       \
    cudaError_t e__ = (expr);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      set_error("deform: %s failed: %s", #expr, cudaGetErrorString(e__));                    \
      return ED3DGS_ERR_CUDA;                                                                \
    }                                                                                        \
  } while (0)

// Event recorded right after the LAST writer of d_embedding in ed3dgs_deform_backward (ed3dgs_deform_set_emb_ready_event):
// the embedding gradient is final one weight-gradient launch (1.2 ms at 1 M Gaussians) before the call's last kernel, and
// a multi-GPU caller starts its exchange from there.
// One event per device (the device current at the call, like the SM reserves in tc_launch.cuh): an event belongs to the
// device it was created on, and recording it on another device's stream fails.
static void* g_emb_ready_event[TC_MAX_DEVICES] = {};
static inline cudaError_t record_emb_ready(bool last_stage, cudaStream_t stream) {
  void* ev = last_stage ? g_emb_ready_event[tc_device_slot()] : nullptr;
  return ev ? cudaEventRecord(static_cast<cudaEvent_t>(ev), stream) : cudaSuccess;
}

// precision of the call in flight on this thread (ed3dgs_deform_weights::precision): 0 = 3xTF32, 1 = single tf32 pass
static thread_local int g_single = 0;

static TcArgs tc_args(long lda, long ldb, int M, int K, int NT, int epi) {
  TcArgs a{};
  a.single = g_single;
  a.lda = lda; a.ldb = ldb; a.M = M; a.K = K; a.NT = NT; a.epilogue = epi;
  a.k_per_cta = K > 0 ? K : 1; a.ones_col = -1;
  return a;
}

static int active_heads(const ed3dgs_deform_weights* w, int* list) {
  int n = 0;
  for (int h = 0; h < 5; ++h)
    if (w->head_w1[h]) list[n++] = h;
  return n;
}

// The stage-resident kernels cover the shape every dataset config uses; anything else takes the layered path.
// ED3DGS_DEFORM_FUSED=0 forces the layered path (A/B timing, tests of the layered kernels).
static bool fused_shape(const ed3dgs_deform_weights* w) {
  static const int enabled = [] { const char* e = getenv("ED3DGS_DEFORM_FUSED"); return e ? atoi(e) : 1; }();
  return enabled && w && w->W == 128 && w->depth == 1 && w->gemb_dim == 32;
}

static cudaError_t fused_func_attr(const void* fn, size_t smem) {
  // per-device attribute: set on every launch (a cached flag would be wrong on the second GPU of a process)
  return cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

static int current_sm_count() { return tc_sm_count(); }

// Forward of one stage in one launch (fused_mlp.cuh).
static int stage_forward_fused(const ed3dgs_deform_weights* w, int s, int N, const float* emb, const float* tau,
                               const float* const base[5], float* const dst[5], const float coef[5], DeformWork& wk,
                               cudaStream_t stream) {
  const int temb = w->temb_dim, gemb = w->gemb_dim;
  const long ld0 = temb + gemb;
  layer0_temporal_bias<<<(128 + 3) / 4, 128, 0, stream>>>(w->feat_w[0], w->feat_b[0], ld0, temb, tau, wk.c1[s], 128);
  int act[5];
  const int na = active_heads(w, act);
  const bool has_rgb = w->head_w1[4] != nullptr;   // the rgb head is the last active slot
  // weight slab images: layer 0 (2), hidden layers (8 per slot), rgb output layer (8, 48 rows)
  PrepArgs pa{};
  FusedFwdArgs fa{};
  int ns = 0;
  uint32_t off = 0;
  auto add = [&](const float* src, long ld, int n_valid, int NT, int K) {
    for (int k0 = 0; k0 < K; k0 += 16) {
      pa.s[ns] = PrepSlab{src, ld, TC_K_MAJOR, n_valid, NT, k0, K, off};
      fa.slab_off[ns] = off;
      off += (uint32_t)NT * 8u;
      ++ns;
    }
  };
  add(w->feat_w[0] + temb, ld0, 128, 128, gemb);
  for (int i = 0; i < na; ++i) add(w->head_w1[act[i]], 128, 128, 128, 128);
  if (has_rgb) add(w->head_w2[4], 128, 48, 48, 128);
  pa.dst = reinterpret_cast<float4*>(wk.wimg[s]);
  {
    ProfScope ps("fused_prepare", stream);
    fused_prepare<<<ns, 128, 0, stream>>>(pa);
  }
  fa.wimg = reinterpret_cast<const float4*>(wk.wimg[s]);
  fa.na = na; fa.has_rgb = has_rgb ? 1 : 0; fa.N = N; fa.npad = wk.npad; fa.single = g_single;
  fa.c1 = wk.c1[s];
  for (int i = 0; i < na; ++i) {
    const int h = act[i];
    fa.b1[i] = w->head_b1[h]; fa.w2[i] = w->head_w2[h]; fa.b2[i] = w->head_b2[h];
    fa.hdim[i] = hHeadDim[h]; fa.base[i] = base[h]; fa.dst[i] = dst[h]; fa.coef[i] = coef[h];
    if (h == 4 && ((reinterpret_cast<uintptr_t>(base[h]) | reinterpret_cast<uintptr_t>(dst[h])) & 15u)) {
      set_error("deform_forward: the rgb rows must be 16-byte aligned");
      return ED3DGS_ERR_ARG;
    }
  }
  fa.zT = wk.zT[s]; fa.hT = wk.hT[s]; fa.Mg = wk.Mg[s]; fa.MgH = wk.MgH[s];
  tc_encode_tiled_fn enc = tc_encode_tiled();
  if (!enc || (reinterpret_cast<uintptr_t>(emb) & 15u)) { set_error("deform_forward: tensor map unavailable"); return ED3DGS_ERR_CUDA; }
  {
    const cuuint64_t dims[2] = {(cuuint64_t)gemb, (cuuint64_t)N};
    const cuuint64_t strides[1] = {(cuuint64_t)gemb * sizeof(float)};
    const cuuint32_t box[2] = {16, (cuuint32_t)FM_TG};
    const cuuint32_t estr[2] = {1, 1};
    if (enc(&fa.emap, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(emb), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
      set_error("deform_forward: cuTensorMapEncodeTiled failed");
      return ED3DGS_ERR_CUDA;
    }
  }
  TC_TRY(fused_func_attr(reinterpret_cast<const void*>(fused_fwd_kernel), FM_FWD_SMEM));
  const int ntiles = (N + FM_TG - 1) / FM_TG;
  const int sms = current_sm_count();
  {
    ProfScope ps("fused_fwd_kernel", stream);
    fused_fwd_kernel<<<ntiles < sms ? ntiles : sms, FM_THREADS, FM_FWD_SMEM, stream>>>(fa);
  }
  TC_TRY(cudaGetLastError());
  for (int h = 0; h < 5; ++h)
    if (!w->head_w1[h] && dst[h] != base[h])
      TC_TRY(cudaMemcpyAsync(dst[h], base[h], (size_t)N * hHeadDim[h] * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  return ED3DGS_OK;
}

// ---- layered path: any net_width up to 256 (multiple of 16), any defor_depth ---------------------------------------------
// tc_rows_kernel puts at most 128 output features on the TMEM lanes, a tile programme holds TC_MAX_STEPS K slabs and a
// launch TC_MAX_BATCH entries: wider layers become several batch entries (128 output features each, the weight / bias /
// output / mask pointers advanced by the chunk) spread over as many launches as the limits ask for.
struct TcEntry {
  TcBatch b;
  int K;   // contraction length of the entry
};

static cudaError_t tc_launch_many(int amaj, int bmaj, const TcArgs& base, const std::vector<TcEntry>& ent, float* scratch,
                                  cudaStream_t stream) {
  size_t i = 0;
  while (i < ent.size()) {
    TcArgs a = base;
    int n = 0, steps = 0, kmax = 0;
    while (i + n < ent.size() && n < TC_MAX_BATCH) {
      const int K = ent[i + n].K;
      const int st = (K + TC_KC - 1) / TC_KC;
      const int km = K > kmax ? K : kmax;
      const size_t floats = (size_t)((km + TC_KC - 1) / TC_KC) * 128 * (2 * TC_PC) * 4 * (size_t)(n + 1);
      if (n > 0 && (steps + st > TC_MAX_STEPS || floats > tc_prep_scratch_floats())) break;
      a.b[n] = ent[i + n].b;
      a.b[n].K = K;
      steps += st;
      kmax = km;
      ++n;
    }
    const cudaError_t e = tc_launch(amaj, bmaj, a, n, scratch, stream);
    if (e != cudaSuccess) return e;
    i += n;
  }
  return cudaSuccess;
}

// dst[g][f] = mask bit (g, f) ? dst[g][f] : 0 ; bias_grad[f] += column sums. One block per 32 Gaussians (one mask word
// per feature), thread = feature. Only the wide layered path needs it (the input gradient of a trunk wider than the tile
// programme is accumulated over several launches and masked afterwards).
__global__ void __launch_bounds__(256) relu_mask_colsum(float* __restrict__ dst, long ld, int W, long N,
                                                        const uint32_t* __restrict__ mask, long mask_ld, int mask_col,
                                                        float* __restrict__ bias_grad) {
  const long g0 = (long)blockIdx.x * 32;
  const int jmax = (int)min(32L, N - g0);
  for (int f = threadIdx.x; f < W; f += blockDim.x) {
    const uint32_t word = mask[(g0 >> 5) * mask_ld + mask_col + f];
    float cs = 0.f;
    for (int j = 0; j < jmax; ++j) {
      float* q = dst + (g0 + j) * ld + f;
      const float r = ((word >> j) & 1u) ? *q : 0.f;
      *q = r;
      cs += r;
    }
    if (bias_grad != nullptr && cs != 0.f) atomicAdd(bias_grad + f, cs);
  }
}

static inline int chunk_cols(int W, int c) { return W - c * 128 < 128 ? W - c * 128 : 128; }

// One stage forward: base[h] -> dst[h] = base[h] + delta_h * coef[h].
// acts_only: stop after the hidden layers (the activations and ReLU masks the layered backward reads).
static int stage_forward(const ed3dgs_deform_weights* w, int s, int N, const float* emb, const float* tau,
                         const float* const base[5], float* const dst[5], const float coef[5], DeformWork& wk,
                         cudaStream_t stream, bool acts_only = false) {
  const int W = w->W, D = w->depth, temb = w->temb_dim, gemb = w->gemb_dim;
  const long ld0 = temb + gemb;
  const int nch = (W + 127) / 128;
  layer0_temporal_bias<<<(W + 3) / 4, 128, 0, stream>>>(w->feat_w[0], w->feat_b[0], ld0, temb, tau, wk.c1[s], W);
  {  // layer 0 on the per-Gaussian embedding
    TcArgs a = tc_args(gemb, ld0, N, gemb, W, TC_EPI_BIAS_ACT);
    a.relu = 1;
    std::vector<TcEntry> ent;
    for (int c = 0; c < nch; ++c) {
      TcBatch b{emb, w->feat_w[0] + temb + (long)c * 128 * ld0, wk.R[s][0] + c * 128, wk.c1[s] + c * 128, nullptr, nullptr,
                W, 0, chunk_cols(W, c), 0, 0, 1.f};
      b.mask_out = wk.Mr[s][0]; b.mask_ld = W; b.mask_col = c * 128;
      ent.push_back({b, gemb});
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_K_MAJOR, a, ent, wk.Wprep, stream));
  }
  for (int l = 1; l < D; ++l) {
    TcArgs a = tc_args(W, W, N, W, W, TC_EPI_BIAS_ACT);
    a.relu = 1;
    std::vector<TcEntry> ent;
    for (int c = 0; c < nch; ++c) {
      TcBatch b{wk.R[s][l - 1], w->feat_w[l] + (long)c * 128 * W, wk.R[s][l] + c * 128, w->feat_b[l] + c * 128, nullptr,
                nullptr, W, 0, chunk_cols(W, c), 0, 0, 1.f};
      b.mask_out = wk.Mr[s][l]; b.mask_ld = W; b.mask_col = c * 128;
      ent.push_back({b, W});
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_K_MAJOR, a, ent, wk.Wprep, stream));
  }
  int act[5];
  const int na = active_heads(w, act);
  {  // head hidden layers, all active heads
    TcArgs a = tc_args(W, W, N, W, W, TC_EPI_BIAS_ACT);
    a.relu = 1;
    std::vector<TcEntry> ent;
    for (int i = 0; i < na; ++i) {
      const int h = act[i];
      for (int c = 0; c < nch; ++c) {
        TcBatch b{wk.R[s][D - 1], w->head_w1[h] + (long)c * 128 * W, wk.Rz[s] + (long)i * W + c * 128, w->head_b1[h] + c * 128,
                  nullptr, nullptr, 5L * W, 0, chunk_cols(W, c), 0, 0, 1.f};
        b.mask_out = wk.Mz[s]; b.mask_ld = 5L * W; b.mask_col = i * W + c * 128;
        ent.push_back({b, W});
      }
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_K_MAJOR, a, ent, wk.Wprep, stream));
  }
  if (acts_only) return ED3DGS_OK;
  {  // head outputs fused with  x = x + delta * coef
    TcArgs a = tc_args(5L * W, W, N, W, 48, TC_EPI_HEAD);
    std::vector<TcEntry> ent;
    for (int i = 0; i < na; ++i) {
      const int h = act[i];
      ent.push_back({TcBatch{wk.Rz[s] + (long)i * W, w->head_w2[h], dst[h], w->head_b2[h], base[h], nullptr,
                             hHeadDim[h], hHeadDim[h], hHeadDim[h], 0, 0, coef[h]}, W});
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_K_MAJOR, a, ent, wk.Wprep, stream));
  }
  for (int h = 0; h < 5; ++h)
    if (!w->head_w1[h] && dst[h] != base[h])
      TC_TRY(cudaMemcpyAsync(dst[h], base[h], (size_t)N * hHeadDim[h] * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  return ED3DGS_OK;
}

// Weight-gradient groups are collected and flushed WG_MAX_GROUPS at a time.
struct WgBatcher {
  WgArgs a{};
  cudaStream_t stream;
  WgBatcher(int N, cudaStream_t st) : stream(st) { a.single = g_single; a.N = N; }
  cudaError_t add(const WgGroup& g) {
    if (g.nseg <= 0) return cudaSuccess;
    a.g[a.ngroups++] = g;
    return a.ngroups == WG_MAX_GROUPS ? flush() : cudaSuccess;
  }
  cudaError_t flush() {
    if (a.ngroups == 0) return cudaSuccess;
    const cudaError_t e = tc_wgrad_launch(a, stream);
    const int N = a.N;
    a = WgArgs{};
    a.single = g_single;
    a.N = N;
    return e;
  }
};

static int stage_backward(const ed3dgs_deform_weights* w, const ed3dgs_deform_grads* g, int s, int N,
                          const float* emb, const float* tau, const float* G, const float* colsum, DeformWork& wk,
                          float* d_emb, int accumulate_emb, float* d_tau, cudaStream_t stream, bool last_stage) {
  const int W = w->W, D = w->depth, temb = w->temb_dim, gemb = w->gemb_dim;
  const long ld0 = temb + gemb;
  const int nch = (W + 127) / 128;
  const int pair = 2 * W <= 256 ? 2 : 1;   // heads side by side on the N = 256 side of a weight-gradient group
  int act[5];
  const int na = active_heads(w, act);

  if (g) {
    for (int i = 0; i < na; ++i) {
      const int h = act[i];
      if (g->head_b2[h]) add_vec<<<1, 64, 0, stream>>>(g->head_b2[h], colsum + hHeadOff[h], hHeadDim[h]);
    }
    // dW2_h[j][f] += sum_n G_h[n][j] * Rz_h[n][f]: all 80 columns of G on the M side, the hidden
    // activations of one or two heads on the N side; only the (head rows x head columns) blocks
    // of the product are kept.
    WgBatcher wb(N, stream);
    for (int i0 = 0; i0 < na; i0 += pair) {
      WgGroup gr{};
      const int cnt = i0 + pair <= na ? pair : na - i0;
      gr.P = G; gr.ldp = GW; gr.pcols = GW;
      gr.Q = wk.Rz[s] + (long)i0 * W; gr.ldq = 5L * W; gr.qcols = cnt * W;
      for (int i = i0; i < i0 + cnt; ++i) {
        const int h = act[i];
        if (!g->head_w2[h]) continue;
        gr.seg[gr.nseg++] = WgSegment{g->head_w2[h], 1, (long)W, (i - i0) * W, W, hHeadOff[h], hHeadDim[h]};
      }
      TC_TRY(wb.add(gr));
    }
    TC_TRY(wb.flush());
  }
  {  // dZ_h = (G_h * W2_h) masked by Rz_h > 0
    TcArgs a = tc_args(GW, W, N, 0, W, TC_EPI_MASK);
    a.K = 48;
    std::vector<TcEntry> ent;
    for (int i = 0; i < na; ++i) {
      const int h = act[i];
      for (int c = 0; c < nch; ++c) {
        TcBatch b{G + hHeadOff[h], w->head_w2[h] + c * 128, wk.dZ + (long)i * W + c * 128, nullptr,
                  wk.Rz[s] + (long)i * W + c * 128, (g && g->head_b1[h]) ? g->head_b1[h] + c * 128 : nullptr, 5L * W, 5L * W,
                  chunk_cols(W, c), hHeadPad[h], hHeadDim[h], 1.f};
        b.mask_in = wk.Mz[s]; b.mask_ld = 5L * W; b.mask_col = i * W + c * 128;
        ent.push_back({b, hHeadPad[h]});
      }
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_MN_MAJOR, a, ent, wk.Wprep, stream));
  }
  if (g) {  // dW1_h[o][i] += sum_n dZ_h[n][o] * R[n][i]: trunk activations (128 at a time) on the M side, dZ on the N side
    WgBatcher wb(N, stream);
    for (int mc = 0; mc < nch; ++mc)
      for (int i0 = 0; i0 < na; i0 += pair) {
        WgGroup gr{};
        const int cnt = i0 + pair <= na ? pair : na - i0;
        gr.P = wk.R[s][D - 1] + mc * 128; gr.ldp = W; gr.pcols = chunk_cols(W, mc);
        gr.Q = wk.dZ + (long)i0 * W; gr.ldq = 5L * W; gr.qcols = cnt * W;
        for (int i = i0; i < i0 + cnt; ++i) {
          const int h = act[i];
          if (!g->head_w1[h]) continue;
          gr.seg[gr.nseg++] = WgSegment{g->head_w1[h] + mc * 128, (long)W, 1, (i - i0) * W, W, 0, gr.pcols};
        }
        TC_TRY(wb.add(gr));
      }
    TC_TRY(wb.flush());
  }
  // dR = (dZ[N, na*W] * W1_all[na*W, W]) masked by R > 0
  for (int i = 0; i < na; ++i)
    TC_TRY(cudaMemcpyAsync(wk.Wpack + (size_t)i * W * W, w->head_w1[act[i]], (size_t)W * W * sizeof(float),
                           cudaMemcpyDeviceToDevice, stream));
  float* dh = wk.dH[0];
  float* dh_other = wk.dH[1];
  // the bias gradient of trunk layer l is the column sum of dh at that layer (MASK epilogue);
  // for layer 0 it is dc1, which also carries the temporal half of the first Linear
  TC_TRY(cudaMemsetAsync(wk.dc1, 0, W * sizeof(float), stream));
  auto trunk_bias_grad = [&](int l) -> float* { return l == 0 ? wk.dc1 : (g ? g->feat_b[l] : nullptr); };
  if ((na * W + TC_KC - 1) / TC_KC <= TC_MAX_STEPS) {
    TcArgs a = tc_args(5L * W, W, N, na * W, W, TC_EPI_MASK);
    std::vector<TcEntry> ent;
    float* bg = trunk_bias_grad(D - 1);
    for (int c = 0; c < nch; ++c) {
      TcBatch b{wk.dZ, wk.Wpack + c * 128, dh + c * 128, nullptr, wk.R[s][D - 1] + c * 128, bg ? bg + c * 128 : nullptr,
                W, W, chunk_cols(W, c), 0, 0, 1.f};
      b.mask_in = wk.Mr[s][D - 1]; b.mask_ld = W; b.mask_col = c * 128;
      ent.push_back({b, na * W});
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_MN_MAJOR, a, ent, wk.Wprep, stream));
  } else {
    // contraction longer than one tile programme: accumulate head by head, then mask and take the column sums
    for (int i = 0; i < na; ++i) {
      TcArgs a = tc_args(5L * W, W, N, W, W, TC_EPI_ACCUM);
      a.accumulate = i > 0;
      std::vector<TcEntry> ent;
      for (int c = 0; c < nch; ++c)
        ent.push_back({TcBatch{wk.dZ + (long)i * W, wk.Wpack + (size_t)i * W * W + c * 128, dh + c * 128, nullptr, nullptr,
                               nullptr, W, 0, chunk_cols(W, c), 0, 0, 1.f}, W});
      TC_TRY(tc_launch_many(TC_K_MAJOR, TC_MN_MAJOR, a, ent, wk.Wprep, stream));
    }
    relu_mask_colsum<<<(N + 31) / 32, 256, 0, stream>>>(dh, W, W, N, wk.Mr[s][D - 1], W, 0, trunk_bias_grad(D - 1));
  }
  for (int l = D - 1; l >= 1; --l) {
    if (g && g->feat_w[l]) {  // dW_l[o][i] += sum_n dh[n][o] * R_{l-1}[n][i]
      WgBatcher wb(N, stream);
      for (int mc = 0; mc < nch; ++mc) {
        WgGroup gr{};
        gr.P = wk.R[s][l - 1] + mc * 128; gr.ldp = W; gr.pcols = chunk_cols(W, mc);
        gr.Q = dh; gr.ldq = W; gr.qcols = W;
        gr.nseg = 1;
        gr.seg[0] = WgSegment{g->feat_w[l] + mc * 128, (long)W, 1, 0, W, 0, gr.pcols};
        TC_TRY(wb.add(gr));
      }
      TC_TRY(wb.flush());
    }
    TcArgs a = tc_args(W, W, N, W, W, TC_EPI_MASK);
    std::vector<TcEntry> ent;
    float* bg = trunk_bias_grad(l - 1);
    for (int c = 0; c < nch; ++c) {
      TcBatch b{dh, w->feat_w[l] + c * 128, dh_other + c * 128, nullptr, wk.R[s][l - 1] + c * 128, bg ? bg + c * 128 : nullptr,
                W, W, chunk_cols(W, c), 0, 0, 1.f};
      b.mask_in = wk.Mr[s][l - 1]; b.mask_ld = W; b.mask_col = c * 128;
      ent.push_back({b, W});
    }
    TC_TRY(tc_launch_many(TC_K_MAJOR, TC_MN_MAJOR, a, ent, wk.Wprep, stream));
    float* t = dh; dh = dh_other; dh_other = t;
  }
  // layer 0: weight gradient of the embedding columns, then dE (dc1 came from the MASK epilogue above)
  if (g && g->feat_w[0]) {  // dW0[o][temb + i] += sum_n dh[n][o] * emb[n][i]
    WgArgs a{};
    a.single = g_single;
    a.N = N; a.ngroups = 1;
    a.g[0].P = emb; a.g[0].ldp = gemb; a.g[0].pcols = gemb;
    a.g[0].Q = dh; a.g[0].ldq = W; a.g[0].qcols = W;
    a.g[0].nseg = 1;
    a.g[0].seg[0] = WgSegment{g->feat_w[0] + temb, ld0, 1, 0, W, 0, gemb};
    TC_TRY(tc_wgrad_launch(a, stream));
  }
  {
    TcArgs a = tc_args(W, ld0, N, W, gemb, TC_EPI_ACCUM);
    a.accumulate = accumulate_emb;
    a.b[0] = TcBatch{dh, w->feat_w[0] + temb, d_emb, nullptr, nullptr, nullptr, gemb, 0, gemb, 0, 0, 1.f};
    TC_TRY(tc_launch(TC_K_MAJOR, TC_MN_MAJOR, a, 1, wk.Wprep, stream));
    TC_TRY(record_emb_ready(last_stage, stream));
  }
  layer0_temporal_backward<<<(temb + 31) / 32, 256, 0, stream>>>(
      w->feat_w[0], ld0, temb, W, tau, wk.dc1, g ? g->feat_w[0] : nullptr, g ? g->feat_b[0] : nullptr, d_tau);
  TC_TRY(cudaGetLastError());
  return ED3DGS_OK;
}

#ifdef ED3DGS_TRACE
static long long* g_trace = nullptr;
#endif
// Backward of one stage on the stage-resident path: one chain kernel (dz, dh, dE, bias gradients, feature-major
// gradient matrices) and three weight-gradient launches that contract those matrices over the Gaussians.
static int stage_backward_fused(const ed3dgs_deform_weights* w, const ed3dgs_deform_grads* g, int s, int N,
                                const float* emb, const float* tau, const float* const d_out[5],
                                const float* const d_sub[5], int use_sub, const float coef[5], DeformWork& wk,
                                float* d_emb, int accumulate_emb, float* d_tau, cudaStream_t stream, bool last_stage) {
  const int temb = w->temb_dim, gemb = w->gemb_dim;
  const long ld0 = temb + gemb;
  int act[5];
  const int na = active_heads(w, act);
  const bool has_rgb = w->head_w1[4] != nullptr;
  TC_TRY(cudaMemsetAsync(wk.dc1, 0, 128 * sizeof(float), stream));

  // slab images in the order the chain kernel's MMA warp walks them (see fused_bwd_kernel)
  PrepArgs pa{};
  FusedBwdArgs ba{};
  int nimg = 0;
  uint32_t off = 0;
  auto add_img = [&](const float* src, long ld, int n_valid, int NT, int k0, int kvalid) -> uint32_t {
    pa.s[nimg] = PrepSlab{src, ld, TC_MN_MAJOR, n_valid, NT, k0, kvalid, off};
    const uint32_t at = off;
    off += (uint32_t)NT * 8u;
    ++nimg;
    return at;
  };
  uint32_t w1t[5][8], w2t[5][2][3], w0t[8];
  for (int i = 0; i < na; ++i)
    for (int sl = 0; sl < 8; ++sl) w1t[i][sl] = add_img(w->head_w1[act[i]], 128, 128, 128, sl * 16, 128);   // rows k, K = f
  for (int i = 0; i < na; ++i) {   // second-layer weights transposed: rows f (a half head), K = o
    const int h = act[i], k = hHeadDim[h];
    for (int hf = 0; hf < 2; ++hf)
      for (int sl = 0; sl * 16 < k; ++sl) w2t[i][hf][sl] = add_img(w->head_w2[h] + hf * 64, 128, 64, 64, sl * 16, k);
  }
  for (int sl = 0; sl < 8; ++sl) w0t[sl] = add_img(w->feat_w[0] + temb, ld0, gemb, 32, sl * 16, 128);           // rows j, K = k
  if (nimg > FM_MAX_SLABS) { set_error("deform_backward: slab table overflow"); return ED3DGS_ERR_ARG; }
  pa.dst = reinterpret_cast<float4*>(wk.wimg_b);
  {
    ProfScope ps("fused_prepare", stream);
    fused_prepare<<<nimg, 128, 0, stream>>>(pa);
  }
  int nseq = 0;
  auto push = [&](uint32_t at, int NT) { ba.seq_off[nseq] = at; ba.seq_bytes[nseq] = (uint16_t)(NT * 128); ++nseq; };
  const int nunits = 2 * na;
  auto push_w2t = [&](int unit) {
    const int i = unit >> 1, k = hHeadDim[act[i]];
    for (int sl = 0; sl * 16 < k; ++sl) push(w2t[i][unit & 1][sl], 64);
  };
  push_w2t(0);
  for (int uu = 0; uu < nunits; ++uu) {
    if (uu + 1 < nunits) push_w2t(uu + 1);
    for (int sl = 0; sl < 4; ++sl) push(w1t[uu >> 1][(uu & 1) * 4 + sl], 128);
  }
  for (int sl = 0; sl < 8; ++sl) push(w0t[sl], 32);
  ba.nseq = nseq;
  ba.wimg = reinterpret_cast<const float4*>(wk.wimg_b);
  ba.na = na; ba.has_rgb = has_rgb ? 1 : 0; ba.N = N;
  ba.use_sub = use_sub; ba.accumulate_emb = accumulate_emb; ba.store_dz = 1; ba.single = g_single;
#ifdef ED3DGS_TRACE
  { static long long* tr = nullptr; if (!tr) { cudaMalloc(&tr, 8 * 8000 * 5); } cudaMemsetAsync(tr, 0, 8 * 8000 * 5, stream); ba.trace = tr; g_trace = tr; }
#endif
  int goff = 0;
  for (int i = 0; i < na; ++i) {
    const int h = act[i];
    ba.w2[i] = w->head_w2[h]; ba.hdim[i] = hHeadDim[h]; ba.goff[i] = goff; ba.coef[i] = coef[h];
    goff += hHeadDim[h];
    ba.d_out[i] = d_out ? d_out[h] : nullptr;
    ba.d_sub[i] = d_sub ? d_sub[h] : nullptr;
    ba.db1[i] = g ? g->head_b1[h] : nullptr;
    ba.db2[i] = g ? g->head_b2[h] : nullptr;
  }
  ba.Mg = wk.Mg[s]; ba.MgH = wk.MgH[s];
  ba.dzT = wk.dzT; ba.dhT = wk.dhT; ba.GT = wk.GT; ba.d_emb = d_emb; ba.dc1 = wk.dc1;
  TC_TRY(fused_func_attr(reinterpret_cast<const void*>(fused_bwd_kernel), FM_BWD_SMEM));
  const int ntiles = (N + FM_TG - 1) / FM_TG;
  const int sms = current_sm_count();
  {
    ProfScope ps("fused_bwd_kernel", stream);
    fused_bwd_kernel<<<ntiles < sms ? ntiles : sms, FB_THREADS, FM_BWD_SMEM, stream>>>(ba);
  }
  TC_TRY(cudaGetLastError());
  TC_TRY(record_emb_ready(last_stage, stream));   // d_emb is complete: the weight gradients below do not touch it

  if (g) {
    // every weight gradient of the stage in one launch; the groups of a slice of Gaussians are adjacent work items, so
    // hT / dzT / zT / GT are shared through L2
    WgArgs a{};
    a.single = g_single;
    a.N = N;
    // dW1_h[o][i] += sum_n dzT_h[o][n] hT[i][n]: one head per group (tensor-memory kernel, N = 128) or two heads side by
    // side (shared-memory kernel, N = 256)
    static const bool wg_ts = [] { const char* e = getenv("ED3DGS_WG_TS"); return e ? atoi(e) != 0 : true; }();
    const int per = wg_ts ? 1 : 2;
    for (int i0 = 0; i0 < na; i0 += per) {
      WgGroup& gr = a.g[a.ngroups];
      gr = WgGroup{};
      gr.P = wk.hT[s]; gr.pk = 1; gr.prow_total = 128; gr.pcols = 128;
      gr.Q = wk.dzT + (size_t)i0 * 128 * 128; gr.qk = 1; gr.qrow_total = (long)na * 128; gr.qcols = (i0 + 1 < na ? per : 1) * 128;
      for (int i = i0; i < i0 + per && i < na; ++i) {
        const int h = act[i];
        if (!g->head_w1[h]) continue;
        gr.seg[gr.nseg++] = WgSegment{g->head_w1[h], 128, 1, (i - i0) * 128, 128, 0, 128};
      }
      if (gr.nseg > 0) ++a.ngroups;
    }
    for (int i = 0; i < na; ++i) {   // dW2_h[o][f] += sum_n zT_h[f][n] GT[goff_h + o][n]: 16 (48 for rgb) gradient rows on the N side
      const int h = act[i];
      if (!g->head_w2[h]) continue;
      WgGroup& gr = a.g[a.ngroups++];
      gr = WgGroup{};
      gr.P = wk.zT[s] + (size_t)i * 128 * 128; gr.pk = 1; gr.prow_total = (long)na * 128; gr.pcols = 128;
      gr.Q = wk.GT + (size_t)ba.goff[i] * 128; gr.qk = 1; gr.qrow_total = FB_GT_ROWS; gr.qcols = (hHeadDim[h] + 15) / 16 * 16;
      gr.nseg = 1;
      gr.seg[0] = WgSegment{g->head_w2[h], 128, 1, 0, hHeadDim[h], 0, 128};
    }
    if (g->feat_w[0]) {  // dW0[o][temb + i] += sum_n dhT[o][n] emb[n][i]: the 32 embedding columns on the N side
      WgGroup& gr = a.g[a.ngroups++];
      gr = WgGroup{};
      gr.P = wk.dhT; gr.pk = 1; gr.prow_total = 128; gr.pcols = 128;
      gr.Q = emb; gr.ldq = gemb; gr.qcols = gemb;
      gr.nseg = 1;
      gr.seg[0] = WgSegment{g->feat_w[0] + temb, 1, ld0, 0, gemb, 0, 128};
    }
#ifdef ED3DGS_TIMING_KNOBS
    {  // timing only (results are wrong): ED3DGS_WG_SKIP bit 0 drops the dW1 groups, bit 1 the dW2 groups, bit 2 dW0
      static const int skip = [] { const char* e = getenv("ED3DGS_WG_SKIP"); return e ? atoi(e) : 0; }();
      if (skip) {
        WgArgs b = a;
        b.ngroups = 0;
        for (int i = 0; i < a.ngroups; ++i) {
          const int kind = a.g[i].P == wk.hT[s] ? 1 : (a.g[i].P == wk.dhT ? 4 : 2);
          if (!(skip & kind)) b.g[b.ngroups++] = a.g[i];
        }
        a = b;
      }
    }
#endif
    const int keep_reserve = tc_sm_reserve();
    if (last_stage && tc_sm_reserve_last() > keep_reserve) tc_sm_reserve() = tc_sm_reserve_last();
    const cudaError_t we = tc_wgrad_launch(a, stream, wg_ts);
    tc_sm_reserve() = keep_reserve;
    TC_TRY(we);
  }
  layer0_temporal_backward<<<(temb + 31) / 32, 256, 0, stream>>>(
      w->feat_w[0], ld0, temb, 128, tau, wk.dc1, g ? g->feat_w[0] : nullptr, g ? g->feat_b[0] : nullptr, d_tau);
  TC_TRY(cudaGetLastError());
  return ED3DGS_OK;
}

}  // namespace ed3dgs

using namespace ed3dgs;

extern "C" {

int ed3dgs_temporal_embed_forward(const float* weight, int max_emb, int dim, int n_rows, const float* t_dev,
                                  float* tau, void* stream) {
  if (!weight || !t_dev || !tau || max_emb < 1 || dim < 2 || n_rows < 1) {
    set_error("temporal_embed_forward: bad argument");
    return ED3DGS_ERR_ARG;
  }
  temb_forward<<<(dim + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(weight, max_emb, dim, n_rows, t_dev, tau);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("temporal_embed_forward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int ed3dgs_temporal_embed_backward(const float* weight, int max_emb, int dim, int n_rows, const float* t_dev,
                                   const float* d_tau, float* d_weight, float* d_t, void* stream) {
  if (!weight || !t_dev || !d_tau || !d_weight || !d_t) {
    set_error("temporal_embed_backward: bad argument");
    return ED3DGS_ERR_ARG;
  }
  temb_backward<<<(dim + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(weight, max_emb, dim, n_rows, t_dev,
                                                                                  d_tau, d_weight, d_t);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("temporal_embed_backward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

void ed3dgs_deform_set_emb_ready_event(void* cuda_event) { g_emb_ready_event[tc_device_slot()] = cuda_event; }

int ed3dgs_deform_set_sm_reserve_last(int n_sms) {
  const int prev = tc_sm_reserve_last();
  tc_sm_reserve_last() = n_sms > 0 ? n_sms : 0;
  return prev;
}

int ed3dgs_deform_set_sm_reserve(int n_sms) {
  const int prev = tc_sm_reserve();
  tc_sm_reserve() = n_sms > 0 ? n_sms : 0;
  return prev;
}

size_t ed3dgs_deform_workspace_bytes(int W, int depth, int N) {
  char* p = nullptr;
  carve_deform(p, W, depth < 1 ? 1 : depth, (size_t)(N > 0 ? N : 0));
  return reinterpret_cast<size_t>(p) + 512;
}

int ed3dgs_deform_forward(const ed3dgs_deform_weights* coarse, const ed3dgs_deform_weights* fine, int N,
                          const float* embedding, const float* tau_c, const float* tau_f,
                          const float* const in[ED3DGS_DEFORM_HEADS], float* const sub[ED3DGS_DEFORM_HEADS],
                          float* const out[ED3DGS_DEFORM_HEADS], const float coef[ED3DGS_DEFORM_HEADS],
                          void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (N <= 0) return ED3DGS_OK;
  int rc;
  if ((rc = check_weights(coarse)) || (rc = check_weights(fine))) return rc;
  const ed3dgs_deform_weights* any = coarse ? coarse : fine;
  static const int hd[5] = {3, 3, 4, 1, 48};
  g_single = (any && any->precision == 1) ? 1 : 0;
  if (!any) {
    for (int h = 0; h < 5; ++h) {
      cudaMemcpyAsync(sub[h], in[h], (size_t)N * hd[h] * 4, cudaMemcpyDeviceToDevice, stream);
      cudaMemcpyAsync(out[h], in[h], (size_t)N * hd[h] * 4, cudaMemcpyDeviceToDevice, stream);
    }
    return ED3DGS_OK;
  }
  if (!workspace || workspace_bytes < ed3dgs_deform_workspace_bytes(any->W, any->depth, N)) {
    set_error("deform_forward: workspace too small");
    return ED3DGS_ERR_ARG;
  }
  char* p = static_cast<char*>(workspace);
  DeformWork wk = carve_deform(p, any->W, any->depth, (size_t)N);
  if (coarse) {
    if (fused_shape(coarse)) rc = stage_forward_fused(coarse, 0, N, embedding, tau_c, in, sub, coef, wk, stream);
    else rc = stage_forward(coarse, 0, N, embedding, tau_c, in, sub, coef, wk, stream);
    if (rc) return rc;
  } else {
    for (int h = 0; h < 5; ++h) cudaMemcpyAsync(sub[h], in[h], (size_t)N * hd[h] * 4, cudaMemcpyDeviceToDevice, stream);
  }
  if (fine) {
    const float* base[5] = {sub[0], sub[1], sub[2], sub[3], sub[4]};
    if (fused_shape(fine)) rc = stage_forward_fused(fine, 1, N, embedding, tau_f, base, out, coef, wk, stream);
    else rc = stage_forward(fine, 1, N, embedding, tau_f, base, out, coef, wk, stream);
    if (rc) return rc;
  } else {
    for (int h = 0; h < 5; ++h) cudaMemcpyAsync(out[h], sub[h], (size_t)N * hd[h] * 4, cudaMemcpyDeviceToDevice, stream);
  }
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("deform_forward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int ed3dgs_deform_backward(const ed3dgs_deform_weights* coarse, const ed3dgs_deform_weights* fine,
                           const ed3dgs_deform_grads* g_coarse, const ed3dgs_deform_grads* g_fine, int N,
                           const float* embedding, const float* tau_c, const float* tau_f,
                           const float* const d_out[ED3DGS_DEFORM_HEADS], const float* const d_sub[ED3DGS_DEFORM_HEADS],
                           const float coef[ED3DGS_DEFORM_HEADS], void* workspace, size_t workspace_bytes,
                           float* d_embedding, float* d_tau_c, float* d_tau_f, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const ed3dgs_deform_weights* any = coarse ? coarse : fine;
  if (N <= 0 || !any) return ED3DGS_OK;
  int rc;
  if ((rc = check_weights(coarse)) || (rc = check_weights(fine))) return rc;
  g_single = any->precision == 1 ? 1 : 0;
  if (!workspace || workspace_bytes < ed3dgs_deform_workspace_bytes(any->W, any->depth, N)) {
    set_error("deform_backward: workspace too small");
    return ED3DGS_ERR_ARG;
  }
  char* p = static_cast<char*>(workspace);
  DeformWork wk = carve_deform(p, any->W, any->depth, (size_t)N);

  GradPtrs gp{};
  bool any_sub = false;
  for (int h = 0; h < 5; ++h) {
    gp.d_out[h] = d_out ? d_out[h] : nullptr;
    gp.d_sub[h] = d_sub ? d_sub[h] : nullptr;
    gp.coef[h] = coef[h];
    any_sub = any_sub || gp.d_sub[h] != nullptr;
    if ((h == 2 || h == 4) && (((uintptr_t)gp.d_out[h] | (uintptr_t)gp.d_sub[h]) & 15u)) {
      set_error("deform_backward: the rotation / rgb upstream gradients must be 16-byte aligned");
      return ED3DGS_ERR_ARG;
    }
  }
  const bool fused = fused_shape(any);
  if (fused) {
    // fine stage sees d_out only; coarse stage sees d_out + d_sub
    int accumulate = 0;
    if (fine) {
      if ((rc = stage_backward_fused(fine, g_fine, 1, N, embedding, tau_f, gp.d_out, gp.d_sub, 0, coef, wk, d_embedding,
                                     accumulate, d_tau_f, stream, coarse == nullptr)))
        return rc;
      accumulate = 1;
    }
    if (coarse) {
      if ((rc = stage_backward_fused(coarse, g_coarse, 0, N, embedding, tau_c, gp.d_out, gp.d_sub, any_sub ? 1 : 0, coef,
                                     wk, d_embedding, accumulate, d_tau_c, stream, true)))
        return rc;
    }
  } else {
  const int gblocks = 148 * 6;
  // fine stage sees d_out only; coarse stage sees d_out + d_sub
  cudaMemsetAsync(wk.colsum[1], 0, GW * sizeof(float), stream);
  { ProfScope ps("build_head_grads", stream);
  build_head_grads<<<gblocks, BHG_THREADS, 0, stream>>>(N, gp, 0, wk.G[1], wk.colsum[1]); }
  const float* Gc = wk.G[1];
  const float* colc = wk.colsum[1];
  if (any_sub && coarse) {
    cudaMemsetAsync(wk.colsum[0], 0, GW * sizeof(float), stream);
    build_head_grads<<<gblocks, BHG_THREADS, 0, stream>>>(N, gp, 1, wk.G[0], wk.colsum[0]);
    Gc = wk.G[0];
    colc = wk.colsum[0];
  }
  int accumulate = 0;
  if (fine) {
    if ((rc = stage_backward(fine, g_fine, 1, N, embedding, tau_f, wk.G[1], wk.colsum[1], wk, d_embedding, accumulate,
                             d_tau_f, stream, coarse == nullptr)))
      return rc;
    accumulate = 1;
  }
  if (coarse) {
    if ((rc = stage_backward(coarse, g_coarse, 0, N, embedding, tau_c, Gc, colc, wk, d_embedding, accumulate, d_tau_c,
                             stream, true)))
      return rc;
  }
  }
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("deform_backward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

#ifdef ED3DGS_TRACE
int ed3dgs_trace_read(long long* host, int n) {
  if (!g_trace) return -1;
  cudaDeviceSynchronize();
  cudaMemcpy(host, g_trace, (size_t)n * 8, cudaMemcpyDeviceToHost);
  return 0;
}
#endif

// Test hook (not in the public header): one GEMM through the tensor-core building block.
int ed3dgs_test_tc_gemm(int amaj, int bmaj, int M, int n_valid, int NT, int K, const float* A, long lda,
                        const float* B, long ldb, float* C, long ldc, int k_per_cta, int ones_col,
                        float* bias_grad, int epilogue, int dbg, void* stream) {
  if (amaj == TC_MN_MAJOR) {
    // weight-gradient form: A stored [K][lda] (P side, M columns), B stored [K][ldb] (Q side, n_valid columns)
    WgArgs wa{};
    wa.N = K; wa.ngroups = 1; wa.single = (dbg & 2) ? 1 : 0;
    wa.g[0].P = A; wa.g[0].ldp = lda; wa.g[0].pcols = (M + 3) / 4 * 4;
    wa.g[0].Q = B; wa.g[0].ldq = ldb; wa.g[0].qcols = (n_valid + 3) / 4 * 4;
    wa.g[0].nseg = 1;
    wa.g[0].seg[0] = (epilogue == TC_EPI_ATOMIC_T) ? WgSegment{C, ldc, 1, 0, n_valid, 0, M} : WgSegment{C, 1, ldc, 0, n_valid, 0, M};
    const cudaError_t e = tc_wgrad_launch(wa, static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) { set_error("tc_wgrad launch failed: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
    return ED3DGS_OK;
  }
  TcArgs a{};
  a.b[0] = TcBatch{A, B, C, nullptr, nullptr, bias_grad, ldc, 0, n_valid, 0, 0, 1.f};
  a.lda = lda; a.ldb = ldb;
  a.M = M; a.K = K; a.NT = NT; a.k_per_cta = k_per_cta > 0 ? k_per_cta : K;
  a.epilogue = epilogue; a.relu = 0; a.accumulate = 0; a.ones_col = ones_col; a.dbg = dbg; a.single = (dbg & 2) ? 1 : 0;
  static float* scratch_of[64] = {};   // test entry: one scratch per device, never freed
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("tc_gemm: no current device"); return ED3DGS_ERR_CUDA; }
  float*& scratch = scratch_of[dev];
  if (!scratch && cudaMalloc(&scratch, tc_prep_scratch_floats() * sizeof(float)) != cudaSuccess) {
    set_error("tc_gemm test scratch allocation failed");
    return ED3DGS_ERR_CUDA;
  }
  const cudaError_t e = tc_launch(amaj, bmaj, a, 1, scratch, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) { set_error("tc_gemm launch failed: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

}  // extern "C"
