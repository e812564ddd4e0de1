// distCUDA2: mean squared distance of every point to its 3 nearest neighbours (exact).
// Replaces SimpleKNN::knn (simple-knn/simple_knn.cu:185-221): Morton order (10 bits / axis),
// 1024-point boxes, +-3 window seed, box-pruned exact scan. Differences from the reference:
// no host round trips (the bounding box stays on the device), no thrust allocations (caller
// workspace), points are gathered into Morton order once so the box scans read contiguous
// float4s, box bounds are staged through shared memory, and the sort is this library's own
// stable radix sort. The result is the exact 3-NN mean, so it is bit-identical to the reference.
#include <cfloat>
#include <mutex>
#include "raster_state.cuh"
#include "raster_api.h"

namespace ed3dgs {

int radix_sort_pairs(BinningState& b, uint32_t n, int end_bit, cudaStream_t stream);

constexpr int KNN_BOX = 1024;

struct KnnWork {
  float* bounds;        // [6] min xyz, max xyz
  float4* sorted_pts;   // [P]  xyz + original index bits
  float* box_min;       // [3 * nboxes] (x,y,z planes interleaved as float4-free SoA)
  float* box_max;
  BinningState bin;
};

static KnnWork carve_knn(char*& p, size_t P) {
  KnnWork w{};
  const size_t nb = (P + KNN_BOX - 1) / KNN_BOX;
  carve(p, w.bounds, 8);
  carve(p, w.sorted_pts, P);
  carve(p, w.box_min, 3 * nb);
  carve(p, w.box_max, 3 * nb);
  w.bin = carve_binning(p, P);
  return w;
}

size_t dist_cuda2_workspace(int P) {
  char* p = nullptr;
  carve_knn(p, (size_t)P);
  return reinterpret_cast<size_t>(p) + 256;
}

__device__ __forceinline__ float atomic_min_f(float* addr, float v) {
  // valid for any sign: ints order like floats for >= 0, reversed for < 0
  return (v >= 0) ? __int_as_float(atomicMin((int*)addr, __float_as_int(v)))
                  : __uint_as_float(atomicMax((unsigned*)addr, __float_as_uint(v)));
}
__device__ __forceinline__ float atomic_max_f(float* addr, float v) {
  return (v >= 0) ? __int_as_float(atomicMax((int*)addr, __float_as_int(v)))
                  : __uint_as_float(atomicMin((unsigned*)addr, __float_as_uint(v)));
}

// QUIRK: the reference reduces with init = (0,0,0) for both min and max
// (simple_knn.cu:192-200), so the box always contains the origin. bounds pre-set to 0.
__global__ void __launch_bounds__(256) knn_bounds(int P, const float* __restrict__ pts, float* __restrict__ bounds) {
  float mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v = pts[3 * i + c];
      mn[c] = fminf(mn[c], v);
      mx[c] = fmaxf(mx[c], v);
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      atomic_min_f(bounds + c, mn[c]);
      atomic_max_f(bounds + 3 + c, mx[c]);
    }
  }
}

__device__ __forceinline__ uint32_t morton_spread(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}

__global__ void __launch_bounds__(256)
knn_morton(int P, const float* __restrict__ pts, const float* __restrict__ bounds, uint64_t* __restrict__ keys,
           uint32_t* __restrict__ vals) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 mn = make_float3(bounds[0], bounds[1], bounds[2]);
  const float3 mx = make_float3(bounds[3], bounds[4], bounds[5]);
  const float3 c = make_float3(pts[3 * idx], pts[3 * idx + 1], pts[3 * idx + 2]);
  const uint32_t x = morton_spread(((c.x - mn.x) / (mx.x - mn.x)) * ((1 << 10) - 1));
  const uint32_t y = morton_spread(((c.y - mn.y) / (mx.y - mn.y)) * ((1 << 10) - 1));
  const uint32_t z = morton_spread(((c.z - mn.z) / (mx.z - mn.z)) * ((1 << 10) - 1));
  keys[idx] = (uint64_t)(x | (y << 1) | (z << 2));
  vals[idx] = (uint32_t)idx;
}

// Gather into Morton order and reduce each 1024-point box to its bounds (simple_knn.cu:78-124).
__global__ void __launch_bounds__(KNN_BOX)
knn_gather_boxes(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                 float4* __restrict__ sorted_pts, float* __restrict__ box_min, float* __restrict__ box_max) {
  __shared__ float s_mn[3][32], s_mx[3][32];
  const int idx = blockIdx.x * KNN_BOX + threadIdx.x;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (idx < P) {
    const uint32_t o = order[idx];
    const float3 p = make_float3(pts[3 * o], pts[3 * o + 1], pts[3 * o + 2]);
    sorted_pts[idx] = make_float4(p.x, p.y, p.z, __uint_as_float(o));
    mn[0] = mx[0] = p.x;
    mn[1] = mx[1] = p.y;
    mn[2] = mx[2] = p.z;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
    if (lane == 0) { s_mn[c][warp] = mn[c]; s_mx[c][warp] = mx[c]; }
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float a = s_mn[c][lane], b = s_mx[c][lane];
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) {
        a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
        b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
      }
      if (lane == 0) { box_min[3 * blockIdx.x + c] = a; box_max[3 * blockIdx.x + c] = b; }
    }
  }
}

__device__ __forceinline__ void knn_push3(const float3 ref, const float4 q, float* best) {
  const float3 d = make_float3(q.x - ref.x, q.y - ref.y, q.z - ref.z);
  float dist = d.x * d.x + d.y * d.y + d.z * d.z;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (best[j] > dist) {
      const float t = best[j];
      best[j] = dist;
      dist = t;
    }
  }
}

__device__ __forceinline__ float knn_box_dist(const float* mn, const float* mx, const float3 p) {
  float3 diff = make_float3(0, 0, 0);
  if (p.x < mn[0] || p.x > mx[0]) diff.x = fminf(fabsf(p.x - mn[0]), fabsf(p.x - mx[0]));
  if (p.y < mn[1] || p.y > mx[1]) diff.y = fminf(fabsf(p.y - mn[1]), fabsf(p.y - mx[1]));
  if (p.z < mn[2] || p.z > mx[2]) diff.z = fminf(fabsf(p.z - mn[2]), fabsf(p.z - mx[2]));
  return diff.x * diff.x + diff.y * diff.y + diff.z * diff.z;
}

constexpr int KNN_THREADS = 256;
constexpr int KNN_BOX_CHUNK = 512;

__global__ void __launch_bounds__(KNN_THREADS)
knn_mean_dist(int P, const float4* __restrict__ sorted_pts, const float* __restrict__ box_min,
              const float* __restrict__ box_max, float* __restrict__ dists) {
  __shared__ float s_mn[KNN_BOX_CHUNK * 3], s_mx[KNN_BOX_CHUNK * 3];
  const int idx = blockIdx.x * KNN_THREADS + threadIdx.x;
  const bool live = idx < P;
  float4 me = make_float4(0, 0, 0, 0);
  float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  float reject = FLT_MAX;
  if (live) {
    me = sorted_pts[idx];
    const float3 p = make_float3(me.x, me.y, me.z);
    for (int i = max(0, idx - 3); i <= min(P - 1, idx + 3); ++i) {
      if (i == idx) continue;
      knn_push3(p, sorted_pts[i], best);
    }
    reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
  }
  const float3 p = make_float3(me.x, me.y, me.z);
  const int nboxes = (P + KNN_BOX - 1) / KNN_BOX;
  for (int b0 = 0; b0 < nboxes; b0 += KNN_BOX_CHUNK) {
    const int nb = min(KNN_BOX_CHUNK, nboxes - b0);
    __syncthreads();
    for (int i = threadIdx.x; i < nb * 3; i += KNN_THREADS) {
      s_mn[i] = box_min[3 * b0 + i];
      s_mx[i] = box_max[3 * b0 + i];
    }
    __syncthreads();
    if (!live) continue;
    for (int b = 0; b < nb; ++b) {
      const float dist = knn_box_dist(s_mn + 3 * b, s_mx + 3 * b, p);
      if (dist > reject || dist > best[2]) continue;
      const int lo = (b0 + b) * KNN_BOX, hi = min(P, lo + KNN_BOX);
      for (int i = lo; i < hi; ++i) {
        if (i == idx) continue;
        knn_push3(p, sorted_pts[i], best);
      }
    }
  }
  if (live) dists[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

// ---------------------------------------------------------------------------------------------
// Exact K nearest neighbours of every point (K <= 32), squared distances in double, ascending, self excluded.
// Replaces utils/extra_utils.py:5-16 (o3d_knn: a python loop of P Open3D/FLANN kd-tree queries on the CPU in
// float64, run by train.py:218-229 for the embedding regulariser every time the number of Gaussians changes).
// Same Morton / box machinery as distCUDA2 above with boxes of 256 points; one thread per point in Morton
// order, so the 32 points of a warp are neighbours in space and visit the same boxes: the box scan is
// warp-uniform (a box is scanned if any lane needs it) and every candidate load is a broadcast. Candidates are
// filtered in fp32 against the current K-th distance (with a 1e-6 relative margin); survivors are measured in
// double -- exact for fp32 inputs up to the final rounding -- and inserted into a sorted list ordered by
// (distance, index). The +-K Morton window only provides the initial pruning bound, as in simple_knn.cu:150-170.
constexpr int KNN2_BOX = 256;
constexpr int KNN2_MAXK = 32;
constexpr int KNN2_THREADS = 128;
constexpr int KNN2_CHUNK = 1024;   // box bounds staged per pass: 1024 boxes x 6 floats = 24 KB

__global__ void __launch_bounds__(KNN2_BOX)
knn_box_bounds256(int P, const float4* __restrict__ sorted_pts, float* __restrict__ box_min, float* __restrict__ box_max) {
  __shared__ float s_mn[3][8], s_mx[3][8];
  const int idx = blockIdx.x * KNN2_BOX + threadIdx.x;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (idx < P) {
    const float4 p = sorted_pts[idx];
    mn[0] = mx[0] = p.x; mn[1] = mx[1] = p.y; mn[2] = mx[2] = p.z;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
    if (lane == 0) { s_mn[c][warp] = mn[c]; s_mx[c][warp] = mx[c]; }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int c = threadIdx.x;
    float a = s_mn[c][0], b = s_mx[c][0];
#pragma unroll
    for (int w = 1; w < 8; ++w) { a = fminf(a, s_mn[c][w]); b = fmaxf(b, s_mx[c][w]); }
    box_min[3 * blockIdx.x + c] = a;
    box_max[3 * blockIdx.x + c] = b;
  }
}

struct TopK {
  double d[KNN2_MAXK];
  uint32_t id[KNN2_MAXK];
  int K;
  __device__ __forceinline__ void reset() {
    for (int j = 0; j < K; ++j) { d[j] = DBL_MAX; id[j] = 0xffffffffu; }
  }
  __device__ __forceinline__ double worst() const { return d[K - 1]; }
  // sorted ascending by (distance, index)
  __device__ __forceinline__ void push(double dist, uint32_t index) {
    if (dist > d[K - 1] || (dist == d[K - 1] && index >= id[K - 1])) return;
    int j = K - 1;
    while (j > 0 && (d[j - 1] > dist || (d[j - 1] == dist && id[j - 1] > index))) {
      d[j] = d[j - 1];
      id[j] = id[j - 1];
      --j;
    }
    d[j] = dist;
    id[j] = index;
  }
};

__device__ __forceinline__ double knn_dist64(const float4 a, const float4 b) {
  const double dx = (double)a.x - (double)b.x, dy = (double)a.y - (double)b.y, dz = (double)a.z - (double)b.z;
  // every product and sum rounded on its own (no fma): the value a float64 CPU implementation produces
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

__global__ void __launch_bounds__(KNN2_THREADS)
knn_topk(int P, int K, const float4* __restrict__ sorted_pts, const float* __restrict__ box_min,
         const float* __restrict__ box_max, double* __restrict__ out_d, int* __restrict__ out_i) {
  __shared__ float s_mn[KNN2_CHUNK * 3], s_mx[KNN2_CHUNK * 3];
  const int idx = blockIdx.x * KNN2_THREADS + threadIdx.x;
  const bool live = idx < P;
  float4 me = make_float4(0, 0, 0, 0);
  TopK top;
  top.K = K;
  top.reset();
  double reject = DBL_MAX;
  if (live) {
    me = sorted_pts[idx];
    for (int i = max(0, idx - K); i <= min(P - 1, idx + K); ++i) {
      if (i == idx) continue;
      const float4 q = sorted_pts[i];
      top.push(knn_dist64(me, q), __float_as_uint(q.w));
    }
    reject = top.worst();
    top.reset();
  }
  const float3 p = make_float3(me.x, me.y, me.z);
  const int nboxes = (P + KNN2_BOX - 1) / KNN2_BOX;
  for (int b0 = 0; b0 < nboxes; b0 += KNN2_CHUNK) {
    const int nb = min(KNN2_CHUNK, nboxes - b0);
    __syncthreads();
    for (int i = threadIdx.x; i < nb * 3; i += KNN2_THREADS) {
      s_mn[i] = box_min[3 * b0 + i];
      s_mx[i] = box_max[3 * b0 + i];
    }
    __syncthreads();
    for (int b = 0; b < nb; ++b) {
      // current pruning bound in fp32, rounded up; a box (or candidate) is skipped only when its fp32 distance,
      // rounded down, is still beyond it
      const double bound = fmin(reject, top.worst());
      const float bound32 = (bound >= (double)FLT_MAX) ? FLT_MAX : (float)bound * 1.000001f + 1e-37f;
      const bool need = live && !(knn_box_dist(s_mn + 3 * b, s_mx + 3 * b, p) * 0.999999f > bound32);
      if (!__any_sync(0xffffffffu, need)) continue;
      const int lo = (b0 + b) * KNN2_BOX, hi = min(P, lo + KNN2_BOX);
      for (int i = lo; i < hi; ++i) {
        const float4 q = sorted_pts[i];   // same address for the whole warp
        const float3 d = make_float3(q.x - p.x, q.y - p.y, q.z - p.z);
        const float d32 = d.x * d.x + d.y * d.y + d.z * d.z;
        const double cur = fmin(reject, top.worst());
        const float cur32 = (cur >= (double)FLT_MAX) ? FLT_MAX : (float)cur * 1.000001f + 1e-37f;
        if (!live || i == idx || d32 * 0.999999f > cur32) continue;
        const double d64 = knn_dist64(me, q);
        if (d64 <= reject) top.push(d64, __float_as_uint(q.w));
      }
    }
  }
  if (live) {
    const size_t o = (size_t)__float_as_uint(me.w) * K;
    for (int j = 0; j < K; ++j) { out_d[o + j] = top.d[j]; out_i[o + j] = (int)top.id[j]; }
  }
}

size_t knn_workspace(int P) {
  char* p = nullptr;
  carve_knn(p, (size_t)P);
  float* f;
  const size_t nb = ((size_t)P + KNN2_BOX - 1) / KNN2_BOX;
  carve(p, f, 3 * nb);
  carve(p, f, 3 * nb);
  return reinterpret_cast<size_t>(p) + 256;
}

int knn_exact(int P, int K, const float* points, double* sq_dists, int* indices, void* workspace,
              size_t workspace_bytes, cudaStream_t stream) {
  if (P <= 0) return ED3DGS_OK;
  if (K < 1 || K > KNN2_MAXK || P <= K) {
    set_error("ed3dgs_knn: need 1 <= K <= %d and more than K points (K = %d, P = %d)", KNN2_MAXK, K, P);
    return ED3DGS_ERR_ARG;
  }
  if (workspace == nullptr || workspace_bytes < knn_workspace(P)) {
    set_error("ed3dgs_knn workspace too small: need %zu bytes", knn_workspace(P));
    return ED3DGS_ERR_ARG;
  }
  char* p = reinterpret_cast<char*>(workspace);
  KnnWork w = carve_knn(p, (size_t)P);
  const int nb256 = (P + KNN2_BOX - 1) / KNN2_BOX;
  float *bmin, *bmax;
  carve(p, bmin, 3 * (size_t)nb256);
  carve(p, bmax, 3 * (size_t)nb256);
  cudaMemsetAsync(w.bounds, 0, 8 * sizeof(float), stream);
  knn_bounds<<<min((P + 255) / 256, 1184), 256, 0, stream>>>(P, points, w.bounds);
  knn_morton<<<(P + 255) / 256, 256, 0, stream>>>(P, points, w.bounds, w.bin.keys[2], w.bin.vals[2]);
  const int sel = radix_sort_pairs(w.bin, (uint32_t)P, 30, stream);
  if (sel < 0) { set_error("radix_sort_pairs: %s", cudaGetErrorString(cudaGetLastError())); return ED3DGS_ERR_CUDA; }
  const int nboxes = (P + KNN_BOX - 1) / KNN_BOX;
  knn_gather_boxes<<<nboxes, KNN_BOX, 0, stream>>>(P, points, w.bin.vals[sel], w.sorted_pts, w.box_min, w.box_max);
  knn_box_bounds256<<<nb256, KNN2_BOX, 0, stream>>>(P, w.sorted_pts, bmin, bmax);
  knn_topk<<<(P + KNN2_THREADS - 1) / KNN2_THREADS, KNN2_THREADS, 0, stream>>>(P, K, w.sorted_pts, bmin, bmax, sq_dists,
                                                                             indices);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) { set_error("ed3dgs_knn launch failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int dist_cuda2(int P, const float* points, float* mean_dists, void* workspace, size_t workspace_bytes,
               cudaStream_t stream) {
  if (P <= 0) return ED3DGS_OK;
  if (workspace == nullptr || workspace_bytes < dist_cuda2_workspace(P)) {
    set_error("distCUDA2 workspace too small: need %zu bytes", dist_cuda2_workspace(P));
    return ED3DGS_ERR_ARG;
  }
  char* p = reinterpret_cast<char*>(workspace);
  KnnWork w = carve_knn(p, (size_t)P);
  cudaMemsetAsync(w.bounds, 0, 8 * sizeof(float), stream);
  knn_bounds<<<min((P + 255) / 256, 1184), 256, 0, stream>>>(P, points, w.bounds);
  knn_morton<<<(P + 255) / 256, 256, 0, stream>>>(P, points, w.bounds, w.bin.keys[2], w.bin.vals[2]);
  const int sel = radix_sort_pairs(w.bin, (uint32_t)P, 30, stream);
  if (sel < 0) { set_error("radix_sort_pairs: %s", cudaGetErrorString(cudaGetLastError())); return ED3DGS_ERR_CUDA; }
  const int nboxes = (P + KNN_BOX - 1) / KNN_BOX;
  knn_gather_boxes<<<nboxes, KNN_BOX, 0, stream>>>(P, points, w.bin.vals[sel], w.sorted_pts, w.box_min, w.box_max);
  knn_mean_dist<<<(P + KNN_THREADS - 1) / KNN_THREADS, KNN_THREADS, 0, stream>>>(P, w.sorted_pts, w.box_min,
                                                                                w.box_max, mean_dists);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) { set_error("distCUDA2 launch failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

}  // namespace ed3dgs

// ---------------------------------------------------------------------------------------------
// 3D smoothing filter size per Gaussian. Replaces GaussianModel.compute_3D_filter
// (scene/gaussian_model.py:538-592): a python loop over every training camera with ~15 torch kernels per
// camera (and a lazy image load) becomes one pass: a thread per point walks the camera table in shared memory.
// Camera record (16 floats): R[9] row-major as stored by the reference (camera.R, used as xyz @ R), T[3],
// focal_x, focal_y, width, height. filter = min over cameras that see the point (depth > 0.2, inside the
// image widened by 15 %) of the clamped depth, the maximum of those for points no camera sees,
// / max focal_x * sqrt(0.2).
namespace ed3dgs {

constexpr int F3D_MAX_CAMS_SMEM = 512;

__global__ void __launch_bounds__(256)
filter3d_min_depth(int P, const float* __restrict__ xyz, int C, const float* __restrict__ cams,
                   float* __restrict__ distance, unsigned int* __restrict__ max_valid_bits) {
  __shared__ float s_cam[F3D_MAX_CAMS_SMEM * 16];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float3 p = make_float3(0, 0, 0);
  if (idx < P) p = make_float3(xyz[3 * idx], xyz[3 * idx + 1], xyz[3 * idx + 2]);
  float dist = 100000.0f;
  bool any = false;
  for (int c0 = 0; c0 < C; c0 += F3D_MAX_CAMS_SMEM) {
    const int nc = min(F3D_MAX_CAMS_SMEM, C - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < nc * 16; i += blockDim.x) s_cam[i] = cams[(size_t)c0 * 16 + i];
    __syncthreads();
    for (int c = 0; c < nc; ++c) {
      const float* m = s_cam + 16 * c;
      // xyz @ R (an fma chain, as an fp32 GEMM with K = 3 computes it), then + T as a separate op (:558)
      const float x = __fadd_rn(__fmaf_rn(p.z, m[6], __fmaf_rn(p.y, m[3], __fmul_rn(p.x, m[0]))), m[9]);
      const float y = __fadd_rn(__fmaf_rn(p.z, m[7], __fmaf_rn(p.y, m[4], __fmul_rn(p.x, m[1]))), m[10]);
      float z = __fadd_rn(__fmaf_rn(p.z, m[8], __fmaf_rn(p.y, m[5], __fmul_rn(p.x, m[2]))), m[11]);
      const bool valid_depth = z > 0.2f;
      z = fmaxf(z, 0.001f);
      const float W = m[14], H = m[15];
      // three separately rounded torch ops in the reference (:566-567)
      const float sx = __fadd_rn(__fmul_rn(__fdiv_rn(x, z), m[12]), W / 2.0f);
      const float sy = __fadd_rn(__fmul_rn(__fdiv_rn(y, z), m[13]), H / 2.0f);
      const bool in_screen = sx >= -0.15f * W && sx <= W * 1.15f && sy >= -0.15f * H && sy <= 1.15f * H;
      if (valid_depth && in_screen && idx < P) {
        dist = fminf(dist, z);
        any = true;
      }
    }
  }
  float mx = any ? dist : 0.f;   // depths are > 0.2, so 0 never wins
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(max_valid_bits, __float_as_uint(mx));
  if (idx < P) distance[idx] = any ? dist : -1.0f;   // -1 marks "seen by no camera"
}

__global__ void __launch_bounds__(256)
filter3d_finish(int P, float* __restrict__ filter, const unsigned int* __restrict__ max_valid_bits, float focal_length) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float d = filter[idx];
  if (d < 0.f) d = __uint_as_float(*max_valid_bits);
  filter[idx] = d / focal_length * 0.4472135954999579f;   // (0.2 ** 0.5)
}

// scratch: at least 4 bytes of device memory; *any_valid (host) = 0 when no camera sees any point (the reference
// raises in that case: max() of an empty tensor, gaussian_model.py:587).
int compute_3d_filter(int P, const float* xyz, int C, const float* cams, float focal_length, float* filter,
                      void* scratch, int* any_valid, cudaStream_t stream) {
  if (any_valid) *any_valid = 1;
  if (P <= 0) return ED3DGS_OK;
  if (C <= 0 || !xyz || !cams || !filter || !scratch) { set_error("compute_3d_filter: bad argument"); return ED3DGS_ERR_ARG; }
  unsigned int* mx = static_cast<unsigned int*>(scratch);
  cudaMemsetAsync(mx, 0, sizeof(unsigned int), stream);
  filter3d_min_depth<<<(P + 255) / 256, 256, 0, stream>>>(P, xyz, C, cams, filter, mx);
  filter3d_finish<<<(P + 255) / 256, 256, 0, stream>>>(P, filter, mx, focal_length);
  unsigned int host_mx = 0;
  cudaError_t err = cudaMemcpyAsync(&host_mx, mx, sizeof(host_mx), cudaMemcpyDeviceToHost, stream);
  if (err == cudaSuccess) err = cudaStreamSynchronize(stream);
  if (err == cudaSuccess) err = cudaGetLastError();
  if (err != cudaSuccess) { set_error("compute_3d_filter failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  if (any_valid) *any_valid = host_mx != 0 ? 1 : 0;
  return ED3DGS_OK;
}

}  // namespace ed3dgs

// ---------------------------------------------------------------------------------------------
// distCUDA2 without a caller workspace (ed3dgs_dist_cuda2_static, include/ed3dgs_b200.h).
//
// Who this is for: code that wants the initial Gaussian scales but is in no position to hand over device
// scratch. A plug-in living inside another program's fixed memory pool (an inference server that carved up the
// whole GPU at start-up, a renderer with a frozen arena), a host written in a language whose FFI only ever sees
// the input and output arrays, or a tool running late in a job whose pool is exhausted: none of them can call
// cudaMalloc -- some because the pool owner forbids it, some because there is nothing left to hand out, and
// all of them would pay a device-wide synchronisation for every cudaMalloc / cudaFree pair anyway. For them
// the Morton keys, the sort's ping-pong buffers and its look-back words live in the array below instead, which
// is part of the library image: the driver places it when the module is loaded on a device and keeps it until
// the process exits, whether or not this entry point is ever called.
//
// Limit: KNN_STATIC_MAX_POINTS points (sparse structure-from-motion clouds, the reference's input at
// scene/gaussian_model.py:161, are well inside it). Larger clouds must use ed3dgs_dist_cuda2 with their own
// workspace.
namespace ed3dgs {

constexpr int KNN_STATIC_MAX_POINTS = 262144;
constexpr size_t KNN_STATIC_BYTES = 16u << 20;   // 52 B per point + 0.6 MB of sort words at the limit, rounded up
constexpr int KNN_STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_knn_static_scratch[KNN_STATIC_BYTES];

// g_knn_static_scratch, written in host code, is only the host-side handle of the symbol -- not somewhere a
// kernel can read or write. The address the kernels need differs per device and comes from the runtime, asked
// once per device while that device is current.
static void* g_knn_static_addr[KNN_STATIC_MAX_DEVICES] = {};
static std::mutex g_knn_static_mu;

static int knn_static_address(void** out) {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) { set_error("ed3dgs_dist_cuda2_static: cudaGetDevice: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  if (dev < 0 || dev >= KNN_STATIC_MAX_DEVICES) {
    set_error("ed3dgs_dist_cuda2_static: device ordinal %d outside [0, %d)", dev, KNN_STATIC_MAX_DEVICES);
    return ED3DGS_ERR_ARG;
  }
  std::lock_guard<std::mutex> lk(g_knn_static_mu);
  if (g_knn_static_addr[dev] == nullptr) {
    void* ptr = nullptr;
    err = cudaGetSymbolAddress(&ptr, g_knn_static_scratch);
    if (err != cudaSuccess || ptr == nullptr) {
      set_error("ed3dgs_dist_cuda2_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(err));
      return ED3DGS_ERR_CUDA;
    }
    g_knn_static_addr[dev] = ptr;
  }
  *out = g_knn_static_addr[dev];
  return ED3DGS_OK;
}

int dist_cuda2_static(int P, const float* points, float* mean_dists, cudaStream_t stream) {
  if (P <= 0) return ED3DGS_OK;
  if (P > KNN_STATIC_MAX_POINTS) {
    set_error("ed3dgs_dist_cuda2_static: %d points exceed the built-in scratch (at most %d); "
              "use ed3dgs_dist_cuda2 with a caller workspace",
              P, KNN_STATIC_MAX_POINTS);
    return ED3DGS_ERR_ARG;
  }
  if (points == nullptr || mean_dists == nullptr) {
    set_error("ed3dgs_dist_cuda2_static: points and mean_dists must be device pointers");
    return ED3DGS_ERR_ARG;
  }
  if (dist_cuda2_workspace(P) > KNN_STATIC_BYTES) {   // cannot happen unless the layout grows past the array
    set_error("ed3dgs_dist_cuda2_static: layout needs %zu bytes, the array holds %zu", dist_cuda2_workspace(P),
              KNN_STATIC_BYTES);
    return ED3DGS_ERR_ARG;
  }
  void* scratch = nullptr;
  const int rc = knn_static_address(&scratch);
  if (rc != ED3DGS_OK) return rc;
  // Same launches as the workspace variant. Nothing from an earlier call is relied on: the bounding box and the
  // sort's tickets / look-back words are cleared on the stream by dist_cuda2 and radix_sort_pairs themselves,
  // every other region is fully overwritten before it is read.
  return dist_cuda2(P, points, mean_dists, scratch, KNN_STATIC_BYTES, stream);
}

}  // namespace ed3dgs

extern "C" {

int ed3dgs_dist_cuda2_static_max_points(void) { return ed3dgs::KNN_STATIC_MAX_POINTS; }

int ed3dgs_dist_cuda2_static(int P, const float* points, float* mean_dists, void* stream) {
  ed3dgs::set_error("%s", "");
  return ed3dgs::dist_cuda2_static(P, points, mean_dists, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
