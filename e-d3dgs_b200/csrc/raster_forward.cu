// Forward half of the splat rasterizer for sm_100a: per-Gaussian preprocess, tile-count scan,
// key emission, (radix sort: radix_sort.cu), tile ranges and the front-to-back blend.
// Behavioural contract = CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:255-432) with
// kernels forward.cu:427-545 (preprocessCUDATongue), rasterizer_impl.cu:70-111
// (duplicateWithKeys), :151-173 (identifyTileRanges), forward.cu:550-822 (renderCUDA).
#include <cstdio>
#include "raster_cull.cuh"
#include "gs_math.cuh"
#include "raster_state.cuh"
#include "raster_api.h"
#include "raster_front.h"
#include "profiler.h"

namespace ed3dgs {

int radix_sort_pairs(BinningState& b, uint32_t n, int end_bit, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------
// SH -> RGB (forward.cu:23-74). sh is [M][3] for this Gaussian.
__device__ __forceinline__ float3 sh_to_rgb(int deg, const float3 pos, const float3 campos,
                                            const float* __restrict__ sh, uint8_t& clamp_bits) {
  float3 dir = f3_sub(pos, campos);
  const float len = sqrtf(dot3(dir, dir));
  dir = make_float3(dir.x / len, dir.y / len, dir.z / len);
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    auto S = [&](int i) { return sh[3 * i + c]; };
    float r = kSH0 * S(0);
    if (deg > 0) {
      const float x = dir.x, y = dir.y, z = dir.z;
      r = r - kSH1 * y * S(1) + kSH1 * z * S(2) - kSH1 * x * S(3);
      if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z;
        const float xy = x * y, yz = y * z, xz = x * z;
        r = r + kSH2[0] * xy * S(4) + kSH2[1] * yz * S(5) + kSH2[2] * (2.0f * zz - xx - yy) * S(6) +
            kSH2[3] * xz * S(7) + kSH2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          r = r + kSH3[0] * y * (3.0f * xx - yy) * S(9) + kSH3[1] * xy * z * S(10) +
              kSH3[2] * y * (4.0f * zz - xx - yy) * S(11) +
              kSH3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
              kSH3[4] * x * (4.0f * zz - xx - yy) * S(13) + kSH3[5] * z * (xx - yy) * S(14) +
              kSH3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
    res[c] = r + 0.5f;
  }
  clamp_bits = (res[0] < 0 ? 1 : 0) | (res[1] < 0 ? 2 : 0) | (res[2] < 0 ? 4 : 0);
  return make_float3(fmaxf(res[0], 0.0f), fmaxf(res[1], 0.0f), fmaxf(res[2], 0.0f));
}

// RaDe-GS geometry of one Gaussian (forward.cu:135-262): ray-space plane, camera-space plane,
// view-space normal (all zero in the degenerate case). Returns the reference's `well_conditioned`
// flag. INTE (computeCov2D<true>, forward.cu:187-236) additionally writes the inverse ray-space
// covariance the point integration evaluates (upper triangle, 6 floats; zeros when degenerate).
// For an ill-conditioned covariance the reference assigns that matrix to a shadowed local
// (forward.cu:222) and stores an uninitialised one; the value it meant to store is written here.
template <bool INTE>
__device__ __forceinline__ bool splat_planes(const Cov2DCtx& c, float fx, float fy, float* cam_plane6,
                                             float2& ray_plane, float3& normal, float* inv6) {
  Mat3 evec;
  float eval[3];
  const int D = eig_sym3(c.Vrk, eval, evec);
  const int min_id = argmin3(eval);
  const bool well = eval[min_id] > 0.00000001;
  Mat3 Vinv;
  if (well) {
    const Mat3 diag = mat3_cols(1 / eval[0], 0, 0, 0, 1 / eval[1], 0, 0, 0, 1 / eval[2]);
    Vinv = mat3_mul(mat3_mul(evec, diag), mat3_t(evec));
  } else {
    const float3 vmin = col3(evec, min_id);
    Vinv = mat3_outer(vmin, vmin);
  }
  const Mat3 cam_inv = mat3_mul(mat3_mul(mat3_t(c.W), Vinv), c.W);
  const float3 uvh = make_float3(c.txtz, c.tytz, 1);
  const float3 uvh_m = mat3_vec(cam_inv, uvh);
  const float3 uvh_mn = f3_normalize(uvh_m);
  if (isnan(uvh_mn.x) || D == 0) {
#pragma unroll
    for (int i = 0; i < 6; ++i) cam_plane6[i] = 0;
    ray_plane = make_float2(0, 0);
    normal = make_float3(0, 0, 0);
    if constexpr (INTE) {
#pragma unroll
      for (int i = 0; i < 6; ++i) inv6[i] = 0;
    }
    return well;
  }
  const float3 t = c.t;
  const float u2 = c.txtz * c.txtz, v2 = c.tytz * c.tytz, uv = c.txtz * c.tytz;
  const float l = sqrtf(t.x * t.x + t.y * t.y + t.z * t.z);
  const Mat3 nJ = mat3_cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z,
                            -(t.y) / (t.z * t.z), t.x / l, t.y / l, t.z / l);
  const Mat3 nJ_inv = mat3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -c.txtz, -c.tytz, 0);
  if constexpr (INTE) {
    Mat3 inv_ray;
    if (well) {
      const float ltz = u2 + v2 + 1;
      const Mat3 full = mat3_scale(mat3_cols(v2 + 1, -uv, c.txtz / l * ltz, -uv, u2 + 1, c.tytz / l * ltz,
                                             -c.txtz, -c.tytz, 1 / l * ltz),
                                   t.z / (u2 + v2 + 1));
      const Mat3 T2 = mat3_mul(c.W, mat3_t(full));
      inv_ray = mat3_mul(mat3_mul(mat3_t(T2), Vinv), T2);
    } else {
      const Mat3 T2 = mat3_mul(c.W, nJ);
      const Mat3 cov_ray = mat3_mul(mat3_mul(mat3_t(T2), Vinv), T2);
      Mat3 rvec;
      float rval[3];
      eig_sym3(cov_ray, rval, rvec);
      const int mn = argmin3(rval);
      const int i1 = (mn + 1) % 3, i2 = (mn + 2) % 3;
      const float l1 = rval[i1], l2 = rval[i2];
      Mat3 nv;
#pragma unroll
      for (int r = 0; r < 3; ++r) { nv.m[0][r] = rvec.m[i1][r]; nv.m[1][r] = rvec.m[i2][r]; nv.m[2][r] = rvec.m[mn][r]; }
      const float3 r3 = make_float3(nv.m[0][2], nv.m[1][2], nv.m[2][2]);
      const Mat3 c2 = mat3_cols(1 / l1, 0, -r3.x / r3.z / l1, 0, 1 / l2, -r3.y / r3.z / l2, -r3.x / r3.z / l1,
                                -r3.y / r3.z / l2, 0);
      inv_ray = mat3_mul(mat3_mul(nv, c2), mat3_t(nv));
    }
    const Mat3 sc = mat3_cols(1 / fx, 0, 0, 0, 1 / fy, 0, 0, 0, 1);
    inv_ray = mat3_mul(mat3_mul(sc, inv_ray), sc);
    inv6[0] = inv_ray.m[0][0]; inv6[1] = inv_ray.m[0][1]; inv6[2] = inv_ray.m[0][2];
    inv6[3] = inv_ray.m[1][1]; inv6[4] = inv_ray.m[1][2]; inv6[5] = inv_ray.m[2][2];
  }
  const float vbn = dot3(uvh_mn, uvh);
  const float factor_normal = l / (u2 + v2 + 1);
  const float vb = fmaxf(vbn, 0.0000001f);
  const float3 plane = mat3_vec(nJ_inv, make_float3(uvh_mn.x / vb, uvh_mn.y / vb, uvh_mn.z / vb));
  const float nl = u2 + v2 + 1;
  cam_plane6[0] = (-(v2 + 1) * t.z + plane.x * t.x) / nl / fx;
  cam_plane6[1] = (uv * t.z + plane.y * t.x) / nl / fy;
  cam_plane6[2] = (uv * t.z + plane.x * t.y) / nl / fx;
  cam_plane6[3] = (-(u2 + 1) * t.z + plane.y * t.y) / nl / fy;
  cam_plane6[4] = (t.x + plane.x * t.z) / nl / fx;
  cam_plane6[5] = (t.y + plane.y * t.z) / nl / fy;
  ray_plane = make_float2(plane.x * l / nl / fx, plane.y * l / nl / fy);
  const float3 ray_n = make_float3(-plane.x * factor_normal, -plane.y * factor_normal, -1);
  normal = f3_normalize(mat3_vec(nJ, ray_n));
  return well;
}

// INTE = the point-integration flavour (preprocessCUDA<C, true>, forward.cu:307-424): no tongue class,
// and two more quads per Gaussian in inte_rec: {inv[0..3]} {inv[4], inv[5], well_conditioned, 0}.
template <bool INTE>
__global__ void __launch_bounds__(256)
preprocess_fwd(int P, int D, int M, const float* __restrict__ means, const float* __restrict__ scales,
               float scale_mod, const float* __restrict__ rots, const float* __restrict__ opac,
               const float* __restrict__ tongue, const float* __restrict__ shs,
               const float* __restrict__ cov3D_pre, const float* __restrict__ colors_pre,
               const float* __restrict__ view, const float* __restrict__ proj,
               const float* __restrict__ campos, int W, int H, float tan_fovx, float tan_fovy,
               float fx, float fy, float kernel_size, int* __restrict__ radii, GeomState g,
               uint32_t gx, uint32_t gy, int write_coord, float4* __restrict__ inte_rec) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  radii[idx] = 0;
  g.tiles_touched[idx] = 0;

  const float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  // near cull only (auxiliary.h:155-180); `prefiltered` is forced false by the reference
  // (rasterizer_impl.cu:348-350), so there is no trap path.
  const float4 hom = xform_4x4(p, proj);
  const float pw = 1.0f / (hom.w + 0.0000001f);
  const float3 pproj = make_float3(hom.x * pw, hom.y * pw, hom.z * pw);
  const float3 pv = xform_4x3(p, view);
  if (pv.z <= 0.2f) return;

  float cov6[6];
  if (cov3D_pre != nullptr) {
#pragma unroll
    for (int i = 0; i < 6; ++i) cov6[i] = cov3D_pre[6 * idx + i];
  } else {
    const float3 s = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
    const float4 q = make_float4(rots[4 * idx], rots[4 * idx + 1], rots[4 * idx + 2], rots[4 * idx + 3]);
    cov3d_from_scale_rot(s, scale_mod, q, cov6);
  }

  Cov2DCtx c;
  cov2d_setup(p, fx, fy, tan_fovx, tan_fovy, cov6, view, c);
  const float cov_x = float(c.cov.m[0][0] + kernel_size);
  const float cov_y = float(c.cov.m[0][1]);
  const float cov_z = float(c.cov.m[1][1] + kernel_size);
  // mip-filter opacity compensation, double literals as in forward.cu:119-124
  const float det_0 = max(1e-6, c.cov.m[0][0] * c.cov.m[1][1] - c.cov.m[0][1] * c.cov.m[0][1]);
  const float det_1 = max(1e-6, (c.cov.m[0][0] + kernel_size) * (c.cov.m[1][1] + kernel_size) -
                                    c.cov.m[0][1] * c.cov.m[0][1]);
  float coef = sqrt(det_0 / (det_1 + 1e-6) + 1e-6);
  if (det_0 <= 1e-6 || det_1 <= 1e-6) coef = 0.0f;

  float cam_plane[6];
  float2 ray_plane;
  float3 normal;
  float inv6[6];
  const bool well = splat_planes<INTE>(c, fx, fy, cam_plane, ray_plane, normal, inv6);
  if constexpr (INTE) {
    inte_rec[2 * (size_t)idx] = make_float4(inv6[0], inv6[1], inv6[2], inv6[3]);
    inte_rec[2 * (size_t)idx + 1] = make_float4(inv6[4], inv6[5], well ? 1.f : 0.f, 0.f);
  }
  // x^2 + y^2 + z^2 as nvcc contracts it in the reference (fma(z, z, fma(x, x, y * y))): the point
  // integration subtracts depths of nearly equal size, so |p_view| has to match to the last bit there
  float ts;
  if constexpr (INTE) ts = sqrtf(__fmaf_rn(pv.z, pv.z, __fmaf_rn(pv.x, pv.x, __fmul_rn(pv.y, pv.y))));
  else ts = sqrtf(pv.x * pv.x + pv.y * pv.y + pv.z * pv.z);

  const float det = (cov_x * cov_z - cov_y * cov_y);
  if (det == 0.0f) return;
  const float det_inv = 1.f / det;
  const float3 conic = make_float3(cov_z * det_inv, -cov_y * det_inv, cov_x * det_inv);

  const float mid = 0.5f * (cov_x + cov_z);
  const float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
  const float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
  const float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
  const float2 pix = make_float2(ndc_to_pix(pproj.x, W), ndc_to_pix(pproj.y, H));
  uint2 rmin, rmax;
  tile_rect(pix, (int)my_radius, rmin, rmax, gx, gy);
  if ((rmax.x - rmin.x) * (rmax.y - rmin.y) == 0) return;

  float3 rgb;
  if (colors_pre == nullptr) {
    uint8_t bits;
    rgb = sh_to_rgb(D, p, make_float3(campos[0], campos[1], campos[2]), shs + (size_t)idx * M * 3, bits);
    g.clamped[idx] = bits;
  } else {
    rgb = make_float3(colors_pre[3 * idx], colors_pre[3 * idx + 1], colors_pre[3 * idx + 2]);
  }

  g.depths[idx] = pv.z;
  radii[idx] = (int)my_radius;
  float4* rec = g.rec + (size_t)idx * REC_QUADS;
  rec[0] = make_float4(pix.x, pix.y, ts, INTE ? 0.f : tongue[idx]);
  rec[1] = make_float4(conic.x, conic.y, conic.z, opac[idx] * coef);
  rec[2] = make_float4(rgb.x, rgb.y, rgb.z, ray_plane.x);
  rec[3] = make_float4(ray_plane.y, normal.x, normal.y, normal.z);
  if (write_coord) {
    float4* rc = g.rec_coord + (size_t)idx * REC_COORD_QUADS;
    rc[0] = make_float4(cam_plane[0], cam_plane[1], cam_plane[2], cam_plane[3]);
    rc[1] = make_float4(cam_plane[4], cam_plane[5], pv.x, pv.y);
    rc[2] = make_float4(pv.z, 0.f, 0.f, 0.f);
  }
  g.tiles_touched[idx] = (rmax.y - rmin.y) * (rmax.x - rmin.x);
}

// ---------------------------------------------------------------------------------------------
// Inclusive scan of tiles_touched (replaces cub::DeviceScan::InclusiveSum, rasterizer_impl.cu:355)
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = SCAN_TILE / SCAN_THREADS;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_local(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, uint32_t n,
                 uint32_t* __restrict__ block_sums) {
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t i0 = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint32_t tsum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    v[k] = (i0 + k < n) ? in[i0 + k] : 0u;
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
  for (int w = 0; w < SCAN_THREADS / 32; ++w)
    if (w < warp) before += s_warp[w];
  uint32_t run = before + inc - tsum;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    run += v[k];
    if (i0 + k < n) out[i0 + k] = run;
  }
  if (threadIdx.x == SCAN_THREADS - 1) block_sums[blockIdx.x] = run;
}

__global__ void scan_exclusive_1cta(uint32_t* __restrict__ data, uint32_t count, uint32_t* __restrict__ total_out);

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tiles_add(uint32_t* __restrict__ out, uint32_t n, const uint32_t* __restrict__ block_prefix) {
  const uint32_t add = block_prefix[blockIdx.x];
  if (add == 0) return;
  const uint32_t i0 = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k)
    if (i0 + k < n) out[i0 + k] += add;
}

// ---------------------------------------------------------------------------------------------
// Key emission: key = tile_id << 32 | float_bits(depth), value = Gaussian index, for every tile
// of the Gaussian's rectangle in row-major order (rasterizer_impl.cu:70-111). A Gaussian with few
// tiles is written by its own thread; a large one is written by the whole warp, coalesced.
constexpr uint32_t DUP_COOP_MIN = 16;

__global__ void __launch_bounds__(256)
emit_keys(int P, const float4* __restrict__ rec, const float* __restrict__ depths,
          const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
          uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t gx, uint32_t gy) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t off = 0, count = 0, w = 0;
  uint2 rmin = make_uint2(0, 0), rmax = make_uint2(0, 0);
  uint32_t dbits = 0;
  if (idx < P) {
    const int r = radii[idx];
    if (r > 0) {
      off = (idx == 0) ? 0u : offsets[idx - 1];
      const float4 q0 = rec[(size_t)idx * REC_QUADS];
      tile_rect(make_float2(q0.x, q0.y), r, rmin, rmax, gx, gy);
      w = rmax.x - rmin.x;
      count = w * (rmax.y - rmin.y);
      dbits = __float_as_uint(depths[idx]);
    }
  }
  const bool big = count >= DUP_COOP_MIN;
  if (count > 0 && !big) {
    uint32_t o = off;
    for (uint32_t y = rmin.y; y < rmax.y; ++y)
      for (uint32_t x = rmin.x; x < rmax.x; ++x) {
        keys[o] = ((uint64_t)(y * gx + x) << 32) | dbits;
        vals[o] = (uint32_t)idx;
        ++o;
      }
  }
  uint32_t todo = __ballot_sync(0xffffffffu, big);
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const uint32_t s_off = __shfl_sync(0xffffffffu, off, src);
    const uint32_t s_cnt = __shfl_sync(0xffffffffu, count, src);
    const uint32_t s_w = __shfl_sync(0xffffffffu, w, src);
    const uint32_t s_x0 = __shfl_sync(0xffffffffu, rmin.x, src);
    const uint32_t s_y0 = __shfl_sync(0xffffffffu, rmin.y, src);
    const uint32_t s_d = __shfl_sync(0xffffffffu, dbits, src);
    const uint32_t s_idx = (uint32_t)(idx - lane + src);
    for (uint32_t t = lane; t < s_cnt; t += 32) {
      const uint32_t y = s_y0 + t / s_w, x = s_x0 + t % s_w;
      keys[s_off + t] = ((uint64_t)(y * gx + x) << 32) | s_d;
      vals[s_off + t] = s_idx;
    }
  }
}

// Tile ranges from the sorted key list (rasterizer_impl.cu:151-173); ranges pre-zeroed.
__global__ void __launch_bounds__(256)
tile_ranges(uint32_t L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t cur = (uint32_t)(keys[idx] >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = idx;
      ranges[cur].x = idx;
    }
  }
  if (idx == L - 1) ranges[cur].y = L;
}

// ---------------------------------------------------------------------------------------------
// Front-to-back blend. One CTA per 16x16 tile, one pixel per thread, a warp covers an 8x4 pixel
// patch. Gaussian records are staged in shared memory with 16-byte cp.async copies, double
// buffered: batch i+1 streams in while batch i is blended; the index list for batch i+2 is
// already in registers. Per-pixel arithmetic and its order follow forward.cu:672-755, so the
// images match the reference to rounding.
constexpr int RENDER_THREADS = 256;
constexpr int RENDER_BATCH = 256;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

template <bool COORD, bool DEPTH>
__global__ void __launch_bounds__(RENDER_THREADS)
render_fwd(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
           const float4* __restrict__ rec, const float4* __restrict__ rec_coord, float fx, float fy,
           const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ out_coord,
           float* __restrict__ out_mcoord, float* __restrict__ out_depth, float* __restrict__ out_mdepth,
           float* __restrict__ out_alpha, float* __restrict__ out_tongue, float* __restrict__ out_normal,
           uint32_t* __restrict__ n_contrib, float* __restrict__ accum_depth,
           float* __restrict__ normal_length, float* __restrict__ accum_coord,
           unsigned long long* __restrict__ pair_counters) {
  constexpr bool NORMAL = COORD || DEPTH;
  constexpr bool GEO = NORMAL;
  constexpr int NQ = GEO ? 4 : 3;
  constexpr int NQC = COORD ? REC_COORD_QUADS : 0;
  extern __shared__ __align__(16) float4 s_rec[];  // [2][NQ + NQC][RENDER_BATCH]
  auto SQ = [&](int buf, int q, int j) -> float4& { return s_rec[(buf * (NQ + NQC) + q) * RENDER_BATCH + j]; };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t px = blockIdx.x * TILE + (warp & 1) * 8 + (lane & 7);
  const uint32_t py = blockIdx.y * TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const uint32_t pix_id = W * py + px;
  const float2 pixf = make_float2((float)px, (float)py);
  const float2 pixnf = make_float2((pixf.x - W / 2.f) / fx, (pixf.y - H / 2.f) / fy);
  const float ln = sqrtf(pixnf.x * pixnf.x + pixnf.y * pixnf.y + 1);
  const float patch_x0 = (float)(blockIdx.x * TILE + (warp & 1) * 8), patch_y0 = (float)(blockIdx.y * TILE + (warp >> 1) * 4);

  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + RENDER_BATCH - 1) / RENDER_BATCH;

  bool done = !inside;
  float T = 1.0f;
  uint32_t contributor = 0, last_contributor = 0, max_contributor = 0xffffffffu;
  float C[3] = {0, 0, 0};
  float tongue = 0, weight = 0;
  float Coord[3] = {0, 0, 0}, mCoord[3] = {0, 0, 0};
  float Depth = 0, mDepth = 0;
  float Normal[3] = {0, 0, 0};

  auto issue = [&](int buf, int id) {
    const float4* src = rec + (size_t)id * REC_QUADS;
#pragma unroll
    for (int q = 0; q < NQ; ++q) cp_async16(&SQ(buf, q, tid), src + q);
    if constexpr (COORD) {
      const float4* src2 = rec_coord + (size_t)id * REC_COORD_QUADS;
#pragma unroll
      for (int q = 0; q < NQC; ++q) cp_async16(&SQ(buf, NQ + q, tid), src2 + q);
    }
  };
  auto load_id = [&](int round) -> int {
    const int k = round * RENDER_BATCH + tid;
    return (k < total) ? (int)point_list[range.x + k] : -1;
  };

  int id_next = load_id(0);
  if (id_next >= 0) issue(0, id_next);
  cp_async_commit();
  id_next = load_id(1);

  int todo = total;
  for (int i = 0; i < rounds; ++i, todo -= RENDER_BATCH) {
    cp_async_wait_all();
    const int num_done = __syncthreads_count(done);
    if (num_done == RENDER_THREADS) break;
    const int buf = i & 1;
    if (id_next >= 0) issue(buf ^ 1, id_next);
    cp_async_commit();
    id_next = load_id(i + 2);

    const int n = min(RENDER_BATCH, todo);
    bool warp_done = false;
#pragma unroll 1
    for (int m = 0; m * 32 < n && !warp_done; ++m) {
    // Conservative cull, 32 Gaussians at a time (one per lane): a Gaussian whose alpha >= 1/255 ellipse
    // misses this warp's 8x4 pixel patch fails the exact per-pixel test below for every lane, so the warp
    // skips it. The contributor index is the list position, so skipped entries still count.
    uint32_t cand;
    {
      const int jl = m * 32 + lane;
      bool touch = false;
      if (jl < n) {
        const float4 c0 = SQ(buf, 0, jl);
        const float4 c1 = SQ(buf, 1, jl);
        touch = patch_may_touch(c0.x, c0.y, c1.x, c1.y, c1.z, c1.w, patch_x0, patch_y0);
      }
      cand = __ballot_sync(0xffffffffu, touch);
    }
#pragma unroll 1
    while (cand) {
      if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
      const int j = m * 32 + __ffs(cand) - 1;
      cand &= cand - 1;
      const float4 q0 = SQ(buf, 0, j);
      const float4 q1 = SQ(buf, 1, j);
      if (!done) contributor = (uint32_t)(i * RENDER_BATCH + j + 1);   // a finished pixel keeps the index it stopped at
      const float2 d = make_float2(q0.x - pixf.x, q0.y - pixf.y);
      const float power = -0.5f * (q1.x * d.x * d.x + q1.z * d.y * d.y) - q1.y * d.x * d.y;
      bool use = !done && !(power > 0.0f);
      float alpha = 0.f;
      if (use) {
        alpha = min(0.99f, q1.w * expf(power));
        use = !(alpha < 1.0f / 255.0f);
      }
      float test_T = T;
      if (use) {
        test_T = T * (1 - alpha);
        if (test_T < 0.0001f) {
          done = true;
          use = false;
        }
      }
      if (!use) continue;

      const float aT = alpha * T;
      const float4 q2 = SQ(buf, 2, j);
      C[0] += q2.x * aT;
      C[1] += q2.y * aT;
      C[2] += q2.z * aT;
      tongue += q0.w * aT;
      const bool before_median = T > 0.5;
      if constexpr (COORD) {
        const float4 c0 = SQ(buf, NQ + 0, j), c1 = SQ(buf, NQ + 1, j), c2 = SQ(buf, NQ + 2, j);
        const float coord[3] = {c1.z + c0.x * d.x + c0.y * d.y, c1.w + c0.z * d.x + c0.w * d.y,
                                c2.x + c1.x * d.x + c1.y * d.y};
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) Coord[ch] += coord[ch] * aT;
        if (before_median) {
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) mCoord[ch] = coord[ch];
        }
      }
      if constexpr (GEO) {
        const float4 q3 = SQ(buf, 3, j);
        if constexpr (DEPTH) {
          const float t = q0.z + (q2.w * d.x + q3.x * d.y);
          Depth += t * aT;
          if (before_median) mDepth = t;
        }
        if constexpr (NORMAL) {
          Normal[0] += q3.y * aT;
          Normal[1] += q3.z * aT;
          Normal[2] += q3.w * aT;
        }
        if (before_median) max_contributor = contributor;
      }
      weight += aT;
      T = test_T;
      last_contributor = contributor;
    }
    }
  }
  cp_async_wait_all();

  if (pair_counters) {   // profiling pass only (bench.py's FP32-ALU roofline): list entries this pixel walked
    unsigned long long f = inside ? (unsigned long long)(done ? contributor : (uint32_t)total) : 0ull;
    unsigned long long b = inside ? (unsigned long long)last_contributor : 0ull;
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
      f += __shfl_xor_sync(0xffffffffu, f, off);
      b += __shfl_xor_sync(0xffffffffu, b, off);
    }
    if (lane == 0) { atomicAdd(pair_counters, f); atomicAdd(pair_counters + 1, b); }
  }
  if (inside) {
    const size_t HW = (size_t)H * W;
    n_contrib[pix_id] = last_contributor;
    n_contrib[pix_id + HW] = max_contributor;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix_id] = C[ch] + T * bg[ch];
    out_tongue[pix_id] = tongue;
    out_alpha[pix_id] = weight;
    if constexpr (COORD) {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) {
        out_coord[ch * HW + pix_id] = last_contributor ? Coord[ch] / weight : 0.f;
        accum_coord[ch * HW + pix_id] = Coord[ch];
        out_mcoord[ch * HW + pix_id] = mCoord[ch];
      }
    }
    if constexpr (DEPTH) {
      const float depth_ln = Depth / ln;
      accum_depth[pix_id] = depth_ln;
      out_depth[pix_id] = last_contributor ? depth_ln / weight : 0.f;
      out_mdepth[pix_id] = mDepth / ln;
    }
    if constexpr (NORMAL) {
      if (last_contributor) {
        float len = sqrtf(Normal[0] * Normal[0] + Normal[1] * Normal[1] + Normal[2] * Normal[2]);
        normal_length[pix_id] = len;
        len = max(len, 1.0E-12F);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) out_normal[ch * HW + pix_id] = Normal[ch] / len;
      } else {
        normal_length[pix_id] = 1;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) out_normal[ch * HW + pix_id] = 0;
      }
    }
  }
}

template <bool COORD, bool DEPTH>
static cudaError_t launch_render_fwd(dim3 grid, cudaStream_t stream, const ImageState& img, const uint32_t* point_list,
                              int W, int H, const GeomState& geom, float fx, float fy, const float* bg,
                              float* out_color, float* out_coord, float* out_mcoord, float* out_depth,
                              float* out_mdepth, float* out_alpha, float* out_tongue, float* out_normal) {
  constexpr int NQ = (COORD || DEPTH) ? 4 : 3;
  constexpr int NQC = COORD ? REC_COORD_QUADS : 0;
  const size_t smem = 2 * (NQ + NQC) * RENDER_BATCH * sizeof(float4);
  // per-device attribute: set on every launch
  const cudaError_t e = cudaFuncSetAttribute(render_fwd<COORD, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  render_fwd<COORD, DEPTH><<<grid, RENDER_THREADS, smem, stream>>>(
      img.ranges, point_list, W, H, geom.rec, geom.rec_coord, fx, fy, bg, out_color, out_coord, out_mcoord,
      out_depth, out_mdepth, out_alpha, out_tongue, out_normal, img.n_contrib, img.accum_depth,
      img.normal_length, img.accum_coord, prof_pair_counters());
  return cudaGetLastError();
}

// Next-highest bit of the MSB of n, i.e. bits needed for tile ids (rasterizer_impl.cu:35-50).
static uint32_t tile_id_bits(uint32_t n) {
  uint32_t msb = sizeof(n) * 4;
  uint32_t step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

int raster_front(const RasterFwdArgs& a, float4* inte_rec, FrontEnd& fe, cudaStream_t stream) {
  const int P = a.P, W = a.width, H = a.height;
  const float focal_y = H / (2.0f * a.tan_fovy);
  const float focal_x = W / (2.0f * a.tan_fovx);
  const bool coord = a.require_coord != 0;
  const uint32_t gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t npix = (size_t)W * H, ntiles = (size_t)gx * gy;
  fe.gx = gx; fe.gy = gy; fe.focal_x = focal_x; fe.focal_y = focal_y;

  char* gp = a.geom_alloc(a.geom_user, geom_bytes(P, coord));
  char* ip = a.image_alloc(a.image_user, image_bytes(npix, ntiles, coord));
  if (!gp || !ip) { set_error("state buffer allocation failed"); return ED3DGS_ERR_ALLOC; }
  GeomState& geom = fe.geom;
  ImageState& img = fe.img;
  geom = carve_geom(gp, P, coord);
  img = carve_image(ip, npix, ntiles, coord);

  {
  ProfScope ps("preprocess_fwd", stream);
  if (inte_rec)
    preprocess_fwd<true><<<(P + 255) / 256, 256, 0, stream>>>(
        P, a.D, a.M, a.means3D, a.scales, a.scale_modifier, a.rotations, a.opacities, nullptr, a.shs,
        a.cov3D_precomp, a.colors_precomp, a.viewmatrix, a.projmatrix, a.cam_pos, W, H, a.tan_fovx, a.tan_fovy,
        focal_x, focal_y, a.kernel_size, a.radii, geom, gx, gy, coord ? 1 : 0, inte_rec);
  else
    preprocess_fwd<false><<<(P + 255) / 256, 256, 0, stream>>>(
        P, a.D, a.M, a.means3D, a.scales, a.scale_modifier, a.rotations, a.opacities, a.tongue_class, a.shs,
        a.cov3D_precomp, a.colors_precomp, a.viewmatrix, a.projmatrix, a.cam_pos, W, H, a.tan_fovx, a.tan_fovy,
        focal_x, focal_y, a.kernel_size, a.radii, geom, gx, gy, coord ? 1 : 0, nullptr);
  }
  ED_DEBUG_SYNC("preprocess");

  const uint32_t nsb = (P + SCAN_TILE - 1) / SCAN_TILE;
  {
  ProfScope ps("scan_tiles", stream);
  scan_tiles_local<<<nsb, SCAN_THREADS, 0, stream>>>(geom.tiles_touched, geom.point_offsets, P, geom.block_sums);
  scan_exclusive_1cta<<<1, 1024, 0, stream>>>(geom.block_sums, nsb, geom.block_sums + nsb);
  scan_tiles_add<<<nsb, SCAN_THREADS, 0, stream>>>(geom.point_offsets, P, geom.block_sums);
  }
  ED_DEBUG_SYNC("scan");

  // The host needs R to size the binning buffers (as rasterizer_impl.cu:359).
  uint32_t R = 0;
  ED_CHECK(cudaMemcpyAsync(&R, geom.block_sums + nsb, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  ED_CHECK(cudaStreamSynchronize(stream));
  fe.R = R;

  char* bp = a.binning_alloc(a.binning_user, binning_bytes(R));
  if (!bp) { set_error("binning buffer allocation failed"); return ED3DGS_ERR_ALLOC; }
  BinningState& bin = fe.bin;
  bin = carve_binning(bp, R);

  int sel = 0;
  if (R > 0) {
    {
    ProfScope ps("emit_keys", stream);
    emit_keys<<<(P + 255) / 256, 256, 0, stream>>>(P, geom.rec, geom.depths, geom.point_offsets, a.radii,
                                                  bin.keys[2], bin.vals[2], gx, gy);
    }
    ED_DEBUG_SYNC("emit_keys");
    const int bit = (int)tile_id_bits(gx * gy);
    {
    ProfScope ps("radix_sort_pairs", stream);
    sel = radix_sort_pairs(bin, R, 32 + bit, stream);
    }
    if (sel < 0) { set_error("radix_sort_pairs: %s", cudaGetErrorString(cudaGetLastError())); return ED3DGS_ERR_CUDA; }
    ED_DEBUG_SYNC("radix_sort");
  }
  fe.sel = sel;
  ED_CHECK(cudaMemsetAsync(img.ranges, 0, ntiles * sizeof(uint2), stream));
  if (R > 0) { ProfScope ps("tile_ranges", stream); tile_ranges<<<(R + 255) / 256, 256, 0, stream>>>(R, bin.keys[sel], img.ranges); }
  ED_DEBUG_SYNC("tile_ranges");
  return ED3DGS_OK;
}

int raster_forward(const RasterFwdArgs& a, cudaStream_t stream) {
  const int P = a.P, W = a.width, H = a.height;
  *a.num_rendered = 0;
  if (P == 0) return ED3DGS_OK;
  if (a.colors_precomp == nullptr && a.shs == nullptr) {
    set_error("For non-RGB, provide precomputed Gaussian colors!");
    return ED3DGS_ERR_ARG;
  }
  const bool coord = a.require_coord != 0, depth = a.require_depth != 0;
  FrontEnd fe;
  const int rc = raster_front(a, nullptr, fe, stream);
  if (rc != ED3DGS_OK) return rc;
  *a.num_rendered = (int)fe.R;
  const GeomState& geom = fe.geom;
  const ImageState& img = fe.img;
  const BinningState& bin = fe.bin;
  const uint32_t gx = fe.gx, gy = fe.gy;
  const float focal_x = fe.focal_x, focal_y = fe.focal_y;
  const int sel = fe.sel;

  const dim3 grid(gx, gy, 1);
  const uint32_t* plist = bin.vals[sel];
#define ED_RENDER(C_, D_)                                                                               \
  launch_render_fwd<C_, D_>(grid, stream, img, plist, W, H, geom, focal_x, focal_y, a.background,        \
                            a.out_color, a.out_coord, a.out_mcoord, a.out_depth, a.out_mdepth, a.out_alpha, \
                            a.out_tongue, a.out_normal)
  cudaError_t render_err;
  {
  ProfScope ps("render_fwd", stream);
  if (coord && depth) render_err = ED_RENDER(true, true);
  else if (coord) render_err = ED_RENDER(true, false);
  else if (depth) render_err = ED_RENDER(false, true);
  else render_err = ED_RENDER(false, false);
  }
#undef ED_RENDER
  ED_CHECK(render_err);
  ED_DEBUG_SYNC("render");
  return ED3DGS_OK;
}

}  // namespace ed3dgs
