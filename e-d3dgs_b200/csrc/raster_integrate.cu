// Point integration for mesh extraction (GOF-style): for every query point, the opacity the Gaussians of
// its pixel accumulate in front of it, evaluated with the 3-D ray-space Gaussians; plus the image the
// reference renders alongside. Behavioural contract = CudaRasterizer::Rasterizer::integrate
// (rasterizer_impl.cu:580-851) with kernels forward.cu:307-424 (preprocessCUDA<C, true>), :1027-1071
// (preprocessPointsCUDA), rasterizer_impl.cu:114-146 (createWithKeys) and forward.cu:1109-1543
// (integrateCUDA).
//
// Redesign. The reference gives every pixel thread a 4 KB list of the list positions that contributed
// (uint16 contributed_ids[2048]) and three 256-entry arrays of the points that fall into the pixel, all in
// local memory, walks the tile list once to fill the first and once more per 256 points to integrate.
// Here the tile list is walked ONCE in batches of 256 Gaussians staged in shared memory:
//   pass A (thread = pixel)  the five-sample visibility test of forward.cu:1218-1281; the outcome is one
//                            bit per (pixel, Gaussian of the batch) in shared memory (8 KB);
//   pass B (thread = point)  the tile's points, binned per tile by a counting sort, are spread over the
//                            256 threads whatever pixel they fall into; a warp walks the OR of its 32
//                            points' bit words, so the Gaussian record is a shared-memory broadcast, and
//                            every point integrates the Gaussians whose bit is set for its pixel, in list
//                            order (forward.cu:1452-1511). The running (alpha, T) pair of a point lives in
//                            an L2-resident scratch array between batches.
// A point's result depends only on its own pixel's bits and on list order, so it is independent of how the
// points are ordered inside the tile (the reference sorts them by depth; nothing reads that order).
// Deviations, all in cases where the reference's behaviour is accidental: tile lists longer than 65 535
// entries (its uint16 ids wrap), pixels holding more than 256 points (its second round re-counts the last
// point into the "number of projected points" plane), ill-conditioned covariances (see splat_planes).
#include <cstdio>
#include "raster_cull.cuh"
#include "gs_math.cuh"
#include "raster_state.cuh"
#include "raster_api.h"
#include "raster_front.h"
#include "profiler.h"

namespace ed3dgs {

__global__ void scan_exclusive_1cta(uint32_t* __restrict__ data, uint32_t count, uint32_t* __restrict__ total_out);

constexpr int MAX_CONTRIB = 512 * 4;  // auxiliary.h:31 MAX_NUM_CONTRIBUTORS * 4 (forward.cu:1189,1297)

// Workspace carved by the host wrapper (bytes from integrate_workspace_bytes).
struct InteWork {
  float4* inte_rec;      // [P][2]
  float4* pts;           // [PN] {x, y, depth, tile | rank packed later}  (projection order)
  uint32_t* pts_rank;    // [PN] rank of the point inside its tile, 0xffffffff = not in the image
  float4* sorted;        // [PN] {x, y, depth, bits(point index)} grouped by tile
  float2* state;         // [PN] {alpha accumulated, T} in `sorted` order
  uint32_t* tile_start;  // [tiles + 1] counts, then exclusive scan
};

static InteWork carve_inte(char*& p, size_t P, size_t PN, size_t ntiles) {
  InteWork w{};
  carve(p, w.inte_rec, 2 * P);
  carve(p, w.pts, PN);
  carve(p, w.pts_rank, PN);
  carve(p, w.sorted, PN);
  carve(p, w.state, PN);
  carve(p, w.tile_start, ntiles + 2);
  return w;
}

size_t integrate_workspace_bytes(int P, int PN, int width, int height) {
  const size_t ntiles = (size_t)((width + TILE - 1) / TILE) * ((height + TILE - 1) / TILE);
  char* p = nullptr;
  carve_inte(p, (size_t)(P > 0 ? P : 0), (size_t)(PN > 0 ? PN : 0), ntiles);
  return reinterpret_cast<size_t>(p) + 256;
}

// ---------------------------------------------------------------------------------------------
// Points: projection (forward.cu:1027-1071), tile of the point (rasterizer_impl.cu:114-146) and the
// per-tile count. Points outside the image keep the values the reference pre-fills its outputs with
// (rasterize_points.cu:316-319), so the caller does not have to fill anything.
__global__ void __launch_bounds__(256)
points_project(int PN, const float* __restrict__ pts3D, const float* __restrict__ view, int W, int H, float fx,
               float fy, uint32_t gx, uint32_t gy, float4* __restrict__ pts, uint32_t* __restrict__ rank,
               uint32_t* __restrict__ tile_count, float* __restrict__ out_alpha, float* __restrict__ out_color,
               float* __restrict__ out_coord, float* __restrict__ out_sdf) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  bool valid = false;
  uint32_t tile = 0xffffffffu;
  float2 pix = make_float2(0.f, 0.f);
  float depth = 0.f;
  if (idx < PN) {
    const float3 p = make_float3(pts3D[3 * idx], pts3D[3 * idx + 1], pts3D[3 * idx + 2]);
    const float3 pv = xform_4x3(p, view);
    if (pv.z > 0.2f) {
      pix = make_float2(fx * pv.x / (pv.z + 0.0000001f) + W / 2., fy * pv.y / (pv.z + 0.0000001f) + H / 2.);
      if (!(pix.x < 0 || pix.x >= W || pix.y < 0 || pix.y >= H)) {
        valid = true;
        depth = sqrtf(__fmaf_rn(pv.z, pv.z, __fmaf_rn(pv.x, pv.x, __fmul_rn(pv.y, pv.y))));  // order: see preprocess_fwd
        const int tx = min((int)gx - 1, max(0, (int)(pix.x / TILE)));
        const int ty = min((int)gy - 1, max(0, (int)(pix.y / TILE)));
        tile = (uint32_t)ty * gx + (uint32_t)tx;
      }
    }
    if (!valid) {
      out_alpha[idx] = 1.0f;
      out_color[3 * idx] = out_color[3 * idx + 1] = out_color[3 * idx + 2] = 0.f;
      out_coord[2 * idx] = out_coord[2 * idx + 1] = 0.f;
      out_sdf[idx] = -1000.0f;
      rank[idx] = 0xffffffffu;
    }
  }
  // one atomic per (warp, tile): lanes that share a tile elect a leader
  const uint32_t peers = __match_any_sync(0xffffffffu, tile);
  if (valid) {
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&tile_count[tile], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    rank[idx] = base + __popc(peers & ((1u << lane) - 1u));
    pts[idx] = make_float4(pix.x, pix.y, depth, __uint_as_float(tile));
  }
}

__global__ void __launch_bounds__(256)
points_scatter(int PN, const float4* __restrict__ pts, const uint32_t* __restrict__ rank,
               const uint32_t* __restrict__ tile_start, float4* __restrict__ sorted, float2* __restrict__ state) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= PN) return;
  const uint32_t r = rank[idx];
  if (r == 0xffffffffu) return;
  const float4 p = pts[idx];
  const uint32_t pos = tile_start[__float_as_uint(p.w)] + r;
  sorted[pos] = make_float4(p.x, p.y, p.z, __uint_as_float((uint32_t)idx));
  state[pos] = make_float2(0.f, 1.f);
}

// ---------------------------------------------------------------------------------------------
constexpr int INT_THREADS = 256;
constexpr int INT_BATCH = 256;
constexpr int INT_QUADS = 6;  // rec q0..q3 + the two integration quads
constexpr int INT_WORDS = INT_BATCH / 32;

__device__ __forceinline__ void cp_async16i(void* smem_dst, const void* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}

struct PixelFinal {  // what a point needs from its pixel once the list has been walked
  float r, g, b, mid_depth_center, mid_plane_x, mid_plane_y, mid_x, mid_y;
};

__global__ void __launch_bounds__(INT_THREADS)
integrate_tiles(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                const uint32_t* __restrict__ tile_start, int W, int H, const float4* __restrict__ rec,
                const float4* __restrict__ inte_rec, const float4* __restrict__ sorted, float2* __restrict__ state,
                const float* __restrict__ bg, float* __restrict__ out_color, float* __restrict__ out_alpha_int,
                float* __restrict__ out_color_int, float* __restrict__ out_coord2d, float* __restrict__ out_sdf) {
  __shared__ __align__(16) float4 s_rec[INT_QUADS][INT_BATCH];
  __shared__ uint32_t s_flags[INT_WORDS][INT_THREADS];
  __shared__ PixelFinal s_fin[INT_THREADS];
  __shared__ uint32_t s_cnt[INT_THREADS];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_x0 = blockIdx.x * TILE, tile_y0 = blockIdx.y * TILE;
  const uint32_t px = tile_x0 + (warp & 1) * 8 + (lane & 7);
  const uint32_t py = tile_y0 + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const uint32_t pix_id = W * py + px;
  const float2 pixf = make_float2((float)px + 0.5f, (float)py + 0.5f);
  // sample positions of this warp's 8x4 patch span [x0, x0 + 8] x [y0, y0 + 4] (centres +- 0.5)
  const float box_x0 = (float)(tile_x0 + (warp & 1) * 8), box_y0 = (float)(tile_y0 + (warp >> 1) * 4);

  const uint32_t tile_id = blockIdx.y * gridDim.x + blockIdx.x;
  const uint2 range = ranges[tile_id];
  const int total = (int)(range.y - range.x);
  const int rounds = (total + INT_BATCH - 1) / INT_BATCH;
  const uint32_t pstart = tile_start[tile_id];
  const int npts = (int)(tile_start[tile_id + 1] - pstart);
  const int npts_pad = (npts + 31) & ~31;

  // pass-A state (forward.cu:1180-1194)
  bool done = !inside;
  float T = 1.0f;
  float cT[5] = {1.0f, 1.0f, 1.0f, 1.0f, 1.0f};
  const float offx[5] = {0.0f, -0.5f, 0.5f, -0.5f, 0.5f};
  const float offy[5] = {0.0f, -0.5f, -0.5f, 0.5f, 0.5f};
  float C[3] = {0, 0, 0};
  float Dacc = 0, mDepth = 0, maxDepth = 0, Asum = 0;
  float mid_depth_center = 0;
  float2 mid_plane = make_float2(0, 0), mid_mean2d = make_float2(0, 0);
  int n_used = 0;
  // a sample whose T * (1 - 1/255) is already below 1e-4 fails the transmittance test (forward.cu:1252-1257)
  // for every alpha that passes the 1/255 test: it can never change again
  const float keep255 = 1.0f - 1.0f / 255.0f;

  int todo = total;
  for (int i = 0; i < rounds; ++i, todo -= INT_BATCH) {
    const int n = min(INT_BATCH, todo);
    __syncthreads();  // previous batch fully consumed (pass B reads s_rec and s_flags)
    if (tid < n) {
      const uint32_t id = point_list[range.x + i * INT_BATCH + tid];
      const float4* src = rec + (size_t)id * REC_QUADS;
#pragma unroll
      for (int q = 0; q < 4; ++q) cp_async16i(&s_rec[q][tid], src + q);
      const float4* src2 = inte_rec + (size_t)id * 2;
      cp_async16i(&s_rec[4][tid], src2);
      cp_async16i(&s_rec[5][tid], src2 + 1);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    // ---- pass A: which Gaussians of the batch touch this pixel ----
#pragma unroll 1
    for (int m = 0; m < INT_WORDS; ++m) {
      uint32_t word = 0;
      if (m * 32 < n && !__all_sync(0xffffffffu, done)) {
        uint32_t cand;
        {
          const int jl = m * 32 + lane;
          bool touch = false;
          if (jl < n) {
            const float4 c0 = s_rec[0][jl];
            const float4 c1 = s_rec[1][jl];
            touch = box_may_touch(c0.x, c0.y, c1.x, c1.y, c1.z, c1.w, box_x0, box_x0 + 8.f, box_y0, box_y0 + 4.f);
          }
          cand = __ballot_sync(0xffffffffu, touch);
        }
#pragma unroll 1
        while (cand) {
          const int jb = __ffs(cand) - 1;
          cand &= cand - 1;
          if (done) continue;
          const int j = m * 32 + jb;
          const float4 q0 = s_rec[0][j];
          const float4 con_o = s_rec[1][j];
          const float4 q2 = s_rec[2][j];
          const float depth_center = q0.z;
          const float2 plane = make_float2(q2.w, s_rec[3][j].x);
          bool used = false;
#pragma unroll
          for (int k = 0; k < 5; ++k) {
            const float2 d = make_float2(q0.x - pixf.x - offx[k], q0.y - pixf.y - offy[k]);
            const float depth = depth_center + (plane.x * d.x + plane.y * d.y);
            const float power = -0.5f * (con_o.x * d.x * d.x + con_o.z * d.y * d.y) - con_o.y * d.x * d.y;
            if (power > 0.0f) continue;
            const float alpha = min(0.99f, con_o.w * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = cT[k] * (1 - alpha);
            if (test_T < 0.0001f) continue;
            if (k == 0) {
              C[0] += q2.x * alpha * T;
              C[1] += q2.y * alpha * T;
              C[2] += q2.z * alpha * T;
            }
            if (depth > maxDepth) maxDepth = depth;
            if (k == 0) {
              Asum += alpha * T;
              Dacc += depth * alpha * T;
              if (T > 0.5) {
                mDepth = depth;
                mid_depth_center = depth_center;
                mid_plane = plane;
                mid_mean2d = make_float2(q0.x, q0.y);
              }
              T = test_T;
            }
            cT[k] = test_T;
            used = true;
          }
          if (used) {
            word |= 1u << jb;
            n_used += 1;
            bool alive = false;
#pragma unroll
            for (int k = 0; k < 5; ++k) alive = alive || !(cT[k] * keep255 < 0.0001f);
            if (!alive || n_used >= MAX_CONTRIB) done = true;
          }
        }
      }
      s_flags[m][tid] = word;
    }
    __syncthreads();

    // ---- pass B: every point of the tile integrates the flagged Gaussians of its pixel ----
#pragma unroll 1
    for (int p = tid; p < npts_pad; p += INT_THREADS) {
      const bool valid = p < npts;
      float4 pt = make_float4(0, 0, 0, 0);
      float2 st = make_float2(0, 1);
      int ptid = 0;
      if (valid) {
        pt = sorted[pstart + p];
        st = state[pstart + p];
        const int lx = (int)pt.x - (int)tile_x0, ly = (int)pt.y - (int)tile_y0;
        ptid = (((ly >> 2) * 2 + (lx >> 3)) << 5) + ((ly & 3) << 3) + (lx & 7);
      }
#pragma unroll 1
      for (int m = 0; m * 32 < n; ++m) {
        const uint32_t bits = valid ? s_flags[m][ptid] : 0u;
        uint32_t any = __reduce_or_sync(0xffffffffu, bits);
#pragma unroll 1
        while (any) {
          const int jb = __ffs(any) - 1;
          any &= any - 1;
          if (!((bits >> jb) & 1u)) continue;
          const int j = m * 32 + jb;
          const float4 q0 = s_rec[0][j];
          const float op = s_rec[1][j].w;
          const float2 plane = make_float2(s_rec[2][j].w, s_rec[3][j].x);
          const float4 i0 = s_rec[4][j];
          const float4 i1 = s_rec[5][j];
          const float depth_center = q0.z;
          // The quadratic form below cancels heavily (entries ~ 1/scale^2 against offsets of a few pixels), so
          // its value depends on where the roundings fall. The operation order is spelled out with intrinsics
          // and is the one nvcc emits for the reference's glm expression (forward.cu:1482-1484):
          // a x + b y + c z -> fma(c, z, fma(a, x, b y)).
          const float2 d = make_float2(q0.x - pt.x, q0.y - pt.y);
          const float depth = depth_center + __fmaf_rn(plane.x, d.x, __fmul_rn(plane.y, d.y));
          float alpha;
          float dz;
          bool zero = false;
          if (i1.z != 0.f) {
            dz = depth_center - min(pt.z, depth);
          } else {
            zero = pt.z < depth;
            dz = depth_center;
          }
          if (zero) {
            alpha = 0;
          } else {
            const float mx = __fmaf_rn(i0.z, dz, __fmaf_rn(i0.x, d.x, __fmul_rn(i0.y, d.y)));
            const float my = __fmaf_rn(i1.x, dz, __fmaf_rn(i0.y, d.x, __fmul_rn(i0.w, d.y)));
            const float mz = __fmaf_rn(i1.y, dz, __fmaf_rn(i0.z, d.x, __fmul_rn(i1.x, d.y)));
            const float quad = __fmaf_rn(dz, mz, __fmaf_rn(d.x, mx, __fmul_rn(d.y, my)));
            const float power = __fmul_rn(quad, -0.5f);
            alpha = min(0.99f, op * expf(power));
          }
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = __fmul_rn(st.y, 1 - alpha);
          st.x = __fmaf_rn(alpha, st.y, st.x);
          st.y = test_T;
        }
      }
      if (valid) state[pstart + p] = st;
    }
  }

  // ---- epilogue: image planes (forward.cu:1305-1320,1539-1542) and per-point outputs (:1519-1536) ----
  __syncthreads();
  PixelFinal f;
  f.r = C[0] + T * bg[0];
  f.g = C[1] + T * bg[1];
  f.b = C[2] + T * bg[2];
  f.mid_depth_center = mid_depth_center;
  f.mid_plane_x = mid_plane.x;
  f.mid_plane_y = mid_plane.y;
  f.mid_x = mid_mean2d.x;
  f.mid_y = mid_mean2d.y;
  s_fin[tid] = f;
  s_cnt[tid] = 0;
  __syncthreads();
  for (int p = tid; p < npts; p += INT_THREADS) {
    const float4 pt = sorted[pstart + p];
    const float2 st = state[pstart + p];
    const uint32_t idx = __float_as_uint(pt.w);
    const int lx = (int)pt.x - (int)tile_x0, ly = (int)pt.y - (int)tile_y0;
    const int ptid = (((ly >> 2) * 2 + (lx >> 3)) << 5) + ((ly & 3) << 3) + (lx & 7);
    const PixelFinal g = s_fin[ptid];
    atomicAdd(&s_cnt[ptid], 1u);
    out_alpha_int[idx] = st.x;
    out_color_int[3 * idx] = g.r;
    out_color_int[3 * idx + 1] = g.g;
    out_color_int[3 * idx + 2] = g.b;
    out_coord2d[2 * idx] = pt.x;
    out_coord2d[2 * idx + 1] = pt.y;
    float sdf = -1000.0f;
    if (pt.z > 0) {
      const float2 d = make_float2(g.mid_x - pt.x, g.mid_y - pt.y);
      const float depth = g.mid_depth_center + (g.mid_plane_x * d.x + g.mid_plane_y * d.y);
      sdf = depth - pt.z;
    }
    out_sdf[idx] = sdf;
  }
  __syncthreads();
  if (inside) {
    const size_t HW = (size_t)H * W;
    out_color[0 * HW + pix_id] = f.r;
    out_color[1 * HW + pix_id] = f.g;
    out_color[2 * HW + pix_id] = f.b;
    out_color[3 * HW + pix_id] = Dacc;
    out_color[4 * HW + pix_id] = mDepth;
    out_color[5 * HW + pix_id] = 0.f;
    out_color[6 * HW + pix_id] = maxDepth;
    out_color[7 * HW + pix_id] = Asum;
    out_color[8 * HW + pix_id] = (float)s_cnt[tid];
  }
}

int raster_integrate(const IntegrateArgs& ia, cudaStream_t stream) {
  const RasterFwdArgs& a = ia.fwd;
  const int P = a.P, PN = ia.PN, W = a.width, H = a.height;
  *a.num_rendered = 0;
  if (P == 0 || PN == 0) return ED3DGS_OK;  // rasterize_points.cu:344: nothing is launched
  if (a.colors_precomp == nullptr && a.shs == nullptr) {
    set_error("For non-RGB, provide precomputed Gaussian colors!");
    return ED3DGS_ERR_ARG;
  }
  const uint32_t gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t ntiles = (size_t)gx * gy;
  if (ia.workspace == nullptr || ia.workspace_bytes < integrate_workspace_bytes(P, PN, W, H)) {
    set_error("ed3dgs_integrate: workspace too small");
    return ED3DGS_ERR_ARG;
  }
  char* wp = static_cast<char*>(ia.workspace);
  InteWork w = carve_inte(wp, P, PN, ntiles);

  FrontEnd fe;
  const int rc = raster_front(a, w.inte_rec, fe, stream);
  if (rc != ED3DGS_OK) return rc;
  *a.num_rendered = (int)fe.R;

  {
    ProfScope ps("points_bin", stream);
    ED_CHECK(cudaMemsetAsync(w.tile_start, 0, (ntiles + 2) * sizeof(uint32_t), stream));
    points_project<<<(PN + 255) / 256, 256, 0, stream>>>(PN, ia.points3D, a.viewmatrix, W, H, fe.focal_x, fe.focal_y,
                                                        gx, gy, w.pts, w.pts_rank, w.tile_start,
                                                        ia.out_alpha_integrated, ia.out_color_integrated,
                                                        ia.out_coordinate2d, ia.out_sdf);
    scan_exclusive_1cta<<<1, 1024, 0, stream>>>(w.tile_start, (uint32_t)ntiles + 1, nullptr);
    points_scatter<<<(PN + 255) / 256, 256, 0, stream>>>(PN, w.pts, w.pts_rank, w.tile_start, w.sorted, w.state);
  }
  ED_DEBUG_SYNC("points_bin");
  {
    ProfScope ps("integrate_tiles", stream);
    integrate_tiles<<<dim3(gx, gy, 1), INT_THREADS, 0, stream>>>(
        fe.img.ranges, fe.bin.vals[fe.sel], w.tile_start, W, H, fe.geom.rec, w.inte_rec, w.sorted, w.state,
        a.background, a.out_color, ia.out_alpha_integrated, ia.out_color_integrated, ia.out_coordinate2d, ia.out_sdf);
  }
  ED_DEBUG_SYNC("integrate");
  ED_CHECK(cudaGetLastError());
  return ED3DGS_OK;
}

}  // namespace ed3dgs
