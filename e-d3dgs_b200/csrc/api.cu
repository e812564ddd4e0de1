// extern "C" surface of libed3dgs_b200.so (declared in include/ed3dgs_b200.h) plus the two
// trivial kernels that do not belong to a larger file (visibility mask, state export).
#include <cstdarg>
#include <cstdio>
#include "gs_math.cuh"
#include "raster_state.cuh"
#include "raster_api.h"
#include "profiler.h"
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace ed3dgs {

static thread_local char g_error[1024] = "";
static thread_local int g_packed_is_input = 0;

// ---- optional kernel timing (profiler.h) ----
struct ProfRec { const char* name; cudaEvent_t a, b; };
static bool g_prof_on = false;
static std::vector<ProfRec> g_prof;
static std::mutex g_prof_mu;
static unsigned long long* g_pair_counters = nullptr;
bool prof_enabled() { return g_prof_on; }
unsigned long long* prof_pair_counters() { return g_prof_on ? g_pair_counters : nullptr; }
void prof_push(const char* name, cudaEvent_t a, cudaEvent_t b) {
  std::lock_guard<std::mutex> lk(g_prof_mu);
  g_prof.push_back(ProfRec{name, a, b});
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof g_error, fmt, ap);
  va_end(ap);
}

// rasterizer_impl.cu:54-66 (checkFrustum) with auxiliary.h:155-180 (in_frustum, near cull only)
__global__ void __launch_bounds__(256)
visible_mask(int P, const float* __restrict__ means, const float* __restrict__ view, uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  const float3 pv = xform_4x3(p, view);
  present[idx] = (pv.z <= 0.2f) ? 0 : 1;
}

int mark_visible(int P, const float* means3D, const float* viewmatrix, const float*, uint8_t* present,
                 cudaStream_t stream) {
  if (P <= 0) return ED3DGS_OK;
  visible_mask<<<(P + 255) / 256, 256, 0, stream>>>(P, means3D, viewmatrix, present);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) { set_error("mark_visible launch failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

__global__ void __launch_bounds__(256)
export_geom(int P, int coord, const float4* __restrict__ rec, const float4* __restrict__ rec_coord,
            float* __restrict__ means2D, float* __restrict__ conic_opacity, float* __restrict__ rgb,
            float* __restrict__ normals, float* __restrict__ ray_planes, float* __restrict__ ts,
            float* __restrict__ camera_planes, float* __restrict__ view_points) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float4 q0 = rec[4 * (size_t)idx], q1 = rec[4 * (size_t)idx + 1], q2 = rec[4 * (size_t)idx + 2],
               q3 = rec[4 * (size_t)idx + 3];
  if (means2D) { means2D[2 * idx] = q0.x; means2D[2 * idx + 1] = q0.y; }
  if (ts) ts[idx] = q0.z;
  if (conic_opacity) { conic_opacity[4 * idx] = q1.x; conic_opacity[4 * idx + 1] = q1.y; conic_opacity[4 * idx + 2] = q1.z; conic_opacity[4 * idx + 3] = q1.w; }
  if (rgb) { rgb[3 * idx] = q2.x; rgb[3 * idx + 1] = q2.y; rgb[3 * idx + 2] = q2.z; }
  if (ray_planes) { ray_planes[2 * idx] = q2.w; ray_planes[2 * idx + 1] = q3.x; }
  if (normals) { normals[3 * idx] = q3.y; normals[3 * idx + 1] = q3.z; normals[3 * idx + 2] = q3.w; }
  if (coord) {
    const float4 c0 = rec_coord[3 * (size_t)idx], c1 = rec_coord[3 * (size_t)idx + 1], c2 = rec_coord[3 * (size_t)idx + 2];
    if (camera_planes) {
      camera_planes[6 * idx] = c0.x; camera_planes[6 * idx + 1] = c0.y; camera_planes[6 * idx + 2] = c0.z;
      camera_planes[6 * idx + 3] = c0.w; camera_planes[6 * idx + 4] = c1.x; camera_planes[6 * idx + 5] = c1.y;
    }
    if (view_points) { view_points[3 * idx] = c1.z; view_points[3 * idx + 1] = c1.w; view_points[3 * idx + 2] = c2.x; }
  }
}

static uint32_t id_bits(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

}  // namespace ed3dgs

using namespace ed3dgs;

extern "C" {

const char* ed3dgs_last_error(void) { return g_error; }
const char* ed3dgs_version(void) { return "0.1.0 (sm_100a)"; }

void ed3dgs_profile_enable(int on) {
  g_prof_on = on != 0;
  if (g_prof_on) {   // pair counters live on the device that is current when profiling is switched on
    if (!g_pair_counters && cudaMalloc(&g_pair_counters, 2 * sizeof(unsigned long long)) != cudaSuccess) g_pair_counters = nullptr;
    if (g_pair_counters) cudaMemset(g_pair_counters, 0, 2 * sizeof(unsigned long long));
  }
}

// (pixel, Gaussian) pairs walked since profiling was switched on: out[0] by the forward blend (list entries up to
// each pixel's early exit), out[1] by the backward blend (up to each pixel's last contributor). Resets the counters.
int ed3dgs_profile_pairs(unsigned long long* out) {
  if (!out || !g_pair_counters) { set_error("profile_pairs: profiling was never enabled"); return ED3DGS_ERR_ARG; }
  if (cudaDeviceSynchronize() != cudaSuccess ||
      cudaMemcpy(out, g_pair_counters, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("profile_pairs: %s", cudaGetErrorString(cudaGetLastError()));
    return ED3DGS_ERR_CUDA;
  }
  cudaMemset(g_pair_counters, 0, 2 * sizeof(unsigned long long));
  return ED3DGS_OK;
}

// Synchronises the device, writes "name total_ms launches\n" lines into buf and clears the log.
int ed3dgs_profile_report(char* buf, size_t nbytes) {
  cudaDeviceSynchronize();
  std::lock_guard<std::mutex> lk(g_prof_mu);
  std::map<std::string, std::pair<double, int>> agg;
  for (auto& r : g_prof) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      auto& e = agg[r.name];
      e.first += ms;
      e.second += 1;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof.clear();
  size_t off = 0;
  if (buf && nbytes) buf[0] = 0;
  for (auto& kv : agg) {
    const int n = snprintf(buf + off, off < nbytes ? nbytes - off : 0, "%s %.6f %d\n", kv.first.c_str(), kv.second.first,
                           kv.second.second);
    if (n < 0 || off + (size_t)n >= nbytes) break;
    off += (size_t)n;
  }
  return (int)agg.size();
}

int ed3dgs_rasterize_forward(
    ed3dgs_alloc_fn geom_alloc, void* geom_user, ed3dgs_alloc_fn binning_alloc, void* binning_user,
    ed3dgs_alloc_fn image_alloc, void* image_user, int P, int D, int M, const float* background,
    int width, int height, const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* tongue_class, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
    const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size,
    int prefiltered, float* out_color, float* out_coord, float* out_mcoord, float* out_depth,
    float* out_mdepth, float* out_alpha, float* out_tongue, float* out_normal, int* radii,
    int require_coord, int require_depth, int debug, void* stream, int* num_rendered) {
  g_error[0] = 0;
  if (P < 0 || width <= 0 || height <= 0 || !num_rendered || !geom_alloc || !binning_alloc || !image_alloc) {
    set_error("ed3dgs_rasterize_forward: bad argument");
    return ED3DGS_ERR_ARG;
  }
  if (P > 0 && cov3D_precomp == nullptr && (scales == nullptr || rotations == nullptr)) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return ED3DGS_ERR_ARG;
  }
  RasterFwdArgs a{geom_alloc, geom_user, binning_alloc, binning_user, image_alloc, image_user, P, D, M,
                  background, width, height, means3D, shs, colors_precomp, opacities, tongue_class, scales,
                  scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                  kernel_size, prefiltered, out_color, out_coord, out_mcoord, out_depth, out_mdepth, out_alpha,
                  out_tongue, out_normal, radii, require_coord, require_depth, debug, num_rendered};
  return raster_forward(a, static_cast<cudaStream_t>(stream));
}

size_t ed3dgs_integrate_workspace_bytes(int P, int PN, int width, int height) {
  return integrate_workspace_bytes(P, PN, width, height);
}

int ed3dgs_integrate(
    ed3dgs_alloc_fn geom_alloc, void* geom_user, ed3dgs_alloc_fn binning_alloc, void* binning_user,
    ed3dgs_alloc_fn image_alloc, void* image_user, int PN, int P, int D, int M, const float* background,
    int width, int height, const float* points3D, const float* means3D, const float* shs,
    const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
    const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size, int prefiltered,
    float* out_color, int* radii, float* out_alpha_integrated, float* out_color_integrated,
    float* out_coordinate2d, float* out_sdf, void* workspace, size_t workspace_bytes, int debug,
    void* stream, int* num_rendered) {
  g_error[0] = 0;
  if (P < 0 || PN < 0 || width <= 0 || height <= 0 || !num_rendered || !geom_alloc || !binning_alloc || !image_alloc) {
    set_error("ed3dgs_integrate: bad argument");
    return ED3DGS_ERR_ARG;
  }
  if (P > 0 && cov3D_precomp == nullptr && (scales == nullptr || rotations == nullptr)) {
    set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
    return ED3DGS_ERR_ARG;
  }
  IntegrateArgs ia{};
  ia.fwd = RasterFwdArgs{geom_alloc, geom_user, binning_alloc, binning_user, image_alloc, image_user, P, D, M,
                         background, width, height, means3D, shs, colors_precomp, opacities, nullptr, scales,
                         scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx,
                         tan_fovy, kernel_size, prefiltered, out_color, nullptr, nullptr, nullptr, nullptr, nullptr,
                         nullptr, nullptr, radii, 0, 0, debug, num_rendered};
  ia.PN = PN;
  ia.points3D = points3D;
  ia.out_alpha_integrated = out_alpha_integrated;
  ia.out_color_integrated = out_color_integrated;
  ia.out_coordinate2d = out_coordinate2d;
  ia.out_sdf = out_sdf;
  ia.workspace = workspace;
  ia.workspace_bytes = workspace_bytes;
  return raster_integrate(ia, static_cast<cudaStream_t>(stream));
}

// Test hook (not in the public header): the NEXT ed3dgs_rasterize_backward call of this thread
// treats grad_scratch as input ([P][16|32] packed rows, raster_state.cuh) and runs only the
// per-Gaussian stage. Lets the tests check that stage on the reference's own intermediates.
void ed3dgs_test_packed_gradients_are_input(int on) { g_packed_is_input = on; }

size_t ed3dgs_rasterize_backward_scratch_bytes(int P, int require_coord) {
  return (size_t)(P > 0 ? P : 0) * (require_coord ? 32 : 16) * sizeof(float) + 256;
}

int ed3dgs_rasterize_backward(
    int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
    const float* shs, const float* colors_precomp, const float* alphas, const float* scales,
    float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
    const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, float kernel_size,
    const int* radii, const float* normalmap, char* geom_buffer, char* binning_buffer,
    char* image_buffer, const float* dL_dpix, const float* dL_dpix_coord, const float* dL_dpix_mcoord,
    const float* dL_dpix_depth, const float* dL_dpix_mdepth, const float* dL_dalphas,
    const float* dL_dpixel_normals, float* dL_dmean2D, float* dL_dcolor, float* dL_dopacity,
    float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
    float* grad_scratch, int require_coord, int require_depth, int debug, void* stream) {
  g_error[0] = 0;
  if (P < 0 || R < 0 || width <= 0 || height <= 0 || (P > 0 && (!geom_buffer || !image_buffer || !grad_scratch))) {
    set_error("ed3dgs_rasterize_backward: bad argument");
    return ED3DGS_ERR_ARG;
  }
  RasterBwdArgs a{P, D, M, R, background, width, height, means3D, shs, colors_precomp, alphas, scales,
                  scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy,
                  kernel_size, radii, normalmap, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_dpix_coord,
                  dL_dpix_mcoord, dL_dpix_depth, dL_dpix_mdepth, dL_dalphas, dL_dpixel_normals, dL_dmean2D,
                  dL_dcolor, dL_dopacity, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, grad_scratch,
                  require_coord, require_depth, debug, g_packed_is_input};
  g_packed_is_input = 0;
  return raster_backward(a, static_cast<cudaStream_t>(stream));
}

int ed3dgs_rasterize_get_state(
    int P, int R, int width, int height, int require_coord, const char* geom_buffer,
    const char* binning_buffer, const char* image_buffer, float* depths, float* means2D,
    uint32_t* tiles_touched, uint32_t* point_offsets, float* conic_opacity, float* rgb,
    uint64_t* keys_unsorted, uint32_t* vals_unsorted, uint64_t* keys_sorted, uint32_t* point_list,
    uint32_t* ranges, uint32_t* n_contrib, float* normals, float* ray_planes, float* ts,
    float* camera_planes, float* view_points, void* stream_) {
  g_error[0] = 0;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (P <= 0) return ED3DGS_OK;
  const bool coord = require_coord != 0;
  const uint32_t gx = (width + TILE - 1) / TILE, gy = (height + TILE - 1) / TILE;
  const size_t npix = (size_t)width * height, ntiles = (size_t)gx * gy;
  char* gp = const_cast<char*>(geom_buffer);
  char* ip = const_cast<char*>(image_buffer);
  GeomState geom = carve_geom(gp, P, coord);
  ImageState img = carve_image(ip, npix, ntiles, coord);
  auto cp = [&](void* dst, const void* src, size_t n) {
    if (dst && n) cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToDevice, stream);
  };
  cp(depths, geom.depths, (size_t)P * 4);
  cp(tiles_touched, geom.tiles_touched, (size_t)P * 4);
  cp(point_offsets, geom.point_offsets, (size_t)P * 4);
  cp(ranges, img.ranges, ntiles * 8);
  cp(n_contrib, img.n_contrib, npix * 8);
  export_geom<<<(P + 255) / 256, 256, 0, stream>>>(P, coord ? 1 : 0, geom.rec, geom.rec_coord, means2D, conic_opacity,
                                                  rgb, normals, ray_planes, ts, camera_planes, view_points);
  if (R > 0 && binning_buffer) {
    char* bp = const_cast<char*>(binning_buffer);
    BinningState bin = carve_binning(bp, (size_t)R);
    const int sel = sorted_half((size_t)R, 32 + (int)id_bits(gx * gy));
    cp(keys_unsorted, bin.keys[2], (size_t)R * 8);
    cp(vals_unsorted, bin.vals[2], (size_t)R * 4);
    cp(keys_sorted, bin.keys[sel], (size_t)R * 8);
    cp(point_list, bin.vals[sel], (size_t)R * 4);
  }
  const cudaError_t err = cudaStreamSynchronize(stream);
  if (err != cudaSuccess) { set_error("get_state failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int ed3dgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                        uint8_t* present, void* stream) {
  g_error[0] = 0;
  return mark_visible(P, means3D, viewmatrix, projmatrix, present, static_cast<cudaStream_t>(stream));
}

size_t ed3dgs_dist_cuda2_workspace_bytes(int P) { return dist_cuda2_workspace(P > 0 ? P : 0); }

int ed3dgs_dist_cuda2(int P, const float* points, float* mean_dists, void* workspace, size_t workspace_bytes,
                      void* stream) {
  g_error[0] = 0;
  return dist_cuda2(P, points, mean_dists, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int ed3dgs_compute_3d_filter(int P, const float* xyz, int num_cameras, const float* cameras, float focal_length,
                             float* filter_3d, void* scratch, int* any_valid, void* stream) {
  g_error[0] = 0;
  return compute_3d_filter(P, xyz, num_cameras, cameras, focal_length, filter_3d, scratch, any_valid,
                           static_cast<cudaStream_t>(stream));
}

size_t ed3dgs_knn_workspace_bytes(int P) { return knn_workspace(P > 0 ? P : 0); }

int ed3dgs_knn(int P, int K, const float* points, double* sq_dists, int* indices, void* workspace,
               size_t workspace_bytes, void* stream) {
  g_error[0] = 0;
  return knn_exact(P, K, points, sq_dists, indices, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
