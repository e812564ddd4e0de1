// TEST INFRASTRUCTURE ONLY -- C-ABI shim around the UNMODIFIED reference CUDA sources.
//
// This file is ours; everything it calls is compiled from the reference sources where
// they lie under /root/reference (see oracle/Makefile, target `ref`):
//   submodules/diff-gaussian-rasterization/cuda_rasterizer/{forward,backward,rasterizer_impl}.cu
//   submodules/simple-knn/simple_knn.cu
// It replaces the torch/pybind layer (rasterize_points.cu:35-250, spatial.cu:16-25) by
// plain pointers so that tests and bench.py can drive the reference kernels through
// ctypes on the GPU box, where /root/reference itself does not exist.
// Only tests/, __graft_entry__.smoke() and bench.py may load the resulting library.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <cuda_runtime.h>
#include "cuda_rasterizer/config.h"
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"
#include "simple_knn.h"
#include "cuda_rasterizer/backward.h"

namespace {
struct Buf {
  char* ptr = nullptr;
  size_t cap = 0;
  size_t size = 0;
  char* resize(size_t n) {
    if (n > cap) {
      if (ptr) cudaFree(ptr);
      size_t want = n + n / 4 + 256;
      if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; throw std::runtime_error("cudaMalloc failed"); }
      cap = want;
    }
    size = n;
    return ptr;
  }
  ~Buf() { if (ptr) cudaFree(ptr); }
};
struct Ctx {
  Buf geom, binning, img;
  int P = 0, R = 0, W = 0, H = 0;
};
thread_local char g_err[512];
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err; }

void* ref_create() { return new Ctx(); }
void ref_destroy(void* c) { delete static_cast<Ctx*>(c); }

// Mirrors RasterizeGaussiansCUDA (rasterize_points.cu:35-137) minus tensor allocation:
// the caller passes zero-filled outputs. Returns num_rendered or -1.
int ref_forward(void* c, int P, int D, int M, const float* background, int W, int H,
                const float* means3D, const float* shs, const float* colors_precomp,
                const float* opacities, const float* tongue_class, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                float tan_fovx, float tan_fovy, float kernel_size, int prefiltered,
                float* out_color, float* out_coord, float* out_mcoord, float* out_depth,
                float* out_mdepth, float* out_alpha, float* out_tongue, float* out_normal,
                int* radii, int require_coord, int require_depth, int debug) {
  Ctx* ctx = static_cast<Ctx*>(c);
  try {
    std::function<char*(size_t)> g = [ctx](size_t n) { return ctx->geom.resize(n); };
    std::function<char*(size_t)> b = [ctx](size_t n) { return ctx->binning.resize(n); };
    std::function<char*(size_t)> i = [ctx](size_t n) { return ctx->img.resize(n); };
    int R = CudaRasterizer::Rasterizer::forward(
        g, b, i, P, D, M, background, W, H, means3D, shs, colors_precomp, opacities,
        tongue_class, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
        cam_pos, tan_fovx, tan_fovy, kernel_size, prefiltered != 0, out_color, out_coord,
        out_mcoord, out_depth, out_mdepth, out_alpha, out_tongue, out_normal, radii,
        require_coord != 0, require_depth != 0, debug != 0);
    ctx->P = P; ctx->R = R; ctx->W = W; ctx->H = H;
    return R;
  } catch (const std::exception& e) {
    snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// Mirrors RasterizeGaussiansBackwardCUDA (rasterize_points.cu:139-250); the caller passes
// zero-filled gradient buffers (the reference zero-allocates them at :184-197).
int ref_backward(void* c, int P, int D, int M, int R, const float* background, int W, int H,
                 const float* means3D, const float* shs, const float* colors_precomp,
                 const float* alphas, const float* scales, float scale_modifier,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                 float kernel_size, const int* radii, const float* normalmap,
                 const float* dL_dpix, const float* dL_dpix_coord, const float* dL_dpix_mcoord,
                 const float* dL_dpix_depth, const float* dL_dpix_mdepth, const float* dL_dalphas,
                 const float* dL_dpixel_normals, float* dL_dmean2D, float* dL_dview_points,
                 float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dts,
                 float* dL_dcamera_planes, float* dL_dray_planes, float* dL_dnormals,
                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                 float* dL_drot, int require_coord, int require_depth, int debug) {
  Ctx* ctx = static_cast<Ctx*>(c);
  try {
    CudaRasterizer::Rasterizer::backward(
        P, D, M, R, background, W, H, means3D, shs, colors_precomp, alphas, scales,
        scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
        tan_fovy, kernel_size, radii, normalmap, ctx->geom.ptr, ctx->binning.ptr, ctx->img.ptr,
        dL_dpix, dL_dpix_coord, dL_dpix_mcoord, dL_dpix_depth, dL_dpix_mdepth, dL_dalphas,
        dL_dpixel_normals, dL_dmean2D, dL_dview_points, dL_dconic, dL_dopacity, dL_dcolor, dL_dts,
        dL_dcamera_planes, dL_dray_planes, dL_dnormals, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
        dL_drot, require_coord != 0, require_depth != 0, debug != 0);
    return 0;
  } catch (const std::exception& e) {
    snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// Copies the reference's intermediate state of the last forward into caller buffers
// (device pointers; any may be null). This is what the bit-exact tests compare.
//   depths[P] means2D[2P] tiles_touched[P] point_offsets[P] conic_opacity[4P] rgb[3P]
//   keys_unsorted[R] vals_unsorted[R] keys_sorted[R] point_list[R] ranges[2*tiles]
//   n_contrib[2*W*H]
int ref_get_state(void* c, float* depths, float* means2D, uint32_t* tiles_touched,
                  uint32_t* point_offsets, float* conic_opacity, float* rgb,
                  uint64_t* keys_unsorted, uint32_t* vals_unsorted, uint64_t* keys_sorted,
                  uint32_t* point_list, uint32_t* ranges, uint32_t* n_contrib,
                  float* normals, float* ray_planes, float* ts, float* camera_planes,
                  float* view_points) {
  Ctx* ctx = static_cast<Ctx*>(c);
  const size_t P = ctx->P, R = ctx->R, N = (size_t)ctx->W * ctx->H;
  char* gp = ctx->geom.ptr;
  CudaRasterizer::GeometryState geom = CudaRasterizer::GeometryState::fromChunk(gp, P);
  char* ip = ctx->img.ptr;
  CudaRasterizer::ImageState img = CudaRasterizer::ImageState::fromChunk(ip, N);
  const size_t tiles = (size_t)((ctx->W + BLOCK_X - 1) / BLOCK_X) * ((ctx->H + BLOCK_Y - 1) / BLOCK_Y);
  auto cp = [](void* dst, const void* src, size_t n) {
    if (dst && n) cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice);
  };
  cp(depths, geom.depths, P * 4);
  cp(means2D, geom.means2D, P * 8);
  cp(tiles_touched, geom.tiles_touched, P * 4);
  cp(point_offsets, geom.point_offsets, P * 4);
  cp(conic_opacity, geom.conic_opacity, P * 16);
  cp(rgb, geom.rgb, P * 12);
  cp(normals, geom.normals, P * 12);
  cp(ray_planes, geom.ray_planes, P * 8);
  cp(ts, geom.ts, P * 4);
  cp(camera_planes, geom.camera_planes, P * 24);
  cp(view_points, geom.view_points, P * 12);
  cp(ranges, img.ranges, tiles * 8);
  cp(n_contrib, img.n_contrib, N * 8);
  if (R > 0) {
    char* bp = ctx->binning.ptr;
    CudaRasterizer::BinningState bin = CudaRasterizer::BinningState::fromChunk(bp, R);
    cp(keys_unsorted, bin.point_list_keys_unsorted, R * 8);
    cp(vals_unsorted, bin.point_list_unsorted, R * 4);
    cp(keys_sorted, bin.point_list_keys, R * 8);
    cp(point_list, bin.point_list, R * 4);
  }
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

// Only the per-Gaussian half of the reference backward, i.e. the BACKWARD::preprocess call of
// Rasterizer::backward (rasterizer_impl.cu:544-577) on caller-supplied intermediate gradients,
// argument for argument -- including dL_dconic travelling in the conic_opacity slot (:576).
// Lets the tests feed identical intermediates to the reference and the product.
int ref_backward_preprocess(void* c, int P, int D, int M, int W, int H, const float* means3D,
                            const float* shs, const float* scales, float scale_modifier,
                            const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                            float kernel_size, const int* radii, float* dL_dmean2D, float* dL_dconic,
                            float* dL_dview_points, float* dL_dopacity, float* dL_dcolor, float* dL_dts,
                            float* dL_dcamera_planes, float* dL_dray_planes, float* dL_dnormals,
                            float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale,
                            float* dL_drot) {
  Ctx* ctx = static_cast<Ctx*>(c);
  char* gp = ctx->geom.ptr;
  CudaRasterizer::GeometryState geomState = CudaRasterizer::GeometryState::fromChunk(gp, P);
  const float focal_y = H / (2.0f * tan_fovy);
  const float focal_x = W / (2.0f * tan_fovx);
  const float* cov3D_ptr = (cov3D_precomp != nullptr) ? cov3D_precomp : geomState.cov3D;
  BACKWARD::preprocess(P, D, M, (float3*)means3D, radii, shs, geomState.clamped, (glm::vec3*)scales,
                       (glm::vec4*)rotations, scale_modifier, cov3D_ptr, viewmatrix, projmatrix, focal_x,
                       focal_y, tan_fovx, tan_fovy, kernel_size, (glm::vec3*)campos, (float3*)dL_dmean2D,
                       dL_dconic, (float3*)dL_dview_points, (glm::vec3*)dL_dmean3D, dL_dcolor, dL_dts,
                       (float2*)dL_dcamera_planes, (float2*)dL_dray_planes, dL_dnormals, dL_dcov3D, dL_dsh,
                       (glm::vec3*)dL_dscale, (glm::vec4*)dL_drot, (float4*)dL_dconic, dL_dopacity);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

// Mirrors IntegrateGaussiansToPointsCUDA (rasterize_points.cu:273-392) minus tensor allocation. The caller
// passes the outputs pre-filled the way that function fills them: out_color[9,H,W] = 0, accum_alpha[H,W] = 0,
// radii = 0, out_alpha_integrated[PN] = 1, out_color_integrated[PN,3] = 0, out_coordinate2d[PN,2] = 0,
// out_sdf[PN] = -1000, invraycov[P,6] = 0, condition[max(P,PN)] = 0 (the reference sizes it PN but indexes it
// per Gaussian, rasterize_points.cu:322 vs forward.cu:379). Returns num_rendered or -1.
int ref_integrate(void* c, int PN, int P, int D, int M, const float* background, int W, int H,
                  const float* points3D, const float* means3D, const float* shs, const float* colors_precomp,
                  const float* opacities, const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                  const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size,
                  const float* subpixel_offset, float* out_color, float* accum_alpha, float* invraycov,
                  int* radii, float* out_alpha_integrated, float* out_color_integrated, float* out_coordinate2d,
                  float* out_sdf, bool* condition) {
  Ctx* ctx = static_cast<Ctx*>(c);
  try {
    Buf pointbuf, pointbin;
    std::function<char*(size_t)> g = [ctx](size_t n) { return ctx->geom.resize(n); };
    std::function<char*(size_t)> b = [ctx](size_t n) { return ctx->binning.resize(n); };
    std::function<char*(size_t)> i = [ctx](size_t n) { return ctx->img.resize(n); };
    std::function<char*(size_t)> pb = [&pointbuf](size_t n) { return pointbuf.resize(n); };
    std::function<char*(size_t)> pbin = [&pointbin](size_t n) { return pointbin.resize(n); };
    int R = CudaRasterizer::Rasterizer::integrate(
        g, b, i, pb, pbin, PN, P, D, M, background, W, H, points3D, means3D, shs, colors_precomp, opacities,
        scales, scale_modifier, rotations, cov3D_precomp, nullptr, viewmatrix, projmatrix, cam_pos, tan_fovx,
        tan_fovy, kernel_size, subpixel_offset, false, out_color, accum_alpha, invraycov, radii,
        out_alpha_integrated, out_color_integrated, out_coordinate2d, out_sdf, condition, false);
    cudaDeviceSynchronize();
    ctx->P = P; ctx->R = R; ctx->W = W; ctx->H = H;
    return R;
  } catch (const std::exception& e) {
    snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

// Mirrors distCUDA2 (simple-knn/spatial.cu:16-25).
void ref_dist_cuda2(int P, const float* points, float* mean_dists) {
  SimpleKNN::knn(P, (float3*)points, mean_dists);
}

}  // extern "C"
