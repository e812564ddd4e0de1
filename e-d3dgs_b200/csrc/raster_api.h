// Internal argument packs behind the flat C-ABI in include/ed3dgs_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ed3dgs_b200.h"

namespace ed3dgs {

void set_error(const char* fmt, ...);

struct RasterFwdArgs {
  ed3dgs_alloc_fn geom_alloc; void* geom_user;
  ed3dgs_alloc_fn binning_alloc; void* binning_user;
  ed3dgs_alloc_fn image_alloc; void* image_user;
  int P, D, M;
  const float* background;
  int width, height;
  const float *means3D, *shs, *colors_precomp, *opacities, *tongue_class, *scales;
  float scale_modifier;
  const float *rotations, *cov3D_precomp, *viewmatrix, *projmatrix, *cam_pos;
  float tan_fovx, tan_fovy, kernel_size;
  int prefiltered;
  float *out_color, *out_coord, *out_mcoord, *out_depth, *out_mdepth, *out_alpha, *out_tongue, *out_normal;
  int* radii;
  int require_coord, require_depth, debug;
  int* num_rendered;
};

struct RasterBwdArgs {
  int P, D, M, R;
  const float* background;
  int width, height;
  const float *means3D, *shs, *colors_precomp, *alphas, *scales;
  float scale_modifier;
  const float *rotations, *cov3D_precomp, *viewmatrix, *projmatrix, *campos;
  float tan_fovx, tan_fovy, kernel_size;
  const int* radii;
  const float* normalmap;
  char *geom_buffer, *binning_buffer, *image_buffer;
  const float *dL_dpix, *dL_dpix_coord, *dL_dpix_mcoord, *dL_dpix_depth, *dL_dpix_mdepth, *dL_dalphas,
      *dL_dpixel_normals;
  float *dL_dmean2D, *dL_dcolor, *dL_dopacity, *dL_dmean3D, *dL_dcov3D, *dL_dsh, *dL_dscale, *dL_drot;
  float* grad_scratch;  // [P][16 or 32] packed per-Gaussian gradients, any contents
  int require_coord, require_depth, debug;
  int packed_is_input;  // test hook: grad_scratch already holds the packed per-Gaussian gradients
};

// Point integration: the forward argument pack (out_color is [9,H,W] here; tongue_class, the depth /
// coordinate planes and require_* are unused) plus the query points and their outputs.
struct IntegrateArgs {
  RasterFwdArgs fwd;
  int PN;
  const float* points3D;
  float *out_alpha_integrated, *out_color_integrated, *out_coordinate2d, *out_sdf;
  void* workspace;
  size_t workspace_bytes;
};

int raster_forward(const RasterFwdArgs& a, cudaStream_t stream);
int raster_integrate(const IntegrateArgs& a, cudaStream_t stream);
size_t integrate_workspace_bytes(int P, int PN, int width, int height);
int raster_backward(const RasterBwdArgs& a, cudaStream_t stream);
int mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                 uint8_t* present, cudaStream_t stream);
int dist_cuda2(int P, const float* points, float* mean_dists, void* workspace, size_t workspace_bytes,
               cudaStream_t stream);
size_t dist_cuda2_workspace(int P);
int knn_exact(int P, int K, const float* points, double* sq_dists, int* indices, void* workspace,
              size_t workspace_bytes, cudaStream_t stream);
size_t knn_workspace(int P);
int compute_3d_filter(int P, const float* xyz, int C, const float* cams, float focal_length, float* filter,
                      void* scratch, int* any_valid, cudaStream_t stream);

}  // namespace ed3dgs
