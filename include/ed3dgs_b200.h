/*
 * ed3dgs_b200 -- C ABI of the B200-native E-D3DGS training hot path.
 *
 * Everything here is plain C: device pointers, sizes, a cudaStream_t passed as void*, int status
 * codes. No torch types. Each entry point names the reference interface it replaces
 * (paths relative to the reference checkout, vladb99/E-D3DGS).
 *
 * All data pointers are DEVICE pointers unless stated otherwise. "Absent" optional inputs are NULL
 * (the reference passes empty tensors whose data pointer is null, rasterize_points.cu:94-98).
 * Every call enqueues on `stream` (NULL = the legacy default stream, which is what the reference
 * uses for everything).
 */
#ifndef ED3DGS_B200_H_
#define ED3DGS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  ED3DGS_OK = 0,
  ED3DGS_ERR_ARG = 1,   /* bad argument (reference: AT_ERROR / std::runtime_error) */
  ED3DGS_ERR_CUDA = 2,  /* a CUDA call or kernel failed */
  ED3DGS_ERR_ALLOC = 3  /* a resize callback returned NULL */
};

/* Last error message of the calling thread ("" when none). */
const char* ed3dgs_last_error(void);
/* "major.minor.patch (sm_100a)" */
const char* ed3dgs_version(void);

/*
 * Measurement hooks (no counterpart in the reference; bench.py's roofline pass). While enabled every kernel group
 * of the library is bracketed by CUDA events on its launching stream; _report synchronises the device, writes
 * "name total_ms launches\n" lines into buf and clears the log; _pairs returns (and resets) the number of
 * (pixel, Gaussian) list entries the forward blend walked (out[0]) and the backward blend walked (out[1]).
 */
void ed3dgs_profile_enable(int on);
int ed3dgs_profile_report(char* buf, size_t nbytes);
int ed3dgs_profile_pairs(unsigned long long* out);

/*
 * Resize callback for the three opaque state buffers (geometry, binning, image).
 * Replaces std::function<char*(size_t)> of CudaRasterizer::Rasterizer::forward
 * (cuda_rasterizer/rasterizer.h:31-34) which the reference binds to torch resize_()
 * (rasterize_points.cu:27-33). Must return a device pointer to at least `bytes` bytes,
 * 256-byte aligned, valid until the matching backward has run.
 */
typedef char* (*ed3dgs_alloc_fn)(void* user, size_t bytes);

/*
 * Forward rasterization. Replaces CudaRasterizer::Rasterizer::forward
 * (cuda_rasterizer/rasterizer.h:30-65, rasterizer_impl.cu:255-432) as called from
 * RasterizeGaussiansCUDA (rasterize_points.cu:35-137; pybind name `rasterize_gaussians`,
 * ext.cpp:16). Same argument meaning and order; outputs need NOT be zero-filled by the caller
 * except `radii` semantics are identical (0 for culled Gaussians).
 *   means3D[P,3] shs[P,M,3] colors_precomp[P,3]|NULL opacities[P] tongue_class[P] scales[P,3]
 *   rotations[P,4] cov3D_precomp[P,6]|NULL viewmatrix[16] projmatrix[16] cam_pos[3] background[3]
 *   out_color[3,H,W] out_coord[3,H,W] out_mcoord[3,H,W] out_depth[H,W] out_mdepth[H,W]
 *   out_alpha[H,W] out_tongue[H,W] out_normal[3,H,W] radii[P] (int32)
 * out_coord/out_mcoord are written only when require_coord, out_depth/out_mdepth only when
 * require_depth, out_normal when either (forward.cu:863-870); otherwise they are left untouched.
 * *num_rendered receives the number of (Gaussian, tile) instances (host int), needed by backward.
 */
int ed3dgs_rasterize_forward(
    ed3dgs_alloc_fn geom_alloc, void* geom_user, ed3dgs_alloc_fn binning_alloc, void* binning_user,
    ed3dgs_alloc_fn image_alloc, void* image_user, int P, int D, int M, const float* background,
    int width, int height, const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* tongue_class, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
    const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size,
    int prefiltered, float* out_color, float* out_coord, float* out_mcoord, float* out_depth,
    float* out_mdepth, float* out_alpha, float* out_tongue, float* out_normal, int* radii,
    int require_coord, int require_depth, int debug, void* stream, int* num_rendered);

/*
 * Backward rasterization. Replaces CudaRasterizer::Rasterizer::backward
 * (cuda_rasterizer/rasterizer.h:67-115, rasterizer_impl.cu:436-578) as called from
 * RasterizeGaussiansBackwardCUDA (rasterize_points.cu:139-250; pybind name
 * `rasterize_gaussians_backward`). The eight returned gradients are the eight tensors that
 * function returns; they are fully written here (no caller zero-fill needed):
 *   dL_dmean2D[P,3] dL_dcolor[P,3] dL_dopacity[P] dL_dmean3D[P,3] dL_dcov3D[P,6] dL_dsh[P,M,3]
 *   dL_dscale[P,3] dL_drot[P,4]
 * The reference's eight intermediate gradient tensors (dL_dview_points, dL_dconic, dL_dts,
 * dL_dcamera_planes, dL_dray_planes, dL_dnormals, ...; rasterize_points.cu:184-197) become one
 * packed scratch buffer: grad_scratch, ed3dgs_rasterize_backward_scratch_bytes(P, require_coord)
 * bytes, contents on entry irrelevant.
 */
size_t ed3dgs_rasterize_backward_scratch_bytes(int P, int require_coord);
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
    float* grad_scratch, int require_coord, int require_depth, int debug, void* stream);

/*
 * Copies the intermediate state of a finished forward out of its opaque buffers, in the
 * reference's own layouts, for bit-exact comparison (tests; the reference has no equivalent
 * because its buffers are raw SoA: rasterizer_impl.h:29-87). Any destination may be NULL.
 *   depths[P] means2D[P,2] tiles_touched[P] point_offsets[P] conic_opacity[P,4] rgb[P,3]
 *   keys_unsorted[R] vals_unsorted[R] keys_sorted[R] point_list[R] ranges[tiles,2]
 *   n_contrib[2,H,W] normals[P,3] ray_planes[P,2] ts[P] camera_planes[P,6] view_points[P,3]
 * Entries of culled Gaussians are unspecified (as in the reference).
 */
int ed3dgs_rasterize_get_state(
    int P, int R, int width, int height, int require_coord, const char* geom_buffer,
    const char* binning_buffer, const char* image_buffer, float* depths, float* means2D,
    uint32_t* tiles_touched, uint32_t* point_offsets, float* conic_opacity, float* rgb,
    uint64_t* keys_unsorted, uint32_t* vals_unsorted, uint64_t* keys_sorted, uint32_t* point_list,
    uint32_t* ranges, uint32_t* n_contrib, float* normals, float* ray_planes, float* ts,
    float* camera_planes, float* view_points, void* stream);

/*
 * Point integration for mesh extraction. Replaces CudaRasterizer::Rasterizer::integrate
 * (cuda_rasterizer/rasterizer.h:114-149, rasterizer_impl.cu:580-851) as called from
 * IntegrateGaussiansToPointsCUDA (rasterize_points.cu:273-392; pybind name
 * `integrate_gaussians_to_points`, Python GaussianRasterizer.integrate, __init__.py:245-312).
 *   points3D[PN,3]  (other inputs as ed3dgs_rasterize_forward, without tongue_class)
 *   out_color[9,H,W]  planes: rgb, sum(depth alpha T), median depth, 0, max depth, sum(alpha T),
 *                     number of points in the pixel (forward.cu:1311-1319,1541)
 *   out_alpha_integrated[PN] out_color_integrated[PN,3] out_coordinate2d[PN,2] out_sdf[PN]
 *   radii[P] (int32)
 * All outputs are fully written when P > 0 and PN > 0 (points outside the image receive the values the
 * reference pre-fills: alpha 1, colour 0, coordinate 0, sdf -1000; rasterize_points.cu:316-319); with
 * P == 0 or PN == 0 nothing is written, as in the reference (:344), and the caller supplies the fill.
 * The reference's invraycov / condition / accum_alpha tensors and its point and point-binning buffers
 * (:320-336) live in `workspace` (ed3dgs_integrate_workspace_bytes bytes of device scratch, contents
 * irrelevant). subpixel_offset and view2gaussian_precomp of the reference are dead arguments
 * (forward.cu:1162 reads the former into an unused local) and are not part of this interface.
 */
size_t ed3dgs_integrate_workspace_bytes(int P, int PN, int width, int height);
int ed3dgs_integrate(
    ed3dgs_alloc_fn geom_alloc, void* geom_user, ed3dgs_alloc_fn binning_alloc, void* binning_user,
    ed3dgs_alloc_fn image_alloc, void* image_user, int PN, int P, int D, int M, const float* background,
    int width, int height, const float* points3D, const float* means3D, const float* shs,
    const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
    const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size, int prefiltered,
    float* out_color, int* radii, float* out_alpha_integrated, float* out_color_integrated,
    float* out_coordinate2d, float* out_sdf, void* workspace, size_t workspace_bytes, int debug,
    void* stream, int* num_rendered);

/*
 * Near-plane visibility mask. Replaces CudaRasterizer::Rasterizer::markVisible
 * (cuda_rasterizer/rasterizer.h:24-29, rasterizer_impl.cu:54-66,176-188; pybind `mark_visible`).
 * present[P] is one byte per Gaussian (torch.bool storage).
 */
int ed3dgs_mark_visible(int P, const float* means3D, const float* viewmatrix,
                        const float* projmatrix, uint8_t* present, void* stream);

/*
 * Mean squared distance to the 3 nearest neighbours. Replaces SimpleKNN::knn / distCUDA2
 * (submodules/simple-knn/simple_knn.cu:185-221, spatial.cu:16-25; pybind `distCUDA2`).
 * workspace: ed3dgs_dist_cuda2_workspace_bytes(P) bytes of device scratch.
 */
size_t ed3dgs_dist_cuda2_workspace_bytes(int P);
int ed3dgs_dist_cuda2(int P, const float* points, float* mean_dists, void* workspace,
                      size_t workspace_bytes, void* stream);

/*
 * ed3dgs_dist_cuda2 for callers that cannot provide device scratch. No counterpart in the reference (its
 * distCUDA2 lets thrust / torch allocate whatever it needs, simple_knn.cu:188-211).
 *
 * Some users of this library never get to allocate GPU memory: a plug-in inside a host application that
 * sized one memory pool at start-up and forbids cudaMalloc afterwards, a binding from a language whose
 * FFI layer can pass the input and output arrays it already owns but has no allocator of its own, a step
 * that runs when the job's pool is already full. Same inputs, same result (bit-identical) as
 * ed3dgs_dist_cuda2, but the Morton keys, sort buffers and look-back words live in a 16 MiB array that is
 * part of the library itself. The call performs no cudaMalloc / cudaFree and calls no allocating library.
 *
 *   Limit    P <= ed3dgs_dist_cuda2_static_max_points() (262144). Larger P returns ED3DGS_ERR_ARG with a
 *            message naming the limit; nothing is launched.
 *   Cost     every GPU on which this library's module gets loaded holds the 16 MiB until the process ends,
 *            used or not.
 *   First    the first call on each device looks up the array's device address (with that device current --
 *   call     the caller must have made the device that owns `stream`, `points` and `mean_dists` current,
 *            as for every entry point here). Make that first call before starting any stream capture.
 *   Sharing  there is one array per device, shared by every call and every thread, and it keeps whatever
 *            the previous call left in it. The call clears what it needs cleared itself, so calls queued
 *            one after the other (same stream, or ordered by events) are fine. Two calls that can
 *            overlap on the same GPU (different streams, no ordering) race on the array and both
 *            results are garbage; serialise them.
 */
int ed3dgs_dist_cuda2_static_max_points(void);
int ed3dgs_dist_cuda2_static(int P, const float* points, float* mean_dists, void* stream);

/*
 * Exact K nearest neighbours of every point (self excluded), K <= 32: sq_dists[P,K] (double, ascending) and
 * indices[P,K] (int32). Replaces utils/extra_utils.py:5-16 (`o3d_knn`: P Open3D KDTreeFlann queries in a python
 * loop on the CPU, float64), which train.py:218-229 runs for the embedding regulariser whenever the number
 * of Gaussians changed. Ties are ordered by index. P must exceed K.
 * workspace: ed3dgs_knn_workspace_bytes(P) bytes of device scratch.
 */
size_t ed3dgs_knn_workspace_bytes(int P);
int ed3dgs_knn(int P, int K, const float* points, double* sq_dists, int* indices, void* workspace,
               size_t workspace_bytes, void* stream);

/*
 * 3D smoothing filter size per Gaussian. Replaces GaussianModel.compute_3D_filter (scene/gaussian_model.py:538-592;
 * called at every densification and every 100 iterations late in training, train.py:115,421,439-442).
 *   cameras[num_cameras,16]: R[9] (camera.R as stored, applied as xyz @ R), T[3], focal_x, focal_y, width, height
 *   focal_length: the largest focal_x of the set (:584-585)      filter_3d[P] (the reference keeps it as [P,1])
 *   scratch: >= 4 bytes of device memory       *any_valid (host): 0 when no camera sees any point (the reference
 *   raises there, :587); filter_3d is then meaningless.
 * Synchronises the stream (to report any_valid).
 */
int ed3dgs_compute_3d_filter(int P, const float* xyz, int num_cameras, const float* cameras, float focal_length,
                             float* filter_3d, void* scratch, int* any_valid, void* stream);

/*
 * Per-Gaussian activations between the deformation network and the rasterizer, fused.
 * Replaces the torch statements of gaussian_renderer/__init__.py:77-83: rotation_activation = F.normalize
 * (eps 1e-12), scaling_activation = exp, opacity_activation = sigmoid (scene/gaussian_model.py:37-45) and, when
 * filter_3d != NULL, GaussianModel.apply_scaling_n_opacity_with_3D_filter (scene/gaussian_model.py:594-603):
 * scales = sqrt(exp(s)^2 + f^2), opacity = sigmoid(o) * sqrt(prod exp(s)^2 / prod (exp(s)^2 + f^2)).
 *   scales[P,3], rotations[P,4], opacity[P] (the reference's [P,1]), filter_3d[P] or NULL: raw (pre-activation)
 *   *_out: activated values, fully written.
 * Backward: g_* = gradients w.r.t. the activated values (any may be NULL = zero), d_* = gradients w.r.t. the raw
 * inputs, fully written; everything is recomputed from the raw inputs (nothing saved by the forward).
 */
int ed3dgs_activations_forward(int P, const float* scales, const float* rotations, const float* opacity,
                               const float* filter_3d, float* scales_out, float* rotations_out, float* opacity_out,
                               void* stream);
int ed3dgs_activations_backward(int P, const float* scales, const float* rotations, const float* opacity,
                                const float* filter_3d, const float* g_scales, const float* g_rotations,
                                const float* g_opacity, float* d_scales, float* d_rotations, float* d_opacity,
                                void* stream);

/*
 * Image losses either side of the rasterizer, fused.
 *
 * L1 + SSIM of utils/loss_utils.py:24-72 (`l1_loss`, `ssim` with the 11x11 sigma-1.5 window and zero padding;
 * train.py:195-205). Forward: sums[0] = sum |img - gt|, sums[1] = sum of the SSIM map (device doubles; the caller
 * divides by C*H*W), maps[3,C,H,W] = per-pixel partial derivatives kept for the backward. Backward:
 * d_img = (*upstream) * (coef_l1 * sign(img - gt) + coef_ssim * d(sum ssim)/d img); upstream may be NULL (= 1).
 */
int ed3dgs_ssim_l1_forward(int C, int H, int W, const float* img, const float* gt, double* sums, float* maps,
                           void* stream);
int ed3dgs_ssim_l1_backward(int C, int H, int W, const float* img, const float* gt, const float* maps,
                            const float* upstream, float coef_l1, float coef_ssim, float* d_img, void* stream);
/*
 * Depth -> normal consistency of train.py:249-267 with utils/graphics_utils.py:240-267
 * (`depth_double_to_normal`): normals from the expected and the median depth map by back-projection and central
 * differences, err_k = 1 - <normal, n_k>. Forward: sums[k] = sum over all H*W pixels of err_k (device doubles),
 * out_normals[2,3,H,W] optional (the tensor depth_double_to_normal returns). Backward of
 * w_expected * mean(err_0) + w_median * mean(err_1): d_depth_*[H,W] and d_normal[3,H,W], fully written
 * (any may be NULL).
 */
int ed3dgs_depth_normal_forward(int H, int W, float focal_x, float focal_y, const float* depth_expected,
                                const float* depth_median, const float* normal, float* out_normals, double* sums,
                                void* stream);
int ed3dgs_depth_normal_backward(int H, int W, float focal_x, float focal_y, const float* depth_expected,
                                 const float* depth_median, const float* normal, const float* upstream, float w_expected,
                                 float w_median, float* d_depth_expected, float* d_depth_median, float* d_normal,
                                 void* stream);

/*
 * Per-Gaussian deformation network (coarse + fine stage), fused. Replaces the torch module
 * scene/deformation.py:15-148 (deform_network.forward :108-141, query_time :72-88, deform :90-106,
 * create_net :38-51). See include/ed3dgs_deform.h.
 */
#include "ed3dgs_deform.h"

#ifdef __cplusplus
}
#endif
#endif /* ED3DGS_B200_H_ */
