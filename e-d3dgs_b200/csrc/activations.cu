// Per-Gaussian activations between the deformation network and the rasterizer, fused:
//   rotation  = normalize(q)                      (gaussian_renderer/__init__.py:77, F.normalize, eps 1e-12)
//   scales    = exp(s), opacity = sigmoid(o)      (:79-80; scene/gaussian_model.py:37-45)
// or, with the 3D smoothing filter (scene/gaussian_model.py:594-603):
//   s2 = exp(s)^2, sa = s2 + f^2, scales = sqrt(sa), opacity = sigmoid(o) * sqrt(prod(s2) / prod(sa))
// The torch graph runs 6-14 elementwise/reduction kernels each way and saves their intermediates; here one
// forward and one backward kernel read 32 (36) bytes and write 32 bytes per Gaussian, and the backward
// recomputes everything from the raw inputs. One thread per Gaussian; rows are 12/16/4 bytes, so a warp
// touches contiguous 384/512/128-byte runs.
#include <cstdio>
#include "raster_api.h"
#include "profiler.h"

namespace ed3dgs {

constexpr float ACT_NORM_EPS = 1e-12f;   // F.normalize default eps

__device__ __forceinline__ float act_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

__global__ void __launch_bounds__(256)
activations_fwd(int P, const float* __restrict__ s_in, const float* __restrict__ q_in, const float* __restrict__ o_in,
                const float* __restrict__ filter, float* __restrict__ s_out, float* __restrict__ q_out,
                float* __restrict__ o_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const float4 q = reinterpret_cast<const float4*>(q_in)[i];
  const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  const float d = fmaxf(n, ACT_NORM_EPS);
  reinterpret_cast<float4*>(q_out)[i] = make_float4(q.x / d, q.y / d, q.z / d, q.w / d);
  const float e0 = expf(s_in[3 * i]), e1 = expf(s_in[3 * i + 1]), e2 = expf(s_in[3 * i + 2]);
  const float sg = act_sigmoid(o_in[i]);
  if (filter == nullptr) {
    s_out[3 * i] = e0; s_out[3 * i + 1] = e1; s_out[3 * i + 2] = e2;
    o_out[i] = sg;
  } else {
    const float f = filter[i], f2 = f * f;
    const float a0 = e0 * e0, a1 = e1 * e1, a2 = e2 * e2;
    const float b0 = a0 + f2, b1 = a1 + f2, b2 = a2 + f2;
    const float det1 = a0 * a1 * a2, det2 = b0 * b1 * b2;
    s_out[3 * i] = sqrtf(b0); s_out[3 * i + 1] = sqrtf(b1); s_out[3 * i + 2] = sqrtf(b2);
    o_out[i] = sg * sqrtf(det1 / det2);
  }
}

// g_* are the gradients w.r.t. the activated values (any may be NULL = zero); d_* w.r.t. the raw inputs.
__global__ void __launch_bounds__(256)
activations_bwd(int P, const float* __restrict__ s_in, const float* __restrict__ q_in, const float* __restrict__ o_in,
                const float* __restrict__ filter, const float* __restrict__ g_s, const float* __restrict__ g_q,
                const float* __restrict__ g_o, float* __restrict__ d_s, float* __restrict__ d_q,
                float* __restrict__ d_o) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  {
    // y = q / max(|q|, eps): dq = (g - u (u.g)) / n above the clamp, g / eps below it
    const float4 q = reinterpret_cast<const float4*>(q_in)[i];
    const float4 g = g_q ? reinterpret_cast<const float4*>(g_q)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float4 r;
    if (n > ACT_NORM_EPS) {
      const float inv = 1.0f / n;
      const float4 u = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
      const float ug = u.x * g.x + u.y * g.y + u.z * g.z + u.w * g.w;
      r = make_float4((g.x - u.x * ug) * inv, (g.y - u.y * ug) * inv, (g.z - u.z * ug) * inv, (g.w - u.w * ug) * inv);
    } else {
      const float inv = 1.0f / ACT_NORM_EPS;
      r = make_float4(g.x * inv, g.y * inv, g.z * inv, g.w * inv);
    }
    reinterpret_cast<float4*>(d_q)[i] = r;
  }
  const float e0 = expf(s_in[3 * i]), e1 = expf(s_in[3 * i + 1]), e2 = expf(s_in[3 * i + 2]);
  const float gs0 = g_s ? g_s[3 * i] : 0.f, gs1 = g_s ? g_s[3 * i + 1] : 0.f, gs2 = g_s ? g_s[3 * i + 2] : 0.f;
  const float go = g_o ? g_o[i] : 0.f;
  const float sg = act_sigmoid(o_in[i]);
  if (filter == nullptr) {
    d_s[3 * i] = gs0 * e0; d_s[3 * i + 1] = gs1 * e1; d_s[3 * i + 2] = gs2 * e2;
    d_o[i] = go * sg * (1.0f - sg);
  } else {
    const float f = filter[i], f2 = f * f;
    const float a0 = e0 * e0, a1 = e1 * e1, a2 = e2 * e2;
    const float b0 = a0 + f2, b1 = a1 + f2, b2 = a2 + f2;
    const float det1 = a0 * a1 * a2, det2 = b0 * b1 * b2;
    const float ratio = det1 / det2;
    const float coef = sqrtf(ratio);
    d_o[i] = go * coef * sg * (1.0f - sg);
    // opacity = sg * sqrt(det1 / det2)
    const float d_ratio = go * sg * 0.5f / coef;
    const float d_det1 = d_ratio / det2;
    const float d_det2 = -d_ratio * ratio / det2;
    // d/d a_k: scales_k = sqrt(b_k); det1 = prod a; det2 = prod b
    const float da0 = gs0 * 0.5f / sqrtf(b0) + d_det1 * a1 * a2 + d_det2 * b1 * b2;
    const float da1 = gs1 * 0.5f / sqrtf(b1) + d_det1 * a0 * a2 + d_det2 * b0 * b2;
    const float da2 = gs2 * 0.5f / sqrtf(b2) + d_det1 * a0 * a1 + d_det2 * b0 * b1;
    // a_k = exp(s_k)^2 -> d a_k / d s_k = 2 a_k
    d_s[3 * i] = da0 * 2.0f * a0; d_s[3 * i + 1] = da1 * 2.0f * a1; d_s[3 * i + 2] = da2 * 2.0f * a2;
  }
}

}  // namespace ed3dgs

extern "C" {

int ed3dgs_activations_forward(int P, const float* scales, const float* rotations, const float* opacity,
                               const float* filter_3d, float* scales_out, float* rotations_out, float* opacity_out,
                               void* stream) {
  using namespace ed3dgs;
  if (P < 0) { set_error("activations_forward: P < 0"); return ED3DGS_ERR_ARG; }
  if (P == 0) return ED3DGS_OK;
  if (!scales || !rotations || !opacity || !scales_out || !rotations_out || !opacity_out) {
    set_error("activations_forward: null pointer");
    return ED3DGS_ERR_ARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps("activations", st);
  activations_fwd<<<(P + 255) / 256, 256, 0, st>>>(P, scales, rotations, opacity, filter_3d, scales_out, rotations_out,
                                                  opacity_out);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("activations_forward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int ed3dgs_activations_backward(int P, const float* scales, const float* rotations, const float* opacity,
                                const float* filter_3d, const float* g_scales, const float* g_rotations,
                                const float* g_opacity, float* d_scales, float* d_rotations, float* d_opacity,
                                void* stream) {
  using namespace ed3dgs;
  if (P < 0) { set_error("activations_backward: P < 0"); return ED3DGS_ERR_ARG; }
  if (P == 0) return ED3DGS_OK;
  if (!scales || !rotations || !opacity || !d_scales || !d_rotations || !d_opacity) {
    set_error("activations_backward: null pointer");
    return ED3DGS_ERR_ARG;
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  ProfScope ps("activations", st);
  activations_bwd<<<(P + 255) / 256, 256, 0, st>>>(P, scales, rotations, opacity, filter_3d, g_scales, g_rotations,
                                                  g_opacity, d_scales, d_rotations, d_opacity);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("activations_backward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

}  // extern "C"
