// Image-space losses either side of the rasterizer, fused:
//   * L1 + SSIM (utils/loss_utils.py:24-72: 11x11 Gaussian window, sigma 1.5, zero padding, five depthwise
//     convolutions in the forward and as many again in autograd's backward) -> one forward kernel that also
//     leaves the three per-pixel partial derivative maps, one backward kernel that convolves them back;
//   * depth -> normal consistency (utils/graphics_utils.py:240-267 depth_double_to_normal + train.py:249-267):
//     back-projection, central differences, cross product, normalisation and the (1 - n_render . n_depth)
//     error of both depth maps in one pass, with a hand-written backward to both depths and the normal map.
// These are HBM-light stencil kernels (5.3 M pixels at C3); the win is launch count and intermediate traffic
// (the torch graph materialises ~25 image-sized tensors per call).
#include <cstdio>
#include "raster_api.h"
#include "profiler.h"

namespace ed3dgs {

constexpr int SS_T = 16;          // output tile
constexpr int SS_R = 5;           // window radius (window_size 11)
constexpr int SS_IN = SS_T + 2 * SS_R;

struct SsimWindow { float g[11]; };

// gaussian(11, 1.5) of loss_utils.py:36-38 in fp32: exp(-(x-5)^2 / (2 sigma^2)) / sum
static SsimWindow ssim_window() {
  SsimWindow w;
  float sum = 0.f;
  for (int i = 0; i < 11; ++i) {
    w.g[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5));
    sum += w.g[i];
  }
  for (int i = 0; i < 11; ++i) w.g[i] /= sum;
  return w;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// sums[0] += sum |x - y|, sums[1] += sum ssim_map; maps[3][C,H,W] = d ssim_map / d (mu1, E[x^2], E[xy]).
__global__ void __launch_bounds__(SS_T * SS_T)
ssim_l1_forward(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                double* __restrict__ sums, float* __restrict__ maps) {
  __shared__ float s_x[SS_IN][SS_IN + 1], s_y[SS_IN][SS_IN + 1];
  __shared__ float s_h[5][SS_IN][SS_T + 1];
  __shared__ float s_red[2][8];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
  const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
  const size_t plane = (size_t)blockIdx.z * H * W;
  for (int i = tid; i < SS_IN * SS_IN; i += SS_T * SS_T) {
    const int r = i / SS_IN, c = i % SS_IN;
    const int y = y0 + r - SS_R, x = x0 + c - SS_R;
    const bool in = y >= 0 && y < H && x >= 0 && x < W;
    s_x[r][c] = in ? img[plane + (size_t)y * W + x] : 0.f;
    s_y[r][c] = in ? gt[plane + (size_t)y * W + x] : 0.f;
  }
  __syncthreads();
  for (int i = tid; i < SS_IN * SS_T; i += SS_T * SS_T) {
    const int r = i / SS_T, c = i % SS_T;
    float a0 = 0, a1 = 0, a2 = 0, a3 = 0, a4 = 0;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      const float w = win.g[k], x = s_x[r][c + k], y = s_y[r][c + k];
      a0 += w * x; a1 += w * y; a2 += w * x * x; a3 += w * y * y; a4 += w * x * y;
    }
    s_h[0][r][c] = a0; s_h[1][r][c] = a1; s_h[2][r][c] = a2; s_h[3][r][c] = a3; s_h[4][r][c] = a4;
  }
  __syncthreads();
  float mu1 = 0, mu2 = 0, ex2 = 0, ey2 = 0, exy = 0;
#pragma unroll
  for (int k = 0; k < 11; ++k) {
    const float w = win.g[k];
    mu1 += w * s_h[0][ty + k][tx]; mu2 += w * s_h[1][ty + k][tx]; ex2 += w * s_h[2][ty + k][tx];
    ey2 += w * s_h[3][ty + k][tx]; exy += w * s_h[4][ty + k][tx];
  }
  const int x = x0 + tx, y = y0 + ty;
  float l1 = 0.f, ss = 0.f;
  if (x < W && y < H) {
    const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
    const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
    const float s1 = ex2 - mu1_sq, s2 = ey2 - mu2_sq, s12 = exy - mu12;
    const float A = 2 * mu12 + C1, B = 2 * s12 + C2, Cc = mu1_sq + mu2_sq + C1, D = s1 + s2 + C2;
    const float m = (A * B) / (Cc * D);
    ss = m;
    l1 = fabsf(s_x[ty + SS_R][tx + SS_R] - s_y[ty + SS_R][tx + SS_R]);
    const size_t o = plane + (size_t)y * W + x, CHW = (size_t)gridDim.z * H * W;
    // total derivative wrt mu1 with E[x^2], E[xy] held fixed (sigma1_sq = E[x^2] - mu1^2, sigma12 = E[xy] - mu1 mu2)
    maps[o] = (2 * mu2 * (B - A)) / (Cc * D) - m * (2 * mu1 / Cc) + m * (2 * mu1 / D);
    maps[CHW + o] = -m / D;
    maps[2 * CHW + o] = (2 * A) / (Cc * D);
  }
  l1 = warp_sum(l1);
  ss = warp_sum(ss);
  if ((tid & 31) == 0) { s_red[0][tid >> 5] = l1; s_red[1][tid >> 5] = ss; }
  __syncthreads();
  if (tid < 2) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_red[tid][w];
    atomicAdd(&sums[tid], (double)t);
  }
}

// d_img = c_l1 * sign(x - y) + c_ssim * [ G*(dmu1) + 2 x G*(dEx2) + y G*(dExy) ]   (G symmetric, zero padded)
__global__ void __launch_bounds__(SS_T * SS_T)
ssim_l1_backward(int H, int W, const float* __restrict__ img, const float* __restrict__ gt, SsimWindow win,
                 const float* __restrict__ maps, const float* __restrict__ upstream, float c_l1, float c_ssim,
                 float* __restrict__ d_img) {
  __shared__ float s_m[3][SS_IN][SS_IN + 1];
  __shared__ float s_h[3][SS_IN][SS_T + 1];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * SS_T + tx;
  const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
  const size_t plane = (size_t)blockIdx.z * H * W, CHW = (size_t)gridDim.z * H * W;
  for (int i = tid; i < SS_IN * SS_IN; i += SS_T * SS_T) {
    const int r = i / SS_IN, c = i % SS_IN;
    const int y = y0 + r - SS_R, x = x0 + c - SS_R;
    const bool in = y >= 0 && y < H && x >= 0 && x < W;
    const size_t o = plane + (size_t)y * W + x;
#pragma unroll
    for (int q = 0; q < 3; ++q) s_m[q][r][c] = in ? maps[q * CHW + o] : 0.f;
  }
  __syncthreads();
  for (int i = tid; i < SS_IN * SS_T; i += SS_T * SS_T) {
    const int r = i / SS_T, c = i % SS_T;
    float a0 = 0, a1 = 0, a2 = 0;
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      const float w = win.g[k];
      a0 += w * s_m[0][r][c + k]; a1 += w * s_m[1][r][c + k]; a2 += w * s_m[2][r][c + k];
    }
    s_h[0][r][c] = a0; s_h[1][r][c] = a1; s_h[2][r][c] = a2;
  }
  __syncthreads();
  const int x = x0 + tx, y = y0 + ty;
  if (x >= W || y >= H) return;
  float g0 = 0, g1 = 0, g2 = 0;
#pragma unroll
  for (int k = 0; k < 11; ++k) {
    const float w = win.g[k];
    g0 += w * s_h[0][ty + k][tx]; g1 += w * s_h[1][ty + k][tx]; g2 += w * s_h[2][ty + k][tx];
  }
  const size_t o = plane + (size_t)y * W + x;
  const float xv = img[o], yv = gt[o];
  const float up = upstream ? upstream[0] : 1.f;
  const float diff = xv - yv;
  const float sgn = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
  d_img[o] = up * (c_l1 * sgn + c_ssim * (g0 + 2.f * xv * g1 + yv * g2));
}

int ssim_l1_fwd(int C, int H, int W, const float* img, const float* gt, double* sums, float* maps, cudaStream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0) return ED3DGS_OK;
  ProfScope ps("ssim_l1_forward", stream);
  cudaMemsetAsync(sums, 0, 2 * sizeof(double), stream);
  const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C), block(SS_T, SS_T);
  ssim_l1_forward<<<grid, block, 0, stream>>>(H, W, img, gt, ssim_window(), sums, maps);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("ssim_l1_forward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int ssim_l1_bwd(int C, int H, int W, const float* img, const float* gt, const float* maps, const float* upstream,
                float c_l1, float c_ssim, float* d_img, cudaStream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0) return ED3DGS_OK;
  ProfScope ps("ssim_l1_backward", stream);
  const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C), block(SS_T, SS_T);
  ssim_l1_backward<<<grid, block, 0, stream>>>(H, W, img, gt, ssim_window(), maps, upstream, c_l1, c_ssim, d_img);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("ssim_l1_backward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

// ---------------------------------------------------------------------------------------------
// Depth -> normal consistency. For k in {expected depth, median depth}: P_k(y,x) = depth_k(y,x) * ray(y,x) with
// ray = ((x + .5)/fx - W/(2 fx), (y + .5)/fy - H/(2 fy), 1) (graphics_utils.py:240-254); interior pixels:
// a = P(y+1,x) - P(y-1,x), b = P(y,x+1) - P(y,x-1), n = normalize(a x b) (eps 1e-12), border n = 0 (:256-263);
// err_k = 1 - <normal_render, n_k> (train.py:264).
struct DnCam { float inv_fx, inv_fy, cx, cy; };   // ray = (x + .5) * inv_fx - cx, (y + .5) * inv_fy - cy

__device__ __forceinline__ float3 dn_ray(const DnCam c, int x, int y) {
  return make_float3(((float)x + 0.5f) * c.inv_fx - c.cx, ((float)y + 0.5f) * c.inv_fy - c.cy, 1.0f);
}
__device__ __forceinline__ float3 dn_point(const DnCam c, const float* __restrict__ d, int W, int x, int y) {
  const float3 r = dn_ray(c, x, y);
  const float dv = d[(size_t)y * W + x];
  return make_float3(dv * r.x, dv * r.y, dv * r.z);
}
__device__ __forceinline__ float3 cross3(const float3 a, const float3 b) {
  return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// out_normals (optional) [2,3,H,W]; sums[k] += sum over all pixels of err_k
__global__ void __launch_bounds__(256)
depth_normal_forward(int H, int W, DnCam cam, const float* __restrict__ depth0, const float* __restrict__ depth1,
                     const float* __restrict__ normal, float* __restrict__ out_normals, double* __restrict__ sums) {
  __shared__ float s_red[2][8];
  const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 16 + (threadIdx.x >> 4);
  const size_t HW = (size_t)H * W;
  float err[2] = {0.f, 0.f};
  if (x < W && y < H) {
    const size_t o = (size_t)y * W + x;
    const float3 nr = make_float3(normal[o], normal[HW + o], normal[2 * HW + o]);
    const bool interior = x >= 1 && x < W - 1 && y >= 1 && y < H - 1;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float* d = k ? depth1 : depth0;
      float3 n = make_float3(0.f, 0.f, 0.f);
      if (interior) {
        const float3 pu = dn_point(cam, d, W, x, y + 1), pd = dn_point(cam, d, W, x, y - 1);
        const float3 pr = dn_point(cam, d, W, x + 1, y), pl = dn_point(cam, d, W, x - 1, y);
        const float3 v = cross3(make_float3(pu.x - pd.x, pu.y - pd.y, pu.z - pd.z),
                                make_float3(pr.x - pl.x, pr.y - pl.y, pr.z - pl.z));
        const float len = fmaxf(sqrtf(v.x * v.x + v.y * v.y + v.z * v.z), 1e-12f);
        n = make_float3(v.x / len, v.y / len, v.z / len);
      }
      if (out_normals) {
        float* on = out_normals + (size_t)k * 3 * HW + o;
        on[0] = n.x; on[HW] = n.y; on[2 * HW] = n.z;
      }
      err[k] = 1.f - (nr.x * n.x + nr.y * n.y + nr.z * n.z);
    }
  }
  const float e0 = warp_sum(err[0]), e1 = warp_sum(err[1]);
  if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = e0; s_red[1][threadIdx.x >> 5] = e1; }
  __syncthreads();
  if (threadIdx.x < 2) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += s_red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], (double)t);
  }
}

// loss = w0 * mean(err_0) + w1 * mean(err_1), scaled by *upstream. d_normal fully written; d_depth0/1 must be
// zero on entry (four neighbours scatter into every depth pixel).
__global__ void __launch_bounds__(256)
depth_normal_backward(int H, int W, DnCam cam, const float* __restrict__ depth0, const float* __restrict__ depth1,
                      const float* __restrict__ normal, const float* __restrict__ upstream, float w0, float w1,
                      float* __restrict__ d_depth0, float* __restrict__ d_depth1, float* __restrict__ d_normal) {
  const int x = blockIdx.x * 16 + (threadIdx.x & 15), y = blockIdx.y * 16 + (threadIdx.x >> 4);
  if (x >= W || y >= H) return;
  const size_t HW = (size_t)H * W, o = (size_t)y * W + x;
  const float up = (upstream ? upstream[0] : 1.f) / (float)HW;
  const float3 nr = make_float3(normal[o], normal[HW + o], normal[2 * HW + o]);
  const bool interior = x >= 1 && x < W - 1 && y >= 1 && y < H - 1;
  float3 dnr = make_float3(0.f, 0.f, 0.f);
  if (interior) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float* d = k ? depth1 : depth0;
      float* dd = k ? d_depth1 : d_depth0;
      const float wk = (k ? w1 : w0) * up;
      const float3 ru = dn_ray(cam, x, y + 1), rd = dn_ray(cam, x, y - 1), rr = dn_ray(cam, x + 1, y), rl = dn_ray(cam, x - 1, y);
      const float du = d[o + W], dd_ = d[o - W], dr = d[o + 1], dl = d[o - 1];
      const float3 a = make_float3(du * ru.x - dd_ * rd.x, du * ru.y - dd_ * rd.y, du * ru.z - dd_ * rd.z);
      const float3 b = make_float3(dr * rr.x - dl * rl.x, dr * rr.y - dl * rl.y, dr * rr.z - dl * rl.z);
      const float3 v = cross3(a, b);
      const float norm = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
      const float len = fmaxf(norm, 1e-12f);
      const float3 n = make_float3(v.x / len, v.y / len, v.z / len);
      dnr.x -= wk * n.x; dnr.y -= wk * n.y; dnr.z -= wk * n.z;
      // g = dL/dn = -wk * nr;  dL/dv = (g - n <n, g>) / len  (the clamp branch passes g / eps through)
      const float3 g = make_float3(-wk * nr.x, -wk * nr.y, -wk * nr.z);
      float3 dv;
      if (norm > 1e-12f) {
        const float ng = n.x * g.x + n.y * g.y + n.z * g.z;
        dv = make_float3((g.x - n.x * ng) / len, (g.y - n.y * ng) / len, (g.z - n.z * ng) / len);
      } else {
        dv = make_float3(g.x / len, g.y / len, g.z / len);
      }
      const float3 da = cross3(b, dv);   // v = a x b  =>  dL/da = b x dv, dL/db = dv x a
      const float3 db = cross3(dv, a);
      if (dd) {
        atomicAdd(dd + o + W, da.x * ru.x + da.y * ru.y + da.z * ru.z);
        atomicAdd(dd + o - W, -(da.x * rd.x + da.y * rd.y + da.z * rd.z));
        atomicAdd(dd + o + 1, db.x * rr.x + db.y * rr.y + db.z * rr.z);
        atomicAdd(dd + o - 1, -(db.x * rl.x + db.y * rl.y + db.z * rl.z));
      }
    }
  }
  if (d_normal) { d_normal[o] = dnr.x; d_normal[HW + o] = dnr.y; d_normal[2 * HW + o] = dnr.z; }
}

static DnCam dn_cam(int H, int W, float fx, float fy) {
  return DnCam{1.f / fx, 1.f / fy, (float)W / (2.f * fx), (float)H / (2.f * fy)};
}

int depth_normal_fwd(int H, int W, float fx, float fy, const float* depth0, const float* depth1, const float* normal,
                     float* out_normals, double* sums, cudaStream_t stream) {
  if (H <= 0 || W <= 0) return ED3DGS_OK;
  ProfScope ps("depth_normal_forward", stream);
  cudaMemsetAsync(sums, 0, 2 * sizeof(double), stream);
  depth_normal_forward<<<dim3((W + 15) / 16, (H + 15) / 16), 256, 0, stream>>>(H, W, dn_cam(H, W, fx, fy), depth0, depth1,
                                                                            normal, out_normals, sums);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("depth_normal_forward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

int depth_normal_bwd(int H, int W, float fx, float fy, const float* depth0, const float* depth1, const float* normal,
                     const float* upstream, float w0, float w1, float* d_depth0, float* d_depth1, float* d_normal,
                     cudaStream_t stream) {
  if (H <= 0 || W <= 0) return ED3DGS_OK;
  ProfScope ps("depth_normal_backward", stream);
  if (d_depth0) cudaMemsetAsync(d_depth0, 0, (size_t)H * W * sizeof(float), stream);
  if (d_depth1) cudaMemsetAsync(d_depth1, 0, (size_t)H * W * sizeof(float), stream);
  depth_normal_backward<<<dim3((W + 15) / 16, (H + 15) / 16), 256, 0, stream>>>(
      H, W, dn_cam(H, W, fx, fy), depth0, depth1, normal, upstream, w0, w1, d_depth0, d_depth1, d_normal);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error("depth_normal_backward: %s", cudaGetErrorString(e)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

}  // namespace ed3dgs

using namespace ed3dgs;

extern "C" {

int ed3dgs_ssim_l1_forward(int C, int H, int W, const float* img, const float* gt, double* sums, float* maps,
                           void* stream) {
  return ssim_l1_fwd(C, H, W, img, gt, sums, maps, static_cast<cudaStream_t>(stream));
}

int ed3dgs_ssim_l1_backward(int C, int H, int W, const float* img, const float* gt, const float* maps,
                            const float* upstream, float coef_l1, float coef_ssim, float* d_img, void* stream) {
  return ssim_l1_bwd(C, H, W, img, gt, maps, upstream, coef_l1, coef_ssim, d_img, static_cast<cudaStream_t>(stream));
}

int ed3dgs_depth_normal_forward(int H, int W, float focal_x, float focal_y, const float* depth_expected,
                                const float* depth_median, const float* normal, float* out_normals, double* sums,
                                void* stream) {
  return depth_normal_fwd(H, W, focal_x, focal_y, depth_expected, depth_median, normal, out_normals, sums,
                          static_cast<cudaStream_t>(stream));
}

int ed3dgs_depth_normal_backward(int H, int W, float focal_x, float focal_y, const float* depth_expected,
                                 const float* depth_median, const float* normal, const float* upstream, float w_expected,
                                 float w_median, float* d_depth_expected, float* d_depth_median, float* d_normal,
                                 void* stream) {
  return depth_normal_bwd(H, W, focal_x, focal_y, depth_expected, depth_median, normal, upstream, w_expected, w_median,
                          d_depth_expected, d_depth_median, d_normal, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
