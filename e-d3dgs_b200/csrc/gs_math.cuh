// Per-Gaussian math shared by the forward and backward splat kernels (device only).
//
// The arithmetic SHAPE of every expression that feeds an integer result (depth bits in the
// sort key, pixel centre, radius, tile rectangle) follows the reference so that nvcc's FMA
// contraction lands on the same operations and keys / radii / tile counts come out
// bit-identical:
//   transformPoint4x3/4x4, ndc2Pix, getRect   auxiliary.h:57-93
//   glm mat3*mat3 / mat3*vec3 expression form  third_party/glm/glm/detail/type_mat3x3.inl
//   computeCov3D                               forward.cu:270-304
//   computeCov2D (EWA + RaDe-GS planes)        forward.cu:77-264
//   symmetric 3x3 eigen-solve (tred2 + tqli)   auxiliary.h:217-401
// (paths relative to submodules/diff-gaussian-rasterization/cuda_rasterizer/ of the reference)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ed3dgs {

constexpr int TILE = 16;          // config.h:16-17 (BLOCK_X, BLOCK_Y)
constexpr int TILE_PIX = TILE * TILE;

// ---- column-major 3x3, m[c][r] ------------------------------------------------------------
struct Mat3 {
  float m[3][3];
  __device__ __forceinline__ float* operator[](int c) { return m[c]; }
  __device__ __forceinline__ const float* operator[](int c) const { return m[c]; }
};

__device__ __forceinline__ Mat3 mat3_cols(float a0, float a1, float a2, float b0, float b1, float b2,
                                          float c0, float c1, float c2) {
  Mat3 r;
  r.m[0][0] = a0; r.m[0][1] = a1; r.m[0][2] = a2;
  r.m[1][0] = b0; r.m[1][1] = b1; r.m[1][2] = b2;
  r.m[2][0] = c0; r.m[2][1] = c1; r.m[2][2] = c2;
  return r;
}

__device__ __forceinline__ Mat3 mat3_zero() { return mat3_cols(0, 0, 0, 0, 0, 0, 0, 0, 0); }

__device__ __forceinline__ Mat3 mat3_mul(const Mat3& a, const Mat3& b) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int row = 0; row < 3; ++row) {
      r.m[c][row] = a.m[0][row] * b.m[c][0] + a.m[1][row] * b.m[c][1] + a.m[2][row] * b.m[c][2];
    }
  }
  return r;
}

__device__ __forceinline__ Mat3 mat3_t(const Mat3& a) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int row = 0; row < 3; ++row) r.m[c][row] = a.m[row][c];
  return r;
}

__device__ __forceinline__ float3 mat3_vec(const Mat3& a, const float3 v) {
  return make_float3(a.m[0][0] * v.x + a.m[1][0] * v.y + a.m[2][0] * v.z,
                     a.m[0][1] * v.x + a.m[1][1] * v.y + a.m[2][1] * v.z,
                     a.m[0][2] * v.x + a.m[1][2] * v.y + a.m[2][2] * v.z);
}

// outer(c, r)[col i][row j] = c[j] * r[i]
__device__ __forceinline__ Mat3 mat3_outer(const float3 c, const float3 r) {
  return mat3_cols(c.x * r.x, c.y * r.x, c.z * r.x, c.x * r.y, c.y * r.y, c.z * r.y, c.x * r.z,
                   c.y * r.z, c.z * r.z);
}

__device__ __forceinline__ Mat3 mat3_scale(const Mat3& a, float s) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int row = 0; row < 3; ++row) r.m[c][row] = a.m[c][row] * s;
  return r;
}

__device__ __forceinline__ Mat3 mat3_add(const Mat3& a, const Mat3& b) {
  Mat3 r;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int row = 0; row < 3; ++row) r.m[c][row] = a.m[c][row] + b.m[c][row];
  return r;
}

__device__ __forceinline__ float dot3(const float3 a, const float3 b) {
  const float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
  return tx + ty + tz;
}
__device__ __forceinline__ float3 col3(const Mat3& a, int c) {
  return make_float3(a.m[c][0], a.m[c][1], a.m[c][2]);
}
__device__ __forceinline__ float3 f3_scale(const float3 a, float s) {
  return make_float3(a.x * s, a.y * s, a.z * s);
}
__device__ __forceinline__ float3 f3_add(const float3 a, const float3 b) {
  return make_float3(a.x + b.x, a.y + b.y, a.z + b.z);
}
__device__ __forceinline__ float3 f3_sub(const float3 a, const float3 b) {
  return make_float3(a.x - b.x, a.y - b.y, a.z - b.z);
}
// v * (1/sqrt(dot(v,v)))
__device__ __forceinline__ float3 f3_normalize(const float3 v) {
  const float inv = 1.0f / sqrtf(dot3(v, v));
  return f3_scale(v, inv);
}

// ---- point transforms (row-vector convention, matrices stored transposed) -----------------
__device__ __forceinline__ float3 xform_4x3(const float3 p, const float* __restrict__ mat) {
  return make_float3(mat[0] * p.x + mat[4] * p.y + mat[8] * p.z + mat[12],
                     mat[1] * p.x + mat[5] * p.y + mat[9] * p.z + mat[13],
                     mat[2] * p.x + mat[6] * p.y + mat[10] * p.z + mat[14]);
}
__device__ __forceinline__ float4 xform_4x4(const float3 p, const float* __restrict__ mat) {
  return make_float4(mat[0] * p.x + mat[4] * p.y + mat[8] * p.z + mat[12],
                     mat[1] * p.x + mat[5] * p.y + mat[9] * p.z + mat[13],
                     mat[2] * p.x + mat[6] * p.y + mat[10] * p.z + mat[14],
                     mat[3] * p.x + mat[7] * p.y + mat[11] * p.z + mat[15]);
}
__device__ __forceinline__ float3 xform_vec_4x3_t(const float3 p, const float* __restrict__ mat) {
  return make_float3(mat[0] * p.x + mat[1] * p.y + mat[2] * p.z,
                     mat[4] * p.x + mat[5] * p.y + mat[6] * p.z,
                     mat[8] * p.x + mat[9] * p.y + mat[10] * p.z);
}

// NDC -> pixel centre; the literals are double on purpose (auxiliary.h:57-60).
__device__ __forceinline__ float ndc_to_pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Tile rectangle touched by a disc of integer radius (auxiliary.h:62-72).
__device__ __forceinline__ void tile_rect(const float2 p, int max_radius, uint2& rmin, uint2& rmax,
                                          const uint32_t gx, const uint32_t gy) {
  rmin.x = min(gx, (uint32_t)max((int)0, (int)((p.x - max_radius) / TILE)));
  rmin.y = min(gy, (uint32_t)max((int)0, (int)((p.y - max_radius) / TILE)));
  rmax.x = min(gx, (uint32_t)max((int)0, (int)((p.x + max_radius + TILE - 1) / TILE)));
  rmax.y = min(gy, (uint32_t)max((int)0, (int)((p.y + max_radius + TILE - 1) / TILE)));
}

// ---- covariance from scale + (un-normalised) quaternion ------------------------------------
__device__ __forceinline__ Mat3 quat_to_R(const float4 q) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  return mat3_cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                   2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                   2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// Sigma = (S R)^T (S R); upper triangle out (forward.cu:270-304).
__device__ __forceinline__ void cov3d_from_scale_rot(const float3 scale, float mod, const float4 q,
                                                     float* cov6) {
  Mat3 S = mat3_cols(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
  S.m[0][0] = mod * scale.x;
  S.m[1][1] = mod * scale.y;
  S.m[2][2] = mod * scale.z;
  const Mat3 R = quat_to_R(q);
  const Mat3 M = mat3_mul(S, R);
  const Mat3 Sigma = mat3_mul(mat3_t(M), M);
  cov6[0] = Sigma.m[0][0];
  cov6[1] = Sigma.m[0][1];
  cov6[2] = Sigma.m[0][2];
  cov6[3] = Sigma.m[1][1];
  cov6[4] = Sigma.m[1][2];
  cov6[5] = Sigma.m[2][2];
}

// ---- symmetric 3x3 eigen-decomposition ------------------------------------------------------
// Householder tridiagonalisation followed by implicit-shift QL, with the reference's absolute
// 1e-7 "is zero" tests and 30-iteration cap (auxiliary.h:217-401). Specialised to 3x3: the
// Householder stage only ever reduces the last row. Returns 3, or 0 when QL did not converge.
// evec[c] is the eigenvector of eval[c].
__device__ __forceinline__ bool near_zero(float x) { return fabsf(x) <= 0.0000001f; }

__device__ __forceinline__ float hypot_nr(float a, float b) {
  float absa = fabsf(a), absb = fabsf(b);
  if (absa > absb) {
    absb /= absa;
    absb *= absb;
    return absa * sqrtf(1.0f + absb);
  }
  if (near_zero(absb)) return 0.0f;
  absa /= absb;
  absa *= absa;
  return absb * sqrtf(1.0f + absa);
}

static __device__ __noinline__ int eig_sym3(const Mat3& A, float* eval, Mat3& evec) {
  float a[3][3];  // a[r][c], overwritten by the accumulated transform
  float d[3], e[3];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) a[r][c] = A.m[c][r];

  // --- stage 1: reduce row 2 ---
  {
    float h = 0.f;
    float scale = 0.f;
    scale += fabsf(a[2][0]);
    scale += fabsf(a[2][1]);
    if (near_zero(scale)) {
      e[2] = a[2][1];
    } else {
      a[2][0] /= scale;
      h += a[2][0] * a[2][0];
      a[2][1] /= scale;
      h += a[2][1] * a[2][1];
      float f = a[2][1];
      float g = (f >= 0) ? -sqrtf(h) : sqrtf(h);
      e[2] = scale * g;
      h -= f * g;
      a[2][1] = f - g;
      f = 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        a[j][2] = a[2][j] / h;
        g = 0.f;
        for (int k = 0; k <= j; ++k) g += a[j][k] * a[2][k];
        for (int k = j + 1; k < 2; ++k) g += a[k][j] * a[2][k];
        e[j] = g / h;
        f += e[j] * a[2][j];
      }
      const float hh = f / (h + h);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        f = a[2][j];
        e[j] = g = e[j] - hh * f;
        for (int k = 0; k <= j; ++k) a[j][k] -= (f * e[k] + g * a[2][k]);
      }
    }
    d[2] = h;
  }
  e[1] = a[1][0];
  d[1] = 0.f;
  d[0] = 0.f;
  e[0] = 0.f;

  // --- accumulate the transform ---
  d[0] = a[0][0];
  a[0][0] = 1.f;
  d[1] = a[1][1];
  a[1][1] = 1.f;
  a[0][1] = a[1][0] = 0.f;
  if (!near_zero(d[2])) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float g = 0.f;
      for (int k = 0; k < 2; ++k) g += a[2][k] * a[k][j];
      for (int k = 0; k < 2; ++k) a[k][j] -= g * a[k][2];
    }
  }
  d[2] = a[2][2];
  a[2][2] = 1.f;
  a[0][2] = a[2][0] = 0.f;
  a[1][2] = a[2][1] = 0.f;

  // --- stage 2: QL with implicit shifts (0-based l, m) ---
  e[0] = e[1];
  e[1] = e[2];
  e[2] = 0.f;
  for (int l = 0; l < 3; ++l) {
    int iter = 0;
    int m;
    do {
      for (m = l; m < 2; ++m) {
        if (near_zero(fabsf(e[m]))) break;
      }
      if (m != l) {
        if (iter++ == 30) return 0;
        float g = (d[l + 1] - d[l]) / (2 * e[l]);
        float r = hypot_nr(g, 1.0f);
        const float sr = (g >= 0) ? fabsf(r) : -fabsf(r);
        g = d[m] - d[l] + e[l] / (g + sr);
        float s = 1.f, c = 1.f, p = 0.f;
        int i;
        for (i = m - 1; i >= l; --i) {
          float f = s * e[i];
          const float b = c * e[i];
          e[i + 1] = r = hypot_nr(f, g);
          if (near_zero(r)) {
            d[i + 1] -= p;
            e[m] = 0.f;
            break;
          }
          s = f / r;
          c = g / r;
          g = d[i + 1] - p;
          r = (d[i] - g) * s + 2 * c * b;
          d[i + 1] = g + (p = s * r);
          g = c * r - b;
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            f = a[k][i + 1];
            a[k][i + 1] = s * a[k][i] + c * f;
            a[k][i] = c * a[k][i] - s * f;
          }
        }
        if (near_zero(r) && (i >= l)) continue;
        d[l] -= p;
        e[l] = g;
        e[m] = 0.f;
      }
    } while (m != l);
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) eval[i] = d[i];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) evec.m[i][j] = a[j][i];
  return 3;
}

__device__ __forceinline__ int argmin3(const float* v) {
  return v[0] > v[1] ? (v[1] > v[2] ? 2 : 1) : (v[0] > v[2] ? 2 : 0);
}

// ---- spherical harmonics (auxiliary.h:35-52) -------------------------------------------------
__device__ constexpr float kSH0 = 0.28209479177387814f;
__device__ constexpr float kSH1 = 0.4886025119029199f;
__device__ constexpr float kSH2[5] = {1.0925484305920792f, -1.0925484305920792f,
                                      0.31539156525252005f, -1.0925484305920792f,
                                      0.5462742152960396f};
__device__ constexpr float kSH3[7] = {-0.5900435899266435f, 2.890611442640554f,
                                      -0.4570457994644658f, 0.3731763325901154f,
                                      -0.4570457994644658f, 1.445305721320277f,
                                      -0.5900435899266435f};

// Everything computeCov2D derives from (mean, cov3D, camera); shared by forward and backward.
struct Cov2DCtx {
  float3 t;           // clamped view-space position
  float txtz, tytz;   // after clamping
  float x_mul, y_mul; // 0 when the clamp was active (backward only)
  Mat3 W, T, Vrk, cov;
};

__device__ __forceinline__ void cov2d_setup(const float3 mean, float fx, float fy, float tan_fovx,
                                            float tan_fovy, const float* cov6,
                                            const float* __restrict__ view, Cov2DCtx& o) {
  float3 t = xform_4x3(mean, view);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  float txtz = t.x / t.z;
  float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  o.x_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  o.y_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  o.txtz = t.x / t.z;
  o.tytz = t.y / t.z;
  o.t = t;
  const Mat3 J = mat3_cols(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z), 0.0f, fy / t.z,
                           -(fy * t.y) / (t.z * t.z), 0, 0, 0);
  o.W = mat3_cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  o.T = mat3_mul(o.W, J);
  o.Vrk = mat3_cols(cov6[0], cov6[1], cov6[2], cov6[1], cov6[3], cov6[4], cov6[2], cov6[4], cov6[5]);
  o.cov = mat3_mul(mat3_mul(mat3_t(o.T), mat3_t(o.Vrk)), o.T);
}

}  // namespace ed3dgs
