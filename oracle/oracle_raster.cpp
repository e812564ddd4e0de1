// TEST INFRASTRUCTURE -- plain C++ (host, OpenMP) restatement of the reference rasterizer.
//
// The reference has no CPU rasterizer (its only implementation is CUDA), so this file restates
// the algorithm of submodules/diff-gaussian-rasterization/cuda_rasterizer one function at a time,
// in scalar fp32, citing the lines it follows. It is the checker of the `-m "not gpu"` tests, of
// __graft_entry__.smoke() and the timed CPU baseline of bench.py. It is pinned against golden
// vectors produced by the UNMODIFIED reference CUDA kernels on a B200 (tests/golden/raster_*.pt,
// script tests/golden/make_raster_golden.py). Differences to the GPU are limited to FMA
// contraction / libm rounding (this file is compiled with -ffp-contract=off).
// Nothing on the product path may include, link or call this file.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

namespace {

constexpr int BLOCK_X = 16, BLOCK_Y = 16;  // config.h:16-17

struct V3 { float x, y, z; };
struct M3 {  // column-major, m[c][r], like glm
  float m[3][3];
};

M3 mul(const M3& a, const M3& b) {  // glm mat3*mat3 (type_mat3x3.inl)
  M3 r;
  for (int c = 0; c < 3; ++c)
    for (int row = 0; row < 3; ++row)
      r.m[c][row] = a.m[0][row] * b.m[c][0] + a.m[1][row] * b.m[c][1] + a.m[2][row] * b.m[c][2];
  return r;
}
M3 tr(const M3& a) {
  M3 r;
  for (int c = 0; c < 3; ++c)
    for (int row = 0; row < 3; ++row) r.m[c][row] = a.m[row][c];
  return r;
}
V3 mulv(const M3& a, V3 v) {
  return {a.m[0][0] * v.x + a.m[1][0] * v.y + a.m[2][0] * v.z, a.m[0][1] * v.x + a.m[1][1] * v.y + a.m[2][1] * v.z,
          a.m[0][2] * v.x + a.m[1][2] * v.y + a.m[2][2] * v.z};
}
M3 cols(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1, float c2) {
  M3 r;
  r.m[0][0] = a0; r.m[0][1] = a1; r.m[0][2] = a2;
  r.m[1][0] = b0; r.m[1][1] = b1; r.m[1][2] = b2;
  r.m[2][0] = c0; r.m[2][1] = c1; r.m[2][2] = c2;
  return r;
}
M3 outer(V3 c, V3 r) {  // glm::outerProduct: m[i] = c * r[i]
  return cols(c.x * r.x, c.y * r.x, c.z * r.x, c.x * r.y, c.y * r.y, c.z * r.y, c.x * r.z, c.y * r.z, c.z * r.z);
}
M3 zero3() { return cols(0, 0, 0, 0, 0, 0, 0, 0, 0); }
M3 add(const M3& a, const M3& b) {
  M3 r;
  for (int c = 0; c < 3; ++c)
    for (int k = 0; k < 3; ++k) r.m[c][k] = a.m[c][k] + b.m[c][k];
  return r;
}
float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3 col(const M3& a, int c) { return {a.m[c][0], a.m[c][1], a.m[c][2]}; }
V3 scl(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
V3 addv(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
V3 subv(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
V3 normalize(V3 v) { return scl(v, 1.0f / std::sqrt(dot(v, v))); }

// auxiliary.h:74-113
V3 transformPoint4x3(V3 p, const float* m) {
  return {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
}
void transformPoint4x4(V3 p, const float* m, float* o) {
  o[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  o[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  o[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  o[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}
V3 transformVec4x3Transpose(V3 p, const float* m) {
  return {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
          m[8] * p.x + m[9] * p.y + m[10] * p.z};
}
float ndc2Pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }  // auxiliary.h:57-60

void getRect(float px, float py, int max_radius, uint32_t* rmin, uint32_t* rmax, uint32_t gx, uint32_t gy) {  // :62-72
  rmin[0] = std::min(gx, (uint32_t)std::max(0, (int)((px - max_radius) / BLOCK_X)));
  rmin[1] = std::min(gy, (uint32_t)std::max(0, (int)((py - max_radius) / BLOCK_Y)));
  rmax[0] = std::min(gx, (uint32_t)std::max(0, (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X)));
  rmax[1] = std::min(gy, (uint32_t)std::max(0, (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y)));
}

// auxiliary.h:189-401 (glm_modification): Householder + QL for a symmetric DxD matrix, D = 3.
bool eq0(float x) { return std::fabs(x - 0.0f) <= 0.0000001f; }
float pythag(float a, float b) {
  float absa = std::fabs(a), absb = std::fabs(b);
  if (absa > absb) { absb /= absa; absb *= absb; return absa * std::sqrt(1.0f + absb); }
  if (eq0(absb)) return 0.0f;
  absa /= absb; absa *= absa;
  return absb * std::sqrt(1.0f + absa);
}
int findEigenvaluesSymReal(const M3& A, float* outVal, M3& outVec) {
  const int D = 3;
  float a[9], d[3], e[3];
  for (int r = 0; r < D; r++)
    for (int c = 0; c < D; c++) a[r * D + c] = A.m[c][r];
  int l, k, j, i;
  float scale, hh, h, g, f;
  for (i = D; i >= 2; i--) {
    l = i - 1;
    h = scale = 0;
    if (l > 1) {
      for (k = 1; k <= l; k++) scale += std::fabs(a[(i - 1) * D + (k - 1)]);
      if (eq0(scale)) {
        e[i - 1] = a[(i - 1) * D + (l - 1)];
      } else {
        for (k = 1; k <= l; k++) {
          a[(i - 1) * D + (k - 1)] /= scale;
          h += a[(i - 1) * D + (k - 1)] * a[(i - 1) * D + (k - 1)];
        }
        f = a[(i - 1) * D + (l - 1)];
        g = ((f >= 0) ? -std::sqrt(h) : std::sqrt(h));
        e[i - 1] = scale * g;
        h -= f * g;
        a[(i - 1) * D + (l - 1)] = f - g;
        f = 0;
        for (j = 1; j <= l; j++) {
          a[(j - 1) * D + (i - 1)] = a[(i - 1) * D + (j - 1)] / h;
          g = 0;
          for (k = 1; k <= j; k++) g += a[(j - 1) * D + (k - 1)] * a[(i - 1) * D + (k - 1)];
          for (k = j + 1; k <= l; k++) g += a[(k - 1) * D + (j - 1)] * a[(i - 1) * D + (k - 1)];
          e[j - 1] = g / h;
          f += e[j - 1] * a[(i - 1) * D + (j - 1)];
        }
        hh = f / (h + h);
        for (j = 1; j <= l; j++) {
          f = a[(i - 1) * D + (j - 1)];
          e[j - 1] = g = e[j - 1] - hh * f;
          for (k = 1; k <= j; k++) a[(j - 1) * D + (k - 1)] -= (f * e[k - 1] + g * a[(i - 1) * D + (k - 1)]);
        }
      }
    } else {
      e[i - 1] = a[(i - 1) * D + (l - 1)];
    }
    d[i - 1] = h;
  }
  d[0] = 0;
  e[0] = 0;
  for (i = 1; i <= D; i++) {
    l = i - 1;
    if (!eq0(d[i - 1])) {
      for (j = 1; j <= l; j++) {
        g = 0;
        for (k = 1; k <= l; k++) g += a[(i - 1) * D + (k - 1)] * a[(k - 1) * D + (j - 1)];
        for (k = 1; k <= l; k++) a[(k - 1) * D + (j - 1)] -= g * a[(k - 1) * D + (i - 1)];
      }
    }
    d[i - 1] = a[(i - 1) * D + (i - 1)];
    a[(i - 1) * D + (i - 1)] = 1;
    for (j = 1; j <= l; j++) a[(j - 1) * D + (i - 1)] = a[(i - 1) * D + (j - 1)] = 0;
  }
  int m, iter;
  float s, r, p, c, b;
  for (i = 2; i <= D; i++) e[i - 2] = e[i - 1];
  e[D - 1] = 0;
  for (l = 1; l <= D; l++) {
    iter = 0;
    do {
      for (m = l; m <= D - 1; m++)
        if (eq0(std::fabs(e[m - 1]))) break;
      if (m != l) {
        if (iter++ == 30) return 0;
        g = (d[l] - d[l - 1]) / (2 * e[l - 1]);
        r = pythag(g, 1.0f);
        g = d[m - 1] - d[l - 1] + e[l - 1] / (g + ((g >= 0) ? std::fabs(r) : -std::fabs(r)));
        s = c = 1;
        p = 0;
        for (i = m - 1; i >= l; i--) {
          f = s * e[i - 1];
          b = c * e[i - 1];
          e[i] = r = pythag(f, g);
          if (eq0(r)) {
            d[i] -= p;
            e[m - 1] = 0;
            break;
          }
          s = f / r;
          c = g / r;
          g = d[i] - p;
          r = (d[i - 1] - g) * s + 2 * c * b;
          d[i] = g + (p = s * r);
          g = c * r - b;
          for (k = 1; k <= D; k++) {
            f = a[(k - 1) * D + i];
            a[(k - 1) * D + i] = s * a[(k - 1) * D + (i - 1)] + c * f;
            a[(k - 1) * D + (i - 1)] = c * a[(k - 1) * D + (i - 1)] - s * f;
          }
        }
        if (eq0(r) && (i >= l)) continue;
        d[l - 1] -= p;
        e[l - 1] = g;
        e[m - 1] = 0;
      }
    } while (m != l);
  }
  for (i = 0; i < D; i++) outVal[i] = d[i];
  for (i = 0; i < D; i++)
    for (j = 0; j < D; j++) outVec.m[i][j] = a[j * D + i];
  return D;
}

const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
const float SH_C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                       0.5462742152960396f};
const float SH_C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                       -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};

// forward.cu:23-74
void computeColorFromSH(int deg, V3 pos, V3 campos, const float* sh, float* rgb, bool* clamped) {
  V3 dir = subv(pos, campos);
  const float len = std::sqrt(dot(dir, dir));
  dir = {dir.x / len, dir.y / len, dir.z / len};
  const float x = dir.x, y = dir.y, z = dir.z;
  for (int c = 0; c < 3; ++c) {
    auto S = [&](int i) { return sh[3 * i + c]; };
    float r = SH_C0 * S(0);
    if (deg > 0) {
      r = r - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
      if (deg > 1) {
        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        r = r + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) + SH_C2[2] * (2.0f * zz - xx - yy) * S(6) +
            SH_C2[3] * xz * S(7) + SH_C2[4] * (xx - yy) * S(8);
        if (deg > 2) {
          r = r + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
              SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) + SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
              SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) + SH_C3[5] * z * (xx - yy) * S(14) +
              SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
        }
      }
    }
    r += 0.5f;
    clamped[c] = (r < 0);
    rgb[c] = std::max(r, 0.0f);
  }
}

// forward.cu:270-304
void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D, M3* Mout = nullptr, M3* Rout = nullptr) {
  M3 S = cols(1, 0, 0, 0, 1, 0, 0, 0, 1);
  S.m[0][0] = mod * scale[0];
  S.m[1][1] = mod * scale[1];
  S.m[2][2] = mod * scale[2];
  const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
  const M3 R = cols(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y), 2.f * (x * y + r * z),
                    1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x), 2.f * (x * z - r * y), 2.f * (y * z + r * x),
                    1.f - 2.f * (x * x + y * y));
  const M3 M = mul(S, R);
  const M3 Sigma = mul(tr(M), M);
  cov3D[0] = Sigma.m[0][0]; cov3D[1] = Sigma.m[0][1]; cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1]; cov3D[4] = Sigma.m[1][2]; cov3D[5] = Sigma.m[2][2];
  if (Mout) *Mout = M;
  if (Rout) *Rout = R;
}

struct Cov2D {  // everything forward.cu:77-124 and backward.cu:182-219 derive
  V3 t;
  float txtz, tytz, xmul, ymul;
  M3 W, T, Vrk, cov;
};
Cov2D cov2dSetup(V3 mean, float fx, float fy, float tan_fovx, float tan_fovy, const float* cov3D, const float* view) {
  Cov2D o;
  V3 t = transformPoint4x3(mean, view);
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = std::min(limx, std::max(-limx, txtz)) * t.z;
  t.y = std::min(limy, std::max(-limy, tytz)) * t.z;
  o.xmul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
  o.ymul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
  o.txtz = t.x / t.z;
  o.tytz = t.y / t.z;
  o.t = t;
  const M3 J = cols(fx / t.z, 0.0f, -(fx * t.x) / (t.z * t.z), 0.0f, fy / t.z, -(fy * t.y) / (t.z * t.z), 0, 0, 0);
  o.W = cols(view[0], view[4], view[8], view[1], view[5], view[9], view[2], view[6], view[10]);
  o.T = mul(o.W, J);
  o.Vrk = cols(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
  o.cov = mul(mul(tr(o.T), tr(o.Vrk)), o.T);
  return o;
}

struct Geom {  // GeometryState (rasterizer_impl.h:29-49), AoS on the host
  float depth, ts, means2D[2], conic_opacity[4], rgb[3], camera_plane[6], ray_plane[2], normal[3], view_point[3], cov3D[6];
  float tongue;
  bool clamped[3];
  uint32_t tiles_touched;
  float invraycov[6];  // integration flavour only (computeCov2D<true>, forward.cu:187-236)
  bool condition;
};

struct Ctx {
  int P = 0, W = 0, H = 0, R = 0;
  std::vector<Geom> geom;
  std::vector<uint64_t> keys_unsorted, keys_sorted;
  std::vector<uint32_t> vals_unsorted, point_list, point_offsets;
  std::vector<uint32_t> ranges;     // [tiles][2]
  std::vector<uint32_t> n_contrib;  // [2][H*W]
  std::vector<float> accum_coord, accum_depth, normal_length;
};

// rasterizer_impl.cu:35-50
uint32_t getHigherMsb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

}  // namespace

extern "C" {

void* orc_create() { return new Ctx(); }
void orc_destroy(void* c) { delete static_cast<Ctx*>(c); }

}  // extern "C"

namespace {
// The part Rasterizer::forward (rasterizer_impl.cu:255-400) and Rasterizer::integrate (:580-712) share:
// per-Gaussian preprocess, scan, key emission, sort, tile ranges. inte = preprocessCUDA<C, true>
// (forward.cu:307-424: no tongue class, invraycov3Ds + condition outputs).
int orc_front(Ctx& ctx, int P, int D, int M, int W, int H, const float* means3D, const float* shs,
              const float* colors_precomp, const float* opacities, const float* tongue_class, const float* scales,
              float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
              const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size,
              int* radii, bool inte) {
  ctx.P = P; ctx.W = W; ctx.H = H;
  const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
  const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  ctx.geom.assign(P, Geom{});
  const V3 campos = {cam_pos[0], cam_pos[1], cam_pos[2]};

  // ---- preprocessCUDATongue (forward.cu:427-545) ----
#pragma omp parallel for schedule(dynamic, 256)
  for (int idx = 0; idx < P; ++idx) {
    Geom& g = ctx.geom[idx];
    radii[idx] = 0;
    g.tiles_touched = 0;
    const V3 p_orig = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    float p_hom[4];
    transformPoint4x4(p_orig, projmatrix, p_hom);
    const float p_w = 1.0f / (p_hom[3] + 0.0000001f);
    const V3 p_proj = {p_hom[0] * p_w, p_hom[1] * p_w, p_hom[2] * p_w};
    const V3 p_view = transformPoint4x3(p_orig, viewmatrix);
    if (p_view.z <= 0.2f) continue;  // in_frustum, auxiliary.h:155-180
    if (cov3D_precomp) std::memcpy(g.cov3D, cov3D_precomp + 6 * idx, 24);
    else computeCov3D(scales + 3 * idx, scale_modifier, rotations + 4 * idx, g.cov3D);

    // computeCov2D (forward.cu:77-264)
    const Cov2D c = cov2dSetup(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, g.cov3D, viewmatrix);
    const float cov_x = float(c.cov.m[0][0] + kernel_size), cov_y = float(c.cov.m[0][1]),
                cov_z = float(c.cov.m[1][1] + kernel_size);
    const float det_0 = (float)std::max(1e-6, (double)(c.cov.m[0][0] * c.cov.m[1][1] - c.cov.m[0][1] * c.cov.m[0][1]));
    const float det_1 = (float)std::max(1e-6, (double)((c.cov.m[0][0] + kernel_size) * (c.cov.m[1][1] + kernel_size) -
                                                       c.cov.m[0][1] * c.cov.m[0][1]));
    float coef = (float)std::sqrt(det_0 / (det_1 + 1e-6) + 1e-6);
    if (det_0 <= 1e-6 || det_1 <= 1e-6) coef = 0.0f;
    M3 evec;
    float eval[3];
    const int Dn = findEigenvaluesSymReal(c.Vrk, eval, evec);
    const int min_id = eval[0] > eval[1] ? (eval[1] > eval[2] ? 2 : 1) : (eval[0] > eval[2] ? 2 : 0);
    M3 Vrk_inv;
    const bool well = eval[min_id] > 0.00000001;
    if (well) {
      const M3 diag = cols(1 / eval[0], 0, 0, 0, 1 / eval[1], 0, 0, 0, 1 / eval[2]);
      Vrk_inv = mul(mul(evec, diag), tr(evec));
    } else {
      const V3 vmin = col(evec, min_id);
      Vrk_inv = outer(vmin, vmin);
    }
    const M3 cov_cam_inv = mul(mul(tr(c.W), Vrk_inv), c.W);
    const V3 uvh = {c.txtz, c.tytz, 1};
    const V3 uvh_m = mulv(cov_cam_inv, uvh);
    const V3 uvh_mn = normalize(uvh_m);
    if (std::isnan(uvh_mn.x) || Dn == 0) {
      for (int ch = 0; ch < 6; ch++) g.camera_plane[ch] = 0;
      g.normal[0] = g.normal[1] = g.normal[2] = 0;
      g.ray_plane[0] = g.ray_plane[1] = 0;
    } else {
      const V3 t = c.t;
      const float u2 = c.txtz * c.txtz, v2 = c.tytz * c.tytz, uv = c.txtz * c.tytz;
      const float l = std::sqrt(t.x * t.x + t.y * t.y + t.z * t.z);
      const M3 nJ = cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z), t.x / l, t.y / l, t.z / l);
      const M3 nJ_inv = cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -c.txtz, -c.tytz, 0);
      if (inte) {  // forward.cu:187-236
        M3 inv_cov_ray;
        if (well) {
          const float ltz = u2 + v2 + 1;
          const float f = t.z / (u2 + v2 + 1);
          M3 full = cols(v2 + 1, -uv, c.txtz / l * ltz, -uv, u2 + 1, c.tytz / l * ltz, -c.txtz, -c.tytz, 1 / l * ltz);
          for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) full.m[a][b] = full.m[a][b] * f;
          const M3 T2 = mul(c.W, tr(full));
          inv_cov_ray = mul(mul(tr(T2), Vrk_inv), T2);
        } else {
          // The reference assigns this branch's result to a shadowed local (:222) and stores an uninitialised
          // matrix; the value it meant to store is restated here. Tests avoid ill-conditioned covariances.
          const M3 T2 = mul(c.W, nJ);
          const M3 cov_ray = mul(mul(tr(T2), Vrk_inv), T2);
          M3 rvec;
          float rval[3];
          findEigenvaluesSymReal(cov_ray, rval, rvec);
          const int mn = rval[0] > rval[1] ? (rval[1] > rval[2] ? 2 : 1) : (rval[0] > rval[2] ? 2 : 0);
          const int i1 = (mn + 1) % 3, i2 = (mn + 2) % 3;
          const float l1 = rval[i1], l2 = rval[i2];
          M3 nv;
          for (int r = 0; r < 3; ++r) { nv.m[0][r] = rvec.m[i1][r]; nv.m[1][r] = rvec.m[i2][r]; nv.m[2][r] = rvec.m[mn][r]; }
          const float r3[3] = {nv.m[0][2], nv.m[1][2], nv.m[2][2]};
          const M3 c2 = cols(1 / l1, 0, -r3[0] / r3[2] / l1, 0, 1 / l2, -r3[1] / r3[2] / l2, -r3[0] / r3[2] / l1,
                             -r3[1] / r3[2] / l2, 0);
          inv_cov_ray = mul(mul(nv, c2), tr(nv));
        }
        const M3 sc = cols(1 / focal_x, 0, 0, 0, 1 / focal_y, 0, 0, 0, 1);
        inv_cov_ray = mul(mul(sc, inv_cov_ray), sc);
        g.invraycov[0] = inv_cov_ray.m[0][0]; g.invraycov[1] = inv_cov_ray.m[0][1]; g.invraycov[2] = inv_cov_ray.m[0][2];
        g.invraycov[3] = inv_cov_ray.m[1][1]; g.invraycov[4] = inv_cov_ray.m[1][2]; g.invraycov[5] = inv_cov_ray.m[2][2];
      }
      const float vbn = dot(uvh_mn, uvh);
      const float factor_normal = l / (u2 + v2 + 1);
      const float vb = std::max(vbn, 0.0000001f);
      const V3 plane = mulv(nJ_inv, V3{uvh_mn.x / vb, uvh_mn.y / vb, uvh_mn.z / vb});
      const float nl = u2 + v2 + 1;
      g.camera_plane[0] = (-(v2 + 1) * t.z + plane.x * t.x) / nl / focal_x;
      g.camera_plane[1] = (uv * t.z + plane.y * t.x) / nl / focal_y;
      g.camera_plane[2] = (uv * t.z + plane.x * t.y) / nl / focal_x;
      g.camera_plane[3] = (-(u2 + 1) * t.z + plane.y * t.y) / nl / focal_y;
      g.camera_plane[4] = (t.x + plane.x * t.z) / nl / focal_x;
      g.camera_plane[5] = (t.y + plane.y * t.z) / nl / focal_y;
      g.ray_plane[0] = plane.x * l / nl / focal_x;
      g.ray_plane[1] = plane.y * l / nl / focal_y;
      const V3 ray_n = {-plane.x * factor_normal, -plane.y * factor_normal, -1};
      const V3 nv = normalize(mulv(nJ, ray_n));
      g.normal[0] = nv.x; g.normal[1] = nv.y; g.normal[2] = nv.z;
    }
    g.condition = well;
    // nvcc contracts x*x+y*y+z*z to fma(z,z,fma(x,x,y*y)); the integration subtracts nearly equal depths, so the
    // oracle reproduces that rounding there (one ulp of ts moves alpha by ~1e-4)
    g.ts = inte ? std::sqrt(std::fmaf(p_view.z, p_view.z, std::fmaf(p_view.x, p_view.x, p_view.y * p_view.y)))
                : std::sqrt(p_view.x * p_view.x + p_view.y * p_view.y + p_view.z * p_view.z);
    const float det = (cov_x * cov_z - cov_y * cov_y);
    if (det == 0.0f) continue;
    const float det_inv = 1.f / det;
    const float conic[3] = {cov_z * det_inv, -cov_y * det_inv, cov_x * det_inv};
    const float mid = 0.5f * (cov_x + cov_z);
    const float lambda1 = mid + std::sqrt(std::max(0.1f, mid * mid - det));
    const float lambda2 = mid - std::sqrt(std::max(0.1f, mid * mid - det));
    const float my_radius = std::ceil(3.f * std::sqrt(std::max(lambda1, lambda2)));
    const float pix[2] = {ndc2Pix(p_proj.x, W), ndc2Pix(p_proj.y, H)};
    uint32_t rmin[2], rmax[2];
    getRect(pix[0], pix[1], (int)my_radius, rmin, rmax, gx, gy);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
    if (!colors_precomp) computeColorFromSH(D, p_orig, campos, shs + (size_t)idx * M * 3, g.rgb, g.clamped);
    else std::memcpy(g.rgb, colors_precomp + 3 * idx, 12);
    g.depth = p_view.z;
    g.view_point[0] = p_view.x; g.view_point[1] = p_view.y; g.view_point[2] = p_view.z;
    radii[idx] = (int)my_radius;
    g.means2D[0] = pix[0]; g.means2D[1] = pix[1];
    g.conic_opacity[0] = conic[0]; g.conic_opacity[1] = conic[1]; g.conic_opacity[2] = conic[2];
    g.conic_opacity[3] = opacities[idx] * coef;
    g.tiles_touched = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
    g.tongue = inte ? 0.f : tongue_class[idx];
  }

  // ---- InclusiveSum (rasterizer_impl.cu:355) + duplicateWithKeys (:70-111) ----
  ctx.point_offsets.resize(P);
  uint32_t run = 0;
  for (int i = 0; i < P; ++i) { run += ctx.geom[i].tiles_touched; ctx.point_offsets[i] = run; }
  const int R = (int)run;
  ctx.R = R;
  ctx.keys_unsorted.assign(R, 0);
  ctx.vals_unsorted.assign(R, 0);
#pragma omp parallel for schedule(dynamic, 1024)
  for (int idx = 0; idx < P; ++idx) {
    if (radii[idx] > 0) {
      uint32_t off = (idx == 0) ? 0 : ctx.point_offsets[idx - 1];
      uint32_t rmin[2], rmax[2];
      const Geom& g = ctx.geom[idx];
      getRect(g.means2D[0], g.means2D[1], radii[idx], rmin, rmax, gx, gy);
      uint32_t dbits;
      std::memcpy(&dbits, &g.depth, 4);
      for (uint32_t y = rmin[1]; y < rmax[1]; y++)
        for (uint32_t x = rmin[0]; x < rmax[0]; x++) {
          uint64_t key = y * gx + x;
          key <<= 32;
          key |= dbits;
          ctx.keys_unsorted[off] = key;
          ctx.vals_unsorted[off] = idx;
          off++;
        }
    }
  }
  // ---- SortPairs over bits [0, 32+bit) (:378-386): stable ----
  const int bit = (int)getHigherMsb(gx * gy);
  const uint64_t mask = (bit + 32 >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
  std::vector<uint32_t> order(R);
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
    return (ctx.keys_unsorted[a] & mask) < (ctx.keys_unsorted[b] & mask);
  });
  ctx.keys_sorted.resize(R);
  ctx.point_list.resize(R);
  for (int i = 0; i < R; ++i) { ctx.keys_sorted[i] = ctx.keys_unsorted[order[i]]; ctx.point_list[i] = ctx.vals_unsorted[order[i]]; }
  // ---- identifyTileRanges (:151-173) ----
  ctx.ranges.assign((size_t)gx * gy * 2, 0);
  for (int idx = 0; idx < R; ++idx) {
    const uint32_t cur = (uint32_t)(ctx.keys_sorted[idx] >> 32);
    if (idx == 0) ctx.ranges[2 * cur] = 0;
    else {
      const uint32_t prev = (uint32_t)(ctx.keys_sorted[idx - 1] >> 32);
      if (cur != prev) { ctx.ranges[2 * prev + 1] = idx; ctx.ranges[2 * cur] = idx; }
    }
    if (idx == R - 1) ctx.ranges[2 * cur + 1] = R;
  }
  return R;
}
}  // namespace

extern "C" {

// CudaRasterizer::Rasterizer::forward (rasterizer_impl.cu:255-432) on host pointers.
int orc_forward(void* cv, int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* tongue_class, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy, float kernel_size,
                float* out_color, float* out_coord, float* out_mcoord, float* out_depth, float* out_mdepth,
                float* out_alpha, float* out_tongue, float* out_normal, int* radii, int require_coord,
                int require_depth) {
  Ctx& ctx = *static_cast<Ctx*>(cv);
  const int R = orc_front(ctx, P, D, M, W, H, means3D, shs, colors_precomp, opacities, tongue_class, scales,
                          scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy,
                          kernel_size, radii, false);
  const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
  const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const size_t HW = (size_t)H * W;

  // ---- renderCUDA forward (forward.cu:550-822), one pixel at a time ----
  const bool COORD = require_coord, DEPTH = require_depth, NORMAL = require_coord || require_depth;
  const bool GEO = NORMAL;
  ctx.n_contrib.assign(2 * HW, 0);
  ctx.accum_coord.assign(3 * HW, 0);
  ctx.accum_depth.assign(HW, 0);
  ctx.normal_length.assign(HW, 0);
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < (int)(gx * gy); ++tile) {
    const uint32_t tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ctx.ranges[2 * tile], r1 = ctx.ranges[2 * tile + 1];
    for (uint32_t ly = 0; ly < BLOCK_Y; ++ly)
      for (uint32_t lx = 0; lx < BLOCK_X; ++lx) {
        const uint32_t px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
        const size_t pix_id = (size_t)W * py + px;
        const float pixf[2] = {(float)px, (float)py};
        const float pixnf[2] = {(pixf[0] - W / 2.f) / focal_x, (pixf[1] - H / 2.f) / focal_y};
        const float ln = std::sqrt(pixnf[0] * pixnf[0] + pixnf[1] * pixnf[1] + 1);
        float T = 1.0f;
        uint32_t contributor = 0, last_contributor = 0, max_contributor = (uint32_t)-1;
        float C[3] = {0, 0, 0}, tongue = 0, weight = 0, Coord[3] = {0, 0, 0}, mCoord[3] = {0, 0, 0}, Depth = 0,
              mDepth = 0, Normal[3] = {0, 0, 0};
        for (uint32_t i = r0; i < r1; ++i) {
          contributor++;
          const Geom& g = ctx.geom[ctx.point_list[i]];
          const float d[2] = {g.means2D[0] - pixf[0], g.means2D[1] - pixf[1]};
          const float* con_o = g.conic_opacity;
          const float power = -0.5f * (con_o[0] * d[0] * d[0] + con_o[2] * d[1] * d[1]) - con_o[1] * d[0] * d[1];
          if (power > 0.0f) continue;
          const float alpha = std::min(0.99f, con_o[3] * std::exp(power));
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = T * (1 - alpha);
          if (test_T < 0.0001f) break;  // done = true
          const float aT = alpha * T;
          for (int ch = 0; ch < 3; ch++) C[ch] += g.rgb[ch] * aT;
          tongue += g.tongue * aT;
          const bool before_median = T > 0.5;
          if (COORD) {
            const float coord[3] = {g.view_point[0] + g.camera_plane[0] * d[0] + g.camera_plane[1] * d[1],
                                    g.view_point[1] + g.camera_plane[2] * d[0] + g.camera_plane[3] * d[1],
                                    g.view_point[2] + g.camera_plane[4] * d[0] + g.camera_plane[5] * d[1]};
            for (int ch = 0; ch < 3; ch++) Coord[ch] += coord[ch] * aT;
            if (before_median)
              for (int ch = 0; ch < 3; ch++) mCoord[ch] = coord[ch];
          }
          if (DEPTH) {
            const float t = g.ts + (g.ray_plane[0] * d[0] + g.ray_plane[1] * d[1]);
            Depth += t * aT;
            if (before_median) mDepth = t;
          }
          if (NORMAL)
            for (int ch = 0; ch < 3; ch++) Normal[ch] += g.normal[ch] * aT;
          if (GEO && before_median) max_contributor = contributor;
          weight += aT;
          T = test_T;
          last_contributor = contributor;
        }
        ctx.n_contrib[pix_id] = last_contributor;
        ctx.n_contrib[pix_id + HW] = max_contributor;
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = C[ch] + T * bg[ch];
        out_tongue[pix_id] = tongue;
        out_alpha[pix_id] = weight;
        if (COORD)
          for (int ch = 0; ch < 3; ch++) {
            out_coord[ch * HW + pix_id] = last_contributor ? Coord[ch] / weight : 0;
            ctx.accum_coord[ch * HW + pix_id] = Coord[ch];
            out_mcoord[ch * HW + pix_id] = mCoord[ch];
          }
        if (DEPTH) {
          const float depth_ln = Depth / ln;
          ctx.accum_depth[pix_id] = depth_ln;
          out_depth[pix_id] = last_contributor ? depth_ln / weight : 0;
          out_mdepth[pix_id] = mDepth / ln;
        }
        if (NORMAL) {
          if (last_contributor) {
            float len_normal = std::sqrt(Normal[0] * Normal[0] + Normal[1] * Normal[1] + Normal[2] * Normal[2]);
            ctx.normal_length[pix_id] = len_normal;
            len_normal = std::max(len_normal, 1.0E-12F);
            for (int ch = 0; ch < 3; ch++) out_normal[ch * HW + pix_id] = Normal[ch] / len_normal;
          } else {
            ctx.normal_length[pix_id] = 1;
            for (int ch = 0; ch < 3; ch++) out_normal[ch * HW + pix_id] = 0;
          }
        }
      }
  }
  return R;
}

// CudaRasterizer::Rasterizer::integrate (rasterizer_impl.cu:580-851) on host pointers, restated pixel by
// pixel the way integrateCUDA (forward.cu:1109-1543) does it: first walk = five-sample visibility test that
// records the contributing list positions (uint16, at most MAX_NUM_CONTRIBUTORS*4 = 2048), then, per round of
// at most MAX_NUM_PROJECTED = 256 points that fall into the pixel, a second walk over the recorded positions.
// Outputs must be pre-filled by the caller like rasterize_points.cu:312-322 does (out_color 0, alpha 1,
// colour 0, coordinate 0, sdf -1000). invraycov[P,6] / condition[P] may be null.
int orc_integrate(void* cv, int PN, int P, int D, int M, const float* bg, int W, int H, const float* points3D,
                  const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
                  const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                  const float* viewmatrix, const float* projmatrix, const float* cam_pos, float tan_fovx,
                  float tan_fovy, float kernel_size, float* out_color, int* radii, float* out_alpha_integrated,
                  float* out_color_integrated, float* out_coordinate2d, float* out_sdf, float* invraycov,
                  uint8_t* condition) {
  Ctx& ctx = *static_cast<Ctx*>(cv);
  const int R = orc_front(ctx, P, D, M, W, H, means3D, shs, colors_precomp, opacities, nullptr, scales, scale_modifier,
                          rotations, cov3D_precomp, viewmatrix, projmatrix, cam_pos, tan_fovx, tan_fovy, kernel_size,
                          radii, true);
  const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
  const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const size_t HW = (size_t)H * W;
  for (int i = 0; i < P; ++i) {
    if (invraycov) std::memcpy(invraycov + 6 * i, ctx.geom[i].invraycov, 24);
    if (condition) condition[i] = ctx.geom[i].condition ? 1 : 0;
  }

  // ---- preprocessPointsCUDA (forward.cu:1027-1071) + createWithKeys (rasterizer_impl.cu:114-146) ----
  struct Pt { float x, y, depth; };
  std::vector<Pt> pt(PN);
  std::vector<uint64_t> pkeys;
  std::vector<uint32_t> pvals;
  for (int idx = 0; idx < PN; ++idx) {
    const V3 p_orig = {points3D[3 * idx], points3D[3 * idx + 1], points3D[3 * idx + 2]};
    const V3 p_view = transformPoint4x3(p_orig, viewmatrix);
    if (p_view.z <= 0.2f) continue;
    const float ix = (float)(focal_x * p_view.x / (p_view.z + 0.0000001f) + W / 2.);
    const float iy = (float)(focal_y * p_view.y / (p_view.z + 0.0000001f) + H / 2.);
    if (ix < 0 || ix >= W || iy < 0 || iy >= H) continue;
    pt[idx] = {ix, iy, std::sqrt(std::fmaf(p_view.z, p_view.z, std::fmaf(p_view.x, p_view.x, p_view.y * p_view.y)))};
    const int x = std::min((int)gx - 1, std::max(0, (int)(ix / BLOCK_X)));
    const int y = std::min((int)gy - 1, std::max(0, (int)(iy / BLOCK_Y)));
    uint64_t key = (uint64_t)(y * gx + x);
    key <<= 32;
    uint32_t dbits;
    std::memcpy(&dbits, &pt[idx].depth, 4);
    key |= dbits;
    pkeys.push_back(key);
    pvals.push_back((uint32_t)idx);
  }
  const int NI = (int)pkeys.size();
  const int bit = (int)getHigherMsb(gx * gy);
  const uint64_t mask = (bit + 32 >= 64) ? ~0ull : ((1ull << (32 + bit)) - 1);
  std::vector<uint32_t> order(NI);
  std::iota(order.begin(), order.end(), 0u);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return (pkeys[a] & mask) < (pkeys[b] & mask); });
  std::vector<uint32_t> plist(NI), pranges((size_t)gx * gy * 2, 0);
  for (int i = 0; i < NI; ++i) plist[i] = pvals[order[i]];
  for (int idx = 0; idx < NI; ++idx) {
    const uint32_t cur = (uint32_t)(pkeys[order[idx]] >> 32);
    if (idx == 0) pranges[2 * cur] = 0;
    else {
      const uint32_t prev = (uint32_t)(pkeys[order[idx - 1]] >> 32);
      if (cur != prev) { pranges[2 * prev + 1] = idx; pranges[2 * cur] = idx; }
    }
    if (idx == NI - 1) pranges[2 * cur + 1] = NI;
  }

  // ---- integrateCUDA (forward.cu:1109-1543) ----
  constexpr int MAXC = 512 * 4, MAXP = 256;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < (int)(gx * gy); ++tile) {
    const uint32_t tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ctx.ranges[2 * tile], r1 = ctx.ranges[2 * tile + 1];
    const uint32_t q0 = pranges[2 * tile], q1 = pranges[2 * tile + 1];
    std::vector<uint16_t> contributed_ids(MAXC);
    for (uint32_t ly = 0; ly < BLOCK_Y; ++ly)
      for (uint32_t lx = 0; lx < BLOCK_X; ++lx) {
        const uint32_t px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
        const size_t pix_id = (size_t)W * py + px;
        const float pixf[2] = {(float)px + 0.5f, (float)py + 0.5f};
        float T = 1.0f;
        uint32_t contributor = 0, last_contributor = 0, n_contrib_local = 0;
        float C[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // CHANNELS*2+2
        float mid_depth_center = 0, mid_plane[2] = {0, 0}, mid_mean2d[2] = {0, 0};
        std::fill(contributed_ids.begin(), contributed_ids.end(), (uint16_t)0);
        float corner_Ts[5] = {1.0f, 1.0f, 1.0f, 1.0f, 1.0f};
        const float offset_xs[5] = {0.0f, -0.5f, 0.5f, -0.5f, 0.5f};
        const float offset_ys[5] = {0.0f, -0.5f, -0.5f, 0.5f, 0.5f};
        for (uint32_t i = r0; i < r1; ++i) {  // :1197-1303
          contributor++;
          const Geom& g = ctx.geom[ctx.point_list[i]];
          const float* con_o = g.conic_opacity;
          const float depth_center = g.ts;
          bool used = false;
          for (int k = 0; k < 5; ++k) {
            const float d[2] = {g.means2D[0] - pixf[0] - offset_xs[k], g.means2D[1] - pixf[1] - offset_ys[k]};
            const float depth = depth_center + (g.ray_plane[0] * d[0] + g.ray_plane[1] * d[1]);
            const float power = -0.5f * (con_o[0] * d[0] * d[0] + con_o[2] * d[1] * d[1]) - con_o[1] * d[0] * d[1];
            if (power > 0.0f) continue;
            const float alpha = std::min(0.99f, con_o[3] * std::exp(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = corner_Ts[k] * (1 - alpha);
            if (test_T < 0.0001f) continue;
            if (k == 0)
              for (int ch = 0; ch < 3; ch++) C[ch] += g.rgb[ch] * alpha * T;
            if (depth > C[6]) C[6] = depth;
            if (k == 0) {
              C[7] += alpha * T;
              C[3] += depth * alpha * T;
              if (T > 0.5) {
                C[4] = depth;
                mid_depth_center = depth_center;
                mid_plane[0] = g.ray_plane[0]; mid_plane[1] = g.ray_plane[1];
                mid_mean2d[0] = g.means2D[0]; mid_mean2d[1] = g.means2D[1];
              }
              T = test_T;
            }
            corner_Ts[k] = test_T;
            used = true;
          }
          if (used) {
            last_contributor = contributor;
            contributed_ids[n_contrib_local] = (uint16_t)contributor;
            n_contrib_local += 1;
            if (n_contrib_local >= (uint32_t)MAXC) break;  // done = true (:1297-1301)
          }
        }
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix_id] = C[ch] + T * bg[ch];
        out_color[3 * HW + pix_id] = C[3];
        out_color[4 * HW + pix_id] = C[4];
        out_color[6 * HW + pix_id] = C[6];
        out_color[7 * HW + pix_id] = C[7];

        // rounds of at most 256 points of this pixel (:1337-1537). A pixel whose points fit one round stops
        // there; the reference keeps such a thread looping while another pixel of the tile needs more rounds,
        // which rewrites identical values and is not restated.
        uint32_t point_counter_last = 0;
        int total_projected = 0;
        bool point_done = false;
        while (!point_done) {
          int num_projected = 0;
          bool exceed = false;
          uint32_t point_counter = 0;
          int projected_ids[MAXP];
          float projected_xy[MAXP][2], projected_depth[MAXP];
          for (uint32_t i = q0; i < q1; ++i) {
            point_counter++;
            if (point_counter <= point_counter_last) continue;
            const Pt& p = pt[plist[i]];
            if ((p.x >= (pixf[0] - 0.5)) && (p.x < (pixf[0] + 0.5)) && (p.y >= (pixf[1] - 0.5)) && (p.y < (pixf[1] + 0.5))) {
              if (num_projected >= MAXP) { exceed = true; break; }
              projected_ids[num_projected] = (int)plist[i];
              projected_xy[num_projected][0] = p.x; projected_xy[num_projected][1] = p.y;
              projected_depth[num_projected] = p.depth;
              num_projected += 1;
            }
          }
          point_counter_last = point_counter - 1;
          point_done = !exceed;
          total_projected += num_projected;
          float point_alphas[MAXP], point_Ts[MAXP];
          for (int k = 0; k < num_projected; ++k) { point_alphas[k] = 0.f; point_Ts[k] = 1.f; }
          uint32_t num_iterated = 0;
          uint16_t num_contributed_second = 0;
          for (uint32_t i = r0; i < r1; ++i) {  // :1452-1511
            num_iterated++;
            if (num_iterated > last_contributor) break;
            if (num_iterated != (uint32_t)contributed_ids[num_contributed_second]) continue;
            num_contributed_second += 1;
            const Geom& g = ctx.geom[ctx.point_list[i]];
            const float* con_o = g.conic_opacity;
            const float depth_center = g.ts;
            const float* iv = g.invraycov;
            for (int k = 0; k < num_projected; ++k) {
              const float d[2] = {g.means2D[0] - projected_xy[k][0], g.means2D[1] - projected_xy[k][1]};
              const float depth = depth_center + std::fmaf(g.ray_plane[0], d[0], g.ray_plane[1] * d[1]);
              float alpha;
              // a x + b y + c z as the GPU evaluates the glm expression: fma(c, z, fma(a, x, b y))
              auto quad = [&](const V3& u) {
                const float mx = std::fmaf(iv[2], u.z, std::fmaf(iv[0], u.x, iv[1] * u.y));
                const float my = std::fmaf(iv[4], u.z, std::fmaf(iv[1], u.x, iv[3] * u.y));
                const float mz = std::fmaf(iv[5], u.z, std::fmaf(iv[2], u.x, iv[4] * u.y));
                return std::fmaf(u.z, mz, std::fmaf(u.x, mx, u.y * my));
              };
              if (g.condition) {
                const V3 du = {d[0], d[1], depth_center - std::min(projected_depth[k], depth)};
                const float power = -0.5f * quad(du);
                alpha = std::min(0.99f, con_o[3] * std::exp(power));
              } else if (projected_depth[k] < depth) {
                alpha = 0;
              } else {
                const V3 du = {d[0], d[1], depth_center};
                const float power = -0.5f * quad(du);
                alpha = std::min(0.99f, con_o[3] * std::exp(power));
              }
              if (alpha < 1.0f / 255.0f) continue;
              const float test_T = point_Ts[k] * (1 - alpha);
              point_alphas[k] += alpha * point_Ts[k];
              point_Ts[k] = test_T;
            }
          }
          for (int k = 0; k < num_projected; ++k) {  // :1519-1536
            const int id = projected_ids[k];
            out_alpha_integrated[id] = point_alphas[k];
            for (int ch = 0; ch < 3; ch++) out_color_integrated[3 * id + ch] = C[ch] + T * bg[ch];
            out_coordinate2d[2 * id] = projected_xy[k][0];
            out_coordinate2d[2 * id + 1] = projected_xy[k][1];
            if (projected_depth[k] > 0) {
              const float d[2] = {mid_mean2d[0] - projected_xy[k][0], mid_mean2d[1] - projected_xy[k][1]};
              const float depth = mid_depth_center + (mid_plane[0] * d[0] + mid_plane[1] * d[1]);
              out_sdf[id] = depth - projected_depth[k];
            }
          }
        }
        out_color[8 * HW + pix_id] = (float)total_projected;
      }
  }
  return R;
}

// Intermediate state in the reference's layouts (any pointer may be null).
void orc_get_state(void* cv, float* depths, float* means2D, uint32_t* tiles_touched, uint32_t* point_offsets,
                   float* conic_opacity, float* rgb, uint64_t* keys_unsorted, uint32_t* vals_unsorted,
                   uint64_t* keys_sorted, uint32_t* point_list, uint32_t* ranges, uint32_t* n_contrib, float* normals,
                   float* ray_planes, float* ts) {
  Ctx& c = *static_cast<Ctx*>(cv);
  for (int i = 0; i < c.P; ++i) {
    const Geom& g = c.geom[i];
    if (depths) depths[i] = g.depth;
    if (means2D) { means2D[2 * i] = g.means2D[0]; means2D[2 * i + 1] = g.means2D[1]; }
    if (tiles_touched) tiles_touched[i] = g.tiles_touched;
    if (point_offsets) point_offsets[i] = c.point_offsets[i];
    if (conic_opacity) std::memcpy(conic_opacity + 4 * i, g.conic_opacity, 16);
    if (rgb) std::memcpy(rgb + 3 * i, g.rgb, 12);
    if (normals) std::memcpy(normals + 3 * i, g.normal, 12);
    if (ray_planes) std::memcpy(ray_planes + 2 * i, g.ray_plane, 8);
    if (ts) ts[i] = g.ts;
  }
  auto cp = [](void* d, const void* s, size_t n) { if (d && n) std::memcpy(d, s, n); };
  cp(keys_unsorted, c.keys_unsorted.data(), c.keys_unsorted.size() * 8);
  cp(vals_unsorted, c.vals_unsorted.data(), c.vals_unsorted.size() * 4);
  cp(keys_sorted, c.keys_sorted.data(), c.keys_sorted.size() * 8);
  cp(point_list, c.point_list.data(), c.point_list.size() * 4);
  cp(ranges, c.ranges.data(), c.ranges.size() * 4);
  cp(n_contrib, c.n_contrib.data(), c.n_contrib.size() * 4);
}

// CudaRasterizer::Rasterizer::backward (rasterizer_impl.cu:436-578) on host pointers. All
// gradient outputs must be zero-filled by the caller (rasterize_points.cu:184-197).
void orc_backward(void* cv, int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                  const float* alphas, const float* scales, float scale_modifier, const float* rotations,
                  const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix, const float* campos_,
                  float tan_fovx, float tan_fovy, float kernel_size, const int* radii, const float* normalmap,
                  const float* dL_dpixels, const float* dL_dpixel_coords, const float* dL_dpixel_mcoords,
                  const float* dL_dpixel_depths, const float* dL_dpixel_mdepths, const float* dL_dalphas,
                  const float* dL_dpixel_normals, float* dL_dmean2D, float* dL_dcolors, float* dL_dopacity,
                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                  int require_coord, int require_depth) {
  Ctx& ctx = *static_cast<Ctx*>(cv);
  const float focal_y = H / (2.0f * tan_fovy), focal_x = W / (2.0f * tan_fovx);
  const uint32_t gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
  const size_t HW = (size_t)H * W;
  const bool COORD = require_coord, DEPTH = require_depth, NORMAL = require_coord || require_depth, GEO = NORMAL;
  std::vector<float> dL_dconic(4 * (size_t)P, 0.f), dL_dview_points(3 * (size_t)P, 0.f), dL_dts(P, 0.f),
      dL_dcamera_planes(6 * (size_t)P, 0.f), dL_dray_planes(2 * (size_t)P, 0.f), dL_dnormals(3 * (size_t)P, 0.f);
  auto aadd = [](float* p, float v) {
#pragma omp atomic
    *p += v;
  };

  // ---- renderCUDA backward (backward.cu:631-1016) ----
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < (int)(gx * gy); ++tile) {
    const uint32_t tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ctx.ranges[2 * tile], r1 = ctx.ranges[2 * tile + 1];
    for (uint32_t ly = 0; ly < BLOCK_Y; ++ly)
      for (uint32_t lx = 0; lx < BLOCK_X; ++lx) {
        const uint32_t px = tx * BLOCK_X + lx, py = ty * BLOCK_Y + ly;
        if (px >= (uint32_t)W || py >= (uint32_t)H) continue;
        const size_t pix_id = (size_t)W * py + px;
        const float pixf[2] = {(float)px, (float)py};
        const float pixnf[2] = {(pixf[0] - W / 2.f) / focal_x, (pixf[1] - H / 2.f) / focal_y};
        const float ln = std::sqrt(pixnf[0] * pixnf[0] + pixnf[1] * pixnf[1] + 1);
        const float T_final = 1 - alphas[pix_id], w_final = alphas[pix_id];
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const int last_contributor = (int)ctx.n_contrib[pix_id];
        const int max_contributor = (int)ctx.n_contrib[pix_id + HW];
        float accum_rec[3] = {0}, dL_dpixel[3], accum_coord_rec[3] = {0}, dL_dpixel_coord[3] = {0}, accum_t_rec = 0,
              dL_dpixel_t = 0, dL_dpixel_mt = 0, accum_alpha_rec = 0, dL_dalpha, accum_normal_rec[3] = {0},
              dL_dpixel_normal[3] = {0}, dL_dpixel_mcoord[3] = {0};
        for (int i = 0; i < 3; i++) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
        dL_dalpha = dL_dalphas[pix_id];
        if (GEO) {
          const float ww = w_final * w_final;
          if (COORD)
            for (int i = 0; i < 3; i++) {
              const float gw = dL_dpixel_coords[i * HW + pix_id];
              dL_dalpha -= gw * ctx.accum_coord[i * HW + pix_id] / ww;
              dL_dpixel_coord[i] = gw / w_final;
              dL_dpixel_mcoord[i] = dL_dpixel_mcoords[i * HW + pix_id];
            }
          if (DEPTH) {
            const float gw = dL_dpixel_depths[pix_id];
            dL_dalpha -= gw * ctx.accum_depth[pix_id] / ww;
            dL_dpixel_t = gw / w_final / ln;
            dL_dpixel_mt = dL_dpixel_mdepths[pix_id] / ln;
          }
          if (NORMAL) {
            const V3 gn = {dL_dpixel_normals[pix_id], dL_dpixel_normals[HW + pix_id], dL_dpixel_normals[2 * HW + pix_id]};
            const V3 nn = {normalmap[pix_id], normalmap[HW + pix_id], normalmap[2 * HW + pix_id]};
            const float nlen = ctx.normal_length[pix_id];
            V3 dL;
            if (nlen < 1.0E-12F) dL = {gn.x / 1.0E-12F, gn.y / 1.0E-12F, gn.z / 1.0E-12F};
            else {
              const float dp = dot(gn, nn);
              dL = {(gn.x - dp * nn.x) / nlen, (gn.y - dp * nn.y) / nlen, (gn.z - dp * nn.z) / nlen};
            }
            dL_dpixel_normal[0] = dL.x; dL_dpixel_normal[1] = dL.y; dL_dpixel_normal[2] = dL.z;
          }
        }
        float last_alpha = 0, last_color[3] = {0}, last_coord[3] = {0}, last_t = 0, last_normal[3] = {0};
        const float ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
        for (uint32_t i = r1; i-- > r0;) {
          contributor--;
          if ((int)contributor >= last_contributor) continue;
          const int global_id = (int)ctx.point_list[i];
          const Geom& g = ctx.geom[global_id];
          const float d[2] = {g.means2D[0] - pixf[0], g.means2D[1] - pixf[1]};
          const float* con_o = g.conic_opacity;
          const float power = -0.5f * (con_o[0] * d[0] * d[0] + con_o[2] * d[1] * d[1]) - con_o[1] * d[0] * d[1];
          if (power > 0.0f) continue;
          const float G = std::exp(power);
          const float alpha = std::min(0.99f, con_o[3] * G);
          if (alpha < 1.0f / 255.0f) continue;
          T = T / (1.f - alpha);
          const float dchannel_dcolor = alpha * T;
          float dL_dopa = 0.0f;
          for (int ch = 0; ch < 3; ch++) {
            const float c = g.rgb[ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = c;
            dL_dopa += (c - accum_rec[ch]) * dL_dpixel[ch];
            aadd(&dL_dcolors[global_id * 3 + ch], dchannel_dcolor * dL_dpixel[ch]);
          }
          float dL_dcoords[3] = {0, 0, 0}, dL_dt = 0;
          if (COORD) {
            const float coord[3] = {g.view_point[0] + g.camera_plane[0] * d[0] + g.camera_plane[1] * d[1],
                                    g.view_point[1] + g.camera_plane[2] * d[0] + g.camera_plane[3] * d[1],
                                    g.view_point[2] + g.camera_plane[4] * d[0] + g.camera_plane[5] * d[1]};
            for (int ch = 0; ch < 3; ch++) {
              accum_coord_rec[ch] = last_alpha * last_coord[ch] + (1.f - last_alpha) * accum_coord_rec[ch];
              last_coord[ch] = coord[ch];
              dL_dopa += (coord[ch] - accum_coord_rec[ch]) * dL_dpixel_coord[ch];
              dL_dcoords[ch] = dchannel_dcolor * dL_dpixel_coord[ch];
              if (contributor == (uint32_t)(max_contributor - 1)) dL_dcoords[ch] += dL_dpixel_mcoord[ch];
              aadd(&dL_dview_points[global_id * 3 + ch], dL_dcoords[ch]);
              aadd(&dL_dcamera_planes[global_id * 6 + 2 * ch], dL_dcoords[ch] * d[0] / focal_x);
              aadd(&dL_dcamera_planes[global_id * 6 + 2 * ch + 1], dL_dcoords[ch] * d[1] / focal_y);
            }
          }
          if (DEPTH) {
            const float t = g.ts + (g.ray_plane[0] * d[0] + g.ray_plane[1] * d[1]);
            accum_t_rec = last_alpha * last_t + (1.f - last_alpha) * accum_t_rec;
            last_t = t;
            dL_dopa += (t - accum_t_rec) * dL_dpixel_t;
            dL_dt = dchannel_dcolor * dL_dpixel_t;
            if (contributor == (uint32_t)(max_contributor - 1)) dL_dt += dL_dpixel_mt;
            aadd(&dL_dts[global_id], dL_dt);
            aadd(&dL_dray_planes[global_id * 2], dL_dt * d[0] / focal_x);
            aadd(&dL_dray_planes[global_id * 2 + 1], dL_dt * d[1] / focal_y);
          }
          if (NORMAL)
            for (int ch = 0; ch < 3; ch++) {
              const float c = g.normal[ch];
              accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
              last_normal[ch] = c;
              dL_dopa += (c - accum_normal_rec[ch]) * dL_dpixel_normal[ch];
              aadd(&dL_dnormals[global_id * 3 + ch], dchannel_dcolor * dL_dpixel_normal[ch]);
            }
          accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
          dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
          dL_dopa *= T;
          last_alpha = alpha;
          float bg_dot_dpixel = 0;
          for (int k = 0; k < 3; k++) bg_dot_dpixel += bg[k] * dL_dpixel[k];
          dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
          const float dL_dG = con_o[3] * dL_dopa;
          const float gdx = G * d[0], gdy = G * d[1];
          const float dG_ddelx = -gdx * con_o[0] - gdy * con_o[1];
          const float dG_ddely = -gdy * con_o[2] - gdx * con_o[1];
          float dL_ddelx = dL_dG * dG_ddelx, dL_ddely = dL_dG * dG_ddely;
          if (COORD) {
            dL_ddelx += dL_dcoords[0] * g.camera_plane[0] + dL_dcoords[1] * g.camera_plane[2] + dL_dcoords[2] * g.camera_plane[4];
            dL_ddely += dL_dcoords[0] * g.camera_plane[1] + dL_dcoords[1] * g.camera_plane[3] + dL_dcoords[2] * g.camera_plane[5];
          }
          if (DEPTH) {
            dL_ddelx += dL_dt * g.ray_plane[0];
            dL_ddely += dL_dt * g.ray_plane[1];
          }
          aadd(&dL_dmean2D[global_id * 3], dL_ddelx * ddelx_dx);
          aadd(&dL_dmean2D[global_id * 3 + 1], dL_ddely * ddely_dy);
          aadd(&dL_dmean2D[global_id * 3 + 2],
               std::fabs(dL_dG * dG_ddelx * ddelx_dx) + std::fabs(dL_dG * dG_ddely * ddely_dy));
          aadd(&dL_dconic[global_id * 4], -0.5f * gdx * d[0] * dL_dG);
          aadd(&dL_dconic[global_id * 4 + 1], -0.5f * gdx * d[1] * dL_dG);
          aadd(&dL_dconic[global_id * 4 + 3], -0.5f * gdy * d[1] * dL_dG);
          aadd(&dL_dopacity[global_id], G * dL_dopa);
        }
      }
  }

  // ---- computeCov2DCUDA (backward.cu:145-488) + preprocessCUDA (:560-628) per Gaussian ----
  const V3 campos = {campos_[0], campos_[1], campos_[2]};
#pragma omp parallel for schedule(dynamic, 256)
  for (int idx = 0; idx < P; ++idx) {
    if (!(radii[idx] > 0)) continue;
    const Geom& gs = ctx.geom[idx];
    const float* cov3D = cov3D_precomp ? cov3D_precomp + 6 * idx : gs.cov3D;
    const V3 mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
    const float dLc[3] = {dL_dconic[4 * idx], dL_dconic[4 * idx + 1], dL_dconic[4 * idx + 3]};
    const V3 dL_dnormal = {dL_dnormals[idx * 3], dL_dnormals[idx * 3 + 1], dL_dnormals[idx * 3 + 2]};
    // QUIRK: conic_opacity slot holds dL_dconic (rasterizer_impl.cu:576) -> .w is the yy term
    const float combined_opacity = dL_dconic[4 * idx + 3];
    const float* cp = &dL_dcamera_planes[idx * 6];
    const float cp0[2] = {cp[0], cp[1]}, cp1[2] = {cp[2], cp[3]}, cp2[2] = {cp[4], cp[5]};
    const float drp[2] = {dL_dray_planes[2 * idx], dL_dray_planes[2 * idx + 1]};
    const Cov2D c = cov2dSetup(mean, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, viewmatrix);
    const V3 t = c.t;
    const float txtz = c.txtz, tytz = c.tytz;
    const M3 &Wm = c.W, &T = c.T, &Vrk = c.Vrk, &cov2D = c.cov;
    const float det_0 = (float)std::max(1e-6, (double)(cov2D.m[0][0] * cov2D.m[1][1] - cov2D.m[0][1] * cov2D.m[0][1]));
    const float det_1 = (float)std::max(1e-6, (double)((cov2D.m[0][0] + kernel_size) * (cov2D.m[1][1] + kernel_size) -
                                                       cov2D.m[0][1] * cov2D.m[0][1]));
    const float coef = (float)std::sqrt(det_0 / (det_1 + 1e-6) + 1e-6);
    M3 evec;
    float eval[3];
    const int Dn = findEigenvaluesSymReal(Vrk, eval, evec);
    const int min_id = eval[0] > eval[1] ? (eval[1] > eval[2] ? 2 : 1) : (eval[0] > eval[2] ? 2 : 0);
    M3 Vrk_inv;
    V3 evmin = {0, 0, 0};
    const bool well = eval[min_id] > 0.00000001;
    if (well) {
      const M3 diag = cols(1 / eval[0], 0, 0, 0, 1 / eval[1], 0, 0, 0, 1 / eval[2]);
      Vrk_inv = mul(mul(evec, diag), tr(evec));
    } else {
      evmin = col(evec, min_id);
      Vrk_inv = outer(evmin, evmin);
    }
    const M3 cov_cam_inv = mul(mul(tr(Wm), Vrk_inv), Wm);
    const V3 uvh = {txtz, tytz, 1};
    const V3 uvh_m = mulv(cov_cam_inv, uvh);
    const V3 uvh_mn = normalize(uvh_m);
    const float u2 = txtz * txtz, v2 = tytz * tytz, uv = txtz * tytz;
    M3 dL_dVrk = zero3(), dL_dnJ = zero3();
    V3 plane = {0, 0, 0};
    float dL_du = 0, dL_dv = 0, dL_dl = 0, l = 1, nl = 1;
    if (!(std::isnan(uvh_mn.x) || Dn == 0)) {
      const float vb = dot(uvh_m, uvh), vbn = dot(uvh_mn, uvh);
      l = std::sqrt(t.x * t.x + t.y * t.y + t.z * t.z);
      const M3 nJ = cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z), t.x / l, t.y / l, t.z / l);
      const M3 nJ_inv = cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -txtz, -tytz, 0);
      const float clamp_vb = std::max(vb, 0.0000001f), clamp_vbn = std::max(vbn, 0.0000001f);
      nl = u2 + v2 + 1;
      const float factor_normal = l / nl;
      const V3 uvh_m_vb = {uvh_mn.x / clamp_vbn, uvh_mn.y / clamp_vbn, uvh_mn.z / clamp_vbn};
      plane = mulv(nJ_inv, uvh_m_vb);
      const float cam0[2] = {(-(v2 + 1) * t.z + plane.x * t.x) / nl, (uv * t.z + plane.y * t.x) / nl};
      const float cam1[2] = {(uv * t.z + plane.x * t.y) / nl, (-(u2 + 1) * t.z + plane.y * t.y) / nl};
      const float cam2[2] = {(t.x + plane.x * t.z) / nl, (t.y + plane.y * t.z) / nl};
      const float ray_plane[2] = {plane.x * factor_normal, plane.y * factor_normal};
      const V3 ray_n = {-plane.x * factor_normal, -plane.y * factor_normal, -1};
      const V3 cam_n = mulv(nJ, ray_n);
      const V3 nvec = normalize(cam_n);
      const float lv = std::sqrt(dot(cam_n, cam_n));
      const V3 dn_lv = {dL_dnormal.x / lv, dL_dnormal.y / lv, dL_dnormal.z / lv};
      const V3 dcam_n = subv(dn_lv, scl(nvec, dot(nvec, dn_lv)));
      const V3 dray_n = mulv(tr(nJ), dcam_n);
      dL_dnJ = outer(dcam_n, ray_n);
      dL_dl = (-plane.x * dray_n.x - plane.y * dray_n.y + plane.x * drp[0] + plane.y * drp[1]) / nl;
      const float dpl[2] = {(t.x * cp0[0] + t.y * cp1[0] + t.z * cp2[0] - l * dray_n.x + drp[0] * l) / nl,
                            (t.x * cp0[1] + t.y * cp1[1] + t.z * cp2[1] - l * dray_n.y + drp[1] * l) / nl};
      const V3 dpl3 = {dpl[0], dpl[1], 0};
      const float dL_dnl = (-cp0[0] * cam0[0] - cp0[1] * cam0[1] - cp1[0] * cam1[0] - cp1[1] * cam1[1] - cp2[0] * cam2[0] -
                            cp2[1] * cam2[1] - dray_n.x * ray_n.x - dray_n.y * ray_n.y - drp[0] * ray_plane[0] -
                            drp[1] * ray_plane[1]) / nl;
      const float tmp = dpl[0] * plane.x + dpl[1] * plane.y;
      const V3 W_uvh = mulv(Wm, uvh);
      if (well) {
        M3 Vd;
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) Vd.m[a][b] = Vrk_inv.m[a][b] / clamp_vb;
        const V3 inner = addv(scl(W_uvh, -tmp), mulv(mul(Wm, tr(nJ_inv)), dpl3));
        const M3 op = outer(mulv(Vrk_inv, W_uvh), mulv(Vd, inner));
        for (int a = 0; a < 3; ++a)
          for (int b = 0; b < 3; ++b) dL_dVrk.m[a][b] = -op.m[a][b];
      } else {
        const float dL_dvb = -tmp / clamp_vb;
        const V3 nJi = mulv(tr(nJ_inv), V3{dpl[0] / clamp_vb, dpl[1] / clamp_vb, 0});
        const M3 dVinv = outer(W_uvh, addv(scl(W_uvh, dL_dvb), mulv(Wm, nJi)));
        const V3 dLdv = mulv(add(dVinv, tr(dVinv)), evmin);
        for (int j = 0; j < 3; j++)
          if (j != min_id) {
            const float sc = dot(col(evec, j), dLdv) / std::min(eval[min_id] - eval[j], -0.0000001f);
            dL_dVrk = add(dL_dVrk, outer(scl(col(evec, j), sc), evmin));
          }
      }
      M3 cci;
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) cci.m[a][b] = cov_cam_inv.m[a][b] / clamp_vb;
      const V3 second = mulv(mul(cci, tr(nJ_inv)), dpl3);
      const float two_tmp = 2 * (-tmp);
      const V3 dL_duvh = {two_tmp * uvh_m_vb.x + second.x, two_tmp * uvh_m_vb.y + second.y, two_tmp * uvh_m_vb.z + second.z};
      const M3 dnJinv = outer(dpl3, uvh_m_vb);
      dL_du = dL_dnl * 2 * txtz + dL_duvh.x + (dnJinv.m[0][1] + dnJinv.m[1][0]) * (-tytz) + 2 * dnJinv.m[1][1] * txtz -
              dnJinv.m[2][0] + (cp0[1] * t.y + cp1[0] * t.y + cp1[1] * (-2 * t.x)) / nl;
      dL_dv = dL_dnl * 2 * tytz + dL_duvh.y + (dnJinv.m[0][1] + dnJinv.m[1][0]) * (-txtz) + 2 * dnJinv.m[0][0] * tytz -
              dnJinv.m[2][1] + (cp0[0] * (-2 * t.y) + cp0[1] * t.x + cp1[0] * t.x) / nl;
    }
    const float opacity = (float)(combined_opacity / (coef + 1e-6));
    const float dL_dcoef = dL_dopacity[idx] * opacity;
    const float dL_dsqrtcoef = (float)(dL_dcoef * 0.5 * 1. / (coef + 1e-6));
    const float dL_ddet0 = (float)(dL_dsqrtcoef / (det_1 + 1e-6));
    const float dL_ddet1 = (float)(dL_dsqrtcoef * det_0 * (-1.f / (det_1 * det_1 + 1e-6)));
    const float dcoef_da = dL_ddet0 * cov2D.m[1][1] + dL_ddet1 * (cov2D.m[1][1] + kernel_size);
    const float dcoef_db = (float)(dL_ddet0 * (-2. * cov2D.m[0][1]) + dL_ddet1 * (-2. * cov2D.m[0][1]));
    const float dcoef_dc = dL_ddet0 * cov2D.m[0][0] + dL_ddet1 * (cov2D.m[0][0] + kernel_size);
    const float a = cov2D.m[0][0] + kernel_size, b = cov2D.m[0][1], cc = cov2D.m[1][1] + kernel_size;
    const float denom = a * cc - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float* dcov = dL_dcov3D + 6 * idx;
    if (denom2inv != 0) {
      dL_da = denom2inv * (-cc * cc * dLc[0] + 2 * b * cc * dLc[1] + (denom - a * cc) * dLc[2]);
      dL_dc = denom2inv * (-a * a * dLc[2] + 2 * a * b * dLc[1] + (denom - a * cc) * dLc[0]);
      dL_db = denom2inv * 2 * (b * cc * dLc[0] - (denom + 2 * b * b) * dLc[1] + a * b * dLc[2]);
      if (det_0 <= 1e-6 || det_1 <= 1e-6) dL_dopacity[idx] = 0;
      else {
        dL_da += dcoef_da; dL_dc += dcoef_dc; dL_db += dcoef_db;
        dL_dopacity[idx] = dL_dopacity[idx] * coef;
      }
      dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
      dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
      dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
      dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][1] * dL_dc;
      dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][2] * dL_dc;
      dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db + 2 * T.m[1][1] * T.m[1][2] * dL_dc;
    } else {
      for (int i = 0; i < 6; i++) dcov[i] = 0;
    }
    dcov[0] += dL_dVrk.m[0][0]; dcov[3] += dL_dVrk.m[1][1]; dcov[5] += dL_dVrk.m[2][2];
    dcov[1] += dL_dVrk.m[0][1] + dL_dVrk.m[1][0];
    dcov[2] += dL_dVrk.m[0][2] + dL_dVrk.m[2][0];
    dcov[4] += dL_dVrk.m[1][2] + dL_dVrk.m[2][1];
    auto TV = [&](int r, int cidx) { return T.m[r][0] * Vrk.m[cidx][0] + T.m[r][1] * Vrk.m[cidx][1] + T.m[r][2] * Vrk.m[cidx][2]; };
    const float dL_dT00 = 2 * TV(0, 0) * dL_da + TV(1, 0) * dL_db, dL_dT01 = 2 * TV(0, 1) * dL_da + TV(1, 1) * dL_db,
                dL_dT02 = 2 * TV(0, 2) * dL_da + TV(1, 2) * dL_db, dL_dT10 = 2 * TV(1, 0) * dL_dc + TV(0, 0) * dL_db,
                dL_dT11 = 2 * TV(1, 1) * dL_dc + TV(0, 1) * dL_db, dL_dT12 = 2 * TV(1, 2) * dL_dc + TV(0, 2) * dL_db;
    const float dL_dJ00 = Wm.m[0][0] * dL_dT00 + Wm.m[0][1] * dL_dT01 + Wm.m[0][2] * dL_dT02;
    const float dL_dJ02 = Wm.m[2][0] * dL_dT00 + Wm.m[2][1] * dL_dT01 + Wm.m[2][2] * dL_dT02;
    const float dL_dJ11 = Wm.m[1][0] * dL_dT10 + Wm.m[1][1] * dL_dT11 + Wm.m[1][2] * dL_dT12;
    const float dL_dJ12 = Wm.m[2][0] * dL_dT10 + Wm.m[2][1] * dL_dT11 + Wm.m[2][2] * dL_dT12;
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz, l3 = l * l * l;
    const float h_x = focal_x, h_y = focal_y;
    const float dL_dtx = c.xmul * (-h_x * tz2 * dL_dJ02 + dL_du * tz - dL_dnJ.m[0][2] * tz2 + dL_dnJ.m[2][0] * (1 / l - t.x * t.x / l3) +
                                   dL_dnJ.m[2][1] * (-t.x * t.y / l3) + dL_dnJ.m[2][2] * (-t.x * t.z / l3) +
                                   (cp0[0] * plane.x + cp0[1] * plane.y + cp2[0]) / nl + dL_dl * t.x / l);
    const float dL_dty = c.ymul * (-h_y * tz2 * dL_dJ12 + dL_dv * tz - dL_dnJ.m[1][2] * tz2 + dL_dnJ.m[2][0] * (-t.x * t.y / l3) +
                                   dL_dnJ.m[2][1] * (1 / l - t.y * t.y / l3) + dL_dnJ.m[2][2] * (-t.y * t.z / l3) +
                                   (cp1[0] * plane.x + cp1[1] * plane.y + cp2[1]) / nl + dL_dl * t.y / l);
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12 -
                         (dL_du * t.x + dL_dv * t.y) * tz2 + (dL_dnJ.m[0][0] + dL_dnJ.m[1][1]) * (-tz2) +
                         dL_dnJ.m[0][2] * (2 * t.x * tz3) + dL_dnJ.m[1][2] * (2 * t.y * tz3) +
                         (dL_dnJ.m[2][0] * t.x + dL_dnJ.m[2][1] * t.y) * (-t.z / l3) + dL_dnJ.m[2][2] * (1 / l - t.z * t.z / l3) +
                         (cp0[0] * (-(v2 + 1)) + cp0[1] * uv + cp1[0] * uv + cp1[1] * (-(u2 + 1)) + cp2[0] * plane.x + cp2[1] * plane.y) / nl +
                         dL_dl * t.z / l;
    V3 dmean = transformVec4x3Transpose({dL_dtx, dL_dty, dL_dtz}, viewmatrix);
    {  // preprocessCUDA backward (:587-619)
      float m_hom[4];
      transformPoint4x4(mean, projmatrix, m_hom);
      const float m_w = 1.0f / (m_hom[3] + 0.0000001f);
      const float* proj = projmatrix;
      const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
      const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
      const float gx2 = dL_dmean2D[3 * idx], gy2 = dL_dmean2D[3 * idx + 1];
      const V3 d1 = {(proj[0] * m_w - proj[3] * mul1) * gx2 + (proj[1] * m_w - proj[3] * mul2) * gy2,
                     (proj[4] * m_w - proj[7] * mul1) * gx2 + (proj[5] * m_w - proj[7] * mul2) * gy2,
                     (proj[8] * m_w - proj[11] * mul1) * gx2 + (proj[9] * m_w - proj[11] * mul2) * gy2};
      const V3 mv = transformPoint4x3(mean, viewmatrix);
      const float tt = std::sqrt(mv.x * mv.x + mv.y * mv.y + mv.z * mv.z);
      const float dLt = dL_dts[idx];
      const V3 d2 = transformVec4x3Transpose({dL_dview_points[3 * idx] + mv.x / tt * dLt, dL_dview_points[3 * idx + 1] + mv.y / tt * dLt,
                                              dL_dview_points[3 * idx + 2] + mv.z / tt * dLt}, viewmatrix);
      dmean.x += d1.x + d2.x; dmean.y += d1.y + d2.y; dmean.z += d1.z + d2.z;
    }
    if (shs) {  // computeColorFromSH backward (:21-140)
      const V3 dir_orig = subv(mean, campos);
      const float len = std::sqrt(dot(dir_orig, dir_orig));
      const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
      const float* sh = shs + (size_t)idx * M * 3;
      float gRGB[3] = {dL_dcolors[3 * idx] * (gs.clamped[0] ? 0.f : 1.f), dL_dcolors[3 * idx + 1] * (gs.clamped[1] ? 0.f : 1.f),
                       dL_dcolors[3 * idx + 2] * (gs.clamped[2] ? 0.f : 1.f)};
      float w[16] = {0};
      float dx[3] = {0}, dy[3] = {0}, dz[3] = {0};
      auto S = [&](int i, int ch) { return sh[3 * i + ch]; };
      w[0] = SH_C0;
      if (D > 0) {
        w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x;
        for (int ch = 0; ch < 3; ++ch) { dx[ch] = -SH_C1 * S(3, ch); dy[ch] = -SH_C1 * S(1, ch); dz[ch] = SH_C1 * S(2, ch); }
        if (D > 1) {
          const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
          w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2.f * zz - xx - yy); w[7] = SH_C2[3] * xz;
          w[8] = SH_C2[4] * (xx - yy);
          for (int ch = 0; ch < 3; ++ch) {
            dx[ch] += SH_C2[0] * y * S(4, ch) + SH_C2[2] * 2.f * -x * S(6, ch) + SH_C2[3] * z * S(7, ch) + SH_C2[4] * 2.f * x * S(8, ch);
            dy[ch] += SH_C2[0] * x * S(4, ch) + SH_C2[1] * z * S(5, ch) + SH_C2[2] * 2.f * -y * S(6, ch) + SH_C2[4] * 2.f * -y * S(8, ch);
            dz[ch] += SH_C2[1] * y * S(5, ch) + SH_C2[2] * 2.f * 2.f * z * S(6, ch) + SH_C2[3] * x * S(7, ch);
          }
          if (D > 2) {
            w[9] = SH_C3[0] * y * (3.f * xx - yy); w[10] = SH_C3[1] * xy * z; w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
            w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
            w[14] = SH_C3[5] * z * (xx - yy); w[15] = SH_C3[6] * x * (xx - 3.f * yy);
            for (int ch = 0; ch < 3; ++ch) {
              dx[ch] += (SH_C3[0] * S(9, ch) * 3.f * 2.f * xy + SH_C3[1] * S(10, ch) * yz + SH_C3[2] * S(11, ch) * -2.f * xy +
                         SH_C3[3] * S(12, ch) * -3.f * 2.f * xz + SH_C3[4] * S(13, ch) * (-3.f * xx + 4.f * zz - yy) +
                         SH_C3[5] * S(14, ch) * 2.f * xz + SH_C3[6] * S(15, ch) * 3.f * (xx - yy));
              dy[ch] += (SH_C3[0] * S(9, ch) * 3.f * (xx - yy) + SH_C3[1] * S(10, ch) * xz + SH_C3[2] * S(11, ch) * (-3.f * yy + 4.f * zz - xx) +
                         SH_C3[3] * S(12, ch) * -3.f * 2.f * yz + SH_C3[4] * S(13, ch) * -2.f * xy + SH_C3[5] * S(14, ch) * -2.f * yz +
                         SH_C3[6] * S(15, ch) * -3.f * 2.f * xy);
              dz[ch] += (SH_C3[1] * S(10, ch) * xy + SH_C3[2] * S(11, ch) * 4.f * 2.f * yz + SH_C3[3] * S(12, ch) * 3.f * (2.f * zz - xx - yy) +
                         SH_C3[4] * S(13, ch) * 4.f * 2.f * xz + SH_C3[5] * S(14, ch) * (xx - yy));
            }
          }
        }
      }
      const int ncoef = (D + 1) * (D + 1);
      for (int i = 0; i < M && i < ncoef && i < 16; ++i)
        for (int ch = 0; ch < 3; ++ch) dL_dsh[((size_t)idx * M + i) * 3 + ch] = w[i] * gRGB[ch];
      const V3 ddir = {dx[0] * gRGB[0] + dx[1] * gRGB[1] + dx[2] * gRGB[2], dy[0] * gRGB[0] + dy[1] * gRGB[1] + dy[2] * gRGB[2],
                       dz[0] * gRGB[0] + dz[1] * gRGB[1] + dz[2] * gRGB[2]};
      const V3 v = dir_orig;  // dnormvdv, auxiliary.h:123-133
      const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
      const float invsum32 = 1.0f / std::sqrt(sum2 * sum2 * sum2);
      dmean.x += ((+sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * invsum32;
      dmean.y += (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * invsum32;
      dmean.z += (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * invsum32;
    }
    dL_dmean3D[3 * idx] = dmean.x; dL_dmean3D[3 * idx + 1] = dmean.y; dL_dmean3D[3 * idx + 2] = dmean.z;
    if (scales) {  // computeCov3D backward (:492-555)
      M3 Mm, R;
      float tmpcov[6];
      computeCov3D(scales + 3 * idx, scale_modifier, rotations + 4 * idx, tmpcov, &Mm, &R);
      const float* rot = rotations + 4 * idx;
      const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
      const float s[3] = {scale_modifier * scales[3 * idx], scale_modifier * scales[3 * idx + 1], scale_modifier * scales[3 * idx + 2]};
      const M3 dSigma = cols(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4], 0.5f * dcov[2],
                             0.5f * dcov[4], dcov[5]);
      M3 M2;
      for (int a2 = 0; a2 < 3; ++a2)
        for (int b2 = 0; b2 < 3; ++b2) M2.m[a2][b2] = Mm.m[a2][b2] * 2.0f;
      const M3 dM = mul(M2, dSigma);
      const M3 Rt = tr(R);
      M3 dMt = tr(dM);
      for (int k = 0; k < 3; ++k) dL_dscale[3 * idx + k] = dot(col(Rt, k), col(dMt, k));
      for (int k = 0; k < 3; ++k) { dMt.m[0][k] *= s[0]; dMt.m[1][k] *= s[1]; dMt.m[2][k] *= s[2]; }
      float* dq = dL_drot + 4 * idx;
      dq[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) + 2 * x * (dMt.m[1][2] - dMt.m[2][1]);
      dq[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) + 2 * r * (dMt.m[1][2] - dMt.m[2][1]) - 4 * x * (dMt.m[2][2] + dMt.m[1][1]);
      dq[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) + 2 * z * (dMt.m[1][2] + dMt.m[2][1]) - 4 * y * (dMt.m[2][2] + dMt.m[0][0]);
      dq[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) + 2 * y * (dMt.m[1][2] + dMt.m[2][1]) - 4 * z * (dMt.m[1][1] + dMt.m[0][0]);
    }
  }
}

}  // extern "C"
