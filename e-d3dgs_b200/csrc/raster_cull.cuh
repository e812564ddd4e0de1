// Conservative (pixel patch, Gaussian) culling shared by the forward and backward blend kernels.
#pragma once

namespace ed3dgs {

// Can a Gaussian (centre g, conic (a, b, c), opacity op) reach alpha >= 1/255 anywhere in the 8x4 pixel
// patch whose first pixel is (x0, y0)? Conservative: alpha = op * exp(-q/2) >= 1/255 needs
// q = a dx^2 + 2 b dx dy + c dy^2 <= 2 ln(255 op); the bounding box of that ellipse has half-widths
// sqrt(t c / det), sqrt(t a / det). Margins (1e-3 in t, 1 % + 0.05 px in the box) dwarf fp32 rounding -- except for
// needle splats: det = a c - b^2 cancels, its relative error is ~eps a c / det, so once det < 256 eps a c (conic
// eigenvalue ratio beyond ~3e4) the 1 % margin no longer covers it and the patch is left to the exact per-pixel test.
__device__ __forceinline__ bool patch_may_touch(float gx, float gy, float a, float b, float c, float op, float x0, float y0) {
  if (op < 1.0f / 255.0f) return false;            // G <= 1, so alpha <= op < 1/255 at every pixel
  const float det = a * c - b * b;
  if (!(det > 256.f * 1.1920929e-7f * a * c) || !(a > 0.f) || !(c > 0.f)) return true;   // not an ellipse / ill-conditioned: exact test
  const float t = 2.f * __logf(255.f * op) * 1.001f + 1e-3f;
  const float inv = t / det;
  const float hx = sqrtf(inv * c) * 1.01f + 0.05f, hy = sqrtf(inv * a) * 1.01f + 0.05f;
  return !(gx + hx < x0 || gx - hx > x0 + 7.f || gy + hy < y0 || gy - hy > y0 + 3.f);
}

// Same test against an arbitrary axis-aligned box of sample positions [xmin, xmax] x [ymin, ymax].
__device__ __forceinline__ bool box_may_touch(float gx, float gy, float a, float b, float c, float op, float xmin,
                                              float xmax, float ymin, float ymax) {
  if (op < 1.0f / 255.0f) return false;
  const float det = a * c - b * b;
  if (!(det > 256.f * 1.1920929e-7f * a * c) || !(a > 0.f) || !(c > 0.f)) return true;
  const float t = 2.f * __logf(255.f * op) * 1.001f + 1e-3f;
  const float inv = t / det;
  const float hx = sqrtf(inv * c) * 1.01f + 0.05f, hy = sqrtf(inv * a) * 1.01f + 0.05f;
  return !(gx + hx < xmin || gx - hx > xmax || gy + hy < ymin || gy - hy > ymax);
}

}  // namespace ed3dgs
