// Backward half of the splat rasterizer for sm_100a.
// Behavioural contract = CudaRasterizer::Rasterizer::backward (rasterizer_impl.cu:436-578):
//   blend_bwd       <- backward.cu:631-1016 (renderCUDA), but instead of 16-25 global float
//                      atomics per (pixel, Gaussian) pair the 16/32 gradient components are
//                      transpose-reduced across the warp with 16/31 shuffles, parked in a
//                      per-warp shared-memory slot, summed over the 8 warps of the tile and
//                      sent to HBM as ONE red.global.add.v4.f32 per quad per Gaussian per tile.
//   gaussian_bwd    <- backward.cu:145-488 (computeCov2DCUDA) + :560-628 (preprocessCUDA) +
//                      :21-140 (SH) + :492-555 (cov3D), fused into one per-Gaussian kernel.
// Reference quirks that are reproduced on purpose are marked QUIRK.
#include "raster_cull.cuh"
#include "gs_math.cuh"
#include "raster_state.cuh"
#include "raster_api.h"
#include "profiler.h"

namespace ed3dgs {

constexpr int BWD_THREADS = 256;
constexpr int BWD_WARPS = BWD_THREADS / 32;

__device__ __forceinline__ void cp_async16_b(void* smem_dst, const void* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_b() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all_b() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void red_add_v4(float* gptr, const float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(gptr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

// Sum v[0..V) over the 32 lanes. On return v[0] of lane L holds the total of component
// (V == 16 ? L >> 1 : L). 16 shuffles for V = 16, 31 for V = 32.
template <int V>
__device__ __forceinline__ void warp_transpose_reduce(float (&v)[V], int lane) {
  static_assert(V == 16 || V == 32, "V");
  int n = V;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    if (n > 1) {
      const int half = n / 2;
      const bool upper = (lane & o) != 0;
#pragma unroll
      for (int i = 0; i < V / 2; ++i) {
        if (i < half) {
          const float send = upper ? v[i] : v[i + half];
          const float keep = upper ? v[i + half] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
      }
      n = half;
    } else {
      v[0] += __shfl_xor_sync(0xffffffffu, v[0], o);
    }
  }
}

template <bool COORD, bool DEPTH>
struct BwdCfg {
  static constexpr bool NORMAL = COORD || DEPTH;
  static constexpr bool GEO = NORMAL;
  static constexpr int V = COORD ? 32 : 16;            // gradient components per pair
  // Gaussians per staged batch. 64 keeps a CTA at 41 KB of shared memory (82 KB with 128), so that three CTAs fit an SM
  // when the registers allow it: measured at C3 (variant <0,1>) 1.46 ms with B = 128 / 2 CTAs (112 registers), 1.49 with
  // B = 64 / 2 CTAs, 1.28 with B = 64 / 3 CTAs (79 registers, no spills), 1.39 with 4 CTAs (64 registers, spills).
  static constexpr int B = 64;
  static constexpr int OCC = COORD ? 2 : 3;            // CTAs per SM asked of the compiler (COORD: 78 KB per CTA)
  static constexpr int NQ = GEO ? 4 : 3;
  static constexpr int NQT = NQ + (COORD ? REC_COORD_QUADS : 0);
  static constexpr size_t rec_bytes = (size_t)2 * NQT * B * sizeof(float4);
  static constexpr size_t acc_bytes = (size_t)BWD_WARPS * B * V * sizeof(float);
  static constexpr size_t id_bytes = (size_t)2 * B * sizeof(int);
  static constexpr size_t mask_bytes = (size_t)BWD_WARPS * (B / 32) * sizeof(uint32_t);
  static constexpr size_t smem = rec_bytes + acc_bytes + id_bytes + mask_bytes + 64;
};

template <bool COORD, bool DEPTH>
__global__ void __launch_bounds__(BWD_THREADS, (BwdCfg<COORD, DEPTH>::OCC))
blend_bwd(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
          const float* __restrict__ bg, const float4* __restrict__ rec, const float4* __restrict__ rec_coord,
          const float* __restrict__ alphas, const float* __restrict__ accum_coord,
          const float* __restrict__ accum_depth, const float* __restrict__ normal_length,
          const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpixels,
          const float* __restrict__ dL_dpixel_coords, const float* __restrict__ dL_dpixel_mcoords,
          const float* __restrict__ dL_dpixel_depths, const float* __restrict__ dL_dpixel_mdepths,
          const float* __restrict__ dL_dalphas, const float* __restrict__ dL_dpixel_normals,
          const float* __restrict__ normalmap, float fx, float fy, float* __restrict__ grad) {
  using Cfg = BwdCfg<COORD, DEPTH>;
  constexpr bool NORMAL = Cfg::NORMAL, GEO = Cfg::GEO;
  constexpr int V = Cfg::V, B = Cfg::B, NQ = Cfg::NQ, NQT = Cfg::NQT;

  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* s_rec = reinterpret_cast<float4*>(smem_raw);                          // [2][NQT][B]
  float* s_acc = reinterpret_cast<float*>(smem_raw + Cfg::rec_bytes);           // [WARPS][B][V]
  int* s_id = reinterpret_cast<int*>(smem_raw + Cfg::rec_bytes + Cfg::acc_bytes);  // [2][B]
  uint32_t* s_mask = reinterpret_cast<uint32_t*>(smem_raw + Cfg::rec_bytes + Cfg::acc_bytes + Cfg::id_bytes);
  __shared__ int s_red[BWD_WARPS];
  auto SQ = [&](int buf, int q, int j) -> float4& { return s_rec[(buf * NQT + q) * B + j]; };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t px = blockIdx.x * TILE + (warp & 1) * 8 + (lane & 7);
  const uint32_t py = blockIdx.y * TILE + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < (uint32_t)W && py < (uint32_t)H;
  const uint32_t pix_id = W * py + px;
  const size_t HW = (size_t)H * W;
  const float2 pixf = make_float2((float)px, (float)py);
  const float2 pixnf = make_float2((pixf.x - W / 2.f) / fx, (pixf.y - H / 2.f) / fy);
  const float ln = sqrtf(pixnf.x * pixnf.x + pixnf.y * pixnf.y + 1);
  const float patch_x0 = (float)(blockIdx.x * TILE + (warp & 1) * 8), patch_y0 = (float)(blockIdx.y * TILE + (warp >> 1) * 4);

  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];

  // QUIRK: T_final is rebuilt from the alpha image (backward.cu:706-707).
  const float T_final = inside ? (1 - alphas[pix_id]) : 0;
  const float w_final = inside ? alphas[pix_id] : 0;
  float T = T_final;
  const int last_contributor = inside ? (int)n_contrib[pix_id] : 0;
  const int max_contributor = inside ? (int)n_contrib[pix_id + HW] : 0;

  // Entries at list positions >= max over the tile of last_contributor touch no pixel: skip.
  int tile_last = last_contributor;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) tile_last = max(tile_last, __shfl_xor_sync(0xffffffffu, tile_last, o));
  if (lane == 0) s_red[warp] = tile_last;
  __syncthreads();
  tile_last = 0;
#pragma unroll
  for (int w = 0; w < BWD_WARPS; ++w) tile_last = max(tile_last, s_red[w]);
  const int total = min(tile_last, (int)(range.y - range.x));
  if (total <= 0) return;
  const int rounds = (total + B - 1) / B;

  float accum_rec[3] = {0, 0, 0}, dL_dpixel[3] = {0, 0, 0};
  float accum_coord_rec[3] = {0, 0, 0}, dL_dpixel_coord[3] = {0, 0, 0}, dL_dpixel_mcoord[3] = {0, 0, 0};
  float accum_t_rec = 0, dL_dpixel_t = 0, dL_dpixel_mt = 0;
  float accum_alpha_rec = 0, dL_dalpha = 0;
  float accum_normal_rec[3] = {0, 0, 0}, dL_dpixel_normal[3] = {0, 0, 0};
  if (inside) {
#pragma unroll
    for (int i = 0; i < 3; ++i) dL_dpixel[i] = dL_dpixels[i * HW + pix_id];
    dL_dalpha = dL_dalphas[pix_id];
    if constexpr (GEO) {
      const float ww = w_final * w_final;
      if constexpr (COORD) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const float gw = dL_dpixel_coords[i * HW + pix_id];
          const float ac = accum_coord[i * HW + pix_id];
          dL_dalpha -= gw * ac / ww;
          dL_dpixel_coord[i] = gw / w_final;
          dL_dpixel_mcoord[i] = dL_dpixel_mcoords[i * HW + pix_id];
        }
      }
      if constexpr (DEPTH) {
        const float gw = dL_dpixel_depths[pix_id];
        const float ad = accum_depth[pix_id];
        dL_dalpha -= gw * ad / ww;
        dL_dpixel_t = gw / w_final / ln;
        dL_dpixel_mt = dL_dpixel_mdepths[pix_id] / ln;
      }
      if constexpr (NORMAL) {
        const float3 gn = make_float3(dL_dpixel_normals[pix_id], dL_dpixel_normals[HW + pix_id],
                                      dL_dpixel_normals[2 * HW + pix_id]);
        const float3 nn = make_float3(normalmap[pix_id], normalmap[HW + pix_id], normalmap[2 * HW + pix_id]);
        const float nlen = normal_length[pix_id];
        float3 dL;
        if (nlen < 1.0E-12F) {
          dL = make_float3(gn.x / 1.0E-12F, gn.y / 1.0E-12F, gn.z / 1.0E-12F);
        } else {
          const float dp = dot3(gn, nn);
          dL = make_float3((gn.x - dp * nn.x) / nlen, (gn.y - dp * nn.y) / nlen, (gn.z - dp * nn.z) / nlen);
        }
        dL_dpixel_normal[0] = dL.x;
        dL_dpixel_normal[1] = dL.y;
        dL_dpixel_normal[2] = dL.z;
      }
    }
  }
  const float bg_dot_dpixel = bg[0] * dL_dpixel[0] + bg[1] * dL_dpixel[1] + bg[2] * dL_dpixel[2];

  float last_alpha = 0, last_color[3] = {0, 0, 0}, last_coord[3] = {0, 0, 0}, last_t = 0;
  float last_normal[3] = {0, 0, 0};
  const float ddelx_dx = 0.5 * W;
  const float ddely_dy = 0.5 * H;

  // Staging: thread -> (record j, quad k) pairs of the batch, walking the list back to front.
  auto batch_id = [&](int round, int j) -> int {
    const int k = round * B + j;  // distance from the back of the (truncated) list
    return (k < total) ? (int)point_list[range.x + (total - 1 - k)] : -1;
  };
  constexpr int ITEMS = (B * NQT + BWD_THREADS - 1) / BWD_THREADS;
  int id_next[ITEMS];
  auto prefetch_ids = [&](int round) {
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
      const int q = it * BWD_THREADS + tid;
      id_next[it] = (q < B * NQT && round < rounds) ? batch_id(round, q / NQT) : -1;
    }
  };
  auto issue = [&](int buf) {
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
      const int q = it * BWD_THREADS + tid;
      if (q < B * NQT) {
        const int j = q / NQT, k = q % NQT, id = id_next[it];
        if (id >= 0) {
          const float4* src = (k < NQ) ? (rec + (size_t)id * REC_QUADS + k)
                                       : (rec_coord + (size_t)id * REC_COORD_QUADS + (k - NQ));
          cp_async16_b(&SQ(buf, k, j), src);
        }
        if (k == 0) s_id[buf * B + j] = id;
      }
    }
  };

  prefetch_ids(0);
  issue(0);
  cp_async_commit_b();
  prefetch_ids(1);

  for (int i = 0; i < rounds; ++i) {
    cp_async_wait_all_b();
    __syncthreads();  // batch i staged; previous batch's cross-warp reduction finished
    const int buf = i & 1;
    issue(buf ^ 1);
    cp_async_commit_b();
    prefetch_ids(i + 2);

    const int n = min(B, total - i * B);
#pragma unroll 1
    for (int m = 0; m < B / 32; ++m) {
    uint32_t wmask = 0;
    const int jend = min(n, (m + 1) * 32);
    // Conservative cull, 32 Gaussians at a time (one per lane): a Gaussian whose alpha >= 1/255 ellipse
    // misses this warp's 8x4 pixel patch cannot pass the exact per-pixel test below, so the warp skips it
    // without staging reads or a vote. Results are unchanged: only pairs the exact test rejects are culled.
    uint32_t cand;
    {
      const int jl = m * 32 + lane;
      bool touch = false;
      if (jl < jend) {
        const float4 c0 = SQ(buf, 0, jl);
        const float4 c1 = SQ(buf, 1, jl);
        touch = patch_may_touch(c0.x, c0.y, c1.x, c1.y, c1.z, c1.w, patch_x0, patch_y0);
      }
      cand = __ballot_sync(0xffffffffu, touch);
    }
#pragma unroll 1
    while (cand) {
      const int j = m * 32 + __ffs(cand) - 1;
      cand &= cand - 1;
      const int contributor = total - 1 - (i * B + j);  // list position from the front
      const float4 q0 = SQ(buf, 0, j);
      const float4 q1 = SQ(buf, 1, j);
      const float2 d = make_float2(q0.x - pixf.x, q0.y - pixf.y);
      const float power = -0.5f * (q1.x * d.x * d.x + q1.z * d.y * d.y) - q1.y * d.x * d.y;
      bool valid = inside && (contributor < last_contributor) && !(power > 0.0f);
      float G = 0.f, alpha = 0.f;
      if (valid) {
        G = expf(power);
        alpha = min(0.99f, q1.w * G);
        valid = !(alpha < 1.0f / 255.0f);
      }
      if (!__any_sync(0xffffffffu, valid)) continue;

      float v[V];
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] = 0.f;
      if (valid) {
        T = T / (1.f - alpha);
        const float wc = alpha * T;  // d(channel)/d(colour)
        float dL_dopa = 0.0f;
        const float4 q2 = SQ(buf, 2, j);
        const float col[3] = {q2.x, q2.y, q2.z};
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
          last_color[ch] = col[ch];
          dL_dopa += (col[ch] - accum_rec[ch]) * dL_dpixel[ch];
          v[8 + ch] = wc * dL_dpixel[ch];
        }
        float dL_dcoords[3] = {0, 0, 0};
        float dL_dt = 0.f;
        float4 c0, c1, c2;
        if constexpr (COORD) {
          c0 = SQ(buf, NQ + 0, j);
          c1 = SQ(buf, NQ + 1, j);
          c2 = SQ(buf, NQ + 2, j);
          const float coord[3] = {c1.z + c0.x * d.x + c0.y * d.y, c1.w + c0.z * d.x + c0.w * d.y,
                                  c2.x + c1.x * d.x + c1.y * d.y};
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            accum_coord_rec[ch] = last_alpha * last_coord[ch] + (1.f - last_alpha) * accum_coord_rec[ch];
            last_coord[ch] = coord[ch];
            dL_dopa += (coord[ch] - accum_coord_rec[ch]) * dL_dpixel_coord[ch];
            dL_dcoords[ch] = wc * dL_dpixel_coord[ch];
            if ((uint32_t)contributor == (uint32_t)(max_contributor - 1)) dL_dcoords[ch] += dL_dpixel_mcoord[ch];
          }
          v[16] = dL_dcoords[0];
          v[17] = dL_dcoords[1];
          v[18] = dL_dcoords[2];
          v[19] = dL_dcoords[0] * d.x / fx;
          v[20] = dL_dcoords[0] * d.y / fy;
          v[21] = dL_dcoords[1] * d.x / fx;
          v[22] = dL_dcoords[1] * d.y / fy;
          v[23] = dL_dcoords[2] * d.x / fx;
          v[24] = dL_dcoords[2] * d.y / fy;
        }
        float4 q3;
        if constexpr (GEO) q3 = SQ(buf, 3, j);
        if constexpr (DEPTH) {
          const float t = q0.z + (q2.w * d.x + q3.x * d.y);
          accum_t_rec = last_alpha * last_t + (1.f - last_alpha) * accum_t_rec;
          last_t = t;
          dL_dopa += (t - accum_t_rec) * dL_dpixel_t;
          dL_dt = wc * dL_dpixel_t;
          if ((uint32_t)contributor == (uint32_t)(max_contributor - 1)) dL_dt += dL_dpixel_mt;
          v[7] = dL_dt;
          v[11] = dL_dt * d.x / fx;
          v[12] = dL_dt * d.y / fy;
        }
        if constexpr (NORMAL) {
          const float nrm[3] = {q3.y, q3.z, q3.w};
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            accum_normal_rec[ch] = last_alpha * last_normal[ch] + (1.f - last_alpha) * accum_normal_rec[ch];
            last_normal[ch] = nrm[ch];
            dL_dopa += (nrm[ch] - accum_normal_rec[ch]) * dL_dpixel_normal[ch];
            v[13 + ch] = wc * dL_dpixel_normal[ch];
          }
        }
        accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
        dL_dopa += (1 - accum_alpha_rec) * dL_dalpha;
        dL_dopa *= T;
        last_alpha = alpha;
        dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;

        const float dL_dG = q1.w * dL_dopa;
        const float gdx = G * d.x;
        const float gdy = G * d.y;
        const float dG_ddelx = -gdx * q1.x - gdy * q1.y;
        const float dG_ddely = -gdy * q1.z - gdx * q1.y;
        float dL_ddelx = dL_dG * dG_ddelx;
        float dL_ddely = dL_dG * dG_ddely;
        if constexpr (COORD) {
          dL_ddelx += dL_dcoords[0] * c0.x + dL_dcoords[1] * c0.z + dL_dcoords[2] * c1.x;
          dL_ddely += dL_dcoords[0] * c0.y + dL_dcoords[1] * c0.w + dL_dcoords[2] * c1.y;
        }
        if constexpr (DEPTH) {
          dL_ddelx += dL_dt * q2.w;
          dL_ddely += dL_dt * q3.x;
        }
        v[0] = dL_ddelx * ddelx_dx;
        v[1] = dL_ddely * ddely_dy;
        v[2] = fabsf(dL_dG * dG_ddelx * ddelx_dx) + fabsf(dL_dG * dG_ddely * ddely_dy);  // GOF abs-grad
        v[3] = G * dL_dopa;
        v[4] = -0.5f * gdx * d.x * dL_dG;
        v[5] = -0.5f * gdx * d.y * dL_dG;
        v[6] = -0.5f * gdy * d.y * dL_dG;
      }
      warp_transpose_reduce<V>(v, lane);
      float* slot = s_acc + ((size_t)warp * B + j) * V;
      if constexpr (V == 16) {
        if ((lane & 1) == 0) slot[lane >> 1] = v[0];
      } else {
        slot[lane] = v[0];
      }
      wmask |= 1u << (j & 31);
    }
    if (lane == 0) s_mask[warp * (B / 32) + m] = wmask;
    }
    __syncthreads();

    // Cross-warp sum and one vector reduction per (Gaussian, quad).
    for (int item = tid; item < n * (V / 4); item += BWD_THREADS) {
      const int j = item / (V / 4), quad = item % (V / 4);
      float4 sum = make_float4(0, 0, 0, 0);
      bool any = false;
#pragma unroll
      for (int w = 0; w < BWD_WARPS; ++w) {
        if (s_mask[w * (B / 32) + (j >> 5)] & (1u << (j & 31))) {
          const float4 t = *reinterpret_cast<const float4*>(s_acc + ((size_t)w * B + j) * V + quad * 4);
          sum.x += t.x; sum.y += t.y; sum.z += t.z; sum.w += t.w;
          any = true;
        }
      }
      if (any) red_add_v4(grad + (size_t)s_id[buf * B + j] * V + quad * 4, sum);
    }
  }
  cp_async_wait_all_b();
}

// ---------------------------------------------------------------------------------------------
// SH backward for one Gaussian (backward.cu:21-140). Writes dL_dsh[M][3] (zero above `deg`) and
// returns the view-direction part of dL/dmean.
__device__ __forceinline__ float3 sh_bwd(int deg, int M, const float3 pos, const float3 campos,
                                         const float* sh, uint8_t clamp_bits, float3 dL_dRGB,
                                         float* dL_dsh) {
  const float3 dir_orig = f3_sub(pos, campos);
  const float len = sqrtf(dot3(dir_orig, dir_orig));
  const float3 dir = make_float3(dir_orig.x / len, dir_orig.y / len, dir_orig.z / len);
  dL_dRGB.x *= (clamp_bits & 1) ? 0 : 1;
  dL_dRGB.y *= (clamp_bits & 2) ? 0 : 1;
  dL_dRGB.z *= (clamp_bits & 4) ? 0 : 1;
  const float g[3] = {dL_dRGB.x, dL_dRGB.y, dL_dRGB.z};
  const float x = dir.x, y = dir.y, z = dir.z;
  float w[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) w[i] = 0.f;
  w[0] = kSH0;
  float dx[3] = {0, 0, 0}, dy[3] = {0, 0, 0}, dz[3] = {0, 0, 0};
  auto S = [&](int i, int c) { return sh[3 * i + c]; };
  if (deg > 0) {
    w[1] = -kSH1 * y;
    w[2] = kSH1 * z;
    w[3] = -kSH1 * x;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      dx[c] = -kSH1 * S(3, c);
      dy[c] = -kSH1 * S(1, c);
      dz[c] = kSH1 * S(2, c);
    }
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z;
      const float xy = x * y, yz = y * z, xz = x * z;
      w[4] = kSH2[0] * xy;
      w[5] = kSH2[1] * yz;
      w[6] = kSH2[2] * (2.f * zz - xx - yy);
      w[7] = kSH2[3] * xz;
      w[8] = kSH2[4] * (xx - yy);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        dx[c] += kSH2[0] * y * S(4, c) + kSH2[2] * 2.f * -x * S(6, c) + kSH2[3] * z * S(7, c) +
                 kSH2[4] * 2.f * x * S(8, c);
        dy[c] += kSH2[0] * x * S(4, c) + kSH2[1] * z * S(5, c) + kSH2[2] * 2.f * -y * S(6, c) +
                 kSH2[4] * 2.f * -y * S(8, c);
        dz[c] += kSH2[1] * y * S(5, c) + kSH2[2] * 2.f * 2.f * z * S(6, c) + kSH2[3] * x * S(7, c);
      }
      if (deg > 2) {
        w[9] = kSH3[0] * y * (3.f * xx - yy);
        w[10] = kSH3[1] * xy * z;
        w[11] = kSH3[2] * y * (4.f * zz - xx - yy);
        w[12] = kSH3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
        w[13] = kSH3[4] * x * (4.f * zz - xx - yy);
        w[14] = kSH3[5] * z * (xx - yy);
        w[15] = kSH3[6] * x * (xx - 3.f * yy);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          dx[c] += (kSH3[0] * S(9, c) * 3.f * 2.f * xy + kSH3[1] * S(10, c) * yz +
                    kSH3[2] * S(11, c) * -2.f * xy + kSH3[3] * S(12, c) * -3.f * 2.f * xz +
                    kSH3[4] * S(13, c) * (-3.f * xx + 4.f * zz - yy) + kSH3[5] * S(14, c) * 2.f * xz +
                    kSH3[6] * S(15, c) * 3.f * (xx - yy));
          dy[c] += (kSH3[0] * S(9, c) * 3.f * (xx - yy) + kSH3[1] * S(10, c) * xz +
                    kSH3[2] * S(11, c) * (-3.f * yy + 4.f * zz - xx) + kSH3[3] * S(12, c) * -3.f * 2.f * yz +
                    kSH3[4] * S(13, c) * -2.f * xy + kSH3[5] * S(14, c) * -2.f * yz +
                    kSH3[6] * S(15, c) * -3.f * 2.f * xy);
          dz[c] += (kSH3[1] * S(10, c) * xy + kSH3[2] * S(11, c) * 4.f * 2.f * yz +
                    kSH3[3] * S(12, c) * 3.f * (2.f * zz - xx - yy) + kSH3[4] * S(13, c) * 4.f * 2.f * xz +
                    kSH3[5] * S(14, c) * (xx - yy));
        }
      }
    }
  }
  const int ncoef = (deg + 1) * (deg + 1);
  for (int i = 0; i < M; ++i) {
    const float wi = (i < ncoef && i < 16) ? w[i] : 0.f;
    dL_dsh[3 * i + 0] = wi * g[0];
    dL_dsh[3 * i + 1] = wi * g[1];
    dL_dsh[3 * i + 2] = wi * g[2];
  }
  const float3 dL_ddir = make_float3(dx[0] * g[0] + dx[1] * g[1] + dx[2] * g[2],
                                     dy[0] * g[0] + dy[1] * g[1] + dy[2] * g[2],
                                     dz[0] * g[0] + dz[1] * g[1] + dz[2] * g[2]);
  // d normalize(v) / dv (auxiliary.h:123-133)
  const float3 vv = dir_orig, dv = dL_ddir;
  const float sum2 = vv.x * vv.x + vv.y * vv.y + vv.z * vv.z;
  const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  float3 r;
  r.x = ((+sum2 - vv.x * vv.x) * dv.x - vv.y * vv.x * dv.y - vv.z * vv.x * dv.z) * invsum32;
  r.y = (-vv.x * vv.y * dv.x + (sum2 - vv.y * vv.y) * dv.y - vv.z * vv.y * dv.z) * invsum32;
  r.z = (-vv.x * vv.z * dv.x - vv.y * vv.z * dv.y + (sum2 - vv.z * vv.z) * dv.z) * invsum32;
  return r;
}

// dL/dcov3D -> dL/dscale, dL/dquat (backward.cu:492-555; quaternion not normalised in-kernel).
__device__ __forceinline__ void cov3d_bwd(const float3 scale, float mod, const float4 q, const float* dcov6,
                                          float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  const Mat3 R = quat_to_R(q);
  Mat3 S = mat3_cols(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
  const float3 s = make_float3(mod * scale.x, mod * scale.y, mod * scale.z);
  S.m[0][0] = s.x;
  S.m[1][1] = s.y;
  S.m[2][2] = s.z;
  const Mat3 M = mat3_mul(S, R);
  const Mat3 dSigma = mat3_cols(dcov6[0], 0.5f * dcov6[1], 0.5f * dcov6[2], 0.5f * dcov6[1], dcov6[3],
                                0.5f * dcov6[4], 0.5f * dcov6[2], 0.5f * dcov6[4], dcov6[5]);
  const Mat3 dM = mat3_mul(mat3_scale(M, 2.0f), dSigma);
  const Mat3 Rt = mat3_t(R);
  Mat3 dMt = mat3_t(dM);
  dL_dscale[0] = dot3(col3(Rt, 0), col3(dMt, 0));
  dL_dscale[1] = dot3(col3(Rt, 1), col3(dMt, 1));
  dL_dscale[2] = dot3(col3(Rt, 2), col3(dMt, 2));
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    dMt.m[0][k] *= s.x;
    dMt.m[1][k] *= s.y;
    dMt.m[2][k] *= s.z;
  }
  dL_drot[0] = 2 * z * (dMt.m[0][1] - dMt.m[1][0]) + 2 * y * (dMt.m[2][0] - dMt.m[0][2]) +
               2 * x * (dMt.m[1][2] - dMt.m[2][1]);
  dL_drot[1] = 2 * y * (dMt.m[1][0] + dMt.m[0][1]) + 2 * z * (dMt.m[2][0] + dMt.m[0][2]) +
               2 * r * (dMt.m[1][2] - dMt.m[2][1]) - 4 * x * (dMt.m[2][2] + dMt.m[1][1]);
  dL_drot[2] = 2 * x * (dMt.m[1][0] + dMt.m[0][1]) + 2 * r * (dMt.m[2][0] - dMt.m[0][2]) +
               2 * z * (dMt.m[1][2] + dMt.m[2][1]) - 4 * y * (dMt.m[2][2] + dMt.m[0][0]);
  dL_drot[3] = 2 * r * (dMt.m[0][1] - dMt.m[1][0]) + 2 * x * (dMt.m[2][0] + dMt.m[0][2]) +
               2 * y * (dMt.m[1][2] + dMt.m[2][1]) - 4 * z * (dMt.m[1][1] + dMt.m[0][0]);
}

// Per-Gaussian chain rule for one visible Gaussian (backward.cu:145-628). sh_in / sh_out are this
// Gaussian's [M][3] rows (they may be the same shared-memory row: every read precedes the writes).
template <bool COORD>
__device__ __forceinline__ void gaussian_bwd_one(
    int idx, int D, int M, const float* __restrict__ means, const float* sh_in, uint8_t clamp_bits,
    const float* __restrict__ scales, const float* __restrict__ rots, float scale_mod,
    const float* __restrict__ cov3D_pre, const float* __restrict__ view, const float* __restrict__ proj, float h_x,
    float h_y, float tan_fovx, float tan_fovy, float kernel_size, const float* __restrict__ campos,
    const float* __restrict__ grad, float* __restrict__ dL_dmean2D, float* __restrict__ dL_dcolor,
    float* __restrict__ dL_dopacity, float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D, float* sh_out,
    float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
  constexpr int V = COORD ? 32 : 16;
  const float4* gq = reinterpret_cast<const float4*>(grad + (size_t)idx * V);
  const float4 g0 = gq[0], g1 = gq[1], g2 = gq[2], g3 = gq[3];
  const float3 dmean2D = make_float3(g0.x, g0.y, g0.z);
  const float dopacity_in = g0.w;
  const float3 dL_dconic = make_float3(g1.x, g1.y, g1.z);  // (xx, xy, yy)
  const float dL_dt = g1.w;
  const float3 dcolor = make_float3(g2.x, g2.y, g2.z);
  const float2 dL_dray_plane = make_float2(g2.w, g3.x);
  const float3 dL_dnormal = make_float3(g3.y, g3.z, g3.w);
  float3 dL_dview_point = make_float3(0, 0, 0);
  float2 dcp0 = make_float2(0, 0), dcp1 = make_float2(0, 0), dcp2 = make_float2(0, 0);
  if constexpr (COORD) {
    const float4 g4 = gq[4], g5 = gq[5], g6 = gq[6];
    dL_dview_point = make_float3(g4.x, g4.y, g4.z);
    dcp0 = make_float2(g4.w, g5.x);
    dcp1 = make_float2(g5.y, g5.z);
    dcp2 = make_float2(g5.w, g6.x);
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) dL_dmean2D[3 * idx + i] = (&dmean2D.x)[i];
  dL_dcolor[3 * idx + 0] = dcolor.x;
  dL_dcolor[3 * idx + 1] = dcolor.y;
  dL_dcolor[3 * idx + 2] = dcolor.z;

  const float3 mean = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  float3 scale = make_float3(0, 0, 0);
  float4 quat = make_float4(0, 0, 0, 0);
  float cov6[6];
  if (cov3D_pre != nullptr) {
#pragma unroll
    for (int i = 0; i < 6; ++i) cov6[i] = cov3D_pre[6 * idx + i];
  } else {
    scale = make_float3(scales[3 * idx], scales[3 * idx + 1], scales[3 * idx + 2]);
    quat = make_float4(rots[4 * idx], rots[4 * idx + 1], rots[4 * idx + 2], rots[4 * idx + 3]);
    cov3d_from_scale_rot(scale, scale_mod, quat, cov6);
  }

  // ---------------- conic / plane / normal chain (backward.cu:166-488) ----------------
  Cov2DCtx c;
  cov2d_setup(mean, h_x, h_y, tan_fovx, tan_fovy, cov6, view, c);
  const float3 t = c.t;
  const float txtz = c.txtz, tytz = c.tytz;
  const Mat3& W = c.W;
  const Mat3& T = c.T;
  const Mat3& Vrk = c.Vrk;
  const Mat3& cov2D = c.cov;
  // QUIRK: the reference hands dL_dconic to this stage in the conic_opacity slot
  // (rasterizer_impl.cu:576), so "combined opacity" is really the yy conic gradient.
  const float combined_opacity = dL_dconic.z;

  const float det_0 = max(1e-6, cov2D.m[0][0] * cov2D.m[1][1] - cov2D.m[0][1] * cov2D.m[0][1]);
  const float det_1 = max(1e-6, (cov2D.m[0][0] + kernel_size) * (cov2D.m[1][1] + kernel_size) -
                                    cov2D.m[0][1] * cov2D.m[0][1]);
  const float coef = sqrt(det_0 / (det_1 + 1e-6) + 1e-6);

  Mat3 evec;
  float eval[3];
  const int Dn = eig_sym3(Vrk, eval, evec);
  const int min_id = argmin3(eval);
  Mat3 Vrk_inv;
  float3 evec_min = make_float3(0, 0, 0);
  const bool well = eval[min_id] > 0.00000001;
  if (well) {
    const Mat3 diag = mat3_cols(1 / eval[0], 0, 0, 0, 1 / eval[1], 0, 0, 0, 1 / eval[2]);
    Vrk_inv = mat3_mul(mat3_mul(evec, diag), mat3_t(evec));
  } else {
    evec_min = col3(evec, min_id);
    Vrk_inv = mat3_outer(evec_min, evec_min);
  }
  const Mat3 cov_cam_inv = mat3_mul(mat3_mul(mat3_t(W), Vrk_inv), W);
  const float3 uvh = make_float3(txtz, tytz, 1);
  const float3 uvh_m = mat3_vec(cov_cam_inv, uvh);
  const float3 uvh_mn = f3_normalize(uvh_m);

  const float u2 = txtz * txtz, v2 = tytz * tytz, uv = txtz * tytz;
  Mat3 dL_dVrk = mat3_zero();
  Mat3 dL_dnJ = mat3_zero();
  float3 plane = make_float3(0, 0, 0);
  float dL_du = 0, dL_dv = 0, dL_dl = 0, l = 1, nl = 1;
  if (!(isnan(uvh_mn.x) || Dn == 0)) {
    const float vb = dot3(uvh_m, uvh);
    const float vbn = dot3(uvh_mn, uvh);
    l = sqrtf(t.x * t.x + t.y * t.y + t.z * t.z);
    const Mat3 nJ = mat3_cols(1 / t.z, 0.0f, -(t.x) / (t.z * t.z), 0.0f, 1 / t.z, -(t.y) / (t.z * t.z),
                              t.x / l, t.y / l, t.z / l);
    const Mat3 nJ_inv = mat3_cols(v2 + 1, -uv, 0, -uv, u2 + 1, 0, -txtz, -tytz, 0);
    const float clamp_vb = max(vb, 0.0000001f);
    const float clamp_vbn = max(vbn, 0.0000001f);
    nl = u2 + v2 + 1;
    const float factor_normal = l / nl;
    const float3 uvh_m_vb = make_float3(uvh_mn.x / clamp_vbn, uvh_mn.y / clamp_vbn, uvh_mn.z / clamp_vbn);
    plane = mat3_vec(nJ_inv, uvh_m_vb);
    const float2 cam0 = make_float2((-(v2 + 1) * t.z + plane.x * t.x) / nl, (uv * t.z + plane.y * t.x) / nl);
    const float2 cam1 = make_float2((uv * t.z + plane.x * t.y) / nl, (-(u2 + 1) * t.z + plane.y * t.y) / nl);
    const float2 cam2 = make_float2((t.x + plane.x * t.z) / nl, (t.y + plane.y * t.z) / nl);
    const float2 ray_plane = make_float2(plane.x * factor_normal, plane.y * factor_normal);
    const float3 ray_n = make_float3(-plane.x * factor_normal, -plane.y * factor_normal, -1);
    const float3 cam_n = mat3_vec(nJ, ray_n);
    const float3 nvec = f3_normalize(cam_n);
    const float lv = sqrtf(dot3(cam_n, cam_n));
    const float3 dn_lv = make_float3(dL_dnormal.x / lv, dL_dnormal.y / lv, dL_dnormal.z / lv);
    const float3 dcam_n = f3_sub(dn_lv, f3_scale(nvec, dot3(nvec, dn_lv)));
    const float3 dray_n = mat3_vec(mat3_t(nJ), dcam_n);
    dL_dnJ = mat3_outer(dcam_n, ray_n);
    dL_dl = (-plane.x * dray_n.x - plane.y * dray_n.y + plane.x * dL_dray_plane.x + plane.y * dL_dray_plane.y) / nl;
    const float2 dL_dplane = make_float2(
        (t.x * dcp0.x + t.y * dcp1.x + t.z * dcp2.x - l * dray_n.x + dL_dray_plane.x * l) / nl,
        (t.x * dcp0.y + t.y * dcp1.y + t.z * dcp2.y - l * dray_n.y + dL_dray_plane.y * l) / nl);
    const float3 dplane3 = make_float3(dL_dplane.x, dL_dplane.y, 0);
    const float dL_dnl = (-dcp0.x * cam0.x - dcp0.y * cam0.y - dcp1.x * cam1.x - dcp1.y * cam1.y -
                          dcp2.x * cam2.x - dcp2.y * cam2.y - dray_n.x * ray_n.x - dray_n.y * ray_n.y -
                          dL_dray_plane.x * ray_plane.x - dL_dray_plane.y * ray_plane.y) / nl;
    const float tmp = dL_dplane.x * plane.x + dL_dplane.y * plane.y;
    const float3 W_uvh = mat3_vec(W, uvh);
    if (well) {
      const float3 lhs = mat3_vec(Vrk_inv, W_uvh);
      // (Vrk_inv / vb) * (W_uvh * (-tmp) + W * nJ_inv^T * dplane3)
      const float3 inner = f3_add(f3_scale(W_uvh, -tmp), mat3_vec(mat3_mul(W, mat3_t(nJ_inv)), dplane3));
      Mat3 Vd;
#pragma unroll
      for (int cc = 0; cc < 3; ++cc)
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) Vd.m[cc][rr] = Vrk_inv.m[cc][rr] / clamp_vb;
      const Mat3 op = mat3_outer(lhs, mat3_vec(Vd, inner));
#pragma unroll
      for (int cc = 0; cc < 3; ++cc)
#pragma unroll
        for (int rr = 0; rr < 3; ++rr) dL_dVrk.m[cc][rr] = -op.m[cc][rr];
    } else {
      const float dL_dvb = -tmp / clamp_vb;
      const float3 nJi_dp = mat3_vec(mat3_t(nJ_inv), make_float3(dL_dplane.x / clamp_vb, dL_dplane.y / clamp_vb, 0));
      const Mat3 dVinv = mat3_outer(W_uvh, f3_add(f3_scale(W_uvh, dL_dvb), mat3_vec(W, nJi_dp)));
      const float3 dLdv = mat3_vec(mat3_add(dVinv, mat3_t(dVinv)), evec_min);
      for (int j = 0; j < 3; ++j) {
        if (j != min_id) {
          const float sc = dot3(col3(evec, j), dLdv) / min(eval[min_id] - eval[j], -0.0000001f);
          dL_dVrk = mat3_add(dL_dVrk, mat3_outer(f3_scale(col3(evec, j), sc), evec_min));
        }
      }
    }
    Mat3 cci;
#pragma unroll
    for (int cc = 0; cc < 3; ++cc)
#pragma unroll
      for (int rr = 0; rr < 3; ++rr) cci.m[cc][rr] = cov_cam_inv.m[cc][rr] / clamp_vb;
    const float3 second = mat3_vec(mat3_mul(cci, mat3_t(nJ_inv)), dplane3);
    const float two_tmp = 2 * (-tmp);
    const float3 dL_duvh = make_float3(two_tmp * uvh_m_vb.x + second.x, two_tmp * uvh_m_vb.y + second.y,
                                       two_tmp * uvh_m_vb.z + second.z);
    const Mat3 dnJinv = mat3_outer(dplane3, uvh_m_vb);
    dL_du = dL_dnl * 2 * txtz + dL_duvh.x + (dnJinv.m[0][1] + dnJinv.m[1][0]) * (-tytz) +
            2 * dnJinv.m[1][1] * txtz - dnJinv.m[2][0] +
            (dcp0.y * t.y + dcp1.x * t.y + dcp1.y * (-2 * t.x)) / nl;
    dL_dv = dL_dnl * 2 * tytz + dL_duvh.y + (dnJinv.m[0][1] + dnJinv.m[1][0]) * (-txtz) +
            2 * dnJinv.m[0][0] * tytz - dnJinv.m[2][1] +
            (dcp0.x * (-2 * t.y) + dcp0.y * t.x + dcp1.x * t.x) / nl;
  }

  const float opacity = combined_opacity / (coef + 1e-6);
  const float dL_dcoef = dopacity_in * opacity;
  const float dL_dsqrtcoef = dL_dcoef * 0.5 * 1. / (coef + 1e-6);
  const float dL_ddet0 = dL_dsqrtcoef / (det_1 + 1e-6);
  const float dL_ddet1 = dL_dsqrtcoef * det_0 * (-1.f / (det_1 * det_1 + 1e-6));
  const float dcoef_da = dL_ddet0 * cov2D.m[1][1] + dL_ddet1 * (cov2D.m[1][1] + kernel_size);
  const float dcoef_db = dL_ddet0 * (-2. * cov2D.m[0][1]) + dL_ddet1 * (-2. * cov2D.m[0][1]);
  const float dcoef_dc = dL_ddet0 * cov2D.m[0][0] + dL_ddet1 * (cov2D.m[0][0] + kernel_size);
  const float a = cov2D.m[0][0] + kernel_size;
  const float b = cov2D.m[0][1];
  const float cc_ = cov2D.m[1][1] + kernel_size;
  const float denom = a * cc_ - b * b;
  float dL_da = 0, dL_db = 0, dL_dc = 0;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float dcov[6];
  float dopacity_out = dopacity_in;
  if (denom2inv != 0) {
    dL_da = denom2inv * (-cc_ * cc_ * dL_dconic.x + 2 * b * cc_ * dL_dconic.y + (denom - a * cc_) * dL_dconic.z);
    dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * cc_) * dL_dconic.x);
    dL_db = denom2inv * 2 * (b * cc_ * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
    if (det_0 <= 1e-6 || det_1 <= 1e-6) {
      dopacity_out = 0;
    } else {
      dL_da += dcoef_da;
      dL_dc += dcoef_dc;
      dL_db += dcoef_db;
      dopacity_out = dopacity_in * coef;
    }
    dcov[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
    dcov[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
    dcov[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
    dcov[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db +
              2 * T.m[1][0] * T.m[1][1] * dL_dc;
    dcov[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db +
              2 * T.m[1][0] * T.m[1][2] * dL_dc;
    dcov[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db +
              2 * T.m[1][1] * T.m[1][2] * dL_dc;
  } else {
#pragma unroll
    for (int i = 0; i < 6; ++i) dcov[i] = 0;
  }
  dcov[0] += dL_dVrk.m[0][0];
  dcov[3] += dL_dVrk.m[1][1];
  dcov[5] += dL_dVrk.m[2][2];
  dcov[1] += dL_dVrk.m[0][1] + dL_dVrk.m[1][0];
  dcov[2] += dL_dVrk.m[0][2] + dL_dVrk.m[2][0];
  dcov[4] += dL_dVrk.m[1][2] + dL_dVrk.m[2][1];
  dL_dopacity[idx] = dopacity_out;
#pragma unroll
  for (int i = 0; i < 6; ++i) dL_dcov3D[6 * idx + i] = dcov[i];

  const float dL_dT00 = 2 * (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_da +
                        (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_db;
  const float dL_dT01 = 2 * (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_da +
                        (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_db;
  const float dL_dT02 = 2 * (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_da +
                        (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_db;
  const float dL_dT10 = 2 * (T.m[1][0] * Vrk.m[0][0] + T.m[1][1] * Vrk.m[0][1] + T.m[1][2] * Vrk.m[0][2]) * dL_dc +
                        (T.m[0][0] * Vrk.m[0][0] + T.m[0][1] * Vrk.m[0][1] + T.m[0][2] * Vrk.m[0][2]) * dL_db;
  const float dL_dT11 = 2 * (T.m[1][0] * Vrk.m[1][0] + T.m[1][1] * Vrk.m[1][1] + T.m[1][2] * Vrk.m[1][2]) * dL_dc +
                        (T.m[0][0] * Vrk.m[1][0] + T.m[0][1] * Vrk.m[1][1] + T.m[0][2] * Vrk.m[1][2]) * dL_db;
  const float dL_dT12 = 2 * (T.m[1][0] * Vrk.m[2][0] + T.m[1][1] * Vrk.m[2][1] + T.m[1][2] * Vrk.m[2][2]) * dL_dc +
                        (T.m[0][0] * Vrk.m[2][0] + T.m[0][1] * Vrk.m[2][1] + T.m[0][2] * Vrk.m[2][2]) * dL_db;
  const float dL_dJ00 = W.m[0][0] * dL_dT00 + W.m[0][1] * dL_dT01 + W.m[0][2] * dL_dT02;
  const float dL_dJ02 = W.m[2][0] * dL_dT00 + W.m[2][1] * dL_dT01 + W.m[2][2] * dL_dT02;
  const float dL_dJ11 = W.m[1][0] * dL_dT10 + W.m[1][1] * dL_dT11 + W.m[1][2] * dL_dT12;
  const float dL_dJ12 = W.m[2][0] * dL_dT10 + W.m[2][1] * dL_dT11 + W.m[2][2] * dL_dT12;

  const float tz = 1.f / t.z;
  const float tz2 = tz * tz;
  const float tz3 = tz2 * tz;
  const float l3 = l * l * l;
  const float dL_dtx =
      c.x_mul * (-h_x * tz2 * dL_dJ02 + dL_du * tz - dL_dnJ.m[0][2] * tz2 + dL_dnJ.m[2][0] * (1 / l - t.x * t.x / l3) +
                 dL_dnJ.m[2][1] * (-t.x * t.y / l3) + dL_dnJ.m[2][2] * (-t.x * t.z / l3) +
                 (dcp0.x * plane.x + dcp0.y * plane.y + dcp2.x) / nl + dL_dl * t.x / l);
  const float dL_dty =
      c.y_mul * (-h_y * tz2 * dL_dJ12 + dL_dv * tz - dL_dnJ.m[1][2] * tz2 + dL_dnJ.m[2][0] * (-t.x * t.y / l3) +
                 dL_dnJ.m[2][1] * (1 / l - t.y * t.y / l3) + dL_dnJ.m[2][2] * (-t.y * t.z / l3) +
                 (dcp1.x * plane.x + dcp1.y * plane.y + dcp2.y) / nl + dL_dl * t.y / l);
  const float dL_dtz =
      -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12 -
      (dL_du * t.x + dL_dv * t.y) * tz2 + (dL_dnJ.m[0][0] + dL_dnJ.m[1][1]) * (-tz2) +
      dL_dnJ.m[0][2] * (2 * t.x * tz3) + dL_dnJ.m[1][2] * (2 * t.y * tz3) +
      (dL_dnJ.m[2][0] * t.x + dL_dnJ.m[2][1] * t.y) * (-t.z / l3) + dL_dnJ.m[2][2] * (1 / l - t.z * t.z / l3) +
      (dcp0.x * (-(v2 + 1)) + dcp0.y * uv + dcp1.x * uv + dcp1.y * (-(u2 + 1)) + dcp2.x * plane.x + dcp2.y * plane.y) / nl +
      dL_dl * t.z / l;
  float3 dmean = xform_vec_4x3_t(make_float3(dL_dtx, dL_dty, dL_dtz), view);

  // ---------------- projection / view-point / t chain (backward.cu:587-619) ----------------
  {
    const float3 m = mean;
    const float4 m_hom = xform_4x4(m, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
    float3 d1;
    d1.x = (proj[0] * m_w - proj[3] * mul1) * dmean2D.x + (proj[1] * m_w - proj[3] * mul2) * dmean2D.y;
    d1.y = (proj[4] * m_w - proj[7] * mul1) * dmean2D.x + (proj[5] * m_w - proj[7] * mul2) * dmean2D.y;
    d1.z = (proj[8] * m_w - proj[11] * mul1) * dmean2D.x + (proj[9] * m_w - proj[11] * mul2) * dmean2D.y;
    const float3 m_view = xform_4x3(m, view);
    const float tt = sqrtf(m_view.x * m_view.x + m_view.y * m_view.y + m_view.z * m_view.z);
    const float3 d2 = xform_vec_4x3_t(make_float3(dL_dview_point.x + m_view.x / tt * dL_dt,
                                                  dL_dview_point.y + m_view.y / tt * dL_dt,
                                                  dL_dview_point.z + m_view.z / tt * dL_dt), view);
    dmean.x += d1.x + d2.x;
    dmean.y += d1.y + d2.y;
    dmean.z += d1.z + d2.z;
  }
  if (sh_in) {
    const float3 dsh_mean = sh_bwd(D, M, mean, make_float3(campos[0], campos[1], campos[2]), sh_in, clamp_bits, dcolor,
                                   sh_out);
    dmean.x += dsh_mean.x;
    dmean.y += dsh_mean.y;
    dmean.z += dsh_mean.z;
  }
  dL_dmean3D[3 * idx + 0] = dmean.x;
  dL_dmean3D[3 * idx + 1] = dmean.y;
  dL_dmean3D[3 * idx + 2] = dmean.z;
  if (scales) cov3d_bwd(scale, scale_mod, quat, dcov, dL_dscale + 3 * idx, dL_drot + 4 * idx);
}

// One thread per Gaussian. The SH rows (48 floats per Gaussian at M = 16: the bulk of this kernel's
// traffic) go through shared memory: a warp loads the rows of its visible Gaussians with coalesced
// 16-byte reads, every thread works in place on its own row (row stride 49 words: conflict-free), and the
// warp writes the 32 gradient rows back as one contiguous 6 KB run.
constexpr int GB_THREADS = 128;
constexpr int GB_SH_ROW = 49;

template <bool COORD>
// (measured: asking for 5 / 6 / 8 CTAs per SM instead of 4 spills 250-420 bytes and costs 0.54 / 0.65 / 0.73 ms against 0.46)
__global__ void __launch_bounds__(GB_THREADS, 4)
gaussian_bwd(int P, int D, int M, const float* __restrict__ means, const int* __restrict__ radii,
             const float* __restrict__ shs, const uint8_t* __restrict__ clamped,
             const float* __restrict__ scales, const float* __restrict__ rots, float scale_mod,
             const float* __restrict__ cov3D_pre, const float* __restrict__ view,
             const float* __restrict__ proj, float h_x, float h_y, float tan_fovx, float tan_fovy,
             float kernel_size, const float* __restrict__ campos, const float* __restrict__ grad,
             float* __restrict__ dL_dmean2D, float* __restrict__ dL_dcolor, float* __restrict__ dL_dopacity,
             float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh,
             float* __restrict__ dL_dscale, float* __restrict__ dL_drot) {
  __shared__ float s_sh[GB_THREADS / 32][32 * GB_SH_ROW];
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const bool inb = idx < P;
  const bool vis = inb && radii[idx] > 0;
  const bool staged = shs != nullptr && M == 16;
  const unsigned vmask = __ballot_sync(0xffffffffu, vis);
  const long wbase = (long)blockIdx.x * blockDim.x + wrp * 32;   // first Gaussian of this warp
  float* const wsh = s_sh[wrp];
  float* const srow = wsh + lane * GB_SH_ROW;
  if (staged && vmask != 0u) {
    const float4* g4 = reinterpret_cast<const float4*>(shs + wbase * 48);
#pragma unroll 4
    for (int q = lane; q < 32 * 12; q += 32) {
      const int row = q / 12, c4 = q - row * 12;
      if ((vmask >> row) & 1u) {
        const float4 v = __ldg(g4 + q);
        float* d = wsh + row * GB_SH_ROW + c4 * 4;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
      }
    }
    __syncwarp();
  }
  if (vis) {
    const float* sh_in = staged ? srow : (shs ? shs + (size_t)idx * M * 3 : nullptr);
    float* sh_out = staged ? srow : (shs ? dL_dsh + (size_t)idx * M * 3 : nullptr);
    gaussian_bwd_one<COORD>(idx, D, M, means, sh_in, clamped[idx], scales, rots, scale_mod, cov3D_pre, view, proj, h_x,
                            h_y, tan_fovx, tan_fovy, kernel_size, campos, grad, dL_dmean2D, dL_dcolor, dL_dopacity,
                            dL_dmean3D, dL_dcov3D, sh_out, dL_dscale, dL_drot);
  } else if (inb) {
    // The reference leaves its zero-initialised outputs untouched for culled Gaussians.
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      dL_dmean2D[3 * idx + i] = 0;
      dL_dcolor[3 * idx + i] = 0;
      dL_dmean3D[3 * idx + i] = 0;
    }
    dL_dopacity[idx] = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i) dL_dcov3D[6 * idx + i] = 0;
    if (staged) {
#pragma unroll
      for (int i = 0; i < 48; ++i) srow[i] = 0.f;
    } else if (shs) {
      for (int i = 0; i < 3 * M; ++i) dL_dsh[(size_t)idx * M * 3 + i] = 0;
    }
    if (scales) {
#pragma unroll
      for (int i = 0; i < 3; ++i) dL_dscale[3 * idx + i] = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) dL_drot[4 * idx + i] = 0;
    }
  }
  if (staged) {
    __syncwarp();
    float4* o4 = reinterpret_cast<float4*>(dL_dsh + wbase * 48);
#pragma unroll 4
    for (int q = lane; q < 32 * 12; q += 32) {
      const int row = q / 12, c4 = q - row * 12;
      if (wbase + row < P) {
        const float* d = wsh + row * GB_SH_ROW + c4 * 4;
        o4[q] = make_float4(d[0], d[1], d[2], d[3]);
      }
    }
  }
}

template <bool COORD, bool DEPTH>
static int launch_blend_bwd(dim3 grid, cudaStream_t stream, const RasterBwdArgs& a, const GeomState& geom,
                            const ImageState& img, const uint32_t* plist, float fx, float fy) {
  using Cfg = BwdCfg<COORD, DEPTH>;
  // per-device attribute: set on every launch
  if (cudaFuncSetAttribute(blend_bwd<COORD, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::smem) !=
      cudaSuccess)
    return 1;
  blend_bwd<COORD, DEPTH><<<grid, BWD_THREADS, Cfg::smem, stream>>>(
      img.ranges, plist, a.width, a.height, a.background, geom.rec, geom.rec_coord, a.alphas, img.accum_coord,
      img.accum_depth, img.normal_length, img.n_contrib, a.dL_dpix, a.dL_dpix_coord, a.dL_dpix_mcoord,
      a.dL_dpix_depth, a.dL_dpix_mdepth, a.dL_dalphas, a.dL_dpixel_normals, a.normalmap, fx, fy, a.grad_scratch);
  return 0;
}

static uint32_t tile_id_bits_b(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

int raster_backward(const RasterBwdArgs& a, cudaStream_t stream) {
  const int P = a.P, W = a.width, H = a.height;
  if (P == 0) return ED3DGS_OK;
  const bool coord = a.require_coord != 0, depth = a.require_depth != 0;
  const uint32_t gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
  const size_t npix = (size_t)W * H, ntiles = (size_t)gx * gy;
  char* gp = a.geom_buffer;
  char* ip = a.image_buffer;
  char* bp = a.binning_buffer;
  GeomState geom = carve_geom(gp, P, coord);
  ImageState img = carve_image(ip, npix, ntiles, coord);
  BinningState bin = carve_binning(bp, (size_t)a.R);
  const int sel = sorted_half((size_t)a.R, 32 + (int)tile_id_bits_b(gx * gy));
  const float focal_y = H / (2.0f * a.tan_fovy);
  const float focal_x = W / (2.0f * a.tan_fovx);
  const int V = coord ? 32 : 16;

  cudaError_t err = a.packed_is_input
                        ? cudaSuccess
                        : cudaMemsetAsync(a.grad_scratch, 0, (size_t)P * V * sizeof(float), stream);
  if (err != cudaSuccess) { set_error("memset of gradient scratch failed: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }

  if (a.R > 0 && !a.packed_is_input) {
    ProfScope ps("blend_bwd", stream);
    const dim3 grid(gx, gy, 1);
    const uint32_t* plist = bin.vals[sel];
    int bad;
    if (coord && depth) bad = launch_blend_bwd<true, true>(grid, stream, a, geom, img, plist, focal_x, focal_y);
    else if (coord) bad = launch_blend_bwd<true, false>(grid, stream, a, geom, img, plist, focal_x, focal_y);
    else if (depth) bad = launch_blend_bwd<false, true>(grid, stream, a, geom, img, plist, focal_x, focal_y);
    else bad = launch_blend_bwd<false, false>(grid, stream, a, geom, img, plist, focal_x, focal_y);
    if (bad) { set_error("blend_bwd: %s", cudaGetErrorString(cudaGetLastError())); return ED3DGS_ERR_CUDA; }
  }
  if (a.debug) {
    err = cudaStreamSynchronize(stream);
    if (err != cudaSuccess) { set_error("[CUDA ERROR] after blend_bwd: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  }
  const int threads = GB_THREADS, blocks = (P + threads - 1) / threads;
#define ED_GBWD(C_)                                                                                          \
  gaussian_bwd<C_><<<blocks, threads, 0, stream>>>(                                                          \
      P, a.D, a.M, a.means3D, a.radii, a.shs, geom.clamped, a.scales, a.rotations, a.scale_modifier,         \
      a.cov3D_precomp, a.viewmatrix, a.projmatrix, focal_x, focal_y, a.tan_fovx, a.tan_fovy, a.kernel_size,  \
      a.campos, a.grad_scratch, a.dL_dmean2D, a.dL_dcolor, a.dL_dopacity, a.dL_dmean3D, a.dL_dcov3D,         \
      a.dL_dsh, a.dL_dscale, a.dL_drot)
  {
  ProfScope ps("gaussian_bwd", stream);
  if (coord) ED_GBWD(true); else ED_GBWD(false);
  }
#undef ED_GBWD
  err = a.debug ? cudaStreamSynchronize(stream) : cudaSuccess;
  if (err == cudaSuccess) err = cudaGetLastError();
  if (err != cudaSuccess) { set_error("[CUDA ERROR] in backward: %s", cudaGetErrorString(err)); return ED3DGS_ERR_CUDA; }
  return ED3DGS_OK;
}

}  // namespace ed3dgs
