// Layout of the three opaque byte buffers the rasterizer keeps between forward and backward.
// The reference carves SoA arrays out of them (rasterizer_impl.cu:190-251, structs in
// rasterizer_impl.h:29-87); the layout is invisible to Python, so it is redesigned here:
// one 64-byte record per Gaussian (four float4) holds everything the blend loops gather, so a
// tile stages a Gaussian with four 16-byte cp.async copies instead of eight scattered loads.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ed3dgs {

// Record quads (per Gaussian, written by preprocess, read by both blend kernels):
//   q0 = {mean2D.x, mean2D.y, t (=|p_view|), tongue}
//   q1 = {conic.x, conic.y, conic.z, opacity * mip_coef}
//   q2 = {r, g, b, ray_plane.x}
//   q3 = {ray_plane.y, normal.x, normal.y, normal.z}
// Extra quads when coordinates are rendered (require_coord):
//   c0 = {camera_plane[0..3]}   c1 = {camera_plane[4], camera_plane[5], p_view.x, p_view.y}
//   c2 = {p_view.z, 0, 0, 0}
constexpr int REC_QUADS = 4;
constexpr int REC_COORD_QUADS = 3;

// Packed per-Gaussian gradient rows produced by the backward blend kernel (one vector
// red.global.add.v4.f32 per quad per tile) and consumed by the per-Gaussian backward kernel.
//   g0 = {dmean2D.x, dmean2D.y, |dmean2D|, dopacity}
//   g1 = {dconic.x (xx), dconic.y (xy), dconic.w (yy), dt}
//   g2 = {dr, dg, db, dray_plane.x}
//   g3 = {dray_plane.y, dnormal.x, dnormal.y, dnormal.z}
//   g4 = {dview_point.xyz, dcamera_plane[0]}  g5 = {dcamera_plane[1..4]}  g6 = {dcamera_plane[5],0,0,0}
constexpr int GRAD_QUADS = 4;
constexpr int GRAD_QUADS_COORD = 8;  // 7 used, padded to 8 for the warp transpose-reduce

struct GeomState {
  float4* rec;             // [P][4]
  float4* rec_coord;       // [P][3]   (only when require_coord)
  float* depths;           // [P]
  uint32_t* tiles_touched; // [P]
  uint32_t* point_offsets; // [P]      inclusive scan of tiles_touched
  uint32_t* block_sums;    // [ceil(P/SCAN_TILE)+1]
  uint8_t* clamped;        // [P]      bit c set when SH colour channel c was clamped at 0
};

struct ImageState {
  uint2* ranges;           // [tiles]
  uint32_t* n_contrib;     // [2][W*H] last contributor, median contributor
  float* accum_depth;      // [W*H]
  float* normal_length;    // [W*H]
  float* accum_coord;      // [3][W*H] (only when require_coord)
};

struct BinningState {
  uint64_t* keys[3];       // [0],[1] = sort ping-pong; [2] = unsorted emission order (kept intact)
  uint32_t* vals[3];
  uint32_t* hist;          // radix sort scratch: digit offsets of every pass, tile tickets, look-back status words
};

constexpr int SCAN_TILE = 2048;
constexpr int SORT_TILE = 4096;  // keys per radix block

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
  p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256));
  out = reinterpret_cast<T*>(p);
  p += count * sizeof(T);
}

inline GeomState carve_geom(char*& p, size_t P, bool coord) {
  GeomState g{};
  carve(p, g.rec, P * REC_QUADS);
  carve(p, g.rec_coord, coord ? P * REC_COORD_QUADS : 0);
  carve(p, g.depths, P);
  carve(p, g.tiles_touched, P);
  carve(p, g.point_offsets, P);
  carve(p, g.block_sums, (P + SCAN_TILE - 1) / SCAN_TILE + 1);
  carve(p, g.clamped, P);
  return g;
}
inline size_t geom_bytes(size_t P, bool coord) {
  char* p = nullptr;
  carve_geom(p, P, coord);
  return reinterpret_cast<size_t>(p) + 256;
}

inline ImageState carve_image(char*& p, size_t npix, size_t ntiles, bool coord) {
  ImageState s{};
  carve(p, s.ranges, ntiles);
  carve(p, s.n_contrib, 2 * npix);
  carve(p, s.accum_depth, npix);
  carve(p, s.normal_length, npix);
  carve(p, s.accum_coord, coord ? 3 * npix : 0);
  return s;
}
inline size_t image_bytes(size_t npix, size_t ntiles, bool coord) {
  char* p = nullptr;
  carve_image(p, npix, ntiles, coord);
  return reinterpret_cast<size_t>(p) + 256;
}

inline size_t sort_blocks(size_t R) { return (R + SORT_TILE - 1) / SORT_TILE; }
constexpr int SORT_MAX_PASSES = 8;   // 64 key bits, 8 per pass
// words of radix sort scratch used by `npasses` passes over R keys (radix_sort.cu: digit_base | ticket | status)
inline size_t sort_aux_words(size_t R, int npasses) {
  return (size_t)SORT_MAX_PASSES * 256 + 256 + (size_t)npasses * sort_blocks(R) * 256;
}

inline BinningState carve_binning(char*& p, size_t R) {
  BinningState b{};
  carve(p, b.keys[0], R);
  carve(p, b.keys[1], R);
  carve(p, b.vals[0], R);
  carve(p, b.vals[1], R);
  carve(p, b.keys[2], R);
  carve(p, b.vals[2], R);
  carve(p, b.hist, sort_aux_words(R, SORT_MAX_PASSES));
  return b;
}
// Which ping-pong half holds the sorted list after sorting key bits [0, end_bit).
// (pass 0 reads the emission buffers [2] and writes [0]; later passes alternate [0] <-> [1].)
inline int sorted_half(size_t R, int end_bit) { return (R == 0 || (((end_bit + 7) / 8) & 1)) ? 0 : 1; }

inline size_t binning_bytes(size_t R) {
  char* p = nullptr;
  carve_binning(p, R);
  return reinterpret_cast<size_t>(p) + 256;
}

}  // namespace ed3dgs
