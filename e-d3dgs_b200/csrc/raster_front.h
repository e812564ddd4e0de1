// The stages the image rasterizer and the point integrator share (rasterizer_impl.cu:255-400 and
// :580-712 are the same code twice in the reference): per-Gaussian preprocess, tile-count scan,
// key emission, radix sort, tile ranges. Host-side helpers and error macros for both callers.
#pragma once
#include "raster_api.h"
#include "raster_state.cuh"

namespace ed3dgs {

struct FrontEnd {
  GeomState geom;
  ImageState img;
  BinningState bin;
  int sel = 0;           // which ping-pong half holds the sorted list
  uint32_t R = 0;        // (Gaussian, tile) instances
  uint32_t gx = 0, gy = 0;
  float focal_x = 0, focal_y = 0;
};

// inte_rec != nullptr selects the integration flavour of the preprocess (no tongue class, extra record).
int raster_front(const RasterFwdArgs& a, float4* inte_rec, FrontEnd& fe, cudaStream_t stream);

#define ED_CHECK(stmt)                                                                        \
  do {                                                                                        \
    cudaError_t err__ = (stmt);                                                               \
    if (err__ != cudaSuccess) {                                                               \
      set_error("%s failed: %s (%s:%d)", #stmt, cudaGetErrorString(err__), __FILE__, __LINE__); \
      return ED3DGS_ERR_CUDA;                                                                 \
    }                                                                                         \
  } while (0)

#define ED_DEBUG_SYNC(what)                                                                   \
  do {                                                                                        \
    if (a.debug) {                                                                            \
      cudaError_t err__ = cudaStreamSynchronize(stream);                                      \
      if (err__ == cudaSuccess) err__ = cudaGetLastError();                                   \
      if (err__ != cudaSuccess) {                                                             \
        set_error("[CUDA ERROR] after %s: %s", what, cudaGetErrorString(err__));              \
        return ED3DGS_ERR_CUDA;                                                               \
      }                                                                                       \
    }                                                                                         \
  } while (0)

}  // namespace ed3dgs
