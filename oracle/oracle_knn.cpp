// TEST INFRASTRUCTURE -- CPU restatement of simple-knn's distCUDA2 (simple_knn.cu:185-221):
// mean of the 3 smallest squared distances to the other points. The reference prunes with Morton
// boxes but is exact, so a brute-force scan with the same distance expression and the same
// K-best insertion (updateKBest, simple_knn.cu:131-145) yields the same numbers.
#include <cfloat>
#include <cstdint>

extern "C" void orc_dist_cuda2(int P, const float* pts, float* out) {
#pragma omp parallel for schedule(dynamic, 64)
  for (int i = 0; i < P; ++i) {
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    const float rx = pts[3 * i], ry = pts[3 * i + 1], rz = pts[3 * i + 2];
    for (int j = 0; j < P; ++j) {
      if (j == i) continue;
      const float dx = pts[3 * j] - rx, dy = pts[3 * j + 1] - ry, dz = pts[3 * j + 2] - rz;
      float dist = __builtin_fmaf(dz, dz, __builtin_fmaf(dy, dy, dx * dx));  // nvcc contracts d.x*d.x+d.y*d.y+d.z*d.z this way
      for (int k = 0; k < 3; k++) {
        if (best[k] > dist) {
          const float t = best[k];
          best[k] = dist;
          dist = t;
        }
      }
    }
    out[i] = (best[0] + best[1] + best[2]) / 3.0f;
  }
}
