// occupancy.cuh -- Occupancy.set / get geometry (decoder/occupancy.py:10-47) shared by the pose
// assembly (assemble.cu) and the detection seed loop (det.cu).  Maps are never rasterised: a seed's
// occupancy value is the number of set boxes of its field that cover its cell, mod 256.
#pragma once
#include "common.cuh"

namespace pp {

// round(np.float32) -> Python int (half to even), clamped like the oracle.
__device__ __forceinline__ long long py_round(float v) {
  double r = rint((double)v);
  if (r > 4e18) r = 4e18;
  if (r < -4e18) r = -4e18;
  return (long long)r;
}

// Occupancy.set box (occupancy.py:31-39 + utils.py:61-66) for a map of (oh, ow) cells.
// min_scale_reduced = min_scale / reduction: 2 for CifCaf (cifcaf.py:84), 1 for CifDet (cifdet.py:36).
__device__ __forceinline__ void occupancy_box(float x, float y, float sigma, long long ow,
                                              long long oh, long long *box,
                                              float min_scale_reduced = 2.0f) {
  long long xi = py_round(__fdiv_rn(x, 2.0f)), yi = py_round(__fdiv_rn(y, 2.0f));
  float shalf = __fdiv_rn(sigma, 2.0f);
  long long si = py_round(shalf > min_scale_reduced ? shalf : min_scale_reduced);
  long long minx = xi - si > 0 ? xi - si : 0;
  long long miny = yi - si > 0 ? yi - si : 0;
  long long t = xi + si + 1 < ow ? xi + si + 1 : ow;
  long long maxx = minx + 1 > t ? minx + 1 : t;
  t = yi + si + 1 < oh ? yi + si + 1 : oh;
  long long maxy = miny + 1 > t ? miny + 1 : t;
  if (maxx > ow) maxx = ow;      // numpy slicing clamps to the array
  if (maxy > oh) maxy = oh;
  box[0] = minx; box[1] = maxx; box[2] = miny; box[3] = maxy;
}

// Occupancy.get cell (occupancy.py:41-47 + functional.pyx:281-286)
__device__ __forceinline__ void occupancy_cell(float x, float y, long long ow, long long oh,
                                               long long *cx, long long *cy) {
  *cx = (long long)clipf(__fdiv_rn(x, 2.0f), 0.0f, (float)(ow - 1));
  *cy = (long long)clipf(__fdiv_rn(y, 2.0f), 0.0f, (float)(oh - 1));
}

}  // namespace pp
