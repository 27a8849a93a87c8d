// inverse.cu -- Preprocess.annotations_inverse (transforms/preprocess.py:35-82) for a batch,
// applied to the decoder's result buffers on the device before they are copied to the host.
//
// Arithmetic follows what numpy does for the reference's statements:
//   rotation      float32 array against Python floats      -> float32 operations
//   += offset[i]  float32 array in place with a float64    -> float64 add, rounded to float32
//   / scale[i]    likewise                                 -> float64 divide, rounded to float32
//   -x + (w - 1)  float32 array plus an int64 scalar       -> float64 add, rounded to float32
// The rotation / offset / scale are applied to every joint row, filled or not, as the
// reference does.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

__global__ void annotations_inverse_kernel(float *ann_data, float *ann_scales, const int *n_ann,
                                           float *pre_data, const int *n_pre, const pp_image_meta *meta,
                                           const int *hflip_target, int F, int ann_cap) {
  const int b = blockIdx.y, a = blockIdx.x, j = threadIdx.x;
  const pp_image_meta m = meta[b];
  const bool live = a < n_ann[b] && j < F;
  float *d = ann_data + ((size_t)b * ann_cap + a) * F * 3;
  float x = 0.f, y = 0.f, v = 0.f;
  if (live) {
    x = d[3 * j]; y = d[3 * j + 1]; v = d[3 * j + 2];
    if (m.rotate) {
      const float xo = __fsub_rn(x, m.rot_cx), yo = __fsub_rn(y, m.rot_cy);
      x = __fadd_rn(__fadd_rn(m.rot_cx, __fmul_rn(m.rot_cos, xo)), __fmul_rn(m.rot_sin, yo));
      y = __fadd_rn(__fsub_rn(m.rot_cy, __fmul_rn(m.rot_sin, xo)), __fmul_rn(m.rot_cos, yo));
    }
    x = (float)((double)x + m.offset_x);
    y = (float)((double)y + m.offset_y);
    x = (float)((double)x / m.scale_x);
    y = (float)((double)y / m.scale_y);
    float *js = ann_scales + ((size_t)b * ann_cap + a) * F + j;
    *js = (float)((double)*js / m.scale_x);
    if (m.hflip) x = (float)((double)(-x) + m.flip_w1);
  }
  __syncthreads();            // all rows are read before the (permuted) rows are written
  if (live) {
    const int t = (m.hflip && hflip_target) ? hflip_target[j] : j;
    d[3 * t] = x; d[3 * t + 1] = y; d[3 * t + 2] = v;
  }
  if (pre_data && a < n_pre[b] && j < F) {
    float *q = pre_data + (((size_t)b * ann_cap + a) * F + j) * 3;
    q[0] = (float)((double)(float)((double)q[0] + m.offset_x) / m.scale_x);
    q[1] = (float)((double)(float)((double)q[1] + m.offset_y) / m.scale_y);
  }
}

// Preprocess.anndet_inverse (transforms/preprocess.py:84-94, utils.rotate_box transforms/utils.py:5-28)
// on detection boxes (x, y, w, h): rotation of the four corners in float32 and their bounding box,
// then offset / scale in float64 rounded to float32 per statement.  The reference applies no hflip to
// detections; neither does this.
__global__ void detections_inverse_kernel(float *box, const int *n_det, const pp_image_meta *meta, int ann_cap) {
  const int b = blockIdx.y, a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= ann_cap || a >= n_det[b]) return;
  const pp_image_meta m = meta[b];
  float *q = box + ((size_t)b * ann_cap + a) * 4;
  float x = q[0], y = q[1], w = q[2], h = q[3];
  if (m.rotate) {
    // rotate_box is called with width = rw - 1, height = rh - 1: centre = (rw - 1) / 2 = rot_cx
    const float cx[4] = {x, __fadd_rn(x, w), x, __fadd_rn(x, w)};
    const float cy[4] = {y, y, __fadd_rn(y, h), __fadd_rn(y, h)};
    float lo_x = 0.f, lo_y = 0.f, hi_x = 0.f, hi_y = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float xo = __fsub_rn(cx[k], m.rot_cx), yo = __fsub_rn(cy[k], m.rot_cy);
      const float rx = __fadd_rn(__fadd_rn(m.rot_cx, __fmul_rn(m.rot_cos, xo)), __fmul_rn(m.rot_sin, yo));
      const float ry = __fadd_rn(__fsub_rn(m.rot_cy, __fmul_rn(m.rot_sin, xo)), __fmul_rn(m.rot_cos, yo));
      lo_x = (k == 0 || rx < lo_x) ? rx : lo_x; hi_x = (k == 0 || rx > hi_x) ? rx : hi_x;
      lo_y = (k == 0 || ry < lo_y) ? ry : lo_y; hi_y = (k == 0 || ry > hi_y) ? ry : hi_y;
    }
    x = lo_x; y = lo_y; w = __fsub_rn(hi_x, lo_x); h = __fsub_rn(hi_y, lo_y);
  }
  x = (float)((double)x + m.offset_x);
  y = (float)((double)y + m.offset_y);
  q[0] = (float)((double)x / m.scale_x);
  q[1] = (float)((double)y / m.scale_y);
  q[2] = (float)((double)w / m.scale_x);
  q[3] = (float)((double)h / m.scale_y);
}

int detections_inverse_launch(float *box, const int *n_det, const pp_inverse &inv, int B, int ann_cap,
                              cudaStream_t st) {
  if (!inv.meta || !box || !n_det || B <= 0 || ann_cap <= 0) return PP_ERR_ARG;
  detections_inverse_kernel<<<dim3((ann_cap + 127) / 128, B), 128, 0, st>>>(box, n_det, inv.meta, ann_cap);
  PP_LAUNCH_CHECK("detections_inverse_kernel");
  return PP_OK;
}

int annotations_inverse_launch(float *ann_data, float *ann_scales, const int *n_ann, float *pre_data,
                               const int *n_pre, const pp_inverse &inv, int B, int F, int ann_cap,
                               cudaStream_t st) {
  if (!inv.meta || B <= 0 || F <= 0 || F > PP_MAX_F || ann_cap <= 0) return PP_ERR_ARG;
  const int threads = (F + 31) / 32 * 32;
  annotations_inverse_kernel<<<dim3(ann_cap, B), threads, 0, st>>>(
      ann_data, ann_scales, n_ann, pre_data, pre_data ? n_pre : nullptr, inv.meta, inv.hflip_target, F,
      ann_cap);
  PP_LAUNCH_CHECK("annotations_inverse_kernel");
  return PP_OK;
}

}  // namespace pp
