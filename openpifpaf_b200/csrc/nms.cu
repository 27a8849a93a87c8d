// nms.cu -- keypoint NMS and final ordering, on sm_100a.
//
// Replaces decoder/nms.py:17-57 (nms.Keypoints.annotations) and
// Annotation.score (annotation.py:24-28,60-71) of the reference.
//
// One CTA per image, one WARP per keypoint type.  The reference walks the
// annotations in score order and, per joint, asks a fresh uint8 occupancy map
// whether the joint's cell is taken.  Joints of different types never
// interact, so each warp replays its own type sequentially over the sorted
// annotations, keeping the boxes it has set in a list; "occupied" is the
// number of listed boxes covering the cell, modulo 256 (the uint8 wrap).
// Lanes split the box list.  Scores are float64 with numpy's pairwise
// summation order so that the two stable sorts see the reference's keys.
#include "common.cuh"
#include "stages.cuh"

#include "nms.cuh"

namespace pp {

__global__ void nms_kernel(NmsArgs a) { nms_image(a, blockIdx.x); }

size_t nms_ws_bytes(const pp_shape &s, const pp_caps &c) {
  size_t bc = (size_t)s.B * c.ann_cap;
  size_t n = 0;
  n += align_up(bc * s.F * 3 * sizeof(float), 256);        // work
  n += align_up(bc * sizeof(double), 256);                 // score
  n += 2 * align_up(bc * sizeof(int), 256);                // order, order2
  n += align_up(bc * s.F * 4 * sizeof(long long), 256);    // boxes
  n += 2 * align_up(bc * s.F * 3 * sizeof(float), 256);    // in-place staging (data, scales)
  n += align_up((size_t)s.B * sizeof(int), 256);           // in-place staging (counts)
  return n;
}

NmsArgs nms_make_args(const float *in_data, const float *in_scales, const int *n_in, float *out_data,
                      float *out_scales, double *out_score, int *out_source, int *n_out, void *d_ws,
                      const pp_shape &s, const pp_caps &c, const pp_params &p) {
  NmsArgs a;
  size_t bc = (size_t)s.B * c.ann_cap;
  char *q = (char *)d_ws;
  a.work = (float *)q;      q += align_up(bc * s.F * 3 * sizeof(float), 256);
  a.score = (double *)q;    q += align_up(bc * sizeof(double), 256);
  a.order = (int *)q;       q += align_up(bc * sizeof(int), 256);
  a.order2 = (int *)q;      q += align_up(bc * sizeof(int), 256);
  a.boxes = (long long *)q; q += align_up(bc * s.F * 4 * sizeof(long long), 256);
  a.in_data = in_data; a.in_scales = in_scales; a.n_in = n_in;
  a.out_data = out_data; a.out_scales = out_scales; a.out_score = out_score;
  a.out_source = out_source; a.n_out = n_out;
  a.ann_cap = c.ann_cap; a.F = s.F; a.do_nms = p.nms;
  a.kp_th = p.nms_keypoint_threshold; a.suppression = p.nms_suppression;
  a.inst_th = p.nms_instance_threshold;
  return a;
}

int nms_launch(const float *in_data, const float *in_scales, const int *n_in, float *out_data,
               float *out_scales, double *out_score, int *out_source, int *n_out, void *d_ws,
               const pp_shape &s, const pp_caps &c, const pp_params &p, cudaStream_t st) {
  const NmsArgs a = nms_make_args(in_data, in_scales, n_in, out_data, out_scales, out_score, out_source,
                                  n_out, d_ws, s, c, p);
  int threads = 32 * s.F;
  if (threads > 1024) threads = 1024;
  if (threads < 128) threads = 128;
  prof_mark(p.profile, kStNms, st);
  nms_kernel<<<s.B, threads, 0, st>>>(a);
  PP_LAUNCH_CHECK("nms_kernel");
  return PP_OK;
}

// offset of the in-place staging area inside the nms workspace
size_t nms_ws_staging_offset(const pp_shape &s, const pp_caps &c) {
  size_t bc = (size_t)s.B * c.ann_cap;
  return align_up(bc * s.F * 3 * sizeof(float), 256) + align_up(bc * sizeof(double), 256) +
         2 * align_up(bc * sizeof(int), 256) + align_up(bc * s.F * 4 * sizeof(long long), 256);
}

}  // namespace pp
