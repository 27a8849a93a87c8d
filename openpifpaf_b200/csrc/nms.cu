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

namespace pp {

// ---------------------------------------------------------------------------------------------
// pp_keypoint_nms_noalloc: pp_keypoint_nms for callers that cannot allocate device memory (CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  The scratch pp_keypoint_nms takes from
// the caller's workspace lives in the file-scope device array below, so the call allocates
// nothing: no cudaMalloc / cudaFree, no allocating library routine.
//   * Every GPU that loads the library holds PP_NMS_NOALLOC_BYTES of this array until exit.
//   * Requests whose scratch (nms_ws_bytes(shape, caps)) exceeds it are rejected with
//     PP_ERR_STATIC_LIMIT before anything is enqueued.
//   * All calls (of every host thread and stream, per GPU) share the array: calls that overlap on
//     the GPU race.  Issue them on one stream, or order them with events.
//   * The array keeps the last call's contents.  Nothing in it has to start at zero: every word
//     the kernel reads (staged input rows and counts, working rows, scores, ranks, occupancy
//     boxes) is written earlier in the same call, so nothing is cleared.
static __device__ __align__(256) unsigned char g_nms_noalloc_scratch[PP_NMS_NOALLOC_BYTES];

// In host code the array's name is a host-side stand-in, not a device address: the address the
// kernel gets comes from cudaGetSymbolAddress, once per device, with that device current.
static int nms_noalloc_scratch(void **out) {
  constexpr int kMaxDev = 64;
  static void *cached[kMaxDev] = {nullptr};
  int dev = 0;
  PP_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) return PP_ERR_ARG;
  void *ptr = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
  if (ptr == nullptr) {
    PP_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_nms_noalloc_scratch));
    if (ptr == nullptr) return PP_ERR_CUDA;
    __atomic_store_n(&cached[dev], ptr, __ATOMIC_RELEASE);     // every thread gets the same value
  }
  *out = ptr;
  return PP_OK;
}

}  // namespace pp

extern "C" {

size_t pp_keypoint_nms_noalloc_max_bytes(void) { return (size_t)PP_NMS_NOALLOC_BYTES; }

size_t pp_keypoint_nms_noalloc_bytes(const pp_shape *s, const pp_caps *c) {
  if (!s || !c || s->B <= 0 || s->F <= 0 || s->F > PP_MAX_F || c->ann_cap <= 0) return 0;
  return pp::nms_ws_bytes(*s, *c);
}

int pp_keypoint_nms_noalloc(float *d_ann_data, float *d_ann_scales, int32_t *d_n_ann, double *d_score,
                            int32_t *d_source, const pp_shape *s, const pp_caps *c, const pp_params *p,
                            void *stream) {
  using namespace pp;
  if (!s || !c || !p || !d_ann_data || !d_ann_scales || !d_n_ann) return PP_ERR_ARG;
  if (s->B <= 0 || s->F <= 0 || s->F > PP_MAX_F || c->ann_cap <= 0) return PP_ERR_ARG;
  if (nms_ws_bytes(*s, *c) > (size_t)PP_NMS_NOALLOC_BYTES) return PP_ERR_STATIC_LIMIT;
  void *ws = nullptr;
  int rc = nms_noalloc_scratch(&ws);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // in place, like pp_keypoint_nms: the input rows are staged inside the scratch first
  const size_t bc = (size_t)s->B * c->ann_cap;
  const size_t data_bytes = align_up(bc * s->F * 3 * sizeof(float), 256);
  char *stage = (char *)ws + nms_ws_staging_offset(*s, *c);
  float *in_data = (float *)stage;
  float *in_scales = (float *)(stage + data_bytes);
  int *n_in = (int *)(stage + 2 * data_bytes);
  PP_CUDA_CHECK(cudaMemcpyAsync(in_data, d_ann_data, bc * s->F * 3 * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  PP_CUDA_CHECK(cudaMemcpyAsync(in_scales, d_ann_scales, bc * s->F * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  PP_CUDA_CHECK(cudaMemcpyAsync(n_in, d_n_ann, (size_t)s->B * sizeof(int), cudaMemcpyDeviceToDevice, st));
  return nms_launch(in_data, in_scales, n_in, d_ann_data, d_ann_scales, d_score, d_source, d_n_ann, ws,
                    *s, *c, *p, st);
}

}  // extern "C"
