// det.cu -- CifDet on sm_100a: the seed loop with occupancy of
// generator/cifdet.py:34-44 and nms.Detection (decoder/nms.py:60-102).
//
// CifDetHr and CifDetSeeds reuse the pose kernels (cifhr.cu / seeds.cu, det = 1): the same splat
// with sigma = max(1, 0.1 min(w, h) stride), the same seed collection + tuple sort with 6-float
// records (v, f, x, y, w, h).
//
//   det_assemble_kernel  one CTA per image.
//     seed loop   Occupancy(shape, 2, min_scale=2.0): as in assemble.cu the maps are never
//                 rasterised -- every seed carries a uint8 counter of the set boxes of its
//                 category that cover its cell; "next free seed" is a CTA-wide ballot search and a
//                 new detection bumps the counters of the seeds behind it.
//     soft NMS    annotations arrive sorted by score (seed order), so sorted(key=-score) is the
//                 identity and the instance-threshold filter keeps a prefix.  Detection i looks at
//                 the (still alive) detections before it: the IoUs are evaluated CTA-wide, the
//                 max reduced, the score scaled; then a stable rank sort of the survivors.
//                 Arithmetic is float32 in numpy's order (nms.py:67-77).
#include "common.cuh"
#include "stages.cuh"
#include "occupancy.cuh"

namespace pp {

constexpr int kDetThreads = 256;
constexpr int kDetWarps = kDetThreads / 32;

// nms.Detection.bbox_iou for one pair (np.maximum / np.minimum propagate NaN)
__device__ __forceinline__ float np_max(float a, float b) { return (a != a) ? a : ((b != b) ? b : fmaxf(a, b)); }
__device__ __forceinline__ float np_min(float a, float b) { return (a != a) ? a : ((b != b) ? b : fminf(a, b)); }

__device__ __forceinline__ float det_iou(const float4 a, const float4 b) {
  const float x1 = np_max(a.x, b.x), y1 = np_max(a.y, b.y);
  const float x2 = np_min(__fadd_rn(a.x, a.z), __fadd_rn(b.x, b.z));
  const float y2 = np_min(__fadd_rn(a.y, a.w), __fadd_rn(b.y, b.w));
  const float inter = __fmul_rn(np_max(0.0f, __fsub_rn(x2, x1)), np_max(0.0f, __fsub_rn(y2, y1)));
  const float area_a = __fmul_rn(a.z, a.w), area_b = __fmul_rn(b.z, b.w);
  const float den = __fadd_rn(__fsub_rn(__fadd_rn(area_a, area_b), inter), 1e-5f);
  return __fdiv_rn(inter, den);
}

struct DetShared {
  int warp_min[kDetWarps];
  float warp_max[kDetWarps];
  int warp_nan[kDetWarps];
  long long box[4];
  int f;
};

// nms.Detection.annotations on n detections sorted by descending score (nms.py:79-102), CTA-wide.
__device__ void detection_nms(DetShared &sh, const int *pre_field, const float *pre_score,
                              const float4 *pre_box, int n, int *det_field, float *det_score,
                              float4 *det_box, int *n_det, float *score, const pp_det_params &p,
                              const int *src_map, int *det_source) {
  // src_map: original index of input row i (NULL = i); det_source: optional, per output row
  if (!p.nms) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      det_field[i] = pre_field[i]; det_score[i] = pre_score[i]; det_box[i] = pre_box[i];
      if (det_source) det_source[i] = src_map ? src_map[i] : i;
    }
    if (threadIdx.x == 0) *n_det = n;
    return;
  }
  // scores arrive non-increasing (seed order), so the threshold filter keeps a prefix (:82-85)
  int m = 0;
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const bool keep = i < n && pre_score[i] >= p.nms_instance_threshold;
    if (i < n) score[i] = pre_score[i];
    m += __syncthreads_count(keep);
  }
  __syncthreads();
  for (int i = 1; i < m; i++) {                                  // :88-96
    const float4 bi = pre_box[i];
    float mx = -1.0f;
    int any_nan = 0;
    for (int j = threadIdx.x; j < i; j += blockDim.x) {
      if (!(score[j] >= p.nms_instance_threshold)) continue;
      const float iou = det_iou(bi, pre_box[j]);
      if (iou != iou) any_nan = 1;
      if (iou > mx) mx = iou;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      any_nan |= __shfl_xor_sync(0xffffffffu, any_nan, o);
    }
    if (lane_id() == 0) { sh.warp_max[warp_id()] = mx; sh.warp_nan[warp_id()] = any_nan; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < kDetWarps; w++) { mx = fmaxf(mx, sh.warp_max[w]); any_nan |= sh.warp_nan[w]; }
      if (!any_nan) {           // np.max is NaN when any IoU is: both comparisons are false
        if (mx > p.nms_iou_threshold) score[i] = __fmul_rn(score[i], p.nms_suppression);
        else if (mx > p.nms_iou_threshold_soft) score[i] = __fmul_rn(score[i], p.nms_suppression_soft);
      }
    }
    __syncthreads();
  }
  // survivors, stable descending sort by score (:98-99): rank = number of survivors before it
  int n_out = 0;
  for (int base = 0; base < m; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const bool keep = i < m && score[i] >= p.nms_instance_threshold;
    if (keep) {
      const float si = score[i];
      int rank = 0;
      for (int j = 0; j < m; j++) {
        const float sj = score[j];
        if (sj >= p.nms_instance_threshold && (sj > si || (sj == si && j < i))) rank++;
      }
      det_field[rank] = pre_field[i]; det_score[rank] = si; det_box[rank] = pre_box[i];
      if (det_source) det_source[rank] = src_map ? src_map[i] : i;
    }
    n_out += __syncthreads_count(keep);
  }
  if (threadIdx.x == 0) *n_det = n_out;
}

__global__ void __launch_bounds__(kDetThreads)
det_assemble_kernel(DetArgs a, pp_det_params p, float *work_score) {
  __shared__ DetShared sh;
  const int b = blockIdx.x;
  const float *seeds = a.seeds + (size_t)b * a.seed_cap * 6;
  const int n_seeds = min(a.n_seeds[b], a.seed_cap);
  uint8_t *occ = a.occ_cnt + (size_t)b * a.seed_cap;
  int2 *rec = a.occ_rec + (size_t)b * a.seed_cap;
  int *pre_field = a.pre_field + (size_t)b * a.ann_cap;
  float *pre_score = a.pre_score + (size_t)b * a.ann_cap;
  float4 *pre_box = (float4 *)a.pre_box + (size_t)b * a.ann_cap;
  // Occupancy(cifhr.accumulated.shape, 2, min_scale=2.0): cifdet.py:36, occupancy.py:21-25
  const long long oh = (long long)((double)a.hh / 2.0), ow = (long long)((double)a.hw / 2.0);

  for (int i = threadIdx.x; i < n_seeds; i += blockDim.x) {
    const float *sd = seeds + (size_t)i * 6;
    long long cx, cy;
    occupancy_cell(__ldg(sd + 2), __ldg(sd + 3), ow, oh, &cx, &cy);
    rec[i] = make_int2((int)cx | ((int)cy << 16), (int)__ldg(sd + 1));
    occ[i] = 0;
  }
  __syncthreads();

  int n = 0, cursor = 0;
  for (;;) {
    // ---- next seed whose cell is free (cifdet.py:39-41) ----
    int found = -1;
    for (int base = cursor; base < n_seeds; base += blockDim.x) {
      const int i = base + threadIdx.x;
      const bool hit = i < n_seeds && occ[i] == 0;
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (lane_id() == 0) sh.warp_min[warp_id()] = m ? (base + warp_id() * 32 + __ffs(m) - 1) : -1;
      __syncthreads();
      for (int w = 0; w < kDetWarps; w++)
        if (sh.warp_min[w] >= 0) { found = sh.warp_min[w]; break; }
      __syncthreads();
      if (found >= 0) break;
    }
    if (found < 0) break;
    cursor = found + 1;
    if (n >= a.ann_cap) {
      if (threadIdx.x == 0) atomicOr(&a.status[b], PP_STATUS_ANN_OVERFLOW);
      break;
    }
    if (threadIdx.x == 0) {                                      // cifdet.py:42-44
      const float *sd = seeds + (size_t)found * 6;
      const float v = sd[0], x = sd[2], y = sd[3], w = sd[4], h = sd[5];
      pre_field[n] = (int)sd[1];
      pre_score[n] = v;
      pre_box[n] = make_float4(__fsub_rn(x, __fdiv_rn(w, 2.0f)), __fsub_rn(y, __fdiv_rn(h, 2.0f)), w, h);
      const float m = (h < w) ? h : w;                           // Python min(w, h)
      occupancy_box(x, y, __fmul_rn(0.1f, m), ow, oh, sh.box, 1.0f);
      sh.f = (int)sd[1];
    }
    n++;
    __syncthreads();
    const int f = sh.f;
    const long long bx0 = sh.box[0], bx1 = sh.box[1], by0 = sh.box[2], by1 = sh.box[3];
    for (int i = cursor + threadIdx.x; i < n_seeds; i += blockDim.x) {
      const int2 r = rec[i];
      const long long cx = r.x & 0xffff, cy = r.x >> 16;
      if (r.y == f && cx >= bx0 && cx < bx1 && cy >= by0 && cy < by1)
        occ[i] = (uint8_t)(occ[i] + 1);                          // uint8 += 1 wraps (utils.py:66)
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) a.n_pre[b] = n;

  __syncthreads();          // pre_* written by thread 0 are read by everyone below
  detection_nms(sh, pre_field, pre_score, pre_box, n, a.det_field + (size_t)b * a.ann_cap,
                a.det_score + (size_t)b * a.ann_cap, (float4 *)a.det_box + (size_t)b * a.ann_cap,
                a.n_det + b, work_score + (size_t)b * a.ann_cap, p, nullptr,
                a.det_source ? a.det_source + (size_t)b * a.ann_cap : nullptr);
}

// nms.Detection.annotations(anns) on its own (nms.py:79-102): inputs in any order.
__global__ void __launch_bounds__(kDetThreads)
det_nms_kernel(const int *in_field, const float *in_score, const float *in_box, const int *n_in,
               int *srt_field, float *srt_score, float *srt_box, int *srt_src, int *det_field,
               float *det_score, float *det_box, int *det_source, int *n_det, float *work_score, int cap,
               pp_det_params p) {
  __shared__ DetShared sh;
  const int b = blockIdx.x;
  const size_t o = (size_t)b * cap;
  const int n = min(n_in[b], cap);
  // sorted(anns, key=-score), stable (:85): rank sort into the scratch copy
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float si = in_score[o + i];
    int rank = 0;
    for (int j = 0; j < n; j++) {
      const float sj = in_score[o + j];
      if (sj > si || (sj == si && j < i)) rank++;
    }
    srt_field[o + rank] = in_field[o + i];
    srt_score[o + rank] = si;
    srt_src[o + rank] = i;
    ((float4 *)srt_box)[o + rank] = ((const float4 *)in_box)[o + i];
  }
  __syncthreads();
  detection_nms(sh, srt_field + o, srt_score + o, (const float4 *)srt_box + o, n, det_field + o,
                det_score + o, (float4 *)det_box + o, n_det + b, work_score + o, p, srt_src + o,
                det_source ? det_source + o : nullptr);
}

int det_nms_launch(const int *in_field, const float *in_score, const float *in_box, const int *n_in,
                   int *det_field, float *det_score, float *det_box, int *det_source, int *n_det,
                   void *d_ws, int B, int cap, const pp_det_params &p, cudaStream_t st) {
  if (((uintptr_t)in_box & 15u) || ((uintptr_t)det_box & 15u) || ((uintptr_t)d_ws & 15u)) return PP_ERR_ARG;
  char *q = (char *)d_ws;
  const size_t n = (size_t)B * cap;
  float *srt_box = (float *)q; q += align_up(n * 4 * sizeof(float), 256);
  float *srt_score = (float *)q; q += align_up(n * sizeof(float), 256);
  float *work = (float *)q; q += align_up(n * sizeof(float), 256);
  int *srt_field = (int *)q; q += align_up(n * sizeof(int), 256);
  int *srt_src = (int *)q;
  det_nms_kernel<<<B, kDetThreads, 0, st>>>(in_field, in_score, in_box, n_in, srt_field, srt_score, srt_box,
                                           srt_src, det_field, det_score, det_box, det_source, n_det, work,
                                           cap, p);
  PP_LAUNCH_CHECK("det_nms_kernel");
  return PP_OK;
}

size_t det_nms_ws_bytes(int B, int cap) {
  const size_t n = (size_t)B * cap;
  return align_up(n * 4 * sizeof(float), 256) + 2 * align_up(n * sizeof(float), 256) +
         2 * align_up(n * sizeof(int), 256);
}

int det_assemble_launch(const DetArgs &a, int B, const pp_det_params &p, float *work_score,
                        cudaStream_t st) {
  if (((uintptr_t)a.pre_box & 15u) || ((uintptr_t)a.det_box & 15u)) return PP_ERR_ARG;
  if (a.hw / 2 >= 65536 || a.hh / 2 >= 32768) return PP_ERR_ARG;     // cell packing in occ_rec
  det_assemble_kernel<<<B, kDetThreads, 0, st>>>(a, p, work_score);
  PP_LAUNCH_CHECK("det_assemble_kernel");
  return PP_OK;
}

}  // namespace pp
