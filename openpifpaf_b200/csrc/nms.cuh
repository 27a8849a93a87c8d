// nms.cuh -- device body of the keypoint NMS (nms.cu), shared with the assembly kernel, which runs it
// as its tail for the image of its CTA (assemble.cu).
//
// Replaces decoder/nms.py:17-57 (nms.Keypoints.annotations) and Annotation.score
// (annotation.py:24-28,60-71).  One CTA per image.  The scratch (working copy of the rows, scores,
// ranks, the boxes set so far per keypoint type) lives in shared memory when the caller hands in
// enough of it for the image's annotation count (the assembly kernel passes its whole dynamic
// region, free by then), else in the global scratch of NmsArgs.
#pragma once
#include "common.cuh"
#include "stages.cuh"

namespace pp {

__device__ __forceinline__ long long py_round_d(float v) {
  double r = rint((double)v);
  if (r > 4e18) r = 4e18;
  if (r < -4e18) r = -4e18;
  return (long long)r;
}

// Annotation.score(): sum(score_weights * sort(v)[::-1]) in float64, numpy pairwise order.  One thread.
__device__ inline double annotation_score(const float *data /* (F,3) */, int F) {
  float v[PP_MAX_F];
  for (int j = 0; j < F; j++) v[j] = data[3 * j + 2];
  for (int i = 1; i < F; i++) {          // ascending insertion sort
    float t = v[i];
    int j = i - 1;
    while (j >= 0 && v[j] > t) { v[j + 1] = v[j]; j--; }
    v[j + 1] = t;
  }
  const double wsum = (double)F + 2.0 * (F < 3 ? F : 3);
  double t[PP_MAX_F];
  for (int j = 0; j < F; j++) t[j] = __dmul_rn((j < 3 ? 3.0 : 1.0) / wsum, (double)v[F - 1 - j]);
  if (F < 8) {
    double r = 0.0;
    for (int j = 0; j < F; j++) r = __dadd_rn(r, t[j]);
    return r;
  }
  double r[8];
  int i;
  for (i = 0; i < 8; i++) r[i] = t[i];
  for (i = 8; i < F - (F % 8); i += 8)
    for (int j = 0; j < 8; j++) r[j] = __dadd_rn(r[j], t[i + j]);
  double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                         __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
  for (; i < F; i++) res = __dadd_rn(res, t[i]);
  return res;
}

// The same value by a whole warp (lane j holds joint j): ranks by F shuffles instead of a sort in
// local memory, the float64 sum in numpy's order (eight running sums, pairwise, then the tail).
// Every lane returns the score.  NaN confidences take the one-thread version (same NaN placement).
__device__ __forceinline__ double warp_annotation_score(const float *data /* (F,3) */, int F) {
  const unsigned full = 0xffffffffu;
  const int lane = lane_id();
  const float v = lane < F ? data[3 * lane + 2] : 0.0f;
  if (__any_sync(full, v != v)) {
    double r = 0.0;
    if (lane == 0) r = annotation_score(data, F);
    return __shfl_sync(full, r, 0);
  }
  // position of joint `lane` in sort(v)[::-1]; equal values keep the order the reversed ascending
  // insertion sort gives them (later joint first)
  int rank = 0;
  for (int k = 0; k < F; k++) {
    const float vk = __shfl_sync(full, v, k);
    rank += (vk > v || (vk == v && k > lane)) ? 1 : 0;
  }
  float sorted = 0.0f;
  for (int k = 0; k < F; k++) {
    const float vk = __shfl_sync(full, v, k);
    const int rk = __shfl_sync(full, rank, k);
    if (rk == lane) sorted = vk;
  }
  const double wsum = (double)F + 2.0 * (F < 3 ? F : 3);
  const double t = lane < F ? __dmul_rn((lane < 3 ? 3.0 : 1.0) / wsum, (double)sorted) : 0.0;
  double res;
  if (F < 8) {
    res = 0.0;
    for (int j = 0; j < F; j++) res = __dadd_rn(res, __shfl_sync(full, t, j));
    return res;
  }
  double r = t;                                       // lanes 0..7: the eight running sums
  int i;
  for (i = 8; i < F - (F % 8); i += 8) {
    const double tt = __shfl_sync(full, t, (i + lane) & 31);
    if (lane < 8) r = __dadd_rn(r, tt);
  }
  r = __dadd_rn(r, __shfl_xor_sync(full, r, 1));      // (r0+r1) ...
  r = __dadd_rn(r, __shfl_xor_sync(full, r, 2));      // (r0+r1)+(r2+r3) ...
  r = __dadd_rn(r, __shfl_xor_sync(full, r, 4));
  res = __shfl_sync(full, r, 0);
  for (; i < F; i++) res = __dadd_rn(res, __shfl_sync(full, t, i));
  return res;
}

__host__ __device__ inline size_t nms_scratch_bytes(int n, int F) {
  return (size_t)n * 8 /* score */ + (size_t)F * n * 32 /* boxes */ + (size_t)n * F * 16 /* work, scales */ +
         (size_t)n * 8 /* order, order2 */ + 16;
}

// nms.Keypoints.annotations for image b, CTA-collective (any block size that is a multiple of 32).
// smem / smem_bytes: optional shared-memory scratch (8-byte aligned).
__device__ __forceinline__ void nms_image(const NmsArgs &a, const int b, void *smem = nullptr,
                                          size_t smem_bytes = 0) {
  __shared__ float s_maxx, s_maxy;
  __shared__ int s_m, s_k;
  const int F = a.F, cap = a.ann_cap;
  const int n = min(a.n_in[b], cap);
  const float *in_data = a.in_data + (size_t)b * cap * F * 3;
  const float *in_scales = a.in_scales + (size_t)b * cap * F;
  float *out_data = a.out_data + (size_t)b * cap * F * 3;
  float *out_scales = a.out_scales + (size_t)b * cap * F;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int warp = tid >> 5, lane = tid & 31, nwarps = nt >> 5;
  const unsigned full = 0xffffffffu;

  float *work; double *score; int *order, *order2; long long *boxes;
  const float *scales = in_scales;
  int box_stride;                 // boxes of keypoint type f start at boxes + 4 * f * box_stride
  const bool in_smem = smem != nullptr && nms_scratch_bytes(n, F) <= smem_bytes;
  if (in_smem) {
    char *q = (char *)smem;
    score = (double *)q;       q += (size_t)n * 8;
    boxes = (long long *)q;    q += (size_t)F * n * 32;
    work = (float *)q;         q += (size_t)n * F * 12;
    float *sc = (float *)q;    q += (size_t)n * F * 4;
    order = (int *)q;          q += (size_t)n * 4;
    order2 = (int *)q;
    box_stride = n;
    for (int i = tid; i < n * F; i += nt) sc[i] = in_scales[i];
    scales = sc;
  } else {
    work = a.work + (size_t)b * cap * F * 3;
    score = a.score + (size_t)b * cap;
    order = a.order + (size_t)b * cap; order2 = a.order2 + (size_t)b * cap;
    boxes = a.boxes + (size_t)b * F * cap * 4;
    box_stride = cap;
  }

  // nms.py:20-21  rows below the keypoint threshold are zeroed (x, y and v)
  for (int i = tid; i < n * F; i += nt) {
    float x = in_data[3 * i], y = in_data[3 * i + 1], v = in_data[3 * i + 2];
    if (a.do_nms && v < a.kp_th) x = y = v = 0.0f;
    work[3 * i] = x; work[3 * i + 1] = y; work[3 * i + 2] = v;
  }
  if (tid == 0) { s_maxx = -INFINITY; s_maxy = -INFINITY; s_m = 0; s_k = 0; }
  __syncthreads();
  for (int i = warp; i < n; i += nwarps) {
    const double sc = warp_annotation_score(work + (size_t)i * F * 3, F);
    if (lane == 0) score[i] = sc;
  }
  __syncthreads();

  if (!a.do_nms) {   // CifCaf(nms=None): creation order, untouched
    for (int i = tid; i < n * F * 3; i += nt) out_data[i] = work[i];
    for (int i = tid; i < n * F; i += nt) out_scales[i] = scales[i];
    for (int i = tid; i < n; i += nt) {
      if (a.out_score) a.out_score[(size_t)b * cap + i] = score[i];
      if (a.out_source) a.out_source[(size_t)b * cap + i] = i;
    }
    if (tid == 0) a.n_out[b] = n;
    return;
  }

  // :22 filter by instance threshold; :33 stable sort by descending score
  for (int i = tid; i < n; i += nt) {
    int rank = -1;
    if (score[i] >= a.inst_th) {
      rank = 0;
      for (int j = 0; j < n; j++) {
        if (!(score[j] >= a.inst_th)) continue;
        if (score[j] > score[i] || (score[j] == score[i] && j < i)) rank++;
      }
      PP_DASSERT(rank >= 0 && rank < n);
      order[rank] = i;
      atomicAdd(&s_m, 1);
    }
  }
  __syncthreads();
  const int m = s_m;
  if (m == 0) {                                                    // :24-25
    if (tid == 0) a.n_out[b] = 0;
    return;
  }
  // :27-31 occupancy extent from the surviving annotations (zero rows included)
  {
    float mx = -INFINITY, my = -INFINITY;
    for (int i = tid; i < m * F; i += nt) {
      int an = order[i / F], j = i % F;
      mx = fmaxf(mx, work[((size_t)an * F + j) * 3]);
      my = fmaxf(my, work[((size_t)an * F + j) * 3 + 1]);
    }
    for (int off = 16; off > 0; off >>= 1) {
      mx = fmaxf(mx, __shfl_xor_sync(full, mx, off));
      my = fmaxf(my, __shfl_xor_sync(full, my, off));
    }
    if (lane == 0) {
      atomicMax((int *)&s_maxx, __float_as_int(fmaxf(mx, 0.0f)));   // maxima are >= 0 here or unused
      atomicMax((int *)&s_maxy, __float_as_int(fmaxf(my, 0.0f)));
    }
  }
  __syncthreads();
  // int(max + 1) then Occupancy's int(shape / reduction)
  const long long sh_h = (long long)__fadd_rn(s_maxy, 1.0f), sh_w = (long long)__fadd_rn(s_maxx, 1.0f);
  const long long oh = (long long)((double)sh_h / 2.0), ow = (long long)((double)sh_w / 2.0);

  // :34-45 per keypoint type, sequential over annotations in score order
  for (int f = warp; f < F; f += nwarps) {
    long long *bxs = boxes + (size_t)f * box_stride * 4;
    int nb = 0;
    for (int oi = 0; oi < m; oi++) {
      const int an = order[oi];
      float *xyv = work + ((size_t)an * F + f) * 3;
      const float v = xyv[2];
      if (v == 0.0f) continue;
      const float x = xyv[0], y = xyv[1];
      const long long cx = (long long)clipf(__fdiv_rn(x, 2.0f), 0.0f, (float)(ow - 1));
      const long long cy = (long long)clipf(__fdiv_rn(y, 2.0f), 0.0f, (float)(oh - 1));
      int hits = 0;
      for (int k = lane; k < nb; k += 32) {
        const long long *bx = bxs + (size_t)k * 4;
        hits += (cx >= bx[0] && cx < bx[1] && cy >= bx[2] && cy < bx[3]);
      }
      hits = __reduce_add_sync(full, hits);
      if ((hits & 255) != 0) {
        if (lane == 0) xyv[2] = __fmul_rn(v, a.suppression);       // :42
      } else {
        if (lane == 0) {                                           // :44 Occupancy.set
          const float sigma = scales[(size_t)an * F + f];
          long long xi = py_round_d(__fdiv_rn(x, 2.0f)), yi = py_round_d(__fdiv_rn(y, 2.0f));
          float shalf = __fdiv_rn(sigma, 2.0f);
          long long si = py_round_d(shalf > 2.0f ? shalf : 2.0f);
          long long minx = xi - si > 0 ? xi - si : 0, miny = yi - si > 0 ? yi - si : 0;
          long long t = xi + si + 1 < ow ? xi + si + 1 : ow;
          long long maxx = minx + 1 > t ? minx + 1 : t;
          t = yi + si + 1 < oh ? yi + si + 1 : oh;
          long long maxy = miny + 1 > t ? miny + 1 : t;
          if (maxx > ow) maxx = ow;
          if (maxy > oh) maxy = oh;
          PP_DASSERT(nb < box_stride);
          long long *bx = bxs + (size_t)nb * 4;
          bx[0] = minx; bx[1] = maxx; bx[2] = miny; bx[3] = maxy;
        }
        nb++;
        __syncwarp();
      }
    }
  }
  __syncthreads();
  // :50-53 re-threshold, re-filter, stable re-sort (ties keep the previous order)
  for (int i = tid; i < m * F; i += nt) {
    int an = order[i / F], j = i % F;
    float *r = work + ((size_t)an * F + j) * 3;
    if (r[2] < a.kp_th) r[0] = r[1] = r[2] = 0.0f;
  }
  __syncthreads();
  for (int oi = warp; oi < m; oi += nwarps) {
    const double sc = warp_annotation_score(work + (size_t)order[oi] * F * 3, F);
    if (lane == 0) score[order[oi]] = sc;
  }
  __syncthreads();
  for (int oi = tid; oi < m; oi += nt) {
    const double si = score[order[oi]];
    int rank = -1;
    if (si >= a.inst_th) {
      rank = 0;
      for (int oj = 0; oj < m; oj++) {
        const double sj = score[order[oj]];
        if (!(sj >= a.inst_th)) continue;
        if (sj > si || (sj == si && oj < oi)) rank++;
      }
      atomicAdd(&s_k, 1);
    }
    order2[oi] = rank;
  }
  __syncthreads();
  if (tid == 0) a.n_out[b] = s_k;
  // one warp per surviving annotation writes its rows
  for (int oi = warp; oi < m; oi += nwarps) {
    const int r = order2[oi];
    if (r < 0) continue;
    const int an = order[oi];
    for (int i = lane; i < F * 3; i += 32) out_data[(size_t)r * F * 3 + i] = work[(size_t)an * F * 3 + i];
    for (int j = lane; j < F; j += 32) out_scales[(size_t)r * F + j] = scales[(size_t)an * F + j];
    if (lane == 0) {
      if (a.out_score) a.out_score[(size_t)b * cap + r] = score[an];
      if (a.out_source) a.out_source[(size_t)b * cap + r] = an;
    }
  }
}

}  // namespace pp
