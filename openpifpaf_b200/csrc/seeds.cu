// seeds.cu -- CifSeeds on sm_100a.
//
// Replaces decoder/cif_seeds.py:23-54 (CifSeeds.fill_cif / get) and the
// scalar_values gather of functional.pyx:231-244.
//
//   K3 seeds_collect_kernel  one thread per CIF cell: threshold, hr gather,
//                            v = 0.9 hr + 0.1 c, threshold, warp-aggregated
//                            append of (v, f, x, y, s) to the image's raw list.
//   K3b seeds_sort_kernel    one CTA per image: bitonic sort of 64-bit keys
//                            [orderable(v):32 | f:8 | raw index:24] in shared
//                            memory (L2-resident global scratch beyond 4 Ki seeds); keys
//                            that tie on (v, f) are ordered by (x, y, s) read
//                            from the raw records, which reproduces Python's
//                            sorted(tuples, reverse=True) exactly (cif_seeds.py:54).
//
// This snapshot of the reference has no 3x3 max-pool seed NMS
// (cif_seeds.py:23-50), so neither has this kernel.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kCollectThreads = 256;
constexpr int kSortThreads = 1024;
constexpr int kSortSmemKeys = 4096;    // 32 KiB of dynamic shared memory; larger lists sort in L2
constexpr int kRankSortMax = 256;      // rank sort up to this many seeds (n^2 work: 1300 seeds took 250 us), bitonic above

__device__ __forceinline__ uint32_t orderable(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// REC = floats per seed record: 5 (v, f, x, y, s) for poses, 6 (v, f, x, y, w, h) for CifDet
// (cif_seeds.py:67-90; fields have 7 channels [v, x, y, _, w, h, _])
template <int REC>
__global__ void __launch_bounds__(kCollectThreads)
seeds_collect_kernel(const float *__restrict__ cif, const float *__restrict__ hr,
                     float *__restrict__ raw, int *__restrict__ n_raw, pp_shape s, int cap, int hh,
                     int hw, float th, float score_scale, float min_scale_over_stride,
                     int use_min_scale, uint32_t seed_mask, int mask_on) {
  const int n = s.H * s.W;
  const int blocks_per_field = (n + kCollectThreads - 1) / kCollectThreads;
  const int bf = blockIdx.x / blocks_per_field;
  const int c = (blockIdx.x - bf * blocks_per_field) * kCollectThreads + threadIdx.x;
  const int b = bf / s.F, f = bf - b * s.F;
  if (mask_on && !((seed_mask >> (f & 31)) & 1u)) return;         // cif_seeds.py:29-30
  const float *p = cif + (size_t)bf * (REC == 6 ? 7 : 5) * n;
  const float stride = (float)s.stride;
  bool keep = false;
  float v = 0.f, x = 0.f, y = 0.f, sc = 0.f, sc2 = 0.f;
  if (c < n) {
    float conf = __ldg(p + c);
    if (conf > th) {                                              // :31
      sc = __ldg(p + 4 * n + c);
      if (REC == 6) sc2 = __ldg(p + 5 * n + c);
      if (!use_min_scale || (sc > min_scale_over_stride &&
                             (REC == 5 || sc2 > min_scale_over_stride))) {   // :32-33, :75-77
        x = __fmul_rn(__ldg(p + n + c), stride);
        y = __fmul_rn(__ldg(p + 2 * n + c), stride);
        float hv = scalar_value_at(hr + (size_t)bf * hh * hw, hh, hw, x, y, 0.0f);   // :37
        v = __fadd_rn(__fmul_rn(0.9f, hv), __fmul_rn(0.1f, conf));                  // :38
        if (score_scale != 1.0f) v = __fmul_rn(v, score_scale);   // :41-42
        keep = v > th;                                            // :43
      }
    }
  }
  // warp-aggregated slot allocation (order is irrelevant: the list is sorted next)
  unsigned m = __ballot_sync(0xffffffffu, keep);
  if (m == 0u) return;
  int lane = lane_id();
  int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(&n_raw[b], __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (keep) {
    int slot = base + __popc(m & ((1u << lane) - 1u));
    if (slot < cap) {
      float *o = raw + ((size_t)b * cap + slot) * REC;
      o[0] = v; o[1] = (float)f; o[2] = x; o[3] = y; o[4] = __fmul_rn(sc, stride);   // :44,47
      if (REC == 6) o[5] = __fmul_rn(sc2, stride);                                   // :84
    }
  }
}

// The same records from the compacted CIF sources of the CifHr stage.  Every source passed
// conf > cif_threshold and the min-scale filter; with seed_threshold >= cif_threshold the seed
// candidates are a subset, and x, y, scale and conf (= 16 * v/16, exact) are what the field scan
// would have computed.  One CTA per (image, field).
template <int REC>
__global__ void __launch_bounds__(128)
seeds_from_sources_kernel(CifSrcView src, const float *__restrict__ hr, float *__restrict__ raw,
                          int *__restrict__ n_raw, int F, int cap, int hh, int hw, float th,
                          float score_scale, uint32_t seed_mask, int mask_on) {
  const int bf = blockIdx.x;
  const int b = bf / F, f = bf - b * F;
  if (mask_on && !((seed_mask >> (f & 31)) & 1u)) return;         // cif_seeds.py:29-30
  const int n = min(src.n[bf], src.src_cap);
  const float4 *sa = src.a + (size_t)bf * src.src_cap;
  const float2 *sc = src.c + (size_t)bf * src.src_cap;
  const int lane = lane_id();
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    bool keep = false;
    float v = 0.f;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n) {
      a = sa[i];
      const float conf = __fmul_rn(a.w, 16.0f);
      if (conf > th) {                                            // :31
        float hv = scalar_value_at(hr + (size_t)bf * hh * hw, hh, hw, a.x, a.y, 0.0f);   // :37
        v = __fadd_rn(__fmul_rn(0.9f, hv), __fmul_rn(0.1f, conf));                      // :38
        if (score_scale != 1.0f) v = __fmul_rn(v, score_scale);   // :41-42
        keep = v > th;                                            // :43
      }
    }
    unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m == 0u) continue;
    int leader = __ffs(m) - 1;
    int at = 0;
    if (lane == leader) at = atomicAdd(&n_raw[b], __popc(m));
    at = __shfl_sync(0xffffffffu, at, leader);
    if (keep) {
      int slot = at + __popc(m & ((1u << lane) - 1u));
      if (slot < cap) {
        const float2 e = sc[i];
        float *o = raw + ((size_t)b * cap + slot) * REC;
        o[0] = v; o[1] = (float)f; o[2] = a.x; o[3] = a.y; o[4] = e.x;   // :44,47
        if (REC == 6) o[5] = e.y;                                        // :84
      }
    }
  }
}

// strict "a sorts before b" for descending tuple order (v, f, x, y, s) / (v, f, x, y, w, h)
template <int REC>
__device__ __forceinline__ bool seed_before(uint64_t a, uint64_t b, const float *__restrict__ raw) {
  uint64_t ha = a >> 24, hb = b >> 24;
  if (ha != hb) return ha > hb;
  if (a == b) return false;
  const float *pa = raw + (size_t)(a & 0xffffffu) * REC, *pb = raw + (size_t)(b & 0xffffffu) * REC;
#pragma unroll
  for (int i = 2; i < REC; i++) {
    float xa = pa[i], xb = pb[i];
    if (xa > xb) return true;
    if (xa < xb) return false;
  }
  return false;
}

template <int REC>
__global__ void __launch_bounds__(kSortThreads)
seeds_sort_kernel(const float *__restrict__ raw_all, const int *__restrict__ n_raw,
                  float *__restrict__ out_all, int *__restrict__ n_out, uint64_t *__restrict__ scratch,
                  int cap, int scratch_stride, int *__restrict__ status) {
  extern __shared__ uint64_t s_keys[];
  const int b = blockIdx.x;
  int n = n_raw[b];
  if (n > cap) {
    if (threadIdx.x == 0) atomicOr(&status[b], PP_STATUS_SEED_OVERFLOW);
    n = cap;
  }
  const float *raw = raw_all + (size_t)b * cap * REC;
  float *out = out_all + (size_t)b * cap * REC;
  int npad = 1;
  while (npad < n) npad <<= 1;
  uint64_t *keys = (npad <= kSortSmemKeys) ? s_keys : scratch + (size_t)b * scratch_stride;

  for (int i = threadIdx.x; i < npad; i += blockDim.x) {
    uint64_t k = 0;
    if (i < n) {
      const float *r = raw + (size_t)i * REC;
      k = ((uint64_t)orderable(r[0]) << 32) | ((uint64_t)((uint32_t)r[1] & 0xffu) << 24) |
          (uint64_t)(uint32_t)i;
    }
    keys[i] = k;
  }
  __syncthreads();
  if (n <= kRankSortMax) {
    // short lists: every record counts the records that
    // sort before it -- one barrier instead of the ~50 of the bitonic network
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const uint64_t me = keys[i];
      int rank = 0;
      for (int j = 0; j < n; j++) {
        const uint64_t other = keys[j];
        if ((other >> 24) != (me >> 24)) {
          rank += (other >> 24) > (me >> 24);
        } else if (j != i) {
          const bool before = seed_before<REC>(other, me, raw);
          // records equal in every field keep their raw order
          rank += before || (j < i && !seed_before<REC>(me, other, raw));
        }
      }
      const float *r = raw + (size_t)i * REC;
      float *o = out + (size_t)rank * REC;
#pragma unroll
      for (int c = 0; c < REC; c++) o[c] = r[c];
    }
    if (threadIdx.x == 0) n_out[b] = n;
    return;
  }
  // bitonic network.  A stage with partner distance j < 32 stays inside aligned groups of 32 keys,
  // and a group is always handled by the same warp (position p belongs to thread p % 1024): between
  // two such stages a warp barrier is enough; a stage with j >= 32 crosses groups on both sides.
  bool prev_wide = false;      // the keys were loaded behind a CTA barrier
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= 32 || prev_wide) __syncthreads(); else __syncwarp();
      prev_wide = j >= 32;
      for (int i = threadIdx.x; i < npad; i += blockDim.x) {
        int l = i ^ j;
        if (l > i) {
          uint64_t a = keys[i], c = keys[l];
          bool desc_block = (i & k) == 0;   // this block ends up descending
          bool swap = desc_block ? seed_before<REC>(c, a, raw) : seed_before<REC>(a, c, raw);
          if (swap) { keys[i] = c; keys[l] = a; }
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float *r = raw + (size_t)(keys[i] & 0xffffffu) * REC;
    float *o = out + (size_t)i * REC;
#pragma unroll
    for (int c = 0; c < REC; c++) o[c] = r[c];
  }
  if (threadIdx.x == 0) n_out[b] = n;
}

// ---------------------------------------------------------------------------------------------
// Segmented seed sort (single field set, seeds from the compacted sources).  sorted(seeds,
// reverse=True) orders by (v, f, x, y, s) descending (cif_seeds.py:54), so
//   K3c seeds_field_sort_kernel  one CTA per (image, field): the field's seeds, sorted descending by
//                                (v, x, y, s) in shared memory, written to the field's segment;
//   K3d seeds_merge_kernel       every seed's final position = its position in its own segment + for
//                                every other field g the number of g's seeds that sort before it
//                                (v_g > v, or v_g == v and g > f): one binary search per other field.
// A crowded image (13 k seeds) no longer waits for one CTA's 100-stage bitonic network (0.42 ms).
// light CTAs (128 threads, ~8 KB of shared memory): several fit on an SM beside the resident CTAs of a
// CifHr splat of another batch in flight
constexpr int kSegThreads = 128;
constexpr int kSegSmemKeys = 1024;
constexpr int kMergeStage = 2048;       // scores of ONE other segment staged at a time for the binary searches

__device__ __forceinline__ float unorderable(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u ^ 0x80000000u) : ~u);
}

// strict "a sorts before b" inside one field: v descending, then (x, y, scale[, h]) descending
template <int REC>
__device__ __forceinline__ bool seg_before(uint64_t a, uint64_t b, const float4 *__restrict__ sa,
                                           const float2 *__restrict__ sc) {
  const uint32_t ha = (uint32_t)(a >> 32), hb = (uint32_t)(b >> 32);
  if (ha != hb) return ha > hb;
  if (a == b) return false;
  const float4 pa = sa[(uint32_t)a], pb = sa[(uint32_t)b];
  if (pa.x != pb.x) return pa.x > pb.x;
  if (pa.y != pb.y) return pa.y > pb.y;
  const float2 ea = sc[(uint32_t)a], eb = sc[(uint32_t)b];
  if (ea.x != eb.x) return ea.x > eb.x;
  if (REC == 6 && ea.y != eb.y) return ea.y > eb.y;
  return false;
}

template <int REC>
__global__ void __launch_bounds__(kSegThreads)
seeds_field_sort_kernel(CifSrcView src, const float *__restrict__ hr, float *__restrict__ seg,
                        int *__restrict__ n_seg, uint64_t *__restrict__ scratch, int scratch_stride, int F,
                        int seed_cap, int hh, int hw, float th, float score_scale, uint32_t seed_mask,
                        int mask_on) {
  __shared__ uint64_t s_keys[kSegSmemKeys];
  __shared__ int wtot[kSegThreads / 32];
  const int bf = blockIdx.x;
  const int b = bf / F, f = bf - b * F;
  int n = min(src.n[bf], src.src_cap);
  if (mask_on && !((seed_mask >> (f & 31)) & 1u)) n = 0;          // cif_seeds.py:29-30
  const float4 *sa = src.a + (size_t)bf * src.src_cap;
  const float2 *sc = src.c + (size_t)bf * src.src_cap;
  int npad_max = 1;
  while (npad_max < n) npad_max <<= 1;
  uint64_t *keys = npad_max <= kSegSmemKeys ? s_keys : scratch + (size_t)bf * scratch_stride;
  // the field's seeds (cif_seeds.py:31-47 on the compacted sources), compacted in source order
  int nk = 0;
  for (int base = 0; base < n; base += kSegThreads) {
    const int i = base + threadIdx.x;
    bool keep = false;
    float v = 0.f;
    if (i < n) {
      const float4 a = sa[i];
      const float conf = __fmul_rn(a.w, 16.0f);
      if (conf > th) {                                            // :31
        float hv = scalar_value_at(hr + (size_t)bf * hh * hw, hh, hw, a.x, a.y, 0.0f);   // :37
        v = __fadd_rn(__fmul_rn(0.9f, hv), __fmul_rn(0.1f, conf));                      // :38
        if (score_scale != 1.0f) v = __fmul_rn(v, score_scale);   // :41-42
        keep = v > th;                                            // :43
      }
    }
    int total;
    const int off = block_flag_scan(keep, wtot, &total);
    PP_DASSERT(!keep || nk + off < (npad_max <= kSegSmemKeys ? kSegSmemKeys : scratch_stride));
    if (keep) keys[nk + off] = ((uint64_t)orderable(v) << 32) | (uint32_t)i;
    nk += total;
  }
  __syncthreads();
  float *out = seg + ((size_t)b * seed_cap + (size_t)f * src.src_cap) * REC;
  auto emit = [&](int pos, uint64_t key) {
    const uint32_t i = (uint32_t)key;
    const float4 a = sa[i];
    const float2 e = sc[i];
    PP_DASSERT(pos >= 0 && pos < src.src_cap && i < (uint32_t)src.src_cap);
    float *o = out + (size_t)pos * REC;
    o[0] = unorderable((uint32_t)(key >> 32)); o[1] = (float)f; o[2] = a.x; o[3] = a.y; o[4] = e.x;   // :44,47
    if (REC == 6) o[5] = e.y;                                                                        // :84
  };
  if (threadIdx.x == 0) n_seg[bf] = nk;
  if (nk <= kRankSortMax) {
    // short segments (the usual case): every seed counts the ones that sort before it
    for (int i = threadIdx.x; i < nk; i += kSegThreads) {
      const uint64_t me = keys[i];
      int rank = 0;
      for (int j = 0; j < nk; j++) {
        const uint64_t other = keys[j];
        if ((other >> 32) != (me >> 32)) {
          rank += (other >> 32) > (me >> 32);
        } else if (j != i) {
          rank += seg_before<REC>(other, me, sa, sc) || (j < i && !seg_before<REC>(me, other, sa, sc));
        }
      }
      emit(rank, me);
    }
    return;
  }
  int npad = 1;
  while (npad < nk) npad <<= 1;
  for (int i = nk + threadIdx.x; i < npad; i += kSegThreads) keys[i] = 0;      // pads sort last
  // bitonic network, descending (see seeds_sort_kernel for the barrier rule)
  bool prev_wide = true;
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= 32 || prev_wide) __syncthreads(); else __syncwarp();
      prev_wide = j >= 32;
      for (int i = threadIdx.x; i < npad; i += kSegThreads) {
        const int l = i ^ j;
        if (l > i) {
          const uint64_t a = keys[i], c = keys[l];
          const bool desc_block = (i & k) == 0;
          const bool a_pad = a == 0, c_pad = c == 0;
          bool swap;
          if (a_pad || c_pad) swap = desc_block ? (a_pad && !c_pad) : (c_pad && !a_pad);
          else swap = desc_block ? seg_before<REC>(c, a, sa, sc) : seg_before<REC>(a, c, sa, sc);
          if (swap) { keys[i] = c; keys[l] = a; }
        }
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nk; i += kSegThreads) emit(i, keys[i]);
}

template <int REC>
__global__ void __launch_bounds__(kSegThreads)
seeds_merge_kernel(const float *__restrict__ seg, const int *__restrict__ n_seg, float *__restrict__ out_all,
                   int *__restrict__ n_out, int F, int seed_cap, int src_cap) {
  __shared__ float s_v[kMergeStage];             // the scores of one other segment
  __shared__ int s_n[PP_MAX_CATEGORIES];
  const int bf = blockIdx.x;
  const int b = bf / F, f = bf - b * F;
  for (int g = threadIdx.x; g < F; g += kSegThreads) s_n[g] = n_seg[b * F + g];
  __syncthreads();
  if (f == 0 && threadIdx.x == 0) {
    int total = 0;
    for (int g = 0; g < F; g++) total += s_n[g];
    n_out[b] = min(total, seed_cap);
  }
  const int nk = s_n[f];
  if (nk == 0) return;                           // uniform: no barrier is skipped by part of the CTA
  const float *base = seg + (size_t)b * seed_cap * REC;
  float *out = out_all + (size_t)b * seed_cap * REC;
  // every thread owns up to kOwn of the segment's seeds (more than kOwn * threads: extra rounds)
  constexpr int kOwn = 4;
  for (int k0 = 0; k0 < nk; k0 += kOwn * kSegThreads) {
    float v[kOwn];
    int rank[kOwn];
#pragma unroll
    for (int q = 0; q < kOwn; q++) {
      const int k = k0 + q * kSegThreads + threadIdx.x;
      v[q] = k < nk ? base[((size_t)f * src_cap + k) * REC] : 0.f;
      rank[q] = k;
    }
    for (int g = 0; g < F; g++) {
      const int ng = s_n[g];
      if (g == f || ng == 0) continue;           // uniform
      const float *vg = base + (size_t)g * src_cap * REC;
      const bool staged = ng <= kMergeStage;
      __syncthreads();                           // the previous segment's searches are done
      if (staged) {
        for (int k = threadIdx.x; k < ng; k += kSegThreads) s_v[k] = vg[(size_t)k * REC];
        __syncthreads();
      }
#pragma unroll
      for (int q = 0; q < kOwn; q++) {
        if (k0 + q * kSegThreads + threadIdx.x >= nk) continue;
        int lo = 0, hi = ng;                     // first index of g's descending list that does NOT sort before us
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          const float x = staged ? s_v[mid] : vg[(size_t)mid * REC];
          if (x > v[q] || (x == v[q] && g > f)) lo = mid + 1; else hi = mid;
        }
        rank[q] += lo;
      }
    }
#pragma unroll
    for (int q = 0; q < kOwn; q++) {
      const int k = k0 + q * kSegThreads + threadIdx.x;
      if (k < nk && rank[q] < seed_cap) {
        const float *rec = base + ((size_t)f * src_cap + k) * REC;
        float *o = out + (size_t)rank[q] * REC;
#pragma unroll
        for (int c = 0; c < REC; c++) o[c] = rec[c];
      }
    }
  }
}

struct SeedsWs {
  float *raw;        // (B,cap,5); segmented sort: field f's sorted seeds at f * src_cap
  int *n_raw;        // (B)
  uint64_t *scratch; // (B,pow2(cap)) only used beyond kSortSmemKeys
  int scratch_stride;
  int *n_seg;        // (B,F) seeds per field (segmented sort)
  uint64_t *seg_scratch;   // (B,F,pow2(src_cap)) only used beyond kSegSmemKeys
  int seg_stride;
};

static int pow2_ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }

size_t seeds_ws_bytes(const pp_shape &s, const pp_caps &c, int det) {
  size_t n = 0;
  n += align_up((size_t)s.B * c.seed_cap * (det ? 6 : 5) * sizeof(float), 256);
  n += align_up((size_t)s.B * sizeof(int), 256);
  int p2 = pow2_ceil(c.seed_cap);
  if (p2 > kSortSmemKeys) n += align_up((size_t)s.B * p2 * sizeof(uint64_t), 256);
  n += align_up((size_t)s.B * s.F * sizeof(int), 256);
  int q2 = pow2_ceil(c.src_cap);
  if (q2 > kSegSmemKeys) n += align_up((size_t)s.B * s.F * q2 * sizeof(uint64_t), 256);
  return n;
}

static SeedsWs seeds_carve(void *d_ws, int B, const pp_caps &c, int det, int F = 0) {
  SeedsWs w;
  char *q = (char *)d_ws;
  w.raw = (float *)q; q += align_up((size_t)B * c.seed_cap * (det ? 6 : 5) * sizeof(float), 256);
  w.n_raw = (int *)q; q += align_up((size_t)B * sizeof(int), 256);
  w.scratch = (uint64_t *)q;
  w.scratch_stride = pow2_ceil(c.seed_cap);
  if (w.scratch_stride > kSortSmemKeys) q += align_up((size_t)B * w.scratch_stride * sizeof(uint64_t), 256);
  w.n_seg = (int *)q; q += align_up((size_t)B * F * sizeof(int), 256);
  w.seg_scratch = (uint64_t *)q;
  w.seg_stride = pow2_ceil(c.src_cap);
  return w;
}

int seeds_collect_launch(const float *d_cif, const float *d_hr, void *d_ws, const pp_shape &s,
                         const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_scale,
                         int reset, cudaStream_t st, int det) {
  if (c.seed_cap >= (1 << 24) || s.F > 256) return PP_ERR_ARG;    // key layout: f:8 | index:24
  SeedsWs w = seeds_carve(d_ws, s.B, c, det);
  const int hh = hr_h(s_hr), hw = hr_w(s_hr), n = s.H * s.W;
  prof_mark(p.profile, kStCollect, st);
  if (reset) PP_CUDA_CHECK(cudaMemsetAsync(w.n_raw, 0, (size_t)s.B * sizeof(int), st));
  int blocks = s.B * s.F * ((n + kCollectThreads - 1) / kCollectThreads);
  if (det)
    seeds_collect_kernel<6><<<blocks, kCollectThreads, 0, st>>>(
        d_cif, d_hr, w.raw, w.n_raw, s, c.seed_cap, hh, hw, p.seed_threshold, p.seed_score_scale,
        min_scale / (float)s.stride, min_scale != 0.0f, 0u, 0);
  else
    seeds_collect_kernel<5><<<blocks, kCollectThreads, 0, st>>>(
        d_cif, d_hr, w.raw, w.n_raw, s, c.seed_cap, hh, hw, p.seed_threshold, p.seed_score_scale,
        min_scale / (float)s.stride, min_scale != 0.0f, p.seed_mask,
        p.seed_mask_enabled || p.seed_mask != 0u);
  PP_LAUNCH_CHECK("seeds_collect_kernel");
  return PP_OK;
}

int seeds_collect_from_sources_launch(const CifSrcView &src, const float *d_hr, void *d_ws, const pp_shape &s,
                                      const pp_caps &c, const pp_params &p, cudaStream_t st, int det) {
  if (c.seed_cap >= (1 << 24) || s.F > 256) return PP_ERR_ARG;
  SeedsWs w = seeds_carve(d_ws, s.B, c, det);
  const int hh = hr_h(s), hw = hr_w(s);
  prof_mark(p.profile, kStCollect, st);
  PP_CUDA_CHECK(cudaMemsetAsync(w.n_raw, 0, (size_t)s.B * sizeof(int), st));
  if (det)
    seeds_from_sources_kernel<6><<<s.B * s.F, 128, 0, st>>>(src, d_hr, w.raw, w.n_raw, s.F, c.seed_cap, hh, hw,
                                                           p.seed_threshold, p.seed_score_scale, 0u, 0);
  else
    seeds_from_sources_kernel<5><<<s.B * s.F, 128, 0, st>>>(src, d_hr, w.raw, w.n_raw, s.F, c.seed_cap, hh, hw,
                                                           p.seed_threshold, p.seed_score_scale, p.seed_mask,
                                                           p.seed_mask_enabled || p.seed_mask != 0u);
  PP_LAUNCH_CHECK("seeds_from_sources_kernel");
  return PP_OK;
}

// Both seed stages at once for one field set (seeds from the compacted sources): per-field sort + merge.
bool seeds_segmented_ok(const pp_shape &s, const pp_caps &c) {
  return (long long)c.seed_cap >= (long long)s.F * c.src_cap && s.F <= PP_MAX_CATEGORIES;
}

int seeds_segmented_launch(const CifSrcView &src, const float *d_hr, float *d_seeds, int *d_n_seeds, void *d_ws,
                           const pp_shape &s, const pp_caps &c, const pp_params &p, cudaStream_t st, int det) {
  if (!seeds_segmented_ok(s, c)) return PP_ERR_ARG;
  SeedsWs w = seeds_carve(d_ws, s.B, c, det, s.F);
  const int hh = hr_h(s), hw = hr_w(s);
  const int mask_on = !det && (p.seed_mask_enabled || p.seed_mask != 0u);
  const uint32_t mask = det ? 0u : p.seed_mask;
  prof_mark(p.profile, kStCollect, st);
  if (det)
    seeds_field_sort_kernel<6><<<s.B * s.F, kSegThreads, 0, st>>>(src, d_hr, w.raw, w.n_seg, w.seg_scratch,
                                                                 w.seg_stride, s.F, c.seed_cap, hh, hw,
                                                                 p.seed_threshold, p.seed_score_scale, mask, mask_on);
  else
    seeds_field_sort_kernel<5><<<s.B * s.F, kSegThreads, 0, st>>>(src, d_hr, w.raw, w.n_seg, w.seg_scratch,
                                                                 w.seg_stride, s.F, c.seed_cap, hh, hw,
                                                                 p.seed_threshold, p.seed_score_scale, mask, mask_on);
  PP_LAUNCH_CHECK("seeds_field_sort_kernel");
  prof_mark(p.profile, kStSort, st);
  if (det)
    seeds_merge_kernel<6><<<s.B * s.F, kSegThreads, 0, st>>>(w.raw, w.n_seg, d_seeds, d_n_seeds, s.F, c.seed_cap,
                                                            c.src_cap);
  else
    seeds_merge_kernel<5><<<s.B * s.F, kSegThreads, 0, st>>>(w.raw, w.n_seg, d_seeds, d_n_seeds, s.F, c.seed_cap,
                                                            c.src_cap);
  PP_LAUNCH_CHECK("seeds_merge_kernel");
  return PP_OK;
}

int seeds_sort_launch(float *d_seeds, int *d_n_seeds, void *d_ws, int B, const pp_caps &c,
                      const pp_params &p, int *d_status, cudaStream_t st, int det) {
  SeedsWs w = seeds_carve(d_ws, B, c, det);
  // small capacities only need a small shared array
  size_t need = (size_t)(w.scratch_stride <= kSortSmemKeys ? w.scratch_stride : kSortSmemKeys) *
                sizeof(uint64_t);
  prof_mark(p.profile, kStSort, st);
  if (det)
    seeds_sort_kernel<6><<<B, kSortThreads, need, st>>>(w.raw, w.n_raw, d_seeds, d_n_seeds, w.scratch,
                                                        c.seed_cap, w.scratch_stride, d_status);
  else
    seeds_sort_kernel<5><<<B, kSortThreads, need, st>>>(w.raw, w.n_raw, d_seeds, d_n_seeds, w.scratch,
                                                        c.seed_cap, w.scratch_stride, d_status);
  PP_LAUNCH_CHECK("seeds_sort_kernel");
  return PP_OK;
}

int seeds_launch(const float *d_cif, const float *d_hr, float *d_seeds, int *d_n_seeds, void *d_ws,
                 const pp_shape &s, const pp_caps &c, const pp_params &p, int *d_status,
                 cudaStream_t st, int det) {
  int rc = seeds_collect_launch(d_cif, d_hr, d_ws, s, s, c, p, p.cif_min_scale, 1, st, det);
  if (rc) return rc;
  return seeds_sort_launch(d_seeds, d_n_seeds, d_ws, s.B, c, p, d_status, st, det);
}

}  // namespace pp
