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

struct SeedsWs {
  float *raw;        // (B,cap,5)
  int *n_raw;        // (B)
  uint64_t *scratch; // (B,pow2(cap)) only used beyond kSortSmemKeys
  int scratch_stride;
};

static int pow2_ceil(int v) { int p = 1; while (p < v) p <<= 1; return p; }

size_t seeds_ws_bytes(const pp_shape &s, const pp_caps &c, int det) {
  size_t n = 0;
  n += align_up((size_t)s.B * c.seed_cap * (det ? 6 : 5) * sizeof(float), 256);
  n += align_up((size_t)s.B * sizeof(int), 256);
  int p2 = pow2_ceil(c.seed_cap);
  if (p2 > kSortSmemKeys) n += align_up((size_t)s.B * p2 * sizeof(uint64_t), 256);
  return n;
}

static SeedsWs seeds_carve(void *d_ws, int B, const pp_caps &c, int det) {
  SeedsWs w;
  char *q = (char *)d_ws;
  w.raw = (float *)q; q += align_up((size_t)B * c.seed_cap * (det ? 6 : 5) * sizeof(float), 256);
  w.n_raw = (int *)q; q += align_up((size_t)B * sizeof(int), 256);
  w.scratch = (uint64_t *)q;
  w.scratch_stride = pow2_ceil(c.seed_cap);
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
