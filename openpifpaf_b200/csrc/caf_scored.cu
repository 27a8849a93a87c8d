// caf_scored.cu -- CafScored on sm_100a.
//
// Replaces decoder/caf_scored.py:32-87 (CafScored.fill_caf) of the reference:
// threshold the CAF confidence, scale the 8 geometry rows by the stride,
// re-score both ends against the CifHr map (scalar_values,
// functional.pyx:231-244) and emit two ORDER-PRESERVING compactions per link:
//   forward[l]  rows (score_f, x1, y1, b1, s1, x2, y2, b2, s2)
//   backward[l] rows (score_b, x2, y2, b2, s2, x1, y1, b1, s1)
// Candidate order is row-major cell order, as numpy boolean masking gives it;
// the assembly stage relies on that order only to break exact score ties.
//
// One CTA per (image, link), two phases per slab of cells: (1) the ordered list of the cells above
// the threshold (block_active_cells: coalesced sweep of the confidence plane, two barriers);
// (2) one listed cell per thread: the other eight planes, the two hr gathers and both ordered
// compactions.  The first version interleaved the two per 256 cells and was bound by ~26 serial
// rounds of dependent loads + barriers per CTA (97 us at batch 64).
#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kCafThreads = 256;

constexpr int kCafSlab = 8192;       // cells per slab: the active-cell list of a slab is 16 KiB of smem

__global__ void __launch_bounds__(kCafThreads)
caf_scored_kernel(const float *__restrict__ caf, const float *__restrict__ caf_rest,
                  const float *__restrict__ hr, pp_skeleton sk,
                  float *__restrict__ fwd, int *__restrict__ nfwd, float *__restrict__ bwd,
                  int *__restrict__ nbwd, pp_shape s, int cap, int hh, int hw, float th,
                  float cif_floor, float cif_floor_compl, float min_dist, int use_min, float max_dist,
                  int use_max, int append, int *__restrict__ status) {
  extern __shared__ unsigned short s_cells[];    // min(n, kCafSlab) active cells, relative to the slab
  __shared__ int wtot_f[kCafThreads / 32], wtot_b[kCafThreads / 32];
  const int bl = blockIdx.x;
  const int b = bl / s.L, l = bl - b * s.L;
  const int n = s.H * s.W;
  const float *p = caf + (size_t)bl * 9 * n;
  // caf_rest != NULL (host-mapped input): the eight geometry planes are read from the pinned host
  // buffer, only where a cell is active; the confidence plane was copied to the device
  const float *pr = caf_rest ? caf_rest + (size_t)bl * 9 * n : p;
  const float stride = (float)s.stride;
  const int j1 = sk.joints[l][0] - 1, j2 = sk.joints[l][1] - 1;
  const float *hr_b = hr + (size_t)b * s.F * hh * hw;
  float *fo = fwd + (size_t)bl * 9 * cap;
  float *bo = bwd + (size_t)bl * 9 * cap;
  const int lane = lane_id(), w = warp_id(), nw = blockDim.x >> 5;
  // append: a later field set of a multi-scale configuration continues the lists (caf_scored.py:70,81)
  int cnt_f = append ? nfwd[bl] : 0, cnt_b = append ? nbwd[bl] : 0;

  for (int slab = 0; slab < n; slab += kCafSlab) {
    // phase 1: the cells above the threshold, in row-major order (caf_scored.py:42)
    const int n_act = block_active_cells<kCafThreads>(p, slab, min(n, slab + kCafSlab), th, s_cells, wtot_f);
    // phase 2: one active cell per thread -- eight more planes, two hr gathers, both compactions
    for (int base = 0; base < n_act; base += blockDim.x) {
      const int i = base + threadIdx.x;
      bool kf = false, kb = false;
      float nine[9];
      float sf = 0.f, sb = 0.f;
      if (i < n_act) {
        const int c = slab + s_cells[i];
        nine[0] = __ldg(p + c);
#pragma unroll
        for (int r = 1; r < 9; r++) nine[r] = __ldg(pr + (size_t)r * n + c);
        bool ok = true;
        if (use_min || use_max) {                                 // :47-55 (multi-scale only)
          float dx = __fsub_rn(nine[1], nine[5]), dy = __fsub_rn(nine[2], nine[6]);
          float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
          if (use_min && !(d > min_dist)) ok = false;
          if (use_max && !(d < max_dist)) ok = false;
        }
        if (ok) {
#pragma unroll
          for (int r = 1; r < 9; r++) nine[r] = __fmul_rn(nine[r], stride);   // :58
          sb = sf = nine[0];
          if (cif_floor < 1.0f && j1 < s.F) {                     // :61-64
            float h = scalar_value_at(hr_b + (size_t)j1 * hh * hw, hh, hw, nine[1], nine[2], 0.0f);
            sb = __fmul_rn(nine[0], __fadd_rn(cif_floor, __fmul_rn(cif_floor_compl, h)));
          }
          if (cif_floor < 1.0f && j2 < s.F) {                     // :72-75
            float h = scalar_value_at(hr_b + (size_t)j2 * hh * hw, hh, hw, nine[5], nine[6], 0.0f);
            sf = __fmul_rn(nine[0], __fadd_rn(cif_floor, __fmul_rn(cif_floor_compl, h)));
          }
          kb = sb > th;                                           // :67
          kf = sf > th;                                           // :78
        }
      }
      // two ordered CTA-wide compactions in one barrier pair
      unsigned mf = __ballot_sync(0xffffffffu, kf), mb = __ballot_sync(0xffffffffu, kb);
      if (lane == 0) { wtot_f[w] = __popc(mf); wtot_b[w] = __popc(mb); }
      __syncthreads();
      int base_f = 0, base_b = 0, tot_f = 0, tot_b = 0;
      for (int k = 0; k < nw; k++) {
        int tf = wtot_f[k], tb = wtot_b[k];
        if (k < w) { base_f += tf; base_b += tb; }
        tot_f += tf; tot_b += tb;
      }
      __syncthreads();
      const unsigned below = (1u << lane) - 1u;
      if (kb) {
        int k = cnt_b + base_b + __popc(mb & below);
        if (k < cap) {                                            // :68-69 row permutation
          bo[k] = sb;
          bo[(size_t)1 * cap + k] = nine[5]; bo[(size_t)2 * cap + k] = nine[6];
          bo[(size_t)3 * cap + k] = nine[7]; bo[(size_t)4 * cap + k] = nine[8];
          bo[(size_t)5 * cap + k] = nine[1]; bo[(size_t)6 * cap + k] = nine[2];
          bo[(size_t)7 * cap + k] = nine[3]; bo[(size_t)8 * cap + k] = nine[4];
        }
      }
      if (kf) {
        int k = cnt_f + base_f + __popc(mf & below);
        if (k < cap) {                                            // :79-80
          fo[k] = sf;
#pragma unroll
          for (int r = 1; r < 9; r++) fo[(size_t)r * cap + k] = nine[r];
        }
      }
      cnt_f += tot_f;
      cnt_b += tot_b;
    }
  }
  if (threadIdx.x == 0) {
    if (cnt_f > cap || cnt_b > cap) atomicOr(&status[b], PP_STATUS_CAF_OVERFLOW);
    nfwd[bl] = min(cnt_f, cap);
    nbwd[bl] = min(cnt_b, cap);
  }
}

// ---------------------------------------------------------------------------
// Spatial binning of the candidate lists for the assembly scans.
//
// _grow_connection only wants the candidates whose source end lies in a box
// around the joint (caf_center_s, functional.pyx:346-353); crowded scenes have
// >1000 candidates per link.  One CTA per (image, link, direction) counting-sorts
// the list by the bin of (x_source, y_source).  Order inside a bin is arbitrary:
// the assembly breaks score ties on the ORIGINAL list index, which is kept as
// the seventh row, so results do not depend on it.

BinGeo bin_geo(const pp_shape &s) {
  BinGeo g;
  g.bshift = 5;
  for (;;) {
    g.nbx = (hr_w(s) >> g.bshift) + 1;
    g.nby = (hr_h(s) >> g.bshift) + 1;
    if ((long long)g.nbx * g.nby + 1 <= kMaxBins) break;
    g.bshift++;
  }
  g.nb = g.nbx * g.nby + 1;
  return g;
}

size_t binned_rows_bytes(const pp_shape &s, const pp_caps &c) {
  return (size_t)s.B * s.L * 2 * kBinRows * c.caf_cap * sizeof(float);
}
size_t binned_start_bytes(const pp_shape &s) {
  return (size_t)s.B * s.L * 2 * (bin_geo(s).nb + 1) * sizeof(int);
}

__device__ __forceinline__ int bin_coord(float v, float inv, int n) {
  return (int)fminf(fmaxf(__fmul_rn(v, inv), 0.0f), (float)(n - 1));   // monotonic, clamped
}

__device__ __forceinline__ int bin_of(float x, float y, const BinGeo &g, float inv) {
  if (x != x || y != y) return g.nb - 1;          // NaN passes every box test: "always" bin
  return bin_coord(y, inv, g.nby) * g.nbx + bin_coord(x, inv, g.nbx);
}

// counting sort of one directed list by spatial bin; CTA-collective (kCafThreads threads).
// hist: kMaxBins ints, warp_sum: kCafThreads / 32 ints of shared memory.
__device__ __forceinline__ void bin_one_list(const float *__restrict__ src, const int K, float *__restrict__ dst,
                                             int *__restrict__ st, const int cap, const BinGeo &g, int *hist,
                                             int *warp_sum) {
  const float inv = 1.0f / (float)(1 << g.bshift);
  const int tid = threadIdx.x;
  for (int i = tid; i < g.nb; i += kCafThreads) hist[i] = 0;
  __syncthreads();
  for (int k = tid; k < K; k += kCafThreads)
    atomicAdd(&hist[bin_of(src[(size_t)cap + k], src[2 * (size_t)cap + k], g, inv)], 1);
  __syncthreads();
  // exclusive scan of hist[0..nb): each thread owns a run of consecutive bins
  const int per = (g.nb + kCafThreads - 1) / kCafThreads;
  const int b0 = tid * per, b1 = min(b0 + per, g.nb);
  int local = 0;
  for (int i = b0; i < b1; i++) local += hist[i];
  int incl = local;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    int v = __shfl_up_sync(0xffffffffu, incl, off);
    if ((tid & 31) >= off) incl += v;
  }
  if ((tid & 31) == 31) warp_sum[tid >> 5] = incl;
  __syncthreads();
  int base = incl - local;
  for (int w = 0; w < (tid >> 5); w++) base += warp_sum[w];
  for (int i = b0; i < b1; i++) {
    int h = hist[i];
    hist[i] = base;      // becomes the running insert cursor of the bin
    st[i] = base;
    base += h;
  }
  if (tid == 0) st[g.nb] = K;
  __syncthreads();
  for (int k = tid; k < K; k += kCafThreads) {
    const float xs = src[(size_t)cap + k], ys = src[2 * (size_t)cap + k];
    const int pos = atomicAdd(&hist[bin_of(xs, ys, g, inv)], 1);
    PP_DASSERT(pos >= 0 && pos < K);
    dst[pos] = src[k];
    dst[(size_t)cap + pos] = xs;
    dst[2 * (size_t)cap + pos] = ys;
    dst[3 * (size_t)cap + pos] = src[5 * (size_t)cap + k];
    dst[4 * (size_t)cap + pos] = src[6 * (size_t)cap + k];
    dst[5 * (size_t)cap + pos] = src[8 * (size_t)cap + k];
    dst[6 * (size_t)cap + pos] = __int_as_float(k);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kCafThreads)
caf_bin_kernel(const float *__restrict__ fwd, const int *__restrict__ nfwd,
               const float *__restrict__ bwd, const int *__restrict__ nbwd, float *__restrict__ rows,
               int *__restrict__ start, int cap, BinGeo g) {
  __shared__ int hist[kMaxBins];
  __shared__ int warp_sum[kCafThreads / 32];
  const int bl = blockIdx.x >> 1, dir = blockIdx.x & 1;
  const int K = min(dir ? nbwd[bl] : nfwd[bl], cap);
  bin_one_list((dir ? bwd : fwd) + (size_t)bl * 9 * cap, K, rows + (size_t)blockIdx.x * kBinRows * cap,
               start + (size_t)blockIdx.x * (g.nb + 1), cap, g, hist, warp_sum);
}

// CafScored.fill_caf (caf_scored.py:59-81) from the raw candidates the prescan left in fwd / nfwd
// (threshold, distance filters and stride scaling already applied; cell order), then the spatial
// binning of both lists for the assembly -- one CTA per (image, link).  The forward list is
// compacted IN PLACE: entries are read 256 at a time into registers, a barrier, then written at
// positions that can only be at or below the ones just read.
__global__ void __launch_bounds__(kCafThreads)
caf_score_bin_kernel(const float *__restrict__ hr, pp_skeleton sk, float *__restrict__ fwd,
                     int *__restrict__ nfwd, float *__restrict__ bwd, int *__restrict__ nbwd,
                     float *__restrict__ rows, int *__restrict__ start, pp_shape s, int cap, int hh, int hw,
                     float th, float cif_floor, float cif_floor_compl, BinGeo g) {
  __shared__ int hist[kMaxBins];
  __shared__ int wtot_f[kCafThreads / 32], wtot_b[kCafThreads / 32];
  const int bl = blockIdx.x;
  const int b = bl / s.L, l = bl - b * s.L;
  const int j1 = sk.joints[l][0] - 1, j2 = sk.joints[l][1] - 1;
  const float *hr_b = hr + (size_t)b * s.F * hh * hw;
  float *fo = fwd + (size_t)bl * 9 * cap;
  float *bo = bwd + (size_t)bl * 9 * cap;
  const int lane = lane_id(), w = warp_id(), nw = blockDim.x >> 5;
  const int n_raw = min(nfwd[bl], cap);
  int cnt_f = 0, cnt_b = 0;
  for (int base = 0; base < n_raw; base += kCafThreads) {
    const int i = base + threadIdx.x;
    bool kf = false, kb = false;
    float nine[9];
    float sf = 0.f, sb = 0.f;
    if (i < n_raw) {
#pragma unroll
      for (int r = 0; r < 9; r++) nine[r] = fo[(size_t)r * cap + i];
      sb = sf = nine[0];
      if (cif_floor < 1.0f && j1 < s.F) {                     // :61-64
        float h = scalar_value_at(hr_b + (size_t)j1 * hh * hw, hh, hw, nine[1], nine[2], 0.0f);
        sb = __fmul_rn(nine[0], __fadd_rn(cif_floor, __fmul_rn(cif_floor_compl, h)));
      }
      if (cif_floor < 1.0f && j2 < s.F) {                     // :72-75
        float h = scalar_value_at(hr_b + (size_t)j2 * hh * hw, hh, hw, nine[5], nine[6], 0.0f);
        sf = __fmul_rn(nine[0], __fadd_rn(cif_floor, __fmul_rn(cif_floor_compl, h)));
      }
      kb = sb > th;                                           // :67
      kf = sf > th;                                           // :78
    }
    // two ordered CTA-wide compactions in one barrier pair (the first barrier also orders this
    // round's reads of the forward rows before its in-place writes)
    unsigned mf = __ballot_sync(0xffffffffu, kf), mb = __ballot_sync(0xffffffffu, kb);
    if (lane == 0) { wtot_f[w] = __popc(mf); wtot_b[w] = __popc(mb); }
    __syncthreads();
    int base_f = 0, base_b = 0, tot_f = 0, tot_b = 0;
    for (int k = 0; k < nw; k++) {
      int tf = wtot_f[k], tb = wtot_b[k];
      if (k < w) { base_f += tf; base_b += tb; }
      tot_f += tf; tot_b += tb;
    }
    __syncthreads();
    const unsigned below = (1u << lane) - 1u;
    if (kb) {
      const int k = cnt_b + base_b + __popc(mb & below);      // :68-69 row permutation
      PP_DASSERT(k < cap);
      bo[k] = sb;
      bo[(size_t)1 * cap + k] = nine[5]; bo[(size_t)2 * cap + k] = nine[6];
      bo[(size_t)3 * cap + k] = nine[7]; bo[(size_t)4 * cap + k] = nine[8];
      bo[(size_t)5 * cap + k] = nine[1]; bo[(size_t)6 * cap + k] = nine[2];
      bo[(size_t)7 * cap + k] = nine[3]; bo[(size_t)8 * cap + k] = nine[4];
    }
    if (kf) {
      const int k = cnt_f + base_f + __popc(mf & below);      // :79-80
      PP_DASSERT(k <= i && k < cap);                         // in place: never ahead of the entry just read
      fo[k] = sf;
#pragma unroll
      for (int r = 1; r < 9; r++) fo[(size_t)r * cap + k] = nine[r];
    }
    cnt_f += tot_f;
    cnt_b += tot_b;
  }
  if (threadIdx.x == 0) { nfwd[bl] = cnt_f; nbwd[bl] = cnt_b; }
  __syncthreads();          // the lists this CTA wrote are read back below
  bin_one_list(fo, cnt_f, rows + (size_t)(2 * bl) * kBinRows * cap, start + (size_t)(2 * bl) * (g.nb + 1), cap, g,
               hist, wtot_f);
  bin_one_list(bo, cnt_b, rows + (size_t)(2 * bl + 1) * kBinRows * cap, start + (size_t)(2 * bl + 1) * (g.nb + 1),
               cap, g, hist, wtot_f);
}

int caf_score_bin_launch(const float *d_hr, const pp_skeleton &sk, float score_th, float *d_fwd, int *d_nfwd,
                         float *d_bwd, int *d_nbwd, float *d_rows, int *d_start, const pp_shape &s,
                         const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st) {
  (void)d_status;          // cap >= H * W on this path: nothing can overflow
  prof_mark(p.profile, kStCaf, st);
  caf_score_bin_kernel<<<s.B * s.L, kCafThreads, 0, st>>>(d_hr, sk, d_fwd, d_nfwd, d_bwd, d_nbwd, d_rows, d_start,
                                                         s, c.caf_cap, hr_h(s), hr_w(s), score_th, p.cif_floor,
                                                         p.cif_floor_compl, bin_geo(s));
  PP_LAUNCH_CHECK("caf_score_bin_kernel");
  return PP_OK;
}

int caf_bin_launch(const float *d_fwd, const int *d_nfwd, const float *d_bwd, const int *d_nbwd,
                   float *d_rows, int *d_start, const pp_shape &s, const pp_caps &c, cudaStream_t st) {
  caf_bin_kernel<<<s.B * s.L * 2, kCafThreads, 0, st>>>(d_fwd, d_nfwd, d_bwd, d_nbwd, d_rows, d_start,
                                                       c.caf_cap, bin_geo(s));
  PP_LAUNCH_CHECK("caf_bin_kernel");
  return PP_OK;
}

int caf_scored_launch_ex(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                         float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                         const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_distance,
                         float max_distance, int append, int *d_status, cudaStream_t st,
                         const float *caf_rest) {
  const int hh = hr_h(s_hr), hw = hr_w(s_hr);
  const float stride = (float)s.stride;
  prof_mark(p.profile, kStCaf, st);
  const size_t smem = (size_t)min(s.H * s.W, kCafSlab) * sizeof(unsigned short);
  caf_scored_kernel<<<s.B * s.L, kCafThreads, smem, st>>>(
      d_caf, caf_rest, d_hr, sk, d_fwd, d_nfwd, d_bwd, d_nbwd, s, c.caf_cap, hh, hw, score_th, p.cif_floor,
      p.cif_floor_compl, min_distance / stride, min_distance != 0.0f, max_distance / stride,
      max_distance > 0.0f, append, d_status);
  PP_LAUNCH_CHECK("caf_scored_kernel");
  return PP_OK;
}

int caf_scored_launch(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                      float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                      const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st) {
  return caf_scored_launch_ex(d_caf, d_hr, sk, score_th, d_fwd, d_nfwd, d_bwd, d_nbwd, s, s, c, p,
                              p.caf_min_distance, p.caf_max_distance, 0, d_status, st, nullptr);
}

}  // namespace pp
