// assemble.cu -- greedy pose assembly with occupancy, on sm_100a.
//
// Replaces, for a whole batch, the serial part of CifCaf.__call__
// (decoder/generator/cifcaf.py:84-108), CifCaf._grow / connection_value /
// _grow_connection / _target_with_blend / _target_with_maxscore (:124-307),
// caf_center_s (functional.pyx:338-359) and Occupancy (decoder/occupancy.py,
// decoder/utils.py:61-66, functional.pyx:281-286).
//
// One CTA per image.  What is parallel inside an image:
//   * "next seed that is not occupied" is a CTA-wide search over the sorted
//     seed list -- occupancy only changes after a pose has been grown;
//   * the occupancy map is never rasterised: Occupancy.get(f, x, y) is the
//     number (mod 256, the reference's uint8 wrap) of Occupancy.set boxes of
//     field f covering the cell, so each seed carries a uint8 counter that is
//     bumped, CTA-wide, against the <= F boxes of the pose just finished;
//   * the frontier is evaluated eagerly: when a joint is set, one WARP per
//     outgoing edge scans the link's candidate list (lanes over candidates,
//     per-lane top-2, butterfly merge), applies blend/max and the reverse
//     match.  The reference evaluates lazily through a PriorityQueue whose
//     unevaluated keys are upper bounds of the evaluated ones, so "always take
//     the best evaluated edge whose target is still empty" visits the same
//     edges in the same order (validated by the tests).
//     Ties follow Python's tuple order (-score, (x, y, s, v), start, end).
#include <cuda_pipeline.h>

#include <stdlib.h>

#include "common.cuh"
#include "stages.cuh"
#include "occupancy.cuh"

namespace pp {

constexpr int kAsmThreads = 256;
constexpr int kAsmWarps = kAsmThreads / 32;
constexpr int kMaxEdges = 2 * PP_MAX_L;
// Occupancy bookkeeping.  Up to kHeavySeeds seeds every seed's counter is tested against the boxes of
// each finished pose (a few thousand tests).  Above that (crowded scenes, degenerate fields with 60 k
// seeds and ~1000 poses per image) the seeds are counting-sorted once by (field, band of kOccBands
// occupancy-row bands), a box then only visits the bins it overlaps, and a box that covers the whole
// map just bumps a per-field counter.
constexpr int kHeavySeeds = 4096;
constexpr int kOccBands = 32;
constexpr int kStageBytes = 80 * 1024;        // staged candidate lists, 2 CTAs / SM (batches > #SMs)
constexpr int kStageBytesSolo = 192 * 1024;   // ... 1 CTA / SM

struct Graph {
  int F;
  int8_t deg[PP_MAX_F];
  int8_t nb_end[PP_MAX_F][PP_MAX_DEG];
  int8_t nb_link[PP_MAX_F][PP_MAX_DEG];
  int8_t nb_fwd[PP_MAX_F][PP_MAX_DEG];
};

struct CScale {
  int has;
  float v[PP_MAX_L];
};

// by_source tables of cifcaf.py:62-65 (dict semantics: a repeated (start, end)
// pair overwrites the value and keeps its first position).
int build_graph(const pp_skeleton &sk, int F, Graph *g) {
  if (F > PP_MAX_F || sk.L > PP_MAX_L || sk.L < 0) return PP_ERR_ARG;
  memset(g, 0, sizeof *g);
  g->F = F;
  for (int l = 0; l < sk.L; l++) {
    int a = sk.joints[l][0] - 1, b = sk.joints[l][1] - 1;
    if (a < 0 || b < 0 || a >= F || b >= F) return PP_ERR_ARG;
    for (int dir = 0; dir < 2; dir++) {
      int s = dir ? b : a, e = dir ? a : b, k;
      for (k = 0; k < g->deg[s]; k++)
        if (g->nb_end[s][k] == e) break;
      if (k == g->deg[s]) {
        if (k >= PP_MAX_DEG) return PP_ERR_ARG;
        g->deg[s]++;
      }
      g->nb_end[s][k] = (int8_t)e;
      g->nb_link[s][k] = (int8_t)l;
      g->nb_fwd[s][k] = (int8_t)!dir;
    }
  }
  return PP_OK;
}

struct Cand {      // (score, index) with the tie rule of the connection method
  float s;
  int i;
};

// blend: np.argsort ascending, take the last two => ties prefer the LARGER index
// max:   np.argmax => ties prefer the SMALLER index
__device__ __forceinline__ bool cand_better(Cand a, Cand b, int method) {
  if (a.s > b.s) return true;
  if (a.s < b.s) return false;
  if (a.i < 0 || b.i < 0) return a.i >= 0 && b.i < 0;
  return method == 1 ? a.i < b.i : a.i > b.i;
}

struct Xysv {
  float x, y, s, v;
};

constexpr int kScratch = 256;     // in-box candidate positions per warp and pass
constexpr int kDirectScan = 192;  // lists up to this long are scanned whole (no bin lookup)

struct Cand2 {     // a candidate with its column in the binned store
  float s;
  int i;           // original list index (tie rule)
  int pos;
};

// _grow_connection (cifcaf.py:124-145), warp-collective; every lane returns the
// result.  `base` points at the score row of one directed candidate list in the
// binned store (rows score, source x/y, target x/y/scale, original index;
// `stride` floats apart; K columns sorted by spatial bin, bin b occupying
// columns [start[b], start[b+1])) -- either the shared-memory copy or global.
//
// Three steps keep the serial chain of the one active warp short: (0) only the
// bin rows that intersect the caf_center_s box are visited; (1) a tight loop
// tests the box exactly (functional.pyx:346-353) and compacts the columns of
// the candidates inside it into `scratch`; (2) each lane scores ONE of them
// (sqrt, divide, exp), so the expensive chain runs once, not once per 32
// candidates.  Not inlined: the forward and the reverse match share one copy.
__device__ __noinline__ Xysv warp_grow_connection(float x, float y, float s, const float *base,
                                                  int stride, int K, const int *start, BinGeo geo,
                                                  int *scratch, int method, int exp_mode) {
  const int lane = lane_id();
  const unsigned full = 0xffffffffu;
  const float box = __fmul_rn(2.0f, s);                           // sigma = 2.0 * xy_scale  :129
  const float lox = __fsub_rn(x, box), hix = __fadd_rn(x, box);
  const float loy = __fsub_rn(y, box), hiy = __fadd_rn(y, box);
  const float sigma = __fmul_rn(0.5f, s);                         // :138
  const float sig2 = __fmul_rn(sigma, sigma);
  const float *cs = base, *xs = base + stride, *ys = base + 2 * (size_t)stride;
  const int *oi = reinterpret_cast<const int *>(base + 6 * (size_t)stride);
  Xysv out = {0.f, 0.f, 0.f, 0.f};
  if (K == 0) return out;

  // step 0: column ranges to visit, one per lane (lane 31 = the NaN bin)
  int r_lo = 0, r_hi = 0, n_ranges = 1;
  const bool nanbox = lox != lox || hix != hix || loy != loy || hiy != hiy;
  bool whole = K <= kDirectScan || nanbox;
  if (!whole) {
    const float inv = 1.0f / (float)(1 << geo.bshift);
    const int bx0 = (int)fminf(fmaxf(__fmul_rn(lox, inv), 0.0f), (float)(geo.nbx - 1));
    const int bx1 = (int)fminf(fmaxf(__fmul_rn(hix, inv), 0.0f), (float)(geo.nbx - 1));
    const int by0 = (int)fminf(fmaxf(__fmul_rn(loy, inv), 0.0f), (float)(geo.nby - 1));
    const int by1 = (int)fminf(fmaxf(__fmul_rn(hiy, inv), 0.0f), (float)(geo.nby - 1));
    n_ranges = by1 - by0 + 2;
    if (n_ranges > 32) {
      whole = true;
    } else {
      if (lane < n_ranges - 1) {
        r_lo = start[(by0 + lane) * geo.nbx + bx0];
        r_hi = start[(by0 + lane) * geo.nbx + bx1 + 1];
      } else if (lane == n_ranges - 1) {
        r_lo = start[geo.nb - 1];
        r_hi = K;
      }
    }
  }
  if (whole) { n_ranges = 1; r_lo = 0; r_hi = K; }

  Cand2 t1 = {-INFINITY, -1, 0}, t2 = {-INFINITY, -1, 0};
  int cnt = 0;
  int rr = 0;
  int seg = __shfl_sync(full, r_lo, 0), seg_end = __shfl_sync(full, r_hi, 0);
  while (rr < n_ranges) {
    // step 1: exact box test, compaction of the columns inside it
    int n = 0;
    while (rr < n_ranges && n <= kScratch - 128) {
      if (seg >= seg_end) {
        if (++rr < n_ranges) { seg = __shfl_sync(full, r_lo, rr); seg_end = __shfl_sync(full, r_hi, rr); }
        continue;
      }
      bool in[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int i = seg + 32 * j + lane;
        const float cx = i < seg_end ? xs[i] : 0.f, cy = i < seg_end ? ys[i] : 0.f;
        in[j] = i < seg_end && !(cx < lox || cx > hix || cy < loy || cy > hiy);
      }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const unsigned m = __ballot_sync(full, in[j]);
        if (in[j]) scratch[n + __popc(m & ((1u << lane) - 1u))] = seg + 32 * j + lane;
        n += __popc(m);
      }
      seg += 128;
    }
    __syncwarp();
    cnt += n;
    // step 2: score the candidates inside the box
    for (int t = lane; t < n; t += 32) {
      const int pos = scratch[t];
      const float dx = __fsub_rn(x, xs[pos]), dy = __fsub_rn(y, ys[pos]);
      const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));   // :134
      const float arg = __fdiv_rn(__fmul_rn(-0.5f, __fmul_rn(d, d)), sig2);          // :139
      const float e = exp_mode ? pp_expf(arg) : expf(arg);
      const Cand2 c = {__fmul_rn(e, cs[pos]), oi[pos], pos};
      const Cand ck = {c.s, c.i}, k1 = {t1.s, t1.i}, k2 = {t2.s, t2.i};
      if (cand_better(ck, k1, method)) { t2 = t1; t1 = c; }
      else if (cand_better(ck, k2, method)) { t2 = c; }
    }
    __syncwarp();
  }
  if (cnt == 0) return out;                                       // :130-131
  // warp-wide top-2.  Scores are products of exp() and a confidence above the
  // threshold, i.e. non-negative and ordered like their bit patterns, so two
  // integer warp reductions per rank do it; anything else (NaN from a zero
  // scale, negative thresholds) takes the generic butterfly below.
  const unsigned b1 = __float_as_uint(t1.s), b2 = __float_as_uint(t2.s);
  const bool plain = (t1.i < 0 || b1 <= 0x7f800000u) && (t2.i < 0 || b2 <= 0x7f800000u);
  if (__all_sync(full, plain)) {
    const unsigned k1 = t1.i >= 0 ? b1 + 1u : 0u;
    const unsigned m1 = __reduce_max_sync(full, k1);
    // ties: blend keeps the larger index (argsort), max the smaller (argmax)
    const int w1 = method == 1 ? __reduce_min_sync(full, k1 == m1 ? t1.i : 0x7fffffff)
                               : __reduce_max_sync(full, k1 == m1 ? t1.i : -1);
    const bool mine = t1.i == w1;
    const int p1 = __shfl_sync(full, t1.pos, __ffs(__ballot_sync(full, mine)) - 1);
    const Cand2 o = mine ? t2 : t1;
    const unsigned k2 = o.i >= 0 ? __float_as_uint(o.s) + 1u : 0u;
    const unsigned m2 = __reduce_max_sync(full, k2);
    int w2 = -1, p2 = 0;
    if (m2 != 0u) {
      w2 = method == 1 ? __reduce_min_sync(full, k2 == m2 ? o.i : 0x7fffffff)
                       : __reduce_max_sync(full, k2 == m2 ? o.i : -1);
      p2 = __shfl_sync(full, o.pos, __ffs(__ballot_sync(full, o.i == w2)) - 1);
    }
    t1.s = __uint_as_float(m1 - 1u); t1.i = w1; t1.pos = p1;
    t2.s = m2 ? __uint_as_float(m2 - 1u) : -INFINITY; t2.i = w2; t2.pos = p2;
  } else {
#pragma unroll 1
    for (int off = 16; off > 0; off >>= 1) {
      Cand2 o1 = {__shfl_xor_sync(full, t1.s, off), __shfl_xor_sync(full, t1.i, off),
                  __shfl_xor_sync(full, t1.pos, off)};
      Cand2 o2 = {__shfl_xor_sync(full, t2.s, off), __shfl_xor_sync(full, t2.i, off),
                  __shfl_xor_sync(full, t2.pos, off)};
      const Cand a1 = {t1.s, t1.i}, a2 = {t2.s, t2.i}, c1 = {o1.s, o1.i}, c2 = {o2.s, o2.i};
      Cand2 n1, n2;
      if (cand_better(a1, c1, method)) { n1 = t1; n2 = cand_better(a2, c1, method) ? t2 : o1; }
      else { n1 = o1; n2 = cand_better(a1, c2, method) ? t1 : o2; }
      t1 = n1; t2 = n2;
    }
  }
  const float *xt = base + 3 * (size_t)stride, *yt = base + 4 * (size_t)stride;
  const float *st = base + 5 * (size_t)stride;
  float x1 = xt[t1.pos], y1 = yt[t1.pos], s1 = st[t1.pos];
  if (method == 1) {                                              // _target_with_maxscore :147-155
    out.x = x1; out.y = y1; out.s = s1; out.v = t1.s;
    return out;
  }
  out.x = x1; out.y = y1; out.s = s1; out.v = __fmul_rn(t1.s, 0.5f);
  if (cnt == 1) return out;                                       // :164-170
  if (t2.s < 0.01f || t2.s < __fmul_rn(0.5f, t1.s)) return out;   // :178-179
  float x2 = xt[t2.pos], y2 = yt[t2.pos], s2 = st[t2.pos];
  float ddx = __fsub_rn(x1, x2), ddy = __fsub_rn(y1, y2);
  float d = __fsqrt_rn(__fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy)));   // :182
  if (d > __fdiv_rn(s1, 2.0f)) return out;                        // :183-184
  float den = __fadd_rn(t1.s, t2.s);
  out.x = __fdiv_rn(__fadd_rn(__fmul_rn(t1.s, x1), __fmul_rn(t2.s, x2)), den);   // :186-191
  out.y = __fdiv_rn(__fadd_rn(__fmul_rn(t1.s, y1), __fmul_rn(t2.s, y2)), den);
  out.s = __fdiv_rn(__fadd_rn(__fmul_rn(t1.s, s1), __fmul_rn(t2.s, s2)), den);
  out.v = __fmul_rn(0.5f, den);
  return out;
}

struct CafLists {
  const float *rows;        // (L, 2, kBinRows, cap) binned store of this image
  const int *start;         // (L, 2, nb + 1)
  BinGeo geo;
  int cap;
  const int *n;             // (2L) list lengths in shared memory: [2l] forward, [2l+1] backward
  const float *stage;       // shared-memory copy of the rows of every list (packed), or nullptr
  const int *stage_off;     // (2L) float offset of each list inside `stage`
};

// grow_connection on the directed list (l, fw)
__device__ __forceinline__ Xysv scan_list(const CafLists &cf, int l, bool fw, float x, float y, float s,
                                          int *scratch, int method, int exp_mode) {
  const int i = 2 * l + (fw ? 0 : 1);
  const int K = cf.n[i];
  const int *start = cf.start + (size_t)i * (cf.geo.nb + 1);
  if (cf.stage) return warp_grow_connection(x, y, s, cf.stage + cf.stage_off[i], K, K, start, cf.geo,
                                            scratch, method, exp_mode);
  return warp_grow_connection(x, y, s, cf.rows + (size_t)i * kBinRows * cf.cap, cf.cap, K, start,
                              cf.geo, scratch, method, exp_mode);
}

// connection_value (cifcaf.py:194-217), warp-collective.
__device__ Xysv warp_connection_value(const float (*data)[3], const float *js, const Graph &g,
                                      const CafLists &cf, int start, int k, bool reverse_match,
                                      float keypoint_th, int method, int exp_mode, int *scratch) {
  Xysv zero = {0.f, 0.f, 0.f, 0.f};
  const int l = g.nb_link[start][k];
  const bool fw = g.nb_fwd[start][k] != 0;
  const float sx = data[start][0], sy = data[start][1], sv = data[start][2];
  const float ss = js[start] > 0.0f ? js[start] : 0.0f;           // max(0.0, .)
  Xysv nw = scan_list(cf, l, fw, sx, sy, ss, scratch, method, exp_mode);
  float kp = __fsqrt_rn(__fmul_rn(nw.v, sv));                     // :201
  if (kp < keypoint_th) return zero;
  if (nw.v == 0.0f) return zero;
  float st = nw.s > 0.0f ? nw.s : 0.0f;
  if (reverse_match) {                                            // :208-215
    Xysv rv = scan_list(cf, l, !fw, nw.x, nw.y, st, scratch, method, exp_mode);
    if (rv.s == 0.0f) return zero;
    float dist = __fadd_rn(fabsf(__fsub_rn(sx, rv.x)), fabsf(__fsub_rn(sy, rv.y)));
    if (dist > ss) return zero;
  }
  nw.v = kp;
  return nw;
}

// Phase timing for tuning (build with PP_EXTRA_NVCC=-DPP_ASM_TIMING): thread 0
// accumulates clock64 deltas per phase into g_asm_cycles; read them with
// pp_debug_asm_cycles.  Compiled out otherwise.
#ifdef PP_ASM_TIMING
__device__ unsigned long long g_asm_cycles[8];
#define PP_TICK(slot)                                                              \
  do {                                                                             \
    if (threadIdx.x == 0) {                                                        \
      long long now_ = clock64();                                                  \
      atomicAdd(&g_asm_cycles[slot], (unsigned long long)(now_ - sh.tick));        \
      sh.tick = now_;                                                              \
    }                                                                              \
  } while (0)
#else
#define PP_TICK(slot) do { } while (0)
#endif

struct AsmShared {
#ifdef PP_ASM_TIMING
  long long tick;
#endif
  float data[PP_MAX_F][3];
  float js[PP_MAX_F];
  // evaluated frontier entries (-score, (x, y, s, v), start, end), one array per member
  float p_key[kMaxEdges], p_x[kMaxEdges], p_y[kMaxEdges], p_s[kMaxEdges], p_v[kMaxEdges];
  int p_start[kMaxEdges], p_end[kMaxEdges];
  int n_pend;
  uint32_t queued[PP_MAX_F];
  int new_start[kMaxEdges], new_k[kMaxEdges];
  int n_new;
  int dec[PP_MAX_F][2];
  int n_dec;
  int best;
  long long box[PP_MAX_F][4];
  int box_valid[PP_MAX_F];
  int found;
  int warp_min[kAsmWarps];
  struct Flood { float key, sx, sy, sv, s; int end; } ff[4 * PP_MAX_L];
  // heavy mode (more than kHeavySeeds seeds): seeds are indexed by (field, band of occupancy rows)
  int occ_bin[PP_MAX_F * kOccBands];   // end offset of bin (field, band) in the field-major seed order
  int occ_all[PP_MAX_F];               // Occupancy.set boxes that cover the whole map, per field
  int list_n[2 * PP_MAX_L];      // candidate-list lengths, [2l] forward / [2l+1] backward
  int list_off[2 * PP_MAX_L];    // float offsets into the staged copy
  int stage_total;               // entries over all lists
  int scratch[kAsmWarps][kScratch];
};

// Heap order of the reference's frontier.  Non-greedy: evaluated entries
// (-score, (x, y, s, v), start, end).  Greedy (cifcaf.py:280-281): an entry is
// applied the moment it is evaluated, so only the unevaluated keys
// (-sqrt(v_start) [* confidence_scale], None, start, end) are ever compared;
// `key` then holds that bound.
__device__ __forceinline__ bool pending_less(const AsmShared &sh, int a, int b, int greedy) {
  if (sh.p_key[a] != sh.p_key[b]) return sh.p_key[a] < sh.p_key[b];
  if (!greedy) {
    if (sh.p_x[a] != sh.p_x[b]) return sh.p_x[a] < sh.p_x[b];
    if (sh.p_y[a] != sh.p_y[b]) return sh.p_y[a] < sh.p_y[b];
    if (sh.p_s[a] != sh.p_s[b]) return sh.p_s[a] < sh.p_s[b];
    if (sh.p_v[a] != sh.p_v[b]) return sh.p_v[a] < sh.p_v[b];
  }
  if (sh.p_start[a] != sh.p_start[b]) return sh.p_start[a] < sh.p_start[b];
  return sh.p_end[a] < sh.p_end[b];
}

// ascending float order as ascending unsigned order (NaN excluded by the caller)
__device__ __forceinline__ uint32_t float_orderable(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Load the list lengths of one image into shared memory and, when everything
// fits into `stage_floats` of dynamic shared memory, copy the rows of all 2L
// binned lists there with cp.async (all copies in flight at once).
// CTA-collective.  Returns the CafLists the growth code reads through.
__device__ CafLists load_lists(AsmShared &sh, const float *rows, const int *start, BinGeo geo,
                               const int *nfwd, const int *nbwd, int L, int cap, float *stage,
                               int stage_floats) {
  const int tid = threadIdx.x;
  for (int i = tid; i < 2 * L; i += blockDim.x)
    sh.list_n[i] = min((i & 1) ? nbwd[i >> 1] : nfwd[i >> 1], cap);
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int i = 0; i < 2 * L; i++) { sh.list_off[i] = kBinRows * run; run += sh.list_n[i]; }
    sh.stage_total = run;
  }
  __syncthreads();
  CafLists cf = {rows, start, geo, cap, sh.list_n, nullptr, sh.list_off};
  if (stage == nullptr || kBinRows * sh.stage_total > stage_floats) return cf;
  const int warp = warp_id(), lane = lane_id(), nw = blockDim.x >> 5;
  for (int pr = warp; pr < 2 * kBinRows * L; pr += nw) {      // one warp per (list, row)
    const int i = pr / kBinRows, r = pr - kBinRows * i;
    const int K = sh.list_n[i];
    const float *src = rows + ((size_t)i * kBinRows + r) * cap;
    float *dst = stage + sh.list_off[i] + r * K;
    for (int k = lane; k < K; k += 32) __pipeline_memcpy_async(dst + k, src + k, sizeof(float));
  }
  __pipeline_commit();
  __pipeline_wait_prior(0);
  __syncthreads();
  cf.stage = stage;
  return cf;
}

// add_to_frontier (cifcaf.py:251-263), run by one thread.
__device__ void queue_edges(AsmShared &sh, const Graph &g, int start) {
  for (int k = 0; k < g.deg[start]; k++) {
    int end = g.nb_end[start][k];
    if (sh.data[end][2] > 0.0f) continue;
    if (sh.queued[start] & (1u << end)) continue;
    sh.queued[start] |= 1u << end;
    sh.new_start[sh.n_new] = start;
    sh.new_k[sh.n_new] = k;
    sh.n_new++;
  }
}

// the same for one start joint, by a whole warp (lane = neighbour slot); sets n_new
__device__ __forceinline__ void queue_edges_warp(AsmShared &sh, const Graph &g, int start) {
  const int lane = lane_id();
  bool ok = false;
  int end = 0;
  if (lane < g.deg[start]) {
    end = g.nb_end[start][lane];
    ok = !(sh.data[end][2] > 0.0f) && !(sh.queued[start] & (1u << end));
  }
  const unsigned m = __ballot_sync(0xffffffffu, ok);
  const unsigned bits = __reduce_or_sync(0xffffffffu, ok ? (1u << end) : 0u);
  if (ok) {
    const int pos = __popc(m & ((1u << lane) - 1u));
    sh.new_start[pos] = start;
    sh.new_k[pos] = lane;
  }
  if (lane == 0) {
    sh.queued[start] |= bits;
    sh.n_new = __popc(m);
  }
}

// _grow (cifcaf.py:247-307) in its eager form; CTA-collective.
__device__ void grow(AsmShared &sh, const Graph &g, const CScale &cscale, const CafLists &cf,
                     bool reverse_match, const pp_params &p) {
  const int warp = warp_id(), lane = lane_id();
  const unsigned full = 0xffffffffu;
  if (threadIdx.x == 0) {
    sh.n_pend = 0;
    sh.n_new = 0;
    for (int j = 0; j < g.F; j++) sh.queued[j] = 0u;
    for (int j = 0; j < g.F; j++) {                               // :288-291
      if (sh.data[j][2] == 0.0f) continue;
      queue_edges(sh, g, j);
    }
  }
  __syncthreads();
  PP_TICK(7);
  for (;;) {
    // evaluate the newly queued edges, one warp each
    const int n_new = sh.n_new, base = sh.n_pend;
    for (int e = warp; e < n_new; e += kAsmWarps) {
      int start = sh.new_start[e], k = sh.new_k[e];
      Xysv r = warp_connection_value(sh.data, sh.js, g, cf, start, k, reverse_match,
                                     p.keypoint_threshold, p.connection_method, p.exp_mode,
                                     sh.scratch[warp]);
      if (lane == 0) {
        const int q = base + e;
        float score = r.v;
        if (p.greedy) score = __fsqrt_rn(sh.data[start][2]);                       // :258 (the bound)
        if (cscale.has) score = __fmul_rn(score, cscale.v[g.nb_link[start][k]]);   // :259-260, :282-284
        sh.p_key[q] = -score; sh.p_x[q] = r.x; sh.p_y[q] = r.y; sh.p_s[q] = r.s; sh.p_v[q] = r.v;
        sh.p_start[q] = start;
        sh.p_end[q] = (r.v == 0.0f) ? -1 : g.nb_end[start][k];                     // :277-278
      }
    }
    __syncthreads();
    PP_TICK(3);
    if (warp == 0) {
      // best evaluated entry whose target is still empty (frontier_get + :297-298)
      const int n_pend = base + n_new;
      int best = -1;
      uint32_t bu = 0xffffffffu;
      bool odd = false;                      // NaN key: leave the integer fast path
      for (int i = lane; i < n_pend; i += 32) {
        const int end = sh.p_end[i];
        if (end < 0 || sh.data[end][2] > 0.0f) continue;
        const float key = sh.p_key[i];
        if (key != key) odd = true;
        const uint32_t u = float_orderable(key);
        if (best < 0 || u < bu || (u == bu && pending_less(sh, i, best, p.greedy))) { best = i; bu = u; }
      }
      const unsigned have = __ballot_sync(full, best >= 0);
      if (have == 0u) {
        best = -1;
      } else if (!__any_sync(full, odd)) {
        // primary key by one integer reduction; full tuple order only among exact ties
        const uint32_t m = __reduce_min_sync(full, best >= 0 ? bu : 0xffffffffu);
        unsigned tie = __ballot_sync(full, best >= 0 && bu == m);
        int w = __shfl_sync(full, best, __ffs(tie) - 1);
        for (tie &= tie - 1u; tie; tie &= tie - 1u) {
          const int o = __shfl_sync(full, best, __ffs(tie) - 1);
          if (pending_less(sh, o, w, p.greedy)) w = o;
        }
        best = w;
      } else {
#pragma unroll 1
        for (int off = 16; off > 0; off >>= 1) {
          int o = __shfl_xor_sync(full, best, off);
          if (o >= 0 && (best < 0 || pending_less(sh, o, best, p.greedy))) best = o;
        }
      }
      if (lane == 0) {
        sh.n_pend = n_pend;
        sh.n_new = 0;
        sh.best = best;
      }
      if (best >= 0) {
        const int end = sh.p_end[best];
        if (lane == 0) {
          sh.data[end][0] = sh.p_x[best]; sh.data[end][1] = sh.p_y[best];          // :300-302
          sh.data[end][2] = sh.p_v[best];
          sh.js[end] = sh.p_s[best];
          sh.dec[sh.n_dec][0] = sh.p_start[best]; sh.dec[sh.n_dec][1] = end;       // :303-304
          sh.n_dec++;
          sh.p_end[best] = -1;
        }
        __syncwarp();
        queue_edges_warp(sh, g, end);                                              // :305
      }
    }
    __syncthreads();
    PP_TICK(4);
    if (sh.best < 0) break;
  }
}

__global__ void __launch_bounds__(kAsmThreads)
assemble_kernel(AsmArgs a, Graph g, CScale cscale, pp_params p) {
  __shared__ AsmShared sh;
  const int b = blockIdx.x;
  const int F = g.F;
  const float *seeds = a.seeds + (size_t)b * a.seed_cap * 5;
  const int n_seeds = min(a.n_seeds[b], a.seed_cap);
  uint8_t *occ = a.occ_cnt + (size_t)b * a.seed_cap;
  extern __shared__ __align__(16) float stage_buf[];
#ifdef PP_ASM_TIMING
  if (threadIdx.x == 0) sh.tick = clock64();
#endif
  const CafLists cf = load_lists(sh, a.rows + (size_t)b * a.L * 2 * kBinRows * a.caf_cap,
                                 a.start + (size_t)b * a.L * 2 * (a.geo.nb + 1), a.geo,
                                 a.nfwd + (size_t)b * a.L, a.nbwd + (size_t)b * a.L, a.L, a.caf_cap,
                                 stage_buf, a.stage_floats);
  // Occupancy(cifhr.accumulated.shape, 2, min_scale=4): cifcaf.py:84, occupancy.py:21-25
  const long long oh = (long long)((double)a.hh / 2.0), ow = (long long)((double)a.hw / 2.0);
  float *pre_data = a.pre_data + (size_t)b * a.ann_cap * F * 3;
  float *pre_scales = a.pre_scales + (size_t)b * a.ann_cap * F;
  int *pre_order = a.pre_order ? a.pre_order + (size_t)b * a.ann_cap * F * 2 : nullptr;

  // every seed's Occupancy.get cell is fixed: compute it once (occupancy.py:41-47)
  int2 *rec = a.occ_rec + (size_t)b * a.seed_cap;
  const bool heavy = n_seeds > a.heavy_seeds && n_seeds < (1 << 24) && ow < 65536 && oh < 32768;
  int bshift = 0;
  while (((oh - 1) >> bshift) >= kOccBands) bshift++;
  if (!heavy) {
    for (int i = threadIdx.x; i < n_seeds; i += blockDim.x) {
      const float *sd = seeds + (size_t)i * 5;
      long long cx, cy;
      occupancy_cell(__ldg(sd + 2), __ldg(sd + 3), ow, oh, &cx, &cy);
      rec[i] = make_int2((int)cx, (int)cy | ((int)__ldg(sd + 1) << 24));
      occ[i] = 0;
    }
  } else {
    // counting sort by (field, band): rec[i].x = position | field << 24 of seed i (score order),
    // rec[k].y = cell x | cell y << 16 and occ[k] = counter of the seed at position k
    for (int i = threadIdx.x; i < PP_MAX_F * kOccBands; i += blockDim.x) sh.occ_bin[i] = 0;
    if (threadIdx.x < PP_MAX_F) sh.occ_all[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n_seeds; i += blockDim.x) {
      const float *sd = seeds + (size_t)i * 5;
      long long cx, cy;
      occupancy_cell(__ldg(sd + 2), __ldg(sd + 3), ow, oh, &cx, &cy);
      atomicAdd(&sh.occ_bin[(int)__ldg(sd + 1) * kOccBands + (int)(cy >> bshift)], 1);
    }
    __syncthreads();
    if (warp_id() == 0) {                       // exclusive scan of the PP_MAX_F * kOccBands counts
      constexpr int kPer = PP_MAX_F * kOccBands / 32;
      const int first = lane_id() * kPer;
      int sum = 0;
      for (int q = 0; q < kPer; q++) sum += sh.occ_bin[first + q];
      int incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane_id() >= d) incl += v;
      }
      int run = incl - sum;
      for (int q = 0; q < kPer; q++) {
        const int cnt = sh.occ_bin[first + q];
        sh.occ_bin[first + q] = run;
        run += cnt;
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n_seeds; i += blockDim.x) {
      const float *sd = seeds + (size_t)i * 5;
      long long cx, cy;
      occupancy_cell(__ldg(sd + 2), __ldg(sd + 3), ow, oh, &cx, &cy);
      const int f = (int)__ldg(sd + 1);
      const int pos = atomicAdd(&sh.occ_bin[f * kOccBands + (int)(cy >> bshift)], 1);   // start -> end
      rec[i].x = pos | (f << 24);
      rec[pos].y = (int)cx | ((int)cy << 16);
      occ[pos] = 0;
    }
  }
  __syncthreads();

  // initial_annotations (cifcaf.py:95-98) are grown first, then the seeds
  const int n_init = a.init_n ? min(a.init_n[b], a.init_cap) : 0;
  int n_ann = 0, cursor = 0, ia = 0;
  PP_TICK(0);
  for (;;) {
    const bool from_init = ia < n_init;
    // ---- next seed whose cell is free (cifcaf.py:100-102) -------------------
    int found = -1;
    for (int base = cursor; !from_init && base < n_seeds; base += blockDim.x) {
      int i = base + threadIdx.x;
      bool hit = false;
      if (i < n_seeds) {
        if (!heavy) {
          hit = occ[i] == 0;
        } else {
          const int r = rec[i].x;
          hit = ((occ[r & 0xffffff] + sh.occ_all[r >> 24]) & 255) == 0;
        }
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      if (lane_id() == 0) sh.warp_min[warp_id()] = m ? (base + warp_id() * 32 + __ffs(m) - 1) : -1;
      __syncthreads();
      for (int w = 0; w < kAsmWarps; w++)
        if (sh.warp_min[w] >= 0) { found = sh.warp_min[w]; break; }
      __syncthreads();
      if (found >= 0) break;
    }
    PP_TICK(1);
    if (!from_init) {
      if (found < 0) break;
      cursor = found + 1;
    }
    if (n_ann >= a.ann_cap) {
      if (threadIdx.x == 0) atomicOr(&a.status[b], PP_STATUS_ANN_OVERFLOW);
      break;
    }
    // ---- new annotation from the seed (:104-105), or the next initial annotation ----
    if (from_init) {
      const float *idata = a.init_data + ((size_t)b * a.init_cap + ia) * F * 3;
      const float *iscales = a.init_scales + ((size_t)b * a.init_cap + ia) * F;
      for (int i = threadIdx.x; i < F * 3; i += blockDim.x) (&sh.data[0][0])[i] = idata[i];
      for (int i = threadIdx.x; i < F; i += blockDim.x) sh.js[i] = iscales[i];
      if (threadIdx.x == 0) sh.n_dec = 0;
      ia++;
    } else {
      for (int i = threadIdx.x; i < F * 3; i += blockDim.x) (&sh.data[0][0])[i] = 0.0f;
      for (int i = threadIdx.x; i < F; i += blockDim.x) sh.js[i] = 0.0f;
    }
    __syncthreads();
    if (!from_init && threadIdx.x == 0) {
      const float *sd = seeds + (size_t)found * 5;
      int f = (int)sd[1];
      sh.data[f][0] = sd[2]; sh.data[f][1] = sd[3]; sh.data[f][2] = sd[0];
      sh.js[f] = sd[4];
      sh.n_dec = 0;
    }
    __syncthreads();
    PP_TICK(2);
    grow(sh, g, cscale, cf, true, p);                             // :106
    // ---- store + mark_occupied (:87-93, :107-108) -------------------------
    for (int i = threadIdx.x; i < F * 3; i += blockDim.x)
      pre_data[(size_t)n_ann * F * 3 + i] = (&sh.data[0][0])[i];
    for (int i = threadIdx.x; i < F; i += blockDim.x) {
      pre_scales[(size_t)n_ann * F + i] = sh.js[i];
      if (pre_order) {
        pre_order[((size_t)n_ann * F + i) * 2] = i < sh.n_dec ? sh.dec[i][0] : -1;
        pre_order[((size_t)n_ann * F + i) * 2 + 1] = i < sh.n_dec ? sh.dec[i][1] : -1;
      }
      sh.box_valid[i] = sh.data[i][2] != 0.0f;
      if (sh.box_valid[i]) occupancy_box(sh.data[i][0], sh.data[i][1], sh.js[i], ow, oh, sh.box[i]);
    }
    n_ann++;
    __syncthreads();
    PP_TICK(5);
    if (heavy) {
      // one warp per joint: a box that covers the whole map bumps the field's counter, any other box
      // visits the seeds of its field in the row bands it overlaps
      for (int j = warp_id(); j < F; j += kAsmWarps) {
        if (!sh.box_valid[j]) continue;
        const long long minx = sh.box[j][0], maxx = sh.box[j][1], miny = sh.box[j][2], maxy = sh.box[j][3];
        if (minx >= maxx || miny >= maxy) continue;        // box outside the map: empty numpy slice
        if (minx <= 0 && maxx >= ow && miny <= 0 && maxy >= oh) {
          if (lane_id() == 0) sh.occ_all[j]++;
          continue;
        }
        const int b0 = j * kOccBands + (int)(miny >> bshift), b1 = j * kOccBands + (int)((maxy - 1) >> bshift);
        const int k0 = b0 == 0 ? 0 : sh.occ_bin[b0 - 1], k1 = sh.occ_bin[b1];
        for (int kk = k0 + lane_id(); kk < k1; kk += 4 * 32) {
          int c[4];
#pragma unroll
          for (int q = 0; q < 4; q++) c[q] = kk + q * 32 < k1 ? rec[kk + q * 32].y : -1;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (kk + q * 32 >= k1) continue;
            const long long cx = c[q] & 0xffff, cy = (unsigned)c[q] >> 16;
            if (cx >= minx && cx < maxx && cy >= miny && cy < maxy) occ[kk + q * 32]++;   // uint8 wraps
          }
        }
      }
      __syncthreads();
      PP_TICK(6);
      continue;
    }
    // Occupancy.set of the pose's joints, seen from the seeds: four seeds per thread and
    // trip so that four loads are in flight
    for (int i0 = cursor + threadIdx.x; i0 < n_seeds; i0 += 4 * blockDim.x) {
      int2 r[4];
      uint8_t c[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int i = i0 + j * blockDim.x;
        r[j] = i < n_seeds ? rec[i] : make_int2(0, 0);
        c[j] = i < n_seeds ? occ[i] : 0;
      }
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int i = i0 + j * blockDim.x;
        if (i >= n_seeds) continue;
        const int f = r[j].y >> 24;
        if (!sh.box_valid[f]) continue;
        const long long cx = r[j].x, cy = r[j].y & 0xffffff;
        if (cx >= sh.box[f][0] && cx < sh.box[f][1] && cy >= sh.box[f][2] && cy < sh.box[f][3])
          occ[i] = (uint8_t)(c[j] + 1);                           // uint8 += 1 wraps (utils.py:66)
      }
    }
    __syncthreads();
    PP_TICK(6);
  }
  if (threadIdx.x == 0) a.n_pre[b] = n_ann;
}

// ---------------------------------------------------------------------------
// complete_annotations (cifcaf.py:333-351) + _flood_fill (:309-331).
// Annotations are independent here, so the grid is (annotation, image).

__device__ __forceinline__ bool flood_less(const AsmShared::Flood &a, const AsmShared::Flood &b) {
  // Python tuple order of (-v, end_i, [x, y, v], s)
  if (a.key != b.key) return a.key < b.key;
  if (a.end != b.end) return a.end < b.end;
  if (a.sx != b.sx) return a.sx < b.sx;
  if (a.sy != b.sy) return a.sy < b.sy;
  if (a.sv != b.sv) return a.sv < b.sv;
  return a.s < b.s;
}

// one thread; `prio` is the closure quirk of cifcaf.py:317,325 (see oracle flood_fill)
__device__ void flood_add(AsmShared &sh, const Graph &g, int &n, int start, float prio) {
  for (int k = 0; k < g.deg[start]; k++) {
    int end = g.nb_end[start][k];
    if (sh.data[end][2] > 0.0f) continue;
    if (n >= 4 * PP_MAX_L) return;     // cannot happen: <= 2 pushes per directed edge
    AsmShared::Flood &e = sh.ff[n++];
    e.key = -prio; e.end = end;
    e.sx = sh.data[start][0]; e.sy = sh.data[start][1]; e.sv = sh.data[start][2];
    e.s = sh.js[start];
  }
}

__device__ void flood_fill(AsmShared &sh, const Graph &g) {
  int n = 0;
  for (int j = 0; j < g.F; j++) {
    if (sh.data[j][2] == 0.0f) continue;
    flood_add(sh, g, n, j, sh.data[j][2]);
  }
  while (n > 0) {
    int m = 0;
    for (int i = 1; i < n; i++)
      if (flood_less(sh.ff[i], sh.ff[m])) m = i;
    AsmShared::Flood e = sh.ff[m];
    sh.ff[m] = sh.ff[--n];
    if (sh.data[e.end][2] > 0.0f) continue;
    sh.data[e.end][0] = e.sx; sh.data[e.end][1] = e.sy;
    sh.data[e.end][2] = 0.00001f;
    sh.js[e.end] = e.s;
    flood_add(sh, g, n, e.end, e.sv);
  }
}

__global__ void __launch_bounds__(kAsmThreads)
complete_kernel(AsmArgs a, const float *__restrict__ rows2, const int *__restrict__ start2,
                const int *__restrict__ nfwd2, const int *__restrict__ nbwd2,
                float *__restrict__ cmp_data, float *__restrict__ cmp_scales, Graph g, CScale cscale,
                pp_params p) {
  __shared__ AsmShared sh;
  __shared__ int unfilled[PP_MAX_F];
  const int b = blockIdx.y, an = blockIdx.x;
  const int F = g.F;
  if (an >= min(a.n_pre[b], a.ann_cap)) return;
  // the score_th = 1e-4 lists hold nearly every cell: read them in place
  const CafLists cf = load_lists(sh, rows2 + (size_t)b * a.L * 2 * kBinRows * a.caf_cap,
                                 start2 + (size_t)b * a.L * 2 * (a.geo.nb + 1), a.geo,
                                 nfwd2 + (size_t)b * a.L, nbwd2 + (size_t)b * a.L, a.L, a.caf_cap,
                                 nullptr, 0);
  float *pre_data = a.pre_data + ((size_t)b * a.ann_cap + an) * F * 3;
  float *pre_scales = a.pre_scales + ((size_t)b * a.ann_cap + an) * F;
  int *pre_order = a.pre_order + ((size_t)b * a.ann_cap + an) * F * 2;
  for (int i = threadIdx.x; i < F * 3; i += blockDim.x) (&sh.data[0][0])[i] = pre_data[i];
  for (int i = threadIdx.x; i < F; i += blockDim.x) sh.js[i] = pre_scales[i];
  __syncthreads();
  if (threadIdx.x == 0) {
    int n1 = 0;
    while (n1 < F && pre_order[2 * n1] >= 0) {
      sh.dec[n1][0] = pre_order[2 * n1]; sh.dec[n1][1] = pre_order[2 * n1 + 1];
      n1++;
    }
    sh.n_dec = n1;
    sh.found = n1;     // decoding steps of the first _grow
  }
  for (int i = threadIdx.x; i < F; i += blockDim.x) unfilled[i] = sh.data[i][2] == 0.0f;   // :340
  __syncthreads();
  grow(sh, g, cscale, cf, false, p);                                // :341
  // as-grown state (scores not yet clamped) for the host's decoding_order replay
  for (int i = threadIdx.x; i < F * 3; i += blockDim.x) pre_data[i] = (&sh.data[0][0])[i];
  for (int i = threadIdx.x; i < F; i += blockDim.x) {
    pre_scales[i] = sh.js[i];
    int s0 = i < sh.n_dec ? sh.dec[i][0] : -1, s1 = i < sh.n_dec ? sh.dec[i][1] : -1;
    if (i == F - 1 && sh.n_dec < F) { s0 = -2; s1 = sh.found; }   // marker: steps of the first grow
    pre_order[2 * i] = s0; pre_order[2 * i + 1] = s1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    bool any_empty = false;
    for (int j = 0; j < F; j++) {
      if (unfilled[j] && sh.data[j][2] > 0.0f)                     // :342-344
        sh.data[j][2] = 0.001f < sh.data[j][2] ? 0.001f : sh.data[j][2];
      if (sh.data[j][2] == 0.0f) any_empty = true;
    }
    if (any_empty) flood_fill(sh, g);                              // :347-348
  }
  __syncthreads();
  float *od = cmp_data + ((size_t)b * a.ann_cap + an) * F * 3;
  float *os = cmp_scales + ((size_t)b * a.ann_cap + an) * F;
  for (int i = threadIdx.x; i < F * 3; i += blockDim.x) od[i] = (&sh.data[0][0])[i];
  for (int i = threadIdx.x; i < F; i += blockDim.x) os[i] = sh.js[i];
}

int complete_launch(const AsmArgs &args, const float *rows2, const int *start2, const int *nfwd2,
                    const int *nbwd2, float *cmp_data, float *cmp_scales, const pp_skeleton &sk,
                    int B, int F, const pp_params &p, cudaStream_t st) {
  Graph g;
  int rc = build_graph(sk, F, &g);
  if (rc) return rc;
  CScale cs;
  memset(&cs, 0, sizeof cs);
  cs.has = sk.has_confidence_scales;
  for (int l = 0; l < sk.L; l++) cs.v[l] = sk.confidence_scales[l];
  dim3 grid(args.ann_cap, B);
  complete_kernel<<<grid, kAsmThreads, 0, st>>>(args, rows2, start2, nfwd2, nbwd2, cmp_data, cmp_scales,
                                               g, cs, p);
  PP_LAUNCH_CHECK("complete_kernel");
  return PP_OK;
}

int assemble_launch(const AsmArgs &args, const pp_skeleton &sk, int B, int F, const pp_params &p,
                    cudaStream_t st) {
  Graph g;
  int rc = build_graph(sk, F, &g);
  if (rc) return rc;
  CScale cs;
  memset(&cs, 0, sizeof cs);
  cs.has = sk.has_confidence_scales;
  for (int l = 0; l < sk.L; l++) cs.v[l] = sk.confidence_scales[l];
  AsmArgs a2 = args;
  // one image per SM and nothing else on the GPU: give the lists (almost) the whole shared memory;
  // with other batches in flight (pp_params.in_flight) leave room for their CTAs on the same SM
  int stage_bytes = (B <= num_sms() && p.in_flight <= 1) ? kStageBytesSolo : kStageBytes;
  a2.heavy_seeds = kHeavySeeds;
  if (const char *e = getenv("PP_ASM_HEAVY_SEEDS")) a2.heavy_seeds = atoi(e);   // tests: force either mode
  if (const char *e = getenv("PP_ASM_STAGE_KB")) {     // tuning experiments only
    int kb = atoi(e);
    if (kb >= 0 && kb * 1024 <= kStageBytesSolo) stage_bytes = kb * 1024;
  }
  a2.stage_floats = stage_bytes / (int)sizeof(float);
  PP_CUDA_CHECK(cudaFuncSetAttribute(assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kStageBytesSolo));
  prof_mark(p.profile, kStAssemble, st);
  assemble_kernel<<<B, kAsmThreads, stage_bytes, st>>>(a2, g, cs, p);
  PP_LAUNCH_CHECK("assemble_kernel");
  return PP_OK;
}

}  // namespace pp

#ifdef PP_ASM_TIMING
extern "C" int pp_debug_asm_cycles(unsigned long long *out8, int reset) {
  if (cudaMemcpyFromSymbol(out8, pp::g_asm_cycles, 8 * sizeof(unsigned long long)) != cudaSuccess) return -3;
  if (reset) {
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(pp::g_asm_cycles, z, sizeof z);
  }
  return 0;
}
#endif
