// assemble.cu -- greedy pose assembly with occupancy, on sm_100a.
//
// Replaces, for a whole batch, the serial part of CifCaf.__call__
// (decoder/generator/cifcaf.py:84-108), CifCaf._grow / connection_value /
// _grow_connection / _target_with_blend / _target_with_maxscore (:124-307),
// caf_center_s (functional.pyx:338-359) and Occupancy (decoder/occupancy.py,
// decoder/utils.py:61-66, functional.pyx:281-286).
//
// One CTA per image, one WARP per pose:
//   * _grow(seed) is a pure function of the seed and the CafScored lists -- it never looks at the
//     occupancy map (cifcaf.py:247-307).  Only the DECISION which seeds start a pose is serial: a
//     seed is used iff no pose committed before it covers its cell.  So the kernel grows poses
//     speculatively, one warp each, side by side: when the seed loop reaches a free seed for which
//     no grown pose is at hand, warp 0 picks that seed plus the next free seeds that do not lie in
//     the Occupancy.set box an earlier pick is certain to set (its own joint), every warp grows one
//     of them, and the seed loop then continues exactly as the reference's: next free seed -> take
//     its grown pose if there is one (else start a new wave) -> store, mark occupied.  Speculation
//     changes the time, never the result: a grown pose whose seed gets covered first is dropped.
//   * inside a warp the frontier is evaluated eagerly: when a joint is set the warp scans the
//     candidate list of every outgoing edge (lanes over candidates), applies blend / max and the
//     reverse match, and keeps the scored entries; "take the best evaluated edge whose target is
//     still empty" then visits the same edges in the same order as the reference's lazy
//     PriorityQueue, whose unevaluated keys are upper bounds of the evaluated ones (SURVEY.md
//     appendix A4; validated by the tests).  Ties follow Python's tuple order
//     (-score, (x, y, s, v), start, end).  No CTA barrier inside a pose.
//   * the occupancy map is never rasterised: Occupancy.get(f, x, y) is the number (mod 256, the
//     reference's uint8 wrap) of Occupancy.set boxes of field f covering the cell, so each seed
//     carries a uint8 counter that is bumped, CTA-wide, against the <= F boxes of a committed pose;
//     "next free seed" is a CTA-wide ballot search;
//   * nms.Keypoints (nms.cuh) runs as the tail of the same CTA unless force_complete is on.
#include <cuda_pipeline.h>

#include "common.cuh"
#include "stages.cuh"
#include "occupancy.cuh"
#include "nms.cuh"

namespace pp {

#ifndef PP_ASM_WARPS
#define PP_ASM_WARPS 8
#endif
constexpr int kAsmWarps = PP_ASM_WARPS;       // poses grown side by side per image
constexpr int kAsmThreads = 32 * kAsmWarps;
// Occupancy bookkeeping.  Up to kHeavySeeds seeds every seed's counter is tested against the boxes of
// each finished pose (a few thousand tests).  Above that (crowded scenes, degenerate fields with 60 k
// seeds and ~1000 poses per image) the seeds are counting-sorted once by (field, band of kOccBands
// occupancy-row bands), a box then only visits the bins it overlaps, and a box that covers the whole
// map just bumps a per-field counter.
constexpr int kHeavySeeds = 4096;
constexpr int kOccBands = 32;
constexpr int kStageBytes = 72 * 1024;        // staged candidate lists when other batches are in flight
constexpr int kPickScan = 4096;               // seeds examined beyond the first one when a wave is picked

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

static int directed_edges(const Graph &g) {
  int n = 0;
  for (int j = 0; j < g.F; j++) n += g.deg[j];
  return n;
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
// Three steps keep the serial chain of the warp short: (0) only the
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
        PP_DASSERT(n + 32 <= kScratch);
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

// ---- one pose per warp ---------------------------------------------------------------------------
// A warp's private state, carved from dynamic shared memory (grow_words(F, E) 4-byte words per warp,
// E = directed edges of the skeleton; every directed edge is queued at most once, cifcaf.py:255-262).
struct GrowView {
  float *data;       // (F, 3) x, y, v
  float *js;         // (F)    joint_scales
  uint32_t *queued;  // (F)    in_frontier bits, [start] bit end
  // evaluated frontier entries (-score, (x, y, s, v), start, end), one array per member
  float *p_key, *p_x, *p_y, *p_s, *p_v;
  int *p_start, *p_end;
  int *dec;          // (F, 2) decoding_order (start, end)
  int *scratch;      // (kScratch)
  int *meta;         // [0] decoding steps, [1] a connection score above 1 was seen
  int n_edges;       // capacity of the p_* arrays
};

__host__ __device__ inline int grow_words(int F, int E) {
  return (3 * F + F + F + 7 * E + 2 * F + kScratch + 4 + 3) & ~3;
}

__device__ __forceinline__ GrowView grow_view(float *base, int F, int E) {
  GrowView w;
  w.data = base;                  base += 3 * F;
  w.js = base;                    base += F;
  w.queued = (uint32_t *)base;    base += F;
  w.p_key = base;                 base += E;
  w.p_x = base;                   base += E;
  w.p_y = base;                   base += E;
  w.p_s = base;                   base += E;
  w.p_v = base;                   base += E;
  w.p_start = (int *)base;        base += E;
  w.p_end = (int *)base;          base += E;
  w.dec = (int *)base;            base += 2 * F;
  w.scratch = (int *)base;        base += kScratch;
  w.meta = (int *)base;
  w.n_edges = E;
  return w;
}

// connection_value (cifcaf.py:194-217), warp-collective.
__device__ Xysv warp_connection_value(const GrowView &w, const Graph &g, const CafLists &cf, int start,
                                      int k, bool reverse_match, float keypoint_th, int method,
                                      int exp_mode) {
  Xysv zero = {0.f, 0.f, 0.f, 0.f};
  const int l = g.nb_link[start][k];
  const bool fw = g.nb_fwd[start][k] != 0;
  const float sx = w.data[3 * start], sy = w.data[3 * start + 1], sv = w.data[3 * start + 2];
  const float ss = w.js[start] > 0.0f ? w.js[start] : 0.0f;       // max(0.0, .)
  Xysv nw = scan_list(cf, l, fw, sx, sy, ss, w.scratch, method, exp_mode);
  // the eager frontier equals the reference's lazy one as long as no evaluated score exceeds its
  // bound sqrt(v_start), i.e. as long as connection scores stay <= 1
  if (nw.v > 1.0f && lane_id() == 0) w.meta[1] = 1;
  float kp = __fsqrt_rn(__fmul_rn(nw.v, sv));                     // :201
  if (kp < keypoint_th) return zero;
  if (nw.v == 0.0f) return zero;
  float st = nw.s > 0.0f ? nw.s : 0.0f;
  if (reverse_match) {                                            // :208-215
    Xysv rv = scan_list(cf, l, !fw, nw.x, nw.y, st, w.scratch, method, exp_mode);
    if (rv.s == 0.0f) return zero;
    float dist = __fadd_rn(fabsf(__fsub_rn(sx, rv.x)), fabsf(__fsub_rn(sy, rv.y)));
    if (dist > ss) return zero;
  }
  nw.v = kp;
  return nw;
}

// Heap order of the reference's frontier.  Non-greedy: evaluated entries
// (-score, (x, y, s, v), start, end).  Greedy (cifcaf.py:280-281): an entry is
// applied the moment it is evaluated, so only the unevaluated keys
// (-sqrt(v_start) [* confidence_scale], None, start, end) are ever compared;
// `key` then holds that bound.
__device__ __forceinline__ bool pending_less(const GrowView &w, int a, int b, int greedy) {
  if (w.p_key[a] != w.p_key[b]) return w.p_key[a] < w.p_key[b];
  if (!greedy) {
    if (w.p_x[a] != w.p_x[b]) return w.p_x[a] < w.p_x[b];
    if (w.p_y[a] != w.p_y[b]) return w.p_y[a] < w.p_y[b];
    if (w.p_s[a] != w.p_s[b]) return w.p_s[a] < w.p_s[b];
    if (w.p_v[a] != w.p_v[b]) return w.p_v[a] < w.p_v[b];
  }
  if (w.p_start[a] != w.p_start[b]) return w.p_start[a] < w.p_start[b];
  return w.p_end[a] < w.p_end[b];
}

// ascending float order as ascending unsigned order (NaN excluded by the caller)
__device__ __forceinline__ uint32_t float_orderable(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// add_to_frontier (cifcaf.py:251-263) for one start joint + the eager evaluation of the edges it
// queues, warp-collective (lane = neighbour slot for the queueing).  Returns the new number of
// frontier entries.
__device__ __forceinline__ int queue_and_evaluate(const GrowView &w, const Graph &g, const CScale &cscale,
                                                  const CafLists &cf, int start, int n_pend,
                                                  bool reverse_match, const pp_params &p) {
  const int lane = lane_id();
  const unsigned full = 0xffffffffu;
  bool ok = false;
  int end = 0;
  if (lane < g.deg[start]) {
    end = g.nb_end[start][lane];
    ok = !(w.data[3 * end + 2] > 0.0f) && !(w.queued[start] & (1u << end));           // :253-256
  }
  unsigned m = __ballot_sync(full, ok);
  const unsigned bits = __reduce_or_sync(full, ok ? (1u << end) : 0u);
  if (lane == 0) w.queued[start] |= bits;
  while (m) {
    const int k = __ffs(m) - 1;
    m &= m - 1u;
    const Xysv r = warp_connection_value(w, g, cf, start, k, reverse_match, p.keypoint_threshold,
                                         p.connection_method, p.exp_mode);
    if (lane == 0) {
      const int q = n_pend;
      PP_DASSERT(q < w.n_edges);
      float score = r.v;
      if (p.greedy) score = __fsqrt_rn(w.data[3 * start + 2]);                        // :258 (the bound)
      if (cscale.has) score = __fmul_rn(score, cscale.v[g.nb_link[start][k]]);        // :259-260, :282-284
      w.p_key[q] = -score; w.p_x[q] = r.x; w.p_y[q] = r.y; w.p_s[q] = r.s; w.p_v[q] = r.v;
      w.p_start[q] = start;
      w.p_end[q] = (r.v == 0.0f) ? -1 : g.nb_end[start][k];                           // :277-278
    }
    n_pend++;
  }
  __syncwarp();
  return n_pend;
}

// _grow (cifcaf.py:247-307) in its eager form, by one warp on its own state.  w.data / w.js hold the
// annotation, w.meta[0] the decoding steps recorded so far (appended to).
__device__ void warp_grow(const GrowView &w, const Graph &g, const CScale &cscale, const CafLists &cf,
                          bool reverse_match, const pp_params &p) {
  const int lane = lane_id();
  const unsigned full = 0xffffffffu;
  const int F = g.F;
  if (lane < F) w.queued[lane] = 0u;
  __syncwarp();
  int n_pend = 0;
  for (int j = 0; j < F; j++) {                                   // :288-291
    if (w.data[3 * j + 2] == 0.0f) continue;
    n_pend = queue_and_evaluate(w, g, cscale, cf, j, n_pend, reverse_match, p);
  }
  int n_dec = w.meta[0];
  for (;;) {
    // best evaluated entry whose target is still empty (frontier_get + :297-298)
    int best = -1;
    uint32_t bu = 0xffffffffu;
    bool odd = false;                      // NaN key: leave the integer fast path
    for (int i = lane; i < n_pend; i += 32) {
      const int end = w.p_end[i];
      if (end < 0 || w.data[3 * end + 2] > 0.0f) continue;
      const float key = w.p_key[i];
      if (key != key) odd = true;
      const uint32_t u = float_orderable(key);
      if (best < 0 || u < bu || (u == bu && pending_less(w, i, best, p.greedy))) { best = i; bu = u; }
    }
    const unsigned have = __ballot_sync(full, best >= 0);
    if (have == 0u) break;
    if (!__any_sync(full, odd)) {
      // primary key by one integer reduction; full tuple order only among exact ties
      const uint32_t m = __reduce_min_sync(full, best >= 0 ? bu : 0xffffffffu);
      unsigned tie = __ballot_sync(full, best >= 0 && bu == m);
      int win = __shfl_sync(full, best, __ffs(tie) - 1);
      for (tie &= tie - 1u; tie; tie &= tie - 1u) {
        const int o = __shfl_sync(full, best, __ffs(tie) - 1);
        if (pending_less(w, o, win, p.greedy)) win = o;
      }
      best = win;
    } else {
#pragma unroll 1
      for (int off = 16; off > 0; off >>= 1) {
        int o = __shfl_xor_sync(full, best, off);
        if (o >= 0 && (best < 0 || pending_less(w, o, best, p.greedy))) best = o;
      }
    }
    const int end = w.p_end[best];
    __syncwarp();
    if (lane == 0) {
      w.data[3 * end] = w.p_x[best]; w.data[3 * end + 1] = w.p_y[best];               // :300-302
      w.data[3 * end + 2] = w.p_v[best];
      w.js[end] = w.p_s[best];
      PP_DASSERT(n_dec < F && end >= 0 && end < F);
      w.dec[2 * n_dec] = w.p_start[best]; w.dec[2 * n_dec + 1] = end;                 // :303-304
      w.p_end[best] = -1;
    }
    n_dec++;
    __syncwarp();
    n_pend = queue_and_evaluate(w, g, cscale, cf, end, n_pend, reverse_match, p);     // :305
  }
  if (lane == 0) w.meta[0] = n_dec;
  __syncwarp();
}

// Phase timing for tuning (build with PP_EXTRA_NVCC=-DPP_ASM_TIMING): thread 0
// accumulates clock64 deltas per phase into g_asm_cycles; read them with
// pp_debug_asm_cycles.  Compiled out otherwise.
// slots: 0 setup, 1 seed search, 2 wave pick, 3 wave grow, 5 store, 6 occupancy update, 7 nms tail
#ifdef PP_ASM_TIMING
__device__ unsigned long long g_asm_cycles[12];   // 8 waves, 9 picks, 10 commits
#define PP_COUNT(slot, n) do { if (threadIdx.x == 0) atomicAdd(&g_asm_cycles[slot], (unsigned long long)(n)); } while (0)
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
#define PP_COUNT(slot, n) do { } while (0)
#endif

struct ListShared {
  int list_n[2 * PP_MAX_L];      // candidate-list lengths, [2l] forward / [2l+1] backward
  int list_off[2 * PP_MAX_L];    // float offsets into the staged copy
  int stage_total;               // entries over all lists
};

struct AsmShared {
#ifdef PP_ASM_TIMING
  long long tick;
#endif
  ListShared ls;
  long long box[PP_MAX_F][4];
  int box_valid[PP_MAX_F];
  int warp_min[kAsmWarps];
  // the wave: seeds whose poses the warps have grown (pick k lives in warp k's state)
  int n_pick;
  int pick_seed[kAsmWarps];
  int pick_f[kAsmWarps];
  int pick_box[kAsmWarps][4];
  // heavy mode (more than kHeavySeeds seeds): seeds are indexed by (field, band of occupancy rows)
  int occ_bin[PP_MAX_F * kOccBands];   // end offset of bin (field, band) in the field-major seed order
  int occ_all[PP_MAX_F];               // Occupancy.set boxes that cover the whole map, per field
};

// Load the list lengths of one image into shared memory and, when everything
// fits into `stage_floats` of dynamic shared memory, copy the rows of all 2L
// binned lists there with cp.async (all copies in flight at once).
// CTA-collective.  Returns the CafLists the growth code reads through.
__device__ CafLists load_lists(ListShared &sh, const float *rows, const int *start, BinGeo geo,
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

// Everything the seed loop of one image works on.
struct SeedCtx {
  const float *seeds;
  int n_seeds;
  uint8_t *occ;
  int2 *rec;
  bool heavy;
  int bshift;
  long long ow, oh;
};

// Occupancy.get of seed i (score order) right now
__device__ __forceinline__ bool seed_free(const AsmShared &sh, const SeedCtx &c, int i) {
  if (!c.heavy) return c.occ[i] == 0;
  const int r = c.rec[i].x;
  return ((c.occ[r & 0xffffff] + sh.occ_all[r >> 24]) & 255) == 0;
}

// field and occupancy cell of seed i
__device__ __forceinline__ void seed_cell(const SeedCtx &c, int i, int *f, int *cx, int *cy) {
  if (!c.heavy) {
    const int2 r = c.rec[i];
    *f = r.y >> 24; *cx = r.x; *cy = r.y & 0xffffff;
  } else {
    const int r = c.rec[i].x;
    const int cell = c.rec[r & 0xffffff].y;
    *f = r >> 24; *cx = cell & 0xffff; *cy = (unsigned)cell >> 16;
  }
}

// Warp 0: the seeds of the next wave.  Pick 0 is `found` (free, next in the reference's order); the
// others are the next free seeds, in order, that lie outside the Occupancy.set boxes the earlier
// picks' own joints are certain to set if those picks are committed (occupancy.py:31-39 on the
// seed's (x, y, scale)).  Whatever is picked only decides which poses are grown ahead of time.
// 128 seeds per trip: their counters, cells and positions are loaded together, then examined in order.
__device__ void pick_wave(AsmShared &sh, const SeedCtx &c, int found) {
  const int lane = lane_id();
  const unsigned full = 0xffffffffu;
  constexpr int U = 4;
  int n_pick = 0;
  auto add = [&](int i, int f, float x, float y, float s) {          // one lane
    long long bx[4];
    occupancy_box(x, y, s, c.ow, c.oh, bx);
    sh.pick_seed[n_pick] = i;
    sh.pick_f[n_pick] = f;
    for (int q = 0; q < 4; q++) sh.pick_box[n_pick][q] = (int)bx[q];
  };
  if (lane == 0) {
    const float *sd = c.seeds + (size_t)found * 5;
    add(found, (int)__ldg(sd + 1), __ldg(sd + 2), __ldg(sd + 3), __ldg(sd + 4));
  }
  n_pick = 1;
  __syncwarp();
  const int last = min(c.n_seeds, found + 1 + kPickScan);
  for (int base = found + 1; base < last && n_pick < kAsmWarps; base += 32 * U) {
    bool cand[U];
    int f[U], cx[U], cy[U];
    float sx[U], sy[U], ss[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int i = base + 32 * u + lane;
      cand[u] = i < last && seed_free(sh, c, i);
      f[u] = cx[u] = cy[u] = 0;
      sx[u] = sy[u] = ss[u] = 0.0f;
      if (cand[u]) {
        seed_cell(c, i, &f[u], &cx[u], &cy[u]);
        const float *sd = c.seeds + (size_t)i * 5;
        sx[u] = __ldg(sd + 2); sy[u] = __ldg(sd + 3); ss[u] = __ldg(sd + 4);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (n_pick >= kAsmWarps) break;
      for (int k = 0; k < n_pick && cand[u]; k++)
        if (sh.pick_f[k] == f[u] && cx[u] >= sh.pick_box[k][0] && cx[u] < sh.pick_box[k][1] &&
            cy[u] >= sh.pick_box[k][2] && cy[u] < sh.pick_box[k][3])
          cand[u] = false;
      unsigned m = __ballot_sync(full, cand[u]);
      while (m && n_pick < kAsmWarps) {
        const int l = __ffs(m) - 1;
        if (lane == l) add(base + 32 * u + lane, f[u], sx[u], sy[u], ss[u]);
        __syncwarp();
        const int k = n_pick++;
        if (cand[u] && lane > l && sh.pick_f[k] == f[u] && cx[u] >= sh.pick_box[k][0] &&
            cx[u] < sh.pick_box[k][1] && cy[u] >= sh.pick_box[k][2] && cy[u] < sh.pick_box[k][3])
          cand[u] = false;
        m = __ballot_sync(full, cand[u]) & ~((2u << l) - 1u);
      }
    }
  }
  PP_DASSERT(n_pick >= 1 && n_pick <= kAsmWarps);
  if (lane == 0) sh.n_pick = n_pick;
}

__global__ void __launch_bounds__(kAsmThreads)
assemble_kernel(AsmArgs a, Graph g, CScale cscale, pp_params p, int n_edges) {
  __shared__ AsmShared sh;
  const int b = blockIdx.x;
  const int F = g.F;
  const int warp = warp_id(), lane = lane_id();
  const float *seeds = a.seeds + (size_t)b * a.seed_cap * 5;
  const int n_seeds = min(a.n_seeds[b], a.seed_cap);
  uint8_t *occ = a.occ_cnt + (size_t)b * a.seed_cap;
  extern __shared__ __align__(16) float dyn_smem[];
  const int gw = grow_words(F, n_edges);
  const GrowView w = grow_view(dyn_smem + (size_t)warp * gw, F, n_edges);
  float *stage_buf = dyn_smem + (size_t)kAsmWarps * gw;
#ifdef PP_ASM_TIMING
  if (threadIdx.x == 0) sh.tick = clock64();
#endif
#ifdef PP_POISON_SMEM
  // debug build: shared memory starts as garbage (what another kernel's CTA may have left there), so that a
  // read of a word this CTA never wrote shows up in every decode instead of once in 10^5
  {
    const int n_dyn = kAsmWarps * gw + a.stage_floats;
    for (int i = threadIdx.x; i < n_dyn; i += blockDim.x) dyn_smem[i] = __int_as_float(0x4e6e6b28 + i);   // ~1e9
    int *raw = reinterpret_cast<int *>(&sh);
    for (int i = threadIdx.x; i < (int)(sizeof(AsmShared) / 4); i += blockDim.x) raw[i] = 0x4e6e6b28 + i;
    __syncthreads();
  }
#endif
  const CafLists cf = load_lists(sh.ls, a.rows + (size_t)b * a.L * 2 * kBinRows * a.caf_cap,
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
    if (warp == 0) {                            // exclusive scan of the PP_MAX_F * kOccBands counts
      constexpr int kPer = PP_MAX_F * kOccBands / 32;
      const int first = lane * kPer;
      int sum = 0;
      for (int q = 0; q < kPer; q++) sum += sh.occ_bin[first + q];
      int incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
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
  if (threadIdx.x == 0) sh.n_pick = 0;
  __syncthreads();
  const SeedCtx ctx = {seeds, n_seeds, occ, rec, heavy, bshift, ow, oh};

  int n_ann = 0, cursor = 0;
  bool overflow = false;
  PP_TICK(0);

  // store the pose grown in warp k's state + mark_occupied (:87-93, :107-108); CTA-collective
  auto commit = [&](int k) {
    const GrowView v = grow_view(dyn_smem + (size_t)k * gw, F, n_edges);
    const int n_dec = v.meta[0];
    PP_DASSERT(k >= 0 && k < kAsmWarps && n_ann < a.ann_cap && n_dec <= F);
    if (threadIdx.x == 0 && v.meta[1]) atomicOr(&a.status[b], PP_STATUS_SCORE_ABOVE_ONE);
    for (int i = threadIdx.x; i < F * 3; i += blockDim.x) pre_data[(size_t)n_ann * F * 3 + i] = v.data[i];
    for (int i = threadIdx.x; i < F; i += blockDim.x) {
      pre_scales[(size_t)n_ann * F + i] = v.js[i];
      if (pre_order) {
        pre_order[((size_t)n_ann * F + i) * 2] = i < n_dec ? v.dec[2 * i] : -1;
        pre_order[((size_t)n_ann * F + i) * 2 + 1] = i < n_dec ? v.dec[2 * i + 1] : -1;
      }
      sh.box_valid[i] = v.data[3 * i + 2] != 0.0f;
      if (sh.box_valid[i]) occupancy_box(v.data[3 * i], v.data[3 * i + 1], v.js[i], ow, oh, sh.box[i]);
    }
    n_ann++;
    PP_COUNT(10, 1);
    __syncthreads();
    PP_TICK(5);
    if (heavy) {
      // one warp per joint: a box that covers the whole map bumps the field's counter, any other box
      // visits the seeds of its field in the row bands it overlaps
      for (int j = warp; j < F; j += kAsmWarps) {
        if (!sh.box_valid[j]) continue;
        const long long minx = sh.box[j][0], maxx = sh.box[j][1], miny = sh.box[j][2], maxy = sh.box[j][3];
        if (minx >= maxx || miny >= maxy) continue;        // box outside the map: empty numpy slice
        if (minx <= 0 && maxx >= ow && miny <= 0 && maxy >= oh) {
          if (lane == 0) sh.occ_all[j]++;
          continue;
        }
        const int b0 = j * kOccBands + (int)(miny >> bshift), b1 = j * kOccBands + (int)((maxy - 1) >> bshift);
        const int k0 = b0 == 0 ? 0 : sh.occ_bin[b0 - 1], k1 = sh.occ_bin[b1];
        for (int kk = k0 + lane; kk < k1; kk += 4 * 32) {
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
    } else {
      // Occupancy.set of the pose's joints, seen from the seeds the loop has not passed yet: four
      // seeds per thread and trip so that four loads are in flight
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
    }
    __syncthreads();
    PP_TICK(6);
  };

  // ---- initial_annotations (cifcaf.py:95-98): grown and marked before the seeds, kAsmWarps at a time
  const int n_init = a.init_n ? min(a.init_n[b], a.init_cap) : 0;
  for (int base = 0; base < n_init && !overflow; base += kAsmWarps) {
    const int n = min(kAsmWarps, n_init - base);
    if (warp < n) {
      const float *idata = a.init_data + ((size_t)b * a.init_cap + base + warp) * F * 3;
      const float *iscales = a.init_scales + ((size_t)b * a.init_cap + base + warp) * F;
      for (int i = lane; i < F * 3; i += 32) w.data[i] = idata[i];
      for (int i = lane; i < F; i += 32) w.js[i] = iscales[i];
      if (lane == 0) { w.meta[0] = 0; w.meta[1] = 0; }
      __syncwarp();
      warp_grow(w, g, cscale, cf, true, p);
    }
    __syncthreads();
    PP_TICK(3);
    for (int k = 0; k < n; k++) {
      if (n_ann >= a.ann_cap) { overflow = true; break; }
      commit(k);
    }
  }

  // ---- the seed loop (cifcaf.py:100-108)
  int n_pick = 0;
  while (!overflow) {
    // next seed whose cell is free (:100-102)
    int found = -1;
    for (int base = cursor; base < n_seeds; base += blockDim.x) {
      const int i = base + threadIdx.x;
      const bool hit = i < n_seeds && seed_free(sh, ctx, i);
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) sh.warp_min[warp] = m ? (base + warp * 32 + __ffs(m) - 1) : -1;
      __syncthreads();
      for (int q = 0; q < kAsmWarps; q++)
        if (sh.warp_min[q] >= 0) { found = sh.warp_min[q]; break; }
      __syncthreads();
      if (found >= 0) break;
    }
    PP_TICK(1);
    if (found < 0) break;
    cursor = found + 1;
    if (n_ann >= a.ann_cap) { overflow = true; break; }
    // is its pose at hand?
    int k = -1;
    for (int q = 0; q < n_pick; q++)
      if (sh.pick_seed[q] == found) k = q;
    if (k < 0) {
      // ---- a new wave: `found` and the free seeds after it, one pose per warp
      // Every warp must have finished looking `found` up in the OLD picks before warp 0 overwrites them:
      // without this barrier a warp that is late (other kernels' warps share the SM's schedulers when
      // batches are in flight) could read the new pick_seed[0] == found, take the "pose at hand" branch
      // and leave the CTA's barriers out of step (seen as one illegal address in ~60,000 decodes)
      __syncthreads();
      if (warp == 0) pick_wave(sh, ctx, found);
      __syncthreads();
      PP_TICK(2);
      n_pick = sh.n_pick;
      PP_COUNT(8, 1); PP_COUNT(9, n_pick);
      if (warp < n_pick) {
        // new annotation from the seed (:104-105)
        const float *sd = seeds + (size_t)sh.pick_seed[warp] * 5;
        for (int i = lane; i < F * 3; i += 32) w.data[i] = 0.0f;
        for (int i = lane; i < F; i += 32) w.js[i] = 0.0f;
        __syncwarp();
        if (lane == 0) {
          const int f = (int)sd[1];
          w.data[3 * f] = sd[2]; w.data[3 * f + 1] = sd[3]; w.data[3 * f + 2] = sd[0];
          w.js[f] = sd[4];
          w.meta[0] = 0;
          w.meta[1] = 0;
        }
        __syncwarp();
        warp_grow(w, g, cscale, cf, true, p);                       // :106
      }
      __syncthreads();
      PP_TICK(3);
      k = 0;
    }
    commit(k);
  }
  if (overflow && threadIdx.x == 0) atomicOr(&a.status[b], PP_STATUS_ANN_OVERFLOW);
  if (threadIdx.x == 0) a.n_pre[b] = n_ann;
  if (a.fuse_nms) {
    // nms.Keypoints.annotations (nms.py:17-57) on the poses this CTA just wrote
    __threadfence_block();
    __syncthreads();
    nms_image(a.nms, b, dyn_smem, (size_t)(kAsmWarps * gw + a.stage_floats) * sizeof(float));
    PP_TICK(7);
  }
}

// ---------------------------------------------------------------------------
// complete_annotations (cifcaf.py:333-351) + _flood_fill (:309-331).
// Annotations are independent here: one warp each, grid (annotations / kAsmWarps, image).

struct Flood { float key, sx, sy, sv, s; int end; };

__device__ __forceinline__ bool flood_less(const Flood &a, const Flood &b) {
  // Python tuple order of (-v, end_i, [x, y, v], s)
  if (a.key != b.key) return a.key < b.key;
  if (a.end != b.end) return a.end < b.end;
  if (a.sx != b.sx) return a.sx < b.sx;
  if (a.sy != b.sy) return a.sy < b.sy;
  if (a.sv != b.sv) return a.sv < b.sv;
  return a.s < b.s;
}

// one thread; `prio` is the closure quirk of cifcaf.py:317,325 (see oracle flood_fill)
__device__ void flood_add(const GrowView &w, Flood *ff, int ff_cap, const Graph &g, int &n, int start,
                          float prio) {
  for (int k = 0; k < g.deg[start]; k++) {
    int end = g.nb_end[start][k];
    if (w.data[3 * end + 2] > 0.0f) continue;
    if (n >= ff_cap) return;     // cannot happen: <= 2 pushes per directed edge
    Flood &e = ff[n++];
    e.key = -prio; e.end = end;
    e.sx = w.data[3 * start]; e.sy = w.data[3 * start + 1]; e.sv = w.data[3 * start + 2];
    e.s = w.js[start];
  }
}

__device__ void flood_fill(const GrowView &w, Flood *ff, int ff_cap, const Graph &g) {
  int n = 0;
  for (int j = 0; j < g.F; j++) {
    if (w.data[3 * j + 2] == 0.0f) continue;
    flood_add(w, ff, ff_cap, g, n, j, w.data[3 * j + 2]);
  }
  while (n > 0) {
    int m = 0;
    for (int i = 1; i < n; i++)
      if (flood_less(ff[i], ff[m])) m = i;
    Flood e = ff[m];
    ff[m] = ff[--n];
    if (w.data[3 * e.end + 2] > 0.0f) continue;
    w.data[3 * e.end] = e.sx; w.data[3 * e.end + 1] = e.sy;
    w.data[3 * e.end + 2] = 0.00001f;
    w.js[e.end] = e.s;
    flood_add(w, ff, ff_cap, g, n, e.end, e.sv);
  }
}

__host__ __device__ inline int flood_words(int n_edges) { return 2 * n_edges * (int)(sizeof(Flood) / 4); }

__global__ void __launch_bounds__(kAsmThreads)
complete_kernel(AsmArgs a, const float *__restrict__ rows2, const int *__restrict__ start2,
                const int *__restrict__ nfwd2, const int *__restrict__ nbwd2,
                float *__restrict__ cmp_data, float *__restrict__ cmp_scales, Graph g, CScale cscale,
                pp_params p, int n_edges) {
  __shared__ ListShared ls;
  extern __shared__ __align__(16) float dyn_smem[];
  const int b = blockIdx.y, an = blockIdx.x * kAsmWarps + warp_id();
  const int lane = lane_id();
  const int F = g.F;
  // the score_th = 1e-4 lists hold nearly every cell: read them in place
  const CafLists cf = load_lists(ls, rows2 + (size_t)b * a.L * 2 * kBinRows * a.caf_cap,
                                 start2 + (size_t)b * a.L * 2 * (a.geo.nb + 1), a.geo,
                                 nfwd2 + (size_t)b * a.L, nbwd2 + (size_t)b * a.L, a.L, a.caf_cap,
                                 nullptr, 0);
  if (an >= min(a.n_pre[b], a.ann_cap)) return;       // no CTA barrier below
  const int per_warp = grow_words(F, n_edges) + flood_words(n_edges);
  float *mine = dyn_smem + (size_t)warp_id() * per_warp;
  const GrowView w = grow_view(mine, F, n_edges);
  Flood *ff = reinterpret_cast<Flood *>(mine + grow_words(F, n_edges));
  float *pre_data = a.pre_data + ((size_t)b * a.ann_cap + an) * F * 3;
  float *pre_scales = a.pre_scales + ((size_t)b * a.ann_cap + an) * F;
  int *pre_order = a.pre_order + ((size_t)b * a.ann_cap + an) * F * 2;
  for (int i = lane; i < F * 3; i += 32) w.data[i] = pre_data[i];
  for (int i = lane; i < F; i += 32) w.js[i] = pre_scales[i];
  __syncwarp();
  int n1 = 0;                                          // decoding steps of the first _grow
  if (lane == 0) {
    while (n1 < F && pre_order[2 * n1] >= 0) {
      w.dec[2 * n1] = pre_order[2 * n1]; w.dec[2 * n1 + 1] = pre_order[2 * n1 + 1];
      n1++;
    }
    w.meta[0] = n1;
    w.meta[1] = 0;
  }
  n1 = __shfl_sync(0xffffffffu, n1, 0);
  const unsigned unfilled = __ballot_sync(0xffffffffu, lane < F && w.data[3 * lane + 2] == 0.0f);   // :340
  __syncwarp();
  warp_grow(w, g, cscale, cf, false, p);                            // :341
  const int n_dec = w.meta[0];
  // as-grown state (scores not yet clamped) for the host's decoding_order replay
  for (int i = lane; i < F * 3; i += 32) pre_data[i] = w.data[i];
  for (int i = lane; i < F; i += 32) {
    pre_scales[i] = w.js[i];
    int s0 = i < n_dec ? w.dec[2 * i] : -1, s1 = i < n_dec ? w.dec[2 * i + 1] : -1;
    if (i == F - 1 && n_dec < F) { s0 = -2; s1 = n1; }   // marker: steps of the first grow
    pre_order[2 * i] = s0; pre_order[2 * i + 1] = s1;
  }
  __syncwarp();
  if (lane == 0) {
    bool any_empty = false;
    for (int j = 0; j < F; j++) {
      if (((unfilled >> j) & 1u) && w.data[3 * j + 2] > 0.0f)        // :342-344
        w.data[3 * j + 2] = 0.001f < w.data[3 * j + 2] ? 0.001f : w.data[3 * j + 2];
      if (w.data[3 * j + 2] == 0.0f) any_empty = true;
    }
    if (any_empty) flood_fill(w, ff, 2 * n_edges, g);                // :347-348
  }
  __syncwarp();
  float *od = cmp_data + ((size_t)b * a.ann_cap + an) * F * 3;
  float *os = cmp_scales + ((size_t)b * a.ann_cap + an) * F;
  for (int i = lane; i < F * 3; i += 32) od[i] = w.data[i];
  for (int i = lane; i < F; i += 32) os[i] = w.js[i];
}

static CScale make_cscale(const pp_skeleton &sk) {
  CScale cs;
  memset(&cs, 0, sizeof cs);
  cs.has = sk.has_confidence_scales;
  for (int l = 0; l < sk.L; l++) cs.v[l] = sk.confidence_scales[l];
  return cs;
}

int complete_launch(const AsmArgs &args, const float *rows2, const int *start2, const int *nfwd2,
                    const int *nbwd2, float *cmp_data, float *cmp_scales, const pp_skeleton &sk,
                    int B, int F, const pp_params &p, cudaStream_t st) {
  Graph g;
  int rc = build_graph(sk, F, &g);
  if (rc) return rc;
  const CScale cs = make_cscale(sk);
  const int n_edges = directed_edges(g);
  const size_t dyn = (size_t)kAsmWarps * (grow_words(F, n_edges) + flood_words(n_edges)) * sizeof(float);
  PP_CUDA_CHECK(cudaFuncSetAttribute(complete_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)dyn));
  dim3 grid((args.ann_cap + kAsmWarps - 1) / kAsmWarps, B);
  complete_kernel<<<grid, kAsmThreads, dyn, st>>>(args, rows2, start2, nfwd2, nbwd2, cmp_data, cmp_scales,
                                                 g, cs, p, n_edges);
  PP_LAUNCH_CHECK("complete_kernel");
  return PP_OK;
}

// dynamic shared memory an assembly CTA may use at most (the SM's 227 KB minus the kernel's static
// part and the 1 KB the system reserves per CTA)
static int max_dynamic_smem() {
  static int cached = 0;
  if (cached == 0) {
    cudaFuncAttributes at;
    int v = 227 * 1024 - 1024 - 16 * 1024;
    if (cudaFuncGetAttributes(&at, assemble_kernel) == cudaSuccess)
      v = 227 * 1024 - 1024 - (int)at.sharedSizeBytes;
    cached = v;
  }
  return cached;
}

int assemble_launch(const AsmArgs &args, const pp_skeleton &sk, int B, int F, const pp_params &p,
                    cudaStream_t st) {
  Graph g;
  int rc = build_graph(sk, F, &g);
  if (rc) return rc;
  const CScale cs = make_cscale(sk);
  const int n_edges = directed_edges(g);
  AsmArgs a2 = args;
  const int state_bytes = kAsmWarps * grow_words(F, n_edges) * (int)sizeof(float);
  const int max_dyn = max_dynamic_smem();
  if (state_bytes > max_dyn) return PP_ERR_ARG;
  // one image per SM and nothing else on the GPU: give the lists (almost) the whole shared memory;
  // with other batches in flight (pp_params.in_flight) leave room for their CTAs on the same SM
  int stage_bytes = (B <= num_sms() && p.in_flight <= 1) ? max_dyn - state_bytes : kStageBytes;
  if (p.asm_stage_kb > 0) stage_bytes = p.asm_stage_kb * 1024;        // tuning knob
  if (p.asm_stage_kb < 0) stage_bytes = 0;
  if (stage_bytes > max_dyn - state_bytes) stage_bytes = max_dyn - state_bytes;
  stage_bytes &= ~15;
  a2.heavy_seeds = p.asm_heavy_seeds > 0 ? p.asm_heavy_seeds : (p.asm_heavy_seeds < 0 ? 0 : kHeavySeeds);
  a2.stage_floats = stage_bytes / (int)sizeof(float);
  PP_CUDA_CHECK(cudaFuncSetAttribute(assemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn));
  prof_mark(p.profile, kStAssemble, st);
  assemble_kernel<<<B, kAsmThreads, (size_t)state_bytes + stage_bytes, st>>>(a2, g, cs, p, n_edges);
  PP_LAUNCH_CHECK("assemble_kernel");
  return PP_OK;
}

}  // namespace pp

#ifdef PP_ASM_TIMING
extern "C" int pp_debug_asm_cycles(unsigned long long *out12, int reset) {
  if (cudaMemcpyFromSymbol(out12, pp::g_asm_cycles, 12 * sizeof(unsigned long long)) != cudaSuccess) return -3;
  if (reset) {
    unsigned long long z[12] = {0};
    cudaMemcpyToSymbol(pp::g_asm_cycles, z, sizeof z);
  }
  return 0;
}
#endif
