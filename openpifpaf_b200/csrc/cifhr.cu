// cifhr.cu -- CifHr high-resolution accumulation on sm_100a.
//
// Replaces decoder/cif_hr.py:26-65 (CifHr.accumulate / fill_multiple) and
// functional.pyx:105-141 (scalar_square_add_gauss_with_max) of the reference.
//
// The reference walks the sources of one field sequentially and adds a clipped
// Gaussian box per source into the hr map, clamping at 1 after every add.  All
// increments are >= 0, so clamping at any subset of the steps gives the same
// result and only the ORDER of the float adds per pixel matters (ascending
// source index = row-major cell order).  We therefore run it as a gather in
// which every pixel has exactly one owner lane:
//
//   K1 cif_compact_kernel   one CTA per (image, field): ordered compaction of
//                           the cells with conf > v_threshold into a source
//                           list {cx, cy, sigma^2, v} + integer box, plus a
//                           per-strip source count (difference array in smem).
//   K2 cifhr_splat_kernel   persistent, warp-specialised CTAs (one producer
//                           warp + eight consumer warps).  A strip = R full hr
//                           rows = ONE contiguous run of floats in HBM.
//                             * the producer warp walks strip groups from a
//                               global cursor and writes every EMPTY strip (the
//                               vast majority) with TMA bulk stores
//                               (cp.async.bulk shared->global) from a zeroed
//                               shared-memory buffer; it never waits on the
//                               consumers, so the HBM write stream never stalls
//                               behind splat arithmetic;
//                             * the consumer warps take the strips that have
//                               sources from a compact list K1 built.  The
//                               strip is assembled in a shared-memory tile:
//                               warps own 32-pixel patches (R rows x 32/R
//                               columns, one pixel per lane); per 64-source
//                               chunk a 64-bit mask per patch says which
//                               sources touch it, so a warp only evaluates
//                               sources that overlap its patch, in list order,
//                               accumulating in a register.  The finished tile
//                               leaves with one TMA bulk store (double-buffered).
//
// No atomics on hr, no read-modify-write of hr, no separate memset: every hr
// byte is written exactly once.  Result is bit-identical to the reference.
#include <stdlib.h>

#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kSplatThreads = 256;             // consumer threads
constexpr int kSplatWarps = kSplatThreads / 32;
constexpr int kSplatBlock = kSplatThreads + 32; // + the producer warp
constexpr int kChunk = 64;            // sources per mask chunk (one bit each)
constexpr int kGroupCap = kSplatThreads;   // sources staged per ordered-filter block
constexpr int kGroupStripsMax = 32;
static int group_strips(const pp_params &p) {           // strips per CTA (pp_params.cifhr_group)
  return (p.cifhr_group >= 1 && p.cifhr_group <= kGroupStripsMax) ? p.cifhr_group : 3;
}
constexpr int kMaxPatches = 192;      // patches per strip (hw <= 6136, see strip_rshift)
constexpr int kWidePatches = 8;       // a source box spanning more patches than this makes its block "wide"
constexpr int kZeroFloats = 1024;     // zeroed smem the empty strips are stored from (4 KiB pieces)
constexpr int kStripBytesMax = 24 * 1024;
constexpr int kCompactThreads = 256;
constexpr int kCompactSlab = 8192;     // cells per active-cell list (16 KiB of shared memory)
constexpr int kBoundBlock = 64;        // sources per row-bound record (see CifHrWs::blk)
// strips kernel (one warp per strip, see cifhr_strips_kernel)
constexpr int kStripWarpsDefault = 3;  // consumer warps per CTA (pp_params.cifhr_warps)
constexpr int kStripWarpsMax = 8;
constexpr int kStripStage = 64;        // sources staged per warp between two flushes
constexpr int kStripBatch = 4;         // rounds of 32 source records loaded together
constexpr int kNarrowMaxCols = 48;     // a source box wider than this sends the strips it meets to the CTA-per-strip kernel
constexpr int kNarrowMaxSrc = 160;     // so does a strip that meets more sources than this
constexpr int kDiffWideShift = 20;     // per-strip difference array: sources in the low bits, wide sources above

struct CifHrWs {
  float4 *src_a;    // (B,F,src_cap) cx, cy, sigma^2, v
  int4 *src_b;      // (B,F,src_cap) minx, maxx, miny, maxy
  float2 *src_c;    // (B,F,src_cap) scale in pixels (CifDet: w, h in pixels) for the seed stage
  int *n_src;       // (B,F)
  int2 *blk;        // (B,F,blk_cap) [min miny, max maxy) over each run of kBoundBlock sources: sources are in
                    // row-major cell order, so a strip only meets a short range of runs (crowded scenes)
  int blk_cap;
  int *strip_cnt;   // (B,F,nstrips) sources meeting the strip | wide sources << kDiffWideShift; 0 = empty strip
  int2 *busy_list;  // (B*F*nstrips) {bf * nstrips + strip, sources of that field} of the strips with sources, any
                    // order: the narrow strips (one warp each, cifhr_strips_kernel) from the front, the wide ones
                    // (one CTA each, cifhr_splat_kernel) from the back
  int *cursors;     // [0] narrow strips, [1] producer group cursor, [2] narrow work cursor, [3] wide strips,
                    // [4] wide work cursor
  int nstrips;
  int rshift;       // strip rows R = 1 << rshift
};

// rows per strip: the largest power of two <= 8 whose strip fits the smem budget
static int strip_rshift(int hw, int pref = 0) {    // pref: pp_params.cifhr_rshift (log2 rows + 1), 0 = 8 rows
  int r = (pref >= 1 && pref <= 4) ? pref - 1 : 3;
  while (r > 0 && ((size_t)hw << r) * sizeof(float) > (size_t)kStripBytesMax - 32) r--;
  return r;
}

static int num_strips(int hh, int rshift) { return (hh + (1 << rshift) - 1) >> rshift; }

size_t cifhr_ws_bytes(const pp_shape &s, const pp_caps &c) {
  size_t bf = (size_t)s.B * s.F;
  size_t n = 0;
  n += align_up(bf * c.src_cap * sizeof(float4), 256);
  n += align_up(bf * c.src_cap * sizeof(int4), 256);
  n += align_up(bf * c.src_cap * sizeof(float2), 256);
  n += align_up(bf * sizeof(int), 256);
  n += align_up(bf * ((c.src_cap + kBoundBlock - 1) / kBoundBlock) * sizeof(int2), 256);
  n += 3 * align_up(bf * num_strips(hr_h(s), 0) * sizeof(int), 256);   // worst case R = 1; counts + int2 work list
  n += 256;
  return n;
}

static CifHrWs carve(void *ws, const pp_shape &s, const pp_caps &c, int rshift_pref = 0) {
  CifHrWs w;
  char *p = (char *)ws;
  size_t bf = (size_t)s.B * s.F;
  w.src_a = (float4 *)p; p += align_up(bf * c.src_cap * sizeof(float4), 256);
  w.src_b = (int4 *)p;   p += align_up(bf * c.src_cap * sizeof(int4), 256);
  w.src_c = (float2 *)p; p += align_up(bf * c.src_cap * sizeof(float2), 256);
  w.n_src = (int *)p;    p += align_up(bf * sizeof(int), 256);
  w.blk_cap = (c.src_cap + kBoundBlock - 1) / kBoundBlock;
  w.blk = (int2 *)p;     p += align_up(bf * w.blk_cap * sizeof(int2), 256);
  w.strip_cnt = (int *)p; p += align_up(bf * num_strips(hr_h(s), 0) * sizeof(int), 256);
  w.busy_list = (int2 *)p; p += 2 * align_up(bf * num_strips(hr_h(s), 0) * sizeof(int), 256);
  w.cursors = (int *)p;
  w.rshift = strip_rshift(hr_w(s), rshift_pref);
  w.nstrips = num_strips(hr_h(s), w.rshift);
  return w;
}

// ---------------------------------------------------------------------------
// K1: ordered compaction of CIF sources per (image, field).
struct CompactArgs {
  const float *cif, *cif2, *rest;
  CifHrWs w;
  pp_shape s;
  int src_cap, hh, hw;
  float v_th, min_scale_over_stride;
  int use_min_scale, det;
  int *status;
  float neighbors;       // CifHr.neighbors
  int n_planes;          // B * F of the launch (the work list is shared by all of them)
  int all_wide;          // every busy strip goes to the CTA-per-strip kernel (pp_params.cifhr_warps < 0)
};

// dynamic shared memory of a scan CTA (compact or CAF prescan): [nstrips + 1 ints, padded to 16 bytes]
// [staged confidence values: 4 * slab + 32 bytes][active-cell list: 2 * slab bytes]
__host__ __device__ inline size_t scan_smem_diff_bytes(int nstrips) { return ((size_t)(nstrips + 1) * 4 + 15) & ~(size_t)15; }
__host__ __device__ inline size_t scan_smem_bytes(int nstrips, int cells) {
  const size_t slab = (size_t)(cells < kCompactSlab ? cells : kCompactSlab);
  return scan_smem_diff_bytes(nstrips) + 4 * slab + 32 + 2 * slab;
}

__device__ __forceinline__ void cif_compact_body(const CompactArgs &a, const int bf, int *sm_diff, float *stage,
                                                 unsigned short *s_cells, uint64_t *bar, int *warp_tot) {
  const float *__restrict__ cif = a.cif, *__restrict__ cif2 = a.cif2, *__restrict__ rest = a.rest;
  const CifHrWs &w = a.w;
  const pp_shape &s = a.s;
  const int src_cap = a.src_cap, hh = a.hh, hw = a.hw, use_min_scale = a.use_min_scale, det = a.det;
  const float v_th = a.v_th, min_scale_over_stride = a.min_scale_over_stride;
  int *__restrict__ status = a.status;
  uint32_t phase = 0;
  const int n = s.H * s.W;
  // cif2 != NULL: a multi-scale hflip pair accumulated into one map, first all sources of cif, then
  // of cif2, each with v / neighbors / 2 (cif_hr.py:55-57,66-70)
  const float len_cifs = cif2 ? 2.0f : 1.0f;
  const float stride = (float)s.stride;

  float4 *oa = w.src_a + (size_t)bf * src_cap;
  int4 *ob = w.src_b + (size_t)bf * src_cap;
  float2 *oc = w.src_c + (size_t)bf * src_cap;
  int count = 0;
  for (int pass = 0; pass < (cif2 ? 2 : 1); pass++) {
  const float *p = (pass ? cif2 : cif) + (size_t)bf * (det ? 7 : 5) * n;   // CifDet: [v, x, y, _, w, h, _]
  // rest != NULL (host-mapped input, api.cu): only the confidence planes were copied to the device;
  // the other planes are read where a cell is active, straight from the pinned host buffer
  const float *pr = (rest && !pass) ? rest + (size_t)bf * (det ? 7 : 5) * n : p;
  for (int slab = 0; slab < n; slab += kCompactSlab) {
    // phase 1: cells above the threshold in row-major order (cif_hr.py:27); phase 2: one listed cell
    // per thread.  (Interleaving the two per 256 cells cost ~26 serial load + barrier rounds.)
    // the slab's confidence values arrive in shared memory by one bulk copy (common.cuh)
    const int m = min(n, slab + kCompactSlab) - slab;
    const float *sp = stage_floats(p + slab, m, stage, bar, phase);
    const int n_act = block_active_cells_smem<kCompactThreads>(sp, m, v_th, s_cells, warp_tot);
    for (int base = 0; base < n_act; base += blockDim.x) {
      const int i = base + threadIdx.x;
      bool flag = i < n_act;
      int c = 0;
      float conf = 0.f, c4 = 0.f, c5 = 0.f;
      if (flag) {
        c = slab + s_cells[i];
        conf = sp[s_cells[i]];
        c4 = __ldg(pr + 4 * n + c);
        if (det) c5 = __ldg(pr + 5 * n + c);
        if (use_min_scale) flag = c4 > min_scale_over_stride;               // :28-29
        if (flag && use_min_scale && det) flag = c5 > min_scale_over_stride; // :87-89
      }
      int total = n_act - base < (int)blockDim.x ? n_act - base : (int)blockDim.x;
      int off = i - base;
      if (use_min_scale) off = block_flag_scan(flag, warp_tot, &total);     // a second ordered filter
      if (flag) {
        int k = count + off;
        if (k < src_cap) {
          float cx = __fmul_rn(__ldg(pr + n + c), stride);          // :32
          float cy = __fmul_rn(__ldg(pr + 2 * n + c), stride);      // :33
          float sg;
          if (det) {      // CifDetHr: sigma = max(1, 0.1 * min(w, h) * stride), cif_hr.py:91-94
            const float m = (c4 != c4) ? c4 : ((c5 != c5) ? c5 : fminf(c4, c5));   // np.minimum keeps NaN
            sg = __fmul_rn(__fmul_rn(0.1f, m), stride);
          } else {
            sg = __fmul_rn(__fmul_rn(0.5f, c4), stride);
          }
          float sigma = (sg != sg) ? sg : fmaxf(1.0f, sg);          // np.maximum(1.0, .) :34
          float cv = __fdiv_rn(__fdiv_rn(conf, a.neighbors), len_cifs);   // v / neighbors / len_cifs :40
          // functional.pyx:122-125 with truncate = 1.0
          int minx = (int)clipf(__fsub_rn(cx, sigma), 0.0f, (float)(hw - 1));
          int maxx = (int)clipf(__fadd_rn(__fadd_rn(cx, sigma), 1.0f), (float)(minx + 1), (float)hw);
          int miny = (int)clipf(__fsub_rn(cy, sigma), 0.0f, (float)(hh - 1));
          int maxy = (int)clipf(__fadd_rn(__fadd_rn(cy, sigma), 1.0f), (float)(miny + 1), (float)hh);
          PP_DASSERT(k >= 0 && k < src_cap && (miny >> w.rshift) <= w.nstrips && ((maxy - 1) >> w.rshift) + 1 <= w.nstrips);
          oa[k] = make_float4(cx, cy, __fmul_rn(sigma, sigma), cv);
          ob[k] = make_int4(minx, maxx, miny, maxy);
          // what CifSeeds needs besides (cx, cy, cv): the scale(s) in pixels (cif_seeds.py:44,84)
          oc[k] = make_float2(__fmul_rn(c4, stride), det ? __fmul_rn(c5, stride) : 0.0f);
          const int one = 1 + ((maxx - minx > kNarrowMaxCols) ? (1 << kDiffWideShift) : 0);
          atomicAdd(&sm_diff[miny >> w.rshift], one);
          atomicAdd(&sm_diff[((maxy - 1) >> w.rshift) + 1], -one);
        }
      }
      count += total;
    }
    __syncthreads();        // everyone is done with the staged slab before the next copy lands in it
  }
  }
  __syncthreads();
  // row bounds of every run of kBoundBlock sources (one warp per run)
  {
    const int n_kept = min(count, src_cap);
    int2 *bb = w.blk + (size_t)bf * w.blk_cap;
    for (int blk = warp_id(); blk * kBoundBlock < n_kept; blk += kCompactThreads / 32) {
      int lo = 0x7fffffff, hi = -0x7fffffff;
      const int k1 = min(n_kept, (blk + 1) * kBoundBlock);
      for (int k = blk * kBoundBlock + lane_id(); k < k1; k += 32) {
        const int4 b = ob[k];
        lo = min(lo, b.z);
        hi = max(hi, b.w);
      }
      lo = __reduce_min_sync(0xffffffffu, lo);
      hi = __reduce_max_sync(0xffffffffu, hi);
      if (lane_id() == 0) bb[blk] = make_int2(lo, hi);
    }
  }
  // per-strip counts = prefix sum of the difference array; warp 0 scans it 32 strips at a time and
  // appends the strips that have sources to the consumers' work list (order is irrelevant)
  if (warp_id() == 0) {
    const int lane = lane_id();
    if (lane == 0) {
      if (count > src_cap) {
        atomicOr(&status[bf / s.F], PP_STATUS_SRC_OVERFLOW);
        count = src_cap;
      }
      w.n_src[bf] = count;
    }
    count = __shfl_sync(0xffffffffu, count, 0);
    int *sc = w.strip_cnt + (size_t)bf * w.nstrips;
    int run = 0, n_narrow = 0, n_wide = 0;
    // a strip goes to the one-warp-per-strip kernel unless it meets a wide box or very many sources
    // (either kernel gives the same map; this only decides who is faster)
    const bool all_wide = a.all_wide != 0;
    auto is_wide = [all_wide](int v) {
      return all_wide || ((unsigned)v >> kDiffWideShift) != 0u || (v & ((1 << kDiffWideShift) - 1)) > kNarrowMaxSrc;
    };
    for (int base = 0; base < w.nstrips; base += 32) {
      const int i = base + lane;
      int v = i < w.nstrips ? sm_diff[i] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
      }
      v += run;
      if (i < w.nstrips) { sc[i] = v; sm_diff[i] = v; }
      run = __shfl_sync(0xffffffffu, v, 31);
      const bool busy = i < w.nstrips && v != 0;
      n_wide += __popc(__ballot_sync(0xffffffffu, busy && is_wide(v)));
      n_narrow += __popc(__ballot_sync(0xffffffffu, busy && !is_wide(v)));
    }
    if (n_narrow + n_wide) {
      int at_n = 0, at_w = 0;
      if (lane == 0) {
        if (n_narrow) at_n = atomicAdd(&w.cursors[0], n_narrow);
        if (n_wide) at_w = atomicAdd(&w.cursors[3], n_wide);
      }
      at_n = __shfl_sync(0xffffffffu, at_n, 0);
      at_w = __shfl_sync(0xffffffffu, at_w, 0);
      const int last = a.n_planes * w.nstrips - 1;
      for (int base = 0; base < w.nstrips; base += 32) {
        const int i = base + lane;
        const int v = i < w.nstrips ? sm_diff[i] : 0;
        const bool busy = v != 0, wide = busy && is_wide(v);
        const unsigned mn = __ballot_sync(0xffffffffu, busy && !wide);
        const unsigned mw = __ballot_sync(0xffffffffu, wide);
        const unsigned below = (1u << lane) - 1u;
        const int2 e = make_int2(bf * w.nstrips + i, count);
        if (wide) w.busy_list[last - (at_w + __popc(mw & below))] = e;
        else if (busy) w.busy_list[at_n + __popc(mn & below)] = e;
        at_n += __popc(mn);
        at_w += __popc(mw);
      }
    }
  }
}


}  // namespace pp

#include "caf_prescan.cuh"

namespace pp {

// One launch for every confidence-plane scan of a decode: blocks [0, n_cif) compact the CIF sources
// of one (image, field) each, the others collect the raw CAF candidates of one (image, link) each
// (caf_prescan.cuh; nothing there depends on the hr map, which does not exist yet).
__global__ void __launch_bounds__(kCompactThreads, 4)
fields_scan_kernel(CompactArgs a, CafPreArgs pa, int n_cif) {
  extern __shared__ __align__(16) unsigned char scan_smem[];
  __shared__ int warp_tot[kCompactThreads / 32];
  __shared__ __align__(8) uint64_t bar;
  const int cells = a.s.H * a.s.W;
  const size_t slab = (size_t)min(cells, kCompactSlab);
  int *sm_diff = reinterpret_cast<int *>(scan_smem);
  float *stage = reinterpret_cast<float *>(scan_smem + scan_smem_diff_bytes(a.w.nstrips));
  unsigned short *s_cells = reinterpret_cast<unsigned short *>(scan_smem + scan_smem_diff_bytes(a.w.nstrips) +
                                                               4 * slab + 32);
  if (threadIdx.x == 0) mbar_init(&bar, 1);
  __syncthreads();
  if ((int)blockIdx.x < n_cif) {
    for (int i = threadIdx.x; i <= a.w.nstrips; i += blockDim.x) sm_diff[i] = 0;
    __syncthreads();
    cif_compact_body(a, blockIdx.x, sm_diff, stage, s_cells, &bar, warp_tot);
  } else {
    caf_prescan_body<kCompactThreads, kCompactSlab>(pa, a.s, blockIdx.x - n_cif, stage, s_cells, &bar, warp_tot);
  }
}

// ---------------------------------------------------------------------------
// K2: strips -> HBM.

// one pixel inside the source's box: functional.pyx:127-140
__device__ __forceinline__ void splat_core(float &acc, int px, int py, const float4 a) {
  {
    float dx = __fsub_rn((float)px, a.x);
    float dy = __fsub_rn((float)py, a.y);
    float dx2 = __fmul_rn(dx, dx);
    float dy2 = __fmul_rn(dy, dy);
    float d2 = __fadd_rn(dx2, dy2);
    if (!(d2 > a.z)) {                                            // functional.pyx:131
      float vv;
      if (dx2 < 0.25f && dy2 < 0.25f) {
        vv = a.w;                                                 // :134-136
      } else {
        // (-0.5 * d2) / sigma2 is evaluated in double by the reference and
        // rounded to float; with 53 >= 2*24+2 that equals the correctly
        // rounded float quotient.
        float arg = __fdiv_rn(__fmul_rn(-0.5f, d2), a.z);
        vv = __fmul_rn(a.w, approx_exp(arg));                     // :138
      }
      acc = __fadd_rn(acc, vv);                                   // :140
    }
  }
}

__device__ __forceinline__ void splat_px(float &acc, int px, int py, const float4 a, const int4 b) {
  if (px >= b.x && px < b.y && py >= b.z && py < b.w) splat_core(acc, px, py, a);
}

__device__ __forceinline__ float clamp1(float v) { return (v < 1.0f) ? v : 1.0f; }  // :141

// TMA bulk copy shared -> global (no tensor map: the strip is linear memory).
__device__ __forceinline__ void bulk_store(float *gdst, const float *ssrc, uint32_t bytes) {
  uint32_t s = (uint32_t)__cvta_generic_to_shared(ssrc);
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(gdst), "r"(s), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
// generic-proxy writes to smem -> visible to the async proxy (TMA)
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

struct StripGeo {
  size_t gs, ge;   // [gs, ge) floats of hr
  size_t ga;       // gs rounded down to a 16-byte unit; smem index of float g is g - ga
  size_t m0, m1;   // 16-byte aligned middle [m0, m1)
  int r0, r1;
};

__device__ __forceinline__ StripGeo strip_geo(size_t plane0, int strip, int rshift, int hh, int hw) {
  StripGeo g;
  g.r0 = strip << rshift;
  g.r1 = min(g.r0 + (1 << rshift), hh);
  g.gs = plane0 + (size_t)g.r0 * hw;
  g.ge = plane0 + (size_t)g.r1 * hw;
  g.ga = g.gs & ~(size_t)3;
  g.m0 = (g.gs + 3) & ~(size_t)3;
  g.m1 = g.ge & ~(size_t)3;
  if (g.m0 > g.ge) g.m0 = g.ge;
  if (g.m1 < g.m0) g.m1 = g.m0;
  return g;
}

// Write one finished strip from a smem buffer laid out like the strip.  Called
// by the 256 consumer threads after the buffer is complete, fenced and barriered.
__device__ __forceinline__ void emit_strip(float *__restrict__ hr, const float *buf, const StripGeo &g) {
  const int t = threadIdx.x;
  if (t == 0) {
    if (g.m1 > g.m0) bulk_store(hr + g.m0, buf + (g.m0 - g.ga), (uint32_t)((g.m1 - g.m0) * sizeof(float)));
  } else if (t >= 32 && t < 36) {          // head: < 4 floats before the aligned middle
    size_t i = g.gs + (size_t)(t - 32);
    if (i < g.m0) hr[i] = buf[i - g.ga];
  } else if (t >= 36 && t < 40) {          // tail
    size_t i = g.m1 + (size_t)(t - 36);
    if (i < g.ge) hr[i] = buf[i - g.ga];
  }
}

// An empty strip: zeros, in pieces of the small zero buffer.  One warp.
__device__ __forceinline__ void emit_zero_strip(float *__restrict__ hr, const float *zero, const StripGeo &g,
                                                int lane) {
  if (lane == 0) {
    for (size_t o = g.m0; o < g.m1; o += kZeroFloats) {
      size_t n = g.m1 - o < (size_t)kZeroFloats ? g.m1 - o : (size_t)kZeroFloats;
      bulk_store(hr + o, zero, (uint32_t)(n * sizeof(float)));
    }
  } else if (lane >= 1 && lane < 5) {
    size_t i = g.gs + (size_t)(lane - 1);
    if (i < g.m0) hr[i] = 0.f;
  } else if (lane >= 5 && lane < 9) {
    size_t i = g.m1 + (size_t)(lane - 5);
    if (i < g.ge) hr[i] = 0.f;
  }
}

// barrier of the consumer warps only (the producer warp never joins it)
__device__ __forceinline__ void consumer_sync() {
  asm volatile("bar.sync 1, %0;" :: "n"(kSplatThreads) : "memory");
}

// block_flag_scan for the consumer warps; *any_wide = some thread of the block passed wide = true
// (bit 30 of a warp's total carries its vote)
// `clear` (2 * kMaxPatches words) is zeroed between the two barriers when anything was flagged: every
// consumer has then left the code before the scan, and nobody reads `clear` before the second barrier
__device__ __forceinline__ int consumer_flag_scan(bool flag, bool wide, int *warp_tot, int *total,
                                                  bool *any_wide, uint32_t *clear) {
  unsigned m = __ballot_sync(0xffffffffu, flag);
  const bool wv = __any_sync(0xffffffffu, wide);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int in_warp = __popc(m & ((1u << lane) - 1u));
  if (lane == 0) warp_tot[w] = __popc(m) | (wv ? (1 << 30) : 0);
  consumer_sync();
  int base = 0, tot = 0, votes = 0;
#pragma unroll
  for (int i = 0; i < kSplatWarps; i++) {
    const int t = warp_tot[i];
    if (i < w) base += t & 0xffff;
    tot += t & 0xffff;
    votes |= t >> 30;
  }
  if (tot != 0)
    for (int q = threadIdx.x; q < 2 * kMaxPatches; q += kSplatThreads) clear[q] = 0u;
  consumer_sync();
  *total = tot;
  *any_wide = votes != 0;
  return base + in_warp;
}

__global__ void __launch_bounds__(kSplatBlock, 3)
cifhr_splat_kernel(float *__restrict__ hr, CifHrWs w, int n_planes, int kGroupStrips, int src_cap,
                   int hh, int hw, int buf_floats, int merge, int produce, int hr_off) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float *zero = reinterpret_cast<float *>(smem_raw);
  float *tiles = zero + kZeroFloats;                      // two tile buffers
  __shared__ float4 s_a[kGroupCap];
  __shared__ int4 s_b[kGroupCap];
  __shared__ uint32_t s_mask[2][kMaxPatches];             // bit s of word s/32: source s touches the patch
  __shared__ uint32_t s_sat[kMaxPatches / 32];            // bit p: every pixel of patch p has reached 1
  __shared__ int warp_tot[kSplatWarps];
  __shared__ int s_first[kSplatWarps], s_last[kSplatWarps];
  __shared__ int s_item;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rshift = w.rshift;
  const int n_groups = (w.nstrips + kGroupStrips - 1) / kGroupStrips;
  // the strips this kernel owns: the wide end of the work list (usually none: every CTA leaves at once)
  if (!produce && __ldg(w.cursors + 3) == 0) return;

  for (int i = tid; i < (kZeroFloats >> 2); i += kSplatBlock)
    reinterpret_cast<float4 *>(zero)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_async_smem();
  __syncthreads();

  if (warp == kSplatWarps) {
    // ---- producer warp: every empty strip, as TMA zero stores ---------------------
    // merge (multi-scale, cif_hr.py:62 accumulated = np.maximum(ta, accumulated)): an empty strip
    // of this scale leaves the map as it is
    if (merge || !produce) return;
    const int total = n_planes * n_groups;
    for (;;) {
      int item = 0;
      if (lane == 0) item = atomicAdd(&w.cursors[1], 1);
      item = __shfl_sync(0xffffffffu, item, 0);
      if (item >= total) break;
      const int bf = item / n_groups;
      const int strip0 = (item - bf * n_groups) * kGroupStrips;
      const int nloc = min(kGroupStrips, w.nstrips - strip0);
      const size_t plane0 = (size_t)hr_off + (size_t)bf * hh * hw;
      const int cnt = lane < nloc ? __ldg(w.strip_cnt + (size_t)bf * w.nstrips + strip0 + lane) : 1;
      unsigned empty = __ballot_sync(0xffffffffu, cnt == 0);
      for (; empty; empty &= empty - 1u)
        emit_zero_strip(hr, zero, strip_geo(plane0, strip0 + __ffs(empty) - 1, rshift, hh, hw), lane);
    }
    if (lane == 0) bulk_wait_read<0>();     // smem must outlive the reads of the async proxy
    return;
  }

  // ---- consumer warps: the strips that have sources ------------------------------------
  const int n_busy = __ldg(w.cursors + 3);
  const int cshift = 5 - rshift;            // patch = (1 << rshift) rows x (1 << cshift) cols, 1 px per lane
  const int n_patches = (hw + (1 << cshift) - 1) >> cshift;
  const int row = lane >> (5 - rshift);                   // lane's row inside the patch
  const int cg = lane & ((1 << (5 - rshift)) - 1);        // lane's column inside the patch
  int n_tiles = 0;

  for (;;) {
    if (tid == 0) s_item = atomicAdd(&w.cursors[4], 1);
    // the bulk store that last read the tile buffer about to be reused (two tiles ago) must have drained
    if (n_tiles >= 2 && tid == 0) bulk_wait_read<1>();
    consumer_sync();
    const int item = s_item;
    if (item >= n_busy) break;
    const int id = __ldg(w.busy_list + (n_planes * w.nstrips - 1 - item)).x;
    const int bf = id / w.nstrips;
    const StripGeo g = strip_geo((size_t)hr_off + (size_t)bf * hh * hw, id - bf * w.nstrips, rshift, hh, hw);
    float *tile = tiles + (size_t)(n_tiles & 1) * buf_floats;
    for (int i = tid; i < (buf_floats >> 2); i += kSplatThreads)
      reinterpret_cast<float4 *>(tile)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    // saturation map of the strip: the patches that do not exist count as saturated, so the strip is
    // finished when every word reads all ones (increments are >= 0 and the result is clamped at 1)
    if (tid < kMaxPatches / 32) {
      const int first = tid << 5;
      s_sat[tid] = n_patches >= first + 32 ? 0u : (n_patches <= first ? 0xffffffffu : ~((1u << (n_patches - first)) - 1u));
    }
    bool strip_done = false;
    const int nsrc = min(__ldg(w.n_src + bf), src_cap);
    const float4 *ga_src = w.src_a + (size_t)bf * src_cap;
    const int4 *gb_src = w.src_b + (size_t)bf * src_cap;
    const int py = g.r0 + row;
    const int tile_row = (int)(g.gs - g.ga) + row * hw;

    // long source lists (crowded scenes): the range of kBoundBlock-runs whose rows meet the strip
    int i_lo = 0, i_hi = nsrc;
    if (nsrc > kSplatThreads) {
      const int2 *bounds = w.blk + (size_t)bf * w.blk_cap;
      int first = 0x7fffffff, last = -1;
      for (int blk = tid; blk * kBoundBlock < nsrc; blk += kSplatThreads) {
        const int2 b = bounds[blk];
        if (b.x < g.r1 && b.y > g.r0) { first = min(first, blk); last = max(last, blk); }
      }
      first = __reduce_min_sync(0xffffffffu, first);
      last = __reduce_max_sync(0xffffffffu, last);
      if (lane == 0) { s_first[warp] = first; s_last[warp] = last; }
      consumer_sync();
      first = 0x7fffffff; last = -1;
#pragma unroll
      for (int q = 0; q < kSplatWarps; q++) { first = min(first, s_first[q]); last = max(last, s_last[q]); }
      i_lo = last < 0 ? nsrc : first * kBoundBlock;
      i_hi = last < 0 ? nsrc : min(nsrc, (last + 1) * kBoundBlock);
    }
    for (int base = i_lo; base < i_hi && !strip_done; base += kSplatThreads) {
      // ordered filter of this block of sources by the strip's rows
      const int i = base + tid;
      bool keep = false, wide = false, flat = false;
      int4 bb = make_int4(0, 0, 0, 0);
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < i_hi) {
        bb = gb_src[i];
        keep = bb.z < g.r1 && bb.w > g.r0;
      }
      if (keep) {
        a = ga_src[i];
        // "flat" source: its box covers the whole strip and sigma^2 is so large that for every pixel of
        // the strip |-0.5 d2 / sigma^2| stays below 2^-22, where approx_exp() returns exactly 1
        // (1 + x/8 rounds to 1) and the d2 > sigma^2 cut cannot trigger: the increment is v for every
        // pixel.  Marked by a negative sigma^2 (sigma >= 1 always); NaN coordinates never qualify.
        if (bb.x <= 0 && bb.y >= hw && bb.z <= g.r0 && bb.w >= g.r1) {
          const float dxm = fmaxf(fabsf(a.x), fabsf((float)(hw - 1) - a.x));
          const float dym = fmaxf(fabsf((float)g.r0 - a.y), fabsf((float)(g.r1 - 1) - a.y));
          const float d2m = dxm * dxm + dym * dym;       // >= every pixel's d2 (up to rounding)
          // 2^25 leaves a factor 8 of margin; d2m < 1e38 keeps every pixel's d2 finite when sigma^2 is
          // +inf (inf / inf is NaN in the reference too)
          flat = d2m < 1e38f && d2m * 33554432.0f <= a.z;
        }
        wide = flat || ((bb.y - 1) >> cshift) - (bb.x >> cshift) >= kWidePatches;
      }
      int n_staged;
      bool any_wide;
      const int off = consumer_flag_scan(keep, wide, warp_tot, &n_staged, &any_wide, &s_mask[0][0]);   // two barriers, masks cleared
      if (keep) {
        if (flat) a.z = -a.z;                              // only ever read by the wide path below
        PP_DASSERT(off >= 0 && off < kGroupCap);
        s_a[off] = a;
        s_b[off] = bb;
      }
      if (n_staged == 0) continue;

      if (!any_wide && n_staged <= kChunk) {
        // the usual case, one chunk of narrow boxes: the thread that staged a source also marks the patches
        // it touches (the masks were cleared inside the scan), so one barrier separates staging from the
        // patch loop; the next scan's first barrier ends the patch loop
        if (keep) {
          const int p0 = bb.x >> cshift, p1 = (bb.y - 1) >> cshift;
          uint32_t *word = s_mask[off >> 5];
          const uint32_t bit = 1u << (off & 31);
          for (int p = p0; p <= p1; p++) atomicOr(&word[p], bit);
        }
        consumer_sync();
        for (int p = warp; p < n_patches; p += kSplatWarps) {
          unsigned long long m = (unsigned long long)s_mask[0][p] | ((unsigned long long)s_mask[1][p] << 32);
          if (m == 0ull) continue;
          const int px = (p << cshift) + cg;
          const bool ok = py < g.r1 && px < hw;
          float acc = ok ? tile[tile_row + px] : 0.f;
          int seen = 0;
          while (m) {
            const int s = __ffsll((long long)m) - 1;
            m &= m - 1ull;
            splat_px(acc, px, py, s_a[s], s_b[s]);
            if ((++seen & 7) == 0) {
              // once every pixel of the patch reached 1 nothing can change any more
              if (__all_sync(0xffffffffu, !(acc < 1.f))) break;
            }
          }
          if (ok) tile[tile_row + px] = clamp1(acc);
        }
        continue;
      }
      if (!any_wide) {
        // more than one chunk of narrow boxes: one thread per source marks the few patches it touches
        for (int c0 = 0; c0 < n_staged; c0 += kChunk) {
          const int nchunk = min(kChunk, n_staged - c0);
          for (int p = tid; p < 2 * kMaxPatches; p += kSplatThreads) (&s_mask[0][0])[p] = 0u;
          consumer_sync();                                 // staged list + cleared masks visible
          if (tid < nchunk) {
            const int4 b = s_b[c0 + tid];
            const int p0 = b.x >> cshift, p1 = (b.y - 1) >> cshift;
            uint32_t *word = s_mask[tid >> 5];
            const uint32_t bit = 1u << (tid & 31);
            for (int p = p0; p <= p1; p++) atomicOr(&word[p], bit);
          }
          consumer_sync();
          for (int p = warp; p < n_patches; p += kSplatWarps) {
            unsigned long long m = (unsigned long long)s_mask[0][p] | ((unsigned long long)s_mask[1][p] << 32);
            if (m == 0ull) continue;
            const int px = (p << cshift) + cg;
            const bool ok = py < g.r1 && px < hw;
            float acc = ok ? tile[tile_row + px] : 0.f;
            int seen = 0;
            while (m) {
              const int s = __ffsll((long long)m) - 1;
              m &= m - 1ull;
              splat_px(acc, px, py, s_a[c0 + s], s_b[c0 + s]);
              if ((++seen & 7) == 0) {
                // once every pixel of the patch reached 1 nothing can change any more
                if (__all_sync(0xffffffffu, !(acc < 1.f))) break;
              }
            }
            if (ok) tile[tile_row + px] = clamp1(acc);
          }
          consumer_sync();                                 // patches done before masks are cleared again
        }
        continue;
      }

      // blocks with boxes spanning many patches (whole-map sources of degenerate fields): one thread per
      // patch builds its mask from the chunk (no atomics), patches whose 32 pixels all reached 1 are
      // remembered in s_sat and skipped, and the strip ends as soon as every patch is saturated
      for (int c0 = 0; c0 < n_staged; c0 += kChunk) {
        const int nchunk = min(kChunk, n_staged - c0);
        consumer_sync();                                   // staged list visible
        for (int p = tid; p < n_patches; p += kSplatThreads) {
          unsigned long long m = 0ull;
          if (!((s_sat[p >> 5] >> (p & 31)) & 1u)) {
            for (int q = 0; q < nchunk; q++) {
              const int4 b = s_b[c0 + q];
              if (p >= (b.x >> cshift) && p <= ((b.y - 1) >> cshift)) m |= 1ull << q;
            }
          }
          s_mask[0][p] = (uint32_t)m;
          s_mask[1][p] = (uint32_t)(m >> 32);
        }
        consumer_sync();
        for (int p = warp; p < n_patches; p += kSplatWarps) {
          unsigned long long m = (unsigned long long)s_mask[0][p] | ((unsigned long long)s_mask[1][p] << 32);
          if (m == 0ull) continue;
          const int px = (p << cshift) + cg;
          const bool ok = py < g.r1 && px < hw;
          float acc = ok ? tile[tile_row + px] : 1.f;      // lanes outside the map never hold a patch open
          int seen = 0;
          bool sat = false;
          while (m) {
            const int s = __ffsll((long long)m) - 1;
            m &= m - 1ull;
            const float4 sa = s_a[c0 + s];
            if (sa.z < 0.f) acc = __fadd_rn(acc, sa.w);    // flat source
            else splat_px(acc, px, py, sa, s_b[c0 + s]);
            if ((++seen & 7) == 0) {
              sat = __all_sync(0xffffffffu, !(acc < 1.f));
              if (sat) break;
            }
          }
          if (ok) tile[tile_row + px] = clamp1(acc);
          if (sat && lane == 0) atomicOr(&s_sat[p >> 5], 1u << (p & 31));
        }
        consumer_sync();                                   // patches done before the masks are rebuilt
        bool all = true;
#pragma unroll
        for (int q = 0; q < kMaxPatches / 32; q++) all = all && (s_sat[q] == 0xffffffffu);
        if (all) { strip_done = true; break; }             // uniform: every consumer reads the same words
      }
    }
    if (merge) {          // np.maximum with the map of the scales before (NaN from either side wins)
      consumer_sync();    // the last patch loop may have ended without a barrier
      for (size_t i = g.gs + tid; i < g.ge; i += kSplatThreads) {
        const float a = tile[i - g.ga], o = hr[i];
        tile[i - g.ga] = (a != a) ? a : ((o != o) ? o : fmaxf(a, o));
      }
    }
    fence_async_smem();
    consumer_sync();
    emit_strip(hr, tile, g);
    n_tiles++;
  }
  if (tid == 0) bulk_wait_read<0>();
}

// ---------------------------------------------------------------------------
// K2': the strips with narrow boxes (every strip of an ordinary pose field), ONE WARP PER STRIP and no CTA
// barrier anywhere.  The CTA-per-strip kernel above spends a busy strip's time in five CTA barriers and a
// chain of four dependent global loads (7.6 us per strip measured at one CTA per SM) for ~25 (patch, source)
// evaluations; here a warp owns a tile of shared memory, walks the strip's sources in list order and adds
// each source to the pixels of its box (lanes = pixels of the box, read-modify-write of the tile, clamp
// after every add exactly like functional.pyx:140-141), and sends the tile off with one bulk store.  The
// next work-list entry is fetched one strip ahead by lane 0, so the only load latency on a strip's path is
// that of its source records.  The producer warp writes the empty strips, one bulk store each from a
// strip-sized zero buffer.  Per-pixel add order = list order, as in the reference.
// the increment of one pixel inside the source's box, branch-free so that several pixels' chains interleave:
// the same operations as splat_core (functional.pyx:127-140), selects instead of branches.  dx, dx2: the
// column part, shared by the pixels a lane owns in one pass.  Returns the new value of the pixel.
__device__ __forceinline__ float splat_select(float acc, float dx2, int py, const float4 a) {
  const float dy = __fsub_rn((float)py, a.y);
  const float dy2 = __fmul_rn(dy, dy);
  const float d2 = __fadd_rn(dx2, dy2);
  const float arg = __fdiv_rn(__fmul_rn(-0.5f, d2), a.z);
  float e = __fadd_rn(1.0f, __fmul_rn(arg, 0.125f));               // approx_exp, functional.pyx:57-64
  e = __fmul_rn(e, e);
  e = __fmul_rn(e, e);
  e = __fmul_rn(e, e);
  e = (arg > 2.0f || arg < -2.0f) ? 0.0f : e;
  const float vv = (dx2 < 0.25f && dy2 < 0.25f) ? a.w : __fmul_rn(a.w, e);   // :134-138
  const float sum = __fadd_rn(acc, vv);                             // :140
  return (d2 > a.z) ? acc : sum;                                    // :131
}

constexpr int kStripIlp = 4;      // rows per lane in flight (two when the box has few rows in the strip)

// ILP rows per lane in flight: rows r0 + u * rpr + rl of the box, column cx (x0 + cx in the map)
template <int ILP>
__device__ __forceinline__ void strip_pass(float *tile, int row0, int hw, int r0, int rpr, int rl, int nrows,
                                           int cx, bool col_ok, float dx2, int y0, const float4 a) {
  float acc[ILP];
  int idx[ILP];
  bool on[ILP];
#pragma unroll
  for (int u = 0; u < ILP; u++) {
    const int ry = r0 + u * rpr + rl;
    on[u] = col_ok && ry < nrows;
    idx[u] = row0 + ry * hw + cx;
    acc[u] = on[u] ? tile[idx[u]] : 1.0f;
  }
#pragma unroll
  for (int u = 0; u < ILP; u++) {
    const int ry = r0 + u * rpr + rl;
    // at 1 nothing can change any more (the reference clamps after every add, :141); NaN -> 1
    const float v = clamp1(splat_select(acc[u], dx2, y0 + ry, a));
    if (on[u] && acc[u] < 1.0f) tile[idx[u]] = v;
  }
}

// Lanes form a (rows x columns) grid over the source's box with a power-of-two column count (no division):
// a box of 9..16 columns puts two rows on the warp, and four rows per lane in flight cover the 8 rows of a
// strip in one pass.
__device__ __forceinline__ void strip_flush(float *tile, int tile0, const float4 *sa, const int4 *sb, int n,
                                            const StripGeo &g, int hw, int lane) {
  __syncwarp();                                  // the staged records are visible to every lane
  if (n == 0) return;
  float4 a_next = sa[0];
  int4 b_next = sb[0];
  for (int s = 0; s < n; s++) {
    const float4 a = a_next;
    const int4 b = b_next;
    if (s + 1 < n) { a_next = sa[s + 1]; b_next = sb[s + 1]; }   // the staged records do not change here
    const int x0 = max(b.x, 0), wd = min(b.y, hw) - x0;
    const int y0 = max(b.z, g.r0), nrows = min(b.w, g.r1) - y0;
    const int sh = wd > 16 ? 5 : (wd > 8 ? 4 : (wd > 4 ? 3 : (wd > 2 ? 2 : (wd > 1 ? 1 : 0))));
    const int rx = lane & ((1 << sh) - 1), rl = lane >> sh, rpr = 32 >> sh;
    const int row0 = tile0 + (y0 - g.r0) * hw + x0;
    for (int cb = 0; cb < wd; cb += 32) {          // one pass unless the box is wider than 32 columns
      const int cx = cb + rx;
      const float dx = __fsub_rn((float)(x0 + cx), a.x);
      const float dx2 = __fmul_rn(dx, dx);
      if (nrows <= 2 * rpr) {
        strip_pass<2>(tile, row0, hw, 0, rpr, rl, nrows, cx, cx < wd, dx2, y0, a);
      } else {
        for (int r0 = 0; r0 < nrows; r0 += rpr * kStripIlp)
          strip_pass<kStripIlp>(tile, row0, hw, r0, rpr, rl, nrows, cx, cx < wd, dx2, y0, a);
      }
    }
    __syncwarp();                                // the next source may touch the same pixels from other lanes
  }
}

__global__ void __launch_bounds__(32 * (kStripWarpsMax + 1), 1)
cifhr_strips_kernel(float *__restrict__ hr, CifHrWs w, int n_planes, int kGroupStrips, int src_cap,
                    int hh, int hw, int buf_floats, int merge, int depth, int hr_off) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int n_warps = (int)(blockDim.x >> 5) - 1;         // consumer warps; the last warp is the producer
  float *zero = reinterpret_cast<float *>(smem_raw);      // one strip of zeros
  float *tiles = zero + buf_floats;                       // one tile per consumer warp
  float4 *st_a = reinterpret_cast<float4 *>(tiles + (size_t)n_warps * buf_floats);
  int4 *st_b = reinterpret_cast<int4 *>(st_a + n_warps * kStripStage);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rshift = w.rshift;
  for (int i = tid; i < (buf_floats >> 2); i += blockDim.x)
    reinterpret_cast<float4 *>(zero)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  fence_async_smem();
  __syncthreads();

  if (warp == n_warps) {
    // ---- producer warp: every empty strip as one TMA store of zeros -----------------
    if (merge) return;     // np.maximum merge of a further scale: its empty strips leave the map alone
    const int n_groups = (w.nstrips + kGroupStrips - 1) / kGroupStrips;
    const int total = n_planes * n_groups;
    // the work item after this one and its strip counts are in flight while this one's stores are issued
    auto counts = [&](int item) {
      const int bf = item / n_groups;
      const int strip0 = (item - bf * n_groups) * kGroupStrips;
      const int nloc = min(kGroupStrips, w.nstrips - strip0);
      return (item < total && lane < nloc) ? __ldg(w.strip_cnt + (size_t)bf * w.nstrips + strip0 + lane) : 1;
    };
    int it_b = 0, item = 0;
    if (lane == 0) {
      item = atomicAdd(&w.cursors[1], 1);
      it_b = atomicAdd(&w.cursors[1], 1);
    }
    item = __shfl_sync(0xffffffffu, item, 0);
    int cnt = counts(item);
    for (;;) {
      if (item >= total) break;
      const int item_next = __shfl_sync(0xffffffffu, it_b, 0);     // asked for one round ago
      if (lane == 0) it_b = atomicAdd(&w.cursors[1], 1);
      const int cnt_next = counts(item_next);
      const int bf = item / n_groups;
      const int strip0 = (item - bf * n_groups) * kGroupStrips;
      const size_t plane0 = (size_t)hr_off + (size_t)bf * hh * hw;
      unsigned empty = __ballot_sync(0xffffffffu, cnt == 0);
      for (; empty; empty &= empty - 1u) {
        const StripGeo g = strip_geo(plane0, strip0 + __ffs(empty) - 1, rshift, hh, hw);
        if (lane == 0) {
          if (g.m1 > g.m0) bulk_store(hr + g.m0, zero, (uint32_t)((g.m1 - g.m0) * sizeof(float)));
          // a short queue: the SM's bulk-copy unit serves requests in order, and the consumers' tiles
          // wait behind whatever the producer has queued
          if (depth <= 2) bulk_wait_read<2>(); else if (depth <= 4) bulk_wait_read<4>();
          else if (depth <= 8) bulk_wait_read<8>(); else if (depth <= 16) bulk_wait_read<16>();
        } else if (lane < 5) {
          const size_t i = g.gs + (size_t)(lane - 1);
          if (i < g.m0) hr[i] = 0.f;
        } else if (lane < 9) {
          const size_t i = g.m1 + (size_t)(lane - 5);
          if (i < g.ge) hr[i] = 0.f;
        }
      }
      item = item_next;
      cnt = cnt_next;
    }
    if (lane == 0) bulk_wait_read<0>();     // smem must outlive the reads of the async proxy
    return;
  }

  // ---- consumer warps ----------------------------------------------------------------
  const int n_items = __ldg(w.cursors);
  float *tile = tiles + (size_t)warp * buf_floats;
  float4 *sa = st_a + warp * kStripStage;
  int4 *sb = st_b + warp * kStripStage;
  // lane 0 holds the work list entry of this strip and has the next one in flight
  int it_a = 0, it_b = 0;
  int2 ent = make_int2(0, 0);
  if (lane == 0) {
    it_a = atomicAdd(&w.cursors[2], 1);
    it_b = atomicAdd(&w.cursors[2], 1);
    if (it_a < n_items) ent = __ldg(w.busy_list + it_a);
  }
  for (;;) {
    const int item = __shfl_sync(0xffffffffu, it_a, 0);
    if (item >= n_items) break;
    const int id = __shfl_sync(0xffffffffu, ent.x, 0);
    const int nsrc = min(__shfl_sync(0xffffffffu, ent.y, 0), src_cap);
    if (lane == 0) {
      it_a = it_b;                                              // arrived while the last strip was computed
      if (it_a < n_items) ent = __ldg(w.busy_list + it_a);      // used at the top of the next round
      it_b = atomicAdd(&w.cursors[2], 1);                       // used as an address one round later
    }
    const int bf = id / w.nstrips;
    const StripGeo g = strip_geo((size_t)hr_off + (size_t)bf * hh * hw, id - bf * w.nstrips, rshift, hh, hw);
    const float4 *ga_src = w.src_a + (size_t)bf * src_cap;
    const int4 *gb_src = w.src_b + (size_t)bf * src_cap;
    const int tile0 = (int)(g.gs - g.ga);

    // long source lists (crowded scenes): the range of kBoundBlock-runs whose rows meet the strip
    int i_lo = 0, i_hi = nsrc;
    if (nsrc > 32 * kStripBatch) {
      const int2 *bounds = w.blk + (size_t)bf * w.blk_cap;
      int first = 0x7fffffff, last = -1;
      for (int blk = lane; blk * kBoundBlock < nsrc; blk += 32) {
        const int2 b = bounds[blk];
        if (b.x < g.r1 && b.y > g.r0) { first = min(first, blk); last = max(last, blk); }
      }
      first = __reduce_min_sync(0xffffffffu, first);
      last = __reduce_max_sync(0xffffffffu, last);
      i_lo = last < 0 ? nsrc : first * kBoundBlock;
      i_hi = last < 0 ? nsrc : min(nsrc, (last + 1) * kBoundBlock);
    }
    // kStripBatch x 32 records per round of loads (one load latency per 128 sources); the first round is
    // on its way while the last strip's store drains and the tile is cleared
    int4 bb[kStripBatch];
    float4 aa[kStripBatch];
    auto load_batch = [&](int base) {
#pragma unroll
      for (int j = 0; j < kStripBatch; j++) {
        const int i = base + 32 * j + lane;
        bb[j] = make_int4(0, 0, 0, 0);           // rows [0, 0): never kept
        aa[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < i_hi) { bb[j] = gb_src[i]; aa[j] = ga_src[i]; }
      }
    };
    load_batch(i_lo);
    if (lane == 0) bulk_wait_read<0>();          // the store of this warp's last strip has finished reading the tile
    __syncwarp();
    for (int i = lane; i < (buf_floats >> 2); i += 32)
      reinterpret_cast<float4 *>(tile)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    int n_st = 0;
    for (int base = i_lo; base < i_hi; base += 32 * kStripBatch) {
      if (base != i_lo) load_batch(base);
#pragma unroll
      for (int j = 0; j < kStripBatch; j++) {
        const bool keep = bb[j].z < g.r1 && bb[j].w > g.r0;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m == 0u) continue;
        if (n_st + __popc(m) > kStripStage) {
          strip_flush(tile, tile0, sa, sb, n_st, g, hw, lane);
          n_st = 0;
        }
        if (keep) {
          const int k = n_st + __popc(m & ((1u << lane) - 1u));
          PP_DASSERT(k < kStripStage);
          sa[k] = aa[j];
          sb[k] = bb[j];
        }
        n_st += __popc(m);
      }
    }
    strip_flush(tile, tile0, sa, sb, n_st, g, hw, lane);
    if (merge) {          // np.maximum with the map of the scales before (NaN from either side wins)
      for (size_t i = g.gs + lane; i < g.ge; i += 32) {
        const float a = tile[i - g.ga], o = hr[i];
        tile[i - g.ga] = (a != a) ? a : ((o != o) ? o : fmaxf(a, o));
      }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      if (g.m1 > g.m0) bulk_store(hr + g.m0, tile + (g.m0 - g.ga), (uint32_t)((g.m1 - g.m0) * sizeof(float)));
    } else if (lane < 5) {          // head: < 4 floats before the aligned middle
      const size_t i = g.gs + (size_t)(lane - 1);
      if (i < g.m0) hr[i] = tile[i - g.ga];
    } else if (lane < 9) {          // tail
      const size_t i = g.m1 + (size_t)(lane - 5);
      if (i < g.ge) hr[i] = tile[i - g.ga];
    }
    __syncwarp();                   // the head / tail lanes have read the tile before it is cleared again
  }
  if (lane == 0) bulk_wait_read<0>();
}

// s: shape of the field set (B, F, H, W, stride); s_hr: the shape that defines the hr map (set 0 of a
// multi-scale configuration, else s itself).  d_cif2: second field of an hflip pair or NULL.
// merge: np.maximum into the existing map instead of overwriting it.
CifSrcView cifhr_sources(void *d_ws, const pp_shape &s_hr, const pp_caps &c) {
  CifHrWs w = carve(d_ws, s_hr, c);
  return CifSrcView{w.src_a, w.src_c, w.n_src, c.src_cap};
}

int cifhr_launch_ex(const float *d_cif, const float *d_cif2, float *d_hr, void *d_ws, const pp_shape &s,
                    const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_scale,
                    int *d_status, cudaStream_t st, int det, int merge, const float *cif_rest,
                    const CafPreArgs *caf_pre) {
  CifHrWs w = carve(d_ws, s_hr, c, p.cifhr_rshift);
  const int hh = hr_h(s_hr), hw = hr_w(s_hr);
  const int bf = s.B * s.F;
  if (((size_t)hw << w.rshift) * sizeof(float) > (size_t)kStripBytesMax - 32) return PP_ERR_ARG;
  // bulk stores need 16-byte alignment: the kernels index from the 16-byte boundary below the map (a
  // sub-batch slice of a larger map starts at any multiple of 4 bytes) and never write below the map itself
  if (((uintptr_t)d_hr & 3u) != 0) return PP_ERR_ARG;
  const int hr_off = (int)(((uintptr_t)d_hr & 15u) >> 2);
  d_hr -= hr_off;
  // (the bulk loads of the confidence planes round their range out to 16-byte boundaries: field tensors
  // must live in allocations whose base is 16-byte aligned -- cudaMalloc gives 256 -- see pifpaf_b200.h)
  const size_t smem = scan_smem_bytes(w.nstrips, s.H * s.W);
  // pp_params.cifhr_warps: consumer warps per CTA of the strips kernel; < 0 = the CTA-per-strip kernel
  // does every strip (the round-1 schedule, kept for comparison)
  const bool legacy = p.cifhr_warps < 0;
  // default: 3 warps at 8-row strips (the thinner kernel wins with batches in flight), 4 where the strips are
  // 4 rows or fewer (images wider than 767 px: twice the strips, each meeting as many sources; measured on
  // 16 x 1281 px / 50 persons: 25.7 k -> 27-28 k img/s, profiles/r3_strips_sweeps.md)
  int n_warps = (p.cifhr_warps >= 1 && p.cifhr_warps <= kStripWarpsMax) ? p.cifhr_warps
                : (w.rshift < 3 ? kStripWarpsDefault + 1 : kStripWarpsDefault);
  prof_mark(p.profile, kStCompact, st);
  PP_CUDA_CHECK(cudaMemsetAsync(w.cursors, 0, 8 * sizeof(int), st));
  CompactArgs ca = {d_cif, d_cif2, cif_rest, w, s, c.src_cap, hh, hw, p.cif_threshold,
                    min_scale / (float)s.stride, min_scale != 0.0f, det, d_status,
                    p.cif_neighbors > 0.0f ? p.cif_neighbors : 16.0f, bf, legacy ? 1 : 0};
  CafPreArgs pa = {};
  int n_caf = 0;
  if (caf_pre) {
    pa = *caf_pre;
    n_caf = s.B * pa.L;
  }
  PP_CUDA_CHECK(cudaFuncSetAttribute(fields_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)scan_smem_bytes(4096, kCompactSlab)));
#ifdef PP_DEBUG_SKIP
  // timing / fault-isolation experiments only: 32 = no strips / splat kernels, 64 = no field scan
  const int dbg_skip = getenv("PP_DEBUG_SKIP") ? atoi(getenv("PP_DEBUG_SKIP")) : 0;
#else
  const int dbg_skip = 0;
#endif
  if (!(dbg_skip & 64)) fields_scan_kernel<<<bf + n_caf, kCompactThreads, smem, st>>>(ca, pa, bf);
  PP_LAUNCH_CHECK("fields_scan_kernel");
  if (dbg_skip & 32) return gate_signal(p.splat_gate, st);
  // one strip = R*hw floats, plus up to 3 floats of misalignment, rounded to 16 bytes
  const int buf_floats = (int)(((((size_t)hw << w.rshift) + 3 + 3) & ~(size_t)3));
  prof_mark(p.profile, kStSplat, st);       // crumb 11
  int rc = gate_wait(p.splat_gate, st);       // batches in flight: one splat at a time (pp_gate)
  if (rc) return rc;
  const int sms = num_sms();
  if (!legacy) {
    // the strips kernel: empty strips (producer warp) + narrow strips (one warp each).  Persistent, one
    // CTA per SM by default: n_warps tiles + the zero strip + the staged records
    constexpr size_t kStripsSmemMax = 226 * 1024;      // 227 KB per CTA minus the kernel's (tiny) static part
    auto strips_smem = [&](int nw) {
      return ((size_t)(nw + 1) * buf_floats) * sizeof(float) + (size_t)nw * kStripStage * (sizeof(float4) + sizeof(int4));
    };
    while (n_warps > 1 && strips_smem(n_warps) > kStripsSmemMax) n_warps--;
    const size_t dyn = strips_smem(n_warps);
    if (dyn > kStripsSmemMax) return PP_ERR_ARG;
    PP_CUDA_CHECK(cudaFuncSetAttribute(cifhr_strips_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)kStripsSmemMax));
    const int group = (p.cifhr_group >= 1 && p.cifhr_group <= kGroupStripsMax) ? p.cifhr_group : 8;
    const int n_groups = (w.nstrips + group - 1) / group;
    const int per_sm = p.cifhr_ctas_per_sm > 0 ? p.cifhr_ctas_per_sm : 1;
    int grid = sms * per_sm;
    if (grid > bf * n_groups) grid = bf * n_groups;
    cifhr_strips_kernel<<<grid, 32 * (n_warps + 1), dyn, st>>>(d_hr, w, bf, group, c.src_cap, hh, hw,
                                                                buf_floats, merge, p.cifhr_depth > 0 ? p.cifhr_depth : 8, hr_off);
    PP_LAUNCH_CHECK("cifhr_strips_kernel");
  }
  debug_crumb(st, 30);
  {
    // the CTA-per-strip kernel: the strips that meet wide boxes or very long lists (degenerate fields);
    // with ordinary pose fields its list is empty and every CTA leaves at once
    const size_t dyn = ((size_t)kZeroFloats + 2 * (size_t)buf_floats) * sizeof(float);
    // per device and cheap: set on every launch rather than caching per process
    PP_CUDA_CHECK(cudaFuncSetAttribute(cifhr_splat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       2 * kStripBytesMax + kZeroFloats * (int)sizeof(float)));
    const int kGroupStrips = legacy ? group_strips(p) : 3;
    const int n_groups = (w.nstrips + kGroupStrips - 1) / kGroupStrips;
    const int per_sm = (legacy && p.cifhr_ctas_per_sm > 0) ? p.cifhr_ctas_per_sm : 3;
    int grid = sms * per_sm;
    if (grid > bf * n_groups) grid = bf * n_groups;
    cifhr_splat_kernel<<<grid, kSplatBlock, dyn, st>>>(d_hr, w, bf, kGroupStrips, c.src_cap, hh, hw,
                                                      buf_floats, merge, legacy ? 1 : 0, hr_off);
    PP_LAUNCH_CHECK("cifhr_splat_kernel");
  }
  return gate_signal(p.splat_gate, st);
}

int cifhr_launch(const float *d_cif, float *d_hr, void *d_ws, const pp_shape &s, const pp_caps &c,
                 const pp_params &p, int *d_status, cudaStream_t st, int det) {
  return cifhr_launch_ex(d_cif, nullptr, d_hr, d_ws, s, s, c, p, p.cif_min_scale, d_status, st, det, 0,
                         nullptr, nullptr);
}

}  // namespace pp
