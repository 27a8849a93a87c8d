// cifhr.cu -- CifHr high-resolution accumulation on sm_100a.
//
// Replaces decoder/cif_hr.py:26-65 (CifHr.accumulate / fill_multiple) and
// functional.pyx:105-141 (scalar_square_add_gauss_with_max) of the reference.
//
// The reference walks the sources of one field sequentially and adds a clipped
// Gaussian box per source into the hr map, clamping at 1 after every add.  All
// increments are >= 0, so the per-step clamp equals one final clamp and only
// the ORDER of the float adds per pixel matters (ascending source index =
// row-major cell order).  We therefore run it as a gather:
//
//   K1 cif_compact_kernel   one CTA per (image, field): ordered compaction of
//                           the cells with conf > v_threshold into a source
//                           list {cx, cy, sigma^2, v} + integer box, plus a
//                           per-strip source count (difference array in smem).
//   K2 cifhr_splat_kernel   persistent CTAs over (image, field, strip of R hr
//                           rows).  A strip is one contiguous run of floats in
//                           HBM.  Empty strips (the vast majority) are pure
//                           128-bit zero stores; the others filter the field's
//                           source list into shared memory and every thread
//                           accumulates its own 4 consecutive pixels over that
//                           list in order, in registers, then stores once.
//
// No atomics on hr, no read-modify-write of hr, no separate memset: every hr
// byte is written exactly once.  Result is bit-identical to the reference.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kStripRows = 8;       // hr rows per strip
constexpr int kSplatThreads = 256;
constexpr int kChunk = 512;         // sources staged in smem at a time
constexpr int kCompactThreads = 256;

struct CifHrWs {
  float4 *src_a;    // (B,F,src_cap) cx, cy, sigma^2, v
  int4 *src_b;      // (B,F,src_cap) minx, maxx, miny, maxy
  int *n_src;       // (B,F)
  int *strip_cnt;   // (B,F,nstrips)
  int nstrips;
};

__host__ __device__ inline int num_strips(int hh) { return (hh + kStripRows - 1) / kStripRows; }

size_t cifhr_ws_bytes(const pp_shape &s, const pp_caps &c) {
  size_t bf = (size_t)s.B * s.F;
  size_t n = 0;
  n += align_up(bf * c.src_cap * sizeof(float4), 256);
  n += align_up(bf * c.src_cap * sizeof(int4), 256);
  n += align_up(bf * sizeof(int), 256);
  n += align_up(bf * num_strips(hr_h(s)) * sizeof(int), 256);
  return n;
}

static CifHrWs carve(void *ws, const pp_shape &s, const pp_caps &c) {
  CifHrWs w;
  char *p = (char *)ws;
  size_t bf = (size_t)s.B * s.F;
  w.src_a = (float4 *)p; p += align_up(bf * c.src_cap * sizeof(float4), 256);
  w.src_b = (int4 *)p;   p += align_up(bf * c.src_cap * sizeof(int4), 256);
  w.n_src = (int *)p;    p += align_up(bf * sizeof(int), 256);
  w.strip_cnt = (int *)p;
  w.nstrips = num_strips(hr_h(s));
  return w;
}

// ---------------------------------------------------------------------------
// K1: ordered compaction of CIF sources per (image, field).
__global__ void __launch_bounds__(kCompactThreads)
cif_compact_kernel(const float *__restrict__ cif, CifHrWs w, pp_shape s, int src_cap, int hh, int hw,
                   float v_th, float min_scale_over_stride, int use_min_scale,
                   int *__restrict__ status) {
  extern __shared__ int sm_diff[];   // nstrips + 1 difference array
  __shared__ int warp_tot[kCompactThreads / 32];
  const int bf = blockIdx.x;
  const int n = s.H * s.W;
  const float *p = cif + (size_t)bf * 5 * n;
  const float stride = (float)s.stride;
  for (int i = threadIdx.x; i <= w.nstrips; i += blockDim.x) sm_diff[i] = 0;
  __syncthreads();

  float4 *oa = w.src_a + (size_t)bf * src_cap;
  int4 *ob = w.src_b + (size_t)bf * src_cap;
  int count = 0;
  for (int base = 0; base < n; base += blockDim.x) {
    int c = base + threadIdx.x;
    bool flag = false;
    float conf = 0.f;
    if (c < n) {
      conf = __ldg(p + c);
      flag = conf > v_th;                                         // cif_hr.py:27
      if (flag && use_min_scale) flag = __ldg(p + 4 * n + c) > min_scale_over_stride;  // :28-29
    }
    int total;
    int off = block_flag_scan(flag, warp_tot, &total);
    if (flag) {
      int k = count + off;
      if (k < src_cap) {
        float cx = __fmul_rn(__ldg(p + n + c), stride);           // :32
        float cy = __fmul_rn(__ldg(p + 2 * n + c), stride);       // :33
        float sg = __fmul_rn(__fmul_rn(0.5f, __ldg(p + 4 * n + c)), stride);
        float sigma = (sg != sg) ? sg : fmaxf(1.0f, sg);          // np.maximum(1.0, .) :34
        float cv = __fdiv_rn(conf, 16.0f);                        // v / neighbors / len_cifs :40
        // functional.pyx:122-125 with truncate = 1.0
        int minx = (int)clipf(__fsub_rn(cx, sigma), 0.0f, (float)(hw - 1));
        int maxx = (int)clipf(__fadd_rn(__fadd_rn(cx, sigma), 1.0f), (float)(minx + 1), (float)hw);
        int miny = (int)clipf(__fsub_rn(cy, sigma), 0.0f, (float)(hh - 1));
        int maxy = (int)clipf(__fadd_rn(__fadd_rn(cy, sigma), 1.0f), (float)(miny + 1), (float)hh);
        oa[k] = make_float4(cx, cy, __fmul_rn(sigma, sigma), cv);
        ob[k] = make_int4(minx, maxx, miny, maxy);
        atomicAdd(&sm_diff[miny / kStripRows], 1);
        atomicAdd(&sm_diff[(maxy - 1) / kStripRows + 1], -1);
      }
    }
    count += total;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (count > src_cap) {
      atomicOr(&status[bf / s.F], PP_STATUS_SRC_OVERFLOW);
      count = src_cap;
    }
    w.n_src[bf] = count;
    int run = 0;
    int *sc = w.strip_cnt + (size_t)bf * w.nstrips;
    for (int i = 0; i < w.nstrips; i++) {
      run += sm_diff[i];
      sc[i] = run;
    }
  }
}

// ---------------------------------------------------------------------------
// K2: per-strip gather.

__device__ __forceinline__ void splat_px(float &acc, int px, int py, const float4 a, const int4 b) {
  if (px >= b.x && px < b.y && py >= b.z && py < b.w) {
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

__device__ __forceinline__ float clamp1(float v) { return (v < 1.0f) ? v : 1.0f; }  // :141

__global__ void __launch_bounds__(kSplatThreads)
cifhr_splat_kernel(float *__restrict__ hr, CifHrWs w, int n_work, int src_cap, int hh, int hw) {
  __shared__ float4 s_a[kChunk];
  __shared__ int4 s_b[kChunk];
  __shared__ int warp_tot[kSplatThreads / 32];
  const size_t plane = (size_t)hh * hw;

  for (int work = blockIdx.x; work < n_work; work += gridDim.x) {
    const int bf = work / w.nstrips;
    const int strip = work - bf * w.nstrips;
    const int r0 = strip * kStripRows;
    const int r1 = min(r0 + kStripRows, hh);
    const size_t gp = (size_t)bf * plane;
    const size_t gs = gp + (size_t)r0 * hw;        // first float of the strip
    const size_t ge = gp + (size_t)r1 * hw;        // one past the last
    const size_t ga = gs & ~(size_t)3;             // 16-byte aligned unit grid
    const int n_units = (int)((ge - ga + 3) >> 2);
    const int cnt = __ldg(w.strip_cnt + work);

    if (cnt == 0) {
      // zero fill: the dominant case.  Interior units are 128-bit stores.
      for (int u = threadIdx.x; u < n_units; u += blockDim.x) {
        size_t g = ga + ((size_t)u << 2);
        if (g >= gs && g + 4 <= ge) {
          *reinterpret_cast<float4 *>(hr + g) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
          for (int i = 0; i < 4; i++)
            if (g + i >= gs && g + i < ge) hr[g + i] = 0.f;
        }
      }
      continue;
    }

    const int nsrc = min(__ldg(w.n_src + bf), src_cap);
    const float4 *ga_src = w.src_a + (size_t)bf * src_cap;
    const int4 *gb_src = w.src_b + (size_t)bf * src_cap;
    const bool single = nsrc <= kChunk;
    int n_staged = 0;

    for (int ubase = 0; ubase < n_units; ubase += blockDim.x) {
      const int u = ubase + threadIdx.x;
      const bool live = u < n_units;
      // dead lanes shadow the last unit so that the warp extents below stay valid
      const size_t g = ga + ((size_t)(live ? u : n_units - 1) << 2);
      // in-plane coordinates of the unit's first float (may be < gs for the head unit)
      long long pl = (long long)g - (long long)gp;
      int y0 = (int)(pl / hw);
      int x0 = (int)(pl - (long long)y0 * hw);
      int px[4], py[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        int xx = x0 + i, yy = y0;
        if (xx >= hw) { xx -= hw; yy += 1; }
        px[i] = xx; py[i] = yy;
      }
      // warp-uniform extent of the 32 units this warp owns (<= 128 pixels)
      int wy0 = __shfl_sync(0xffffffffu, py[0], 0);
      int wy1 = __shfl_sync(0xffffffffu, py[3], 31);
      int wx0 = __shfl_sync(0xffffffffu, px[0], 0);
      int wx1 = __shfl_sync(0xffffffffu, px[3], 31);
      float acc[4] = {0.f, 0.f, 0.f, 0.f};

      for (int cbase = 0; cbase < nsrc; cbase += kChunk) {
        if (!(single && ubase > 0)) {
          // stage the sources of this chunk that touch rows [r0, r1), in order
          __syncthreads();
          n_staged = 0;
          const int cend = min(cbase + kChunk, nsrc);
          for (int b0 = cbase; b0 < cend; b0 += blockDim.x) {
            int i = b0 + threadIdx.x;
            bool keep = false;
            int4 bb = make_int4(0, 0, 0, 0);
            if (i < cend) {
              bb = gb_src[i];
              keep = bb.z < r1 && bb.w > r0;
            }
            int tot;
            int off = block_flag_scan(keep, warp_tot, &tot);
            if (keep) {
              s_a[n_staged + off] = ga_src[i];
              s_b[n_staged + off] = bb;
            }
            n_staged += tot;
          }
          __syncthreads();
        }
        if (live) {
          for (int k = 0; k < n_staged; k++) {
            const int4 b = s_b[k];
            if (b.w <= wy0 || b.z > wy1) continue;                 // warp-uniform row reject
            if (wy0 == wy1 && (b.y <= wx0 || b.x > wx1)) continue; // and column reject
            const float4 a = s_a[k];
#pragma unroll
            for (int i = 0; i < 4; i++) splat_px(acc[i], px[i], py[i], a, b);
            if ((k & 7) == 7) {
              // saturation exit: once every pixel of the warp reached 1 nothing can change
              bool sat = !(acc[0] < 1.f) && !(acc[1] < 1.f) && !(acc[2] < 1.f) && !(acc[3] < 1.f);
              if (__all_sync(__activemask(), sat)) break;
            }
          }
        }
      }
      if (live) {
        if (g >= gs && g + 4 <= ge) {
          *reinterpret_cast<float4 *>(hr + g) =
              make_float4(clamp1(acc[0]), clamp1(acc[1]), clamp1(acc[2]), clamp1(acc[3]));
        } else {
#pragma unroll
          for (int i = 0; i < 4; i++)
            if (g + i >= gs && g + i < ge) hr[g + i] = clamp1(acc[i]);
        }
      }
    }
    __syncthreads();
  }
}

int cifhr_launch(const float *d_cif, float *d_hr, void *d_ws, const pp_shape &s, const pp_caps &c,
                 const pp_params &p, int *d_status, cudaStream_t st) {
  CifHrWs w = carve(d_ws, s, c);
  const int hh = hr_h(s), hw = hr_w(s);
  const int bf = s.B * s.F;
  size_t smem = (size_t)(w.nstrips + 1) * sizeof(int);
  prof_mark(p.profile, kStCompact, st);
  cif_compact_kernel<<<bf, kCompactThreads, smem, st>>>(
      d_cif, w, s, c.src_cap, hh, hw, p.cif_threshold, p.cif_min_scale / (float)s.stride,
      p.cif_min_scale != 0.0f, d_status);
  PP_LAUNCH_CHECK("cif_compact_kernel");
  const int n_work = bf * w.nstrips;
  int grid = num_sms() * 8;
  if (grid > n_work) grid = n_work;
  prof_mark(p.profile, kStSplat, st);
  cifhr_splat_kernel<<<grid, kSplatThreads, 0, st>>>(d_hr, w, n_work, c.src_cap, hh, hw);
  PP_LAUNCH_CHECK("cifhr_splat_kernel");
  return PP_OK;
}

}  // namespace pp
