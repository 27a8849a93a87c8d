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
// One CTA per (image, link).  The confidence plane is streamed with coalesced
// loads; the other eight planes are touched only for cells above threshold.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kCafThreads = 256;

__global__ void __launch_bounds__(kCafThreads)
caf_scored_kernel(const float *__restrict__ caf, const float *__restrict__ hr, pp_skeleton sk,
                  float *__restrict__ fwd, int *__restrict__ nfwd, float *__restrict__ bwd,
                  int *__restrict__ nbwd, pp_shape s, int cap, int hh, int hw, float th,
                  float cif_floor, float cif_floor_compl, float min_dist, int use_min, float max_dist,
                  int use_max, int *__restrict__ status) {
  __shared__ int wtot_f[kCafThreads / 32], wtot_b[kCafThreads / 32];
  const int bl = blockIdx.x;
  const int b = bl / s.L, l = bl - b * s.L;
  const int n = s.H * s.W;
  const float *p = caf + (size_t)bl * 9 * n;
  const float stride = (float)s.stride;
  const int j1 = sk.joints[l][0] - 1, j2 = sk.joints[l][1] - 1;
  const float *hr_b = hr + (size_t)b * s.F * hh * hw;
  float *fo = fwd + (size_t)bl * 9 * cap;
  float *bo = bwd + (size_t)bl * 9 * cap;
  const int lane = lane_id(), w = warp_id(), nw = blockDim.x >> 5;
  int cnt_f = 0, cnt_b = 0;

  for (int base = 0; base < n; base += blockDim.x) {
    const int c = base + threadIdx.x;
    bool kf = false, kb = false;
    float nine[9];
    float sf = 0.f, sb = 0.f;
    if (c < n) {
      nine[0] = __ldg(p + c);
      if (nine[0] > th) {                                         // caf_scored.py:42
#pragma unroll
        for (int r = 1; r < 9; r++) nine[r] = __ldg(p + (size_t)r * n + c);
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
    }
    // two ordered CTA-wide compactions in one barrier pair
    unsigned mf = __ballot_sync(0xffffffffu, kf), mb = __ballot_sync(0xffffffffu, kb);
    if (lane == 0) { wtot_f[w] = __popc(mf); wtot_b[w] = __popc(mb); }
    __syncthreads();
    int base_f = 0, base_b = 0, tot_f = 0, tot_b = 0;
    for (int i = 0; i < nw; i++) {
      int tf = wtot_f[i], tb = wtot_b[i];
      if (i < w) { base_f += tf; base_b += tb; }
      tot_f += tf; tot_b += tb;
    }
    __syncthreads();
    const unsigned below = (1u << lane) - 1u;
    if (kb) {
      int k = cnt_b + base_b + __popc(mb & below);
      if (k < cap) {                                              // :68-69 row permutation
        bo[k] = sb;
        bo[(size_t)1 * cap + k] = nine[5]; bo[(size_t)2 * cap + k] = nine[6];
        bo[(size_t)3 * cap + k] = nine[7]; bo[(size_t)4 * cap + k] = nine[8];
        bo[(size_t)5 * cap + k] = nine[1]; bo[(size_t)6 * cap + k] = nine[2];
        bo[(size_t)7 * cap + k] = nine[3]; bo[(size_t)8 * cap + k] = nine[4];
      }
    }
    if (kf) {
      int k = cnt_f + base_f + __popc(mf & below);
      if (k < cap) {                                              // :79-80
        fo[k] = sf;
#pragma unroll
        for (int r = 1; r < 9; r++) fo[(size_t)r * cap + k] = nine[r];
      }
    }
    cnt_f += tot_f;
    cnt_b += tot_b;
  }
  if (threadIdx.x == 0) {
    if (cnt_f > cap || cnt_b > cap) atomicOr(&status[b], PP_STATUS_CAF_OVERFLOW);
    nfwd[bl] = min(cnt_f, cap);
    nbwd[bl] = min(cnt_b, cap);
  }
}

int caf_scored_launch(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                      float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                      const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st) {
  const int hh = hr_h(s), hw = hr_w(s);
  const float stride = (float)s.stride;
  prof_mark(p.profile, kStCaf, st);
  caf_scored_kernel<<<s.B * s.L, kCafThreads, 0, st>>>(
      d_caf, d_hr, sk, d_fwd, d_nfwd, d_bwd, d_nbwd, s, c.caf_cap, hh, hw, score_th, p.cif_floor,
      p.cif_floor_compl, p.caf_min_distance / stride, p.caf_min_distance != 0.0f,
      p.caf_max_distance / stride, p.caf_max_distance > 0.0f, d_status);
  PP_LAUNCH_CHECK("caf_scored_kernel");
  return PP_OK;
}

}  // namespace pp
