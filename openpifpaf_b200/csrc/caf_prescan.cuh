// caf_prescan.cuh -- the hr-independent half of CafScored.fill_caf (decoder/caf_scored.py:42-58) for
// one (image, link): cells above the score threshold in row-major order, the multi-scale distance
// filters, rows 1..8 scaled by the stride.  Runs inside the CIF compact launch (cifhr.cu), before
// the hr map exists; caf_score_bin_kernel (caf_scored.cu) finishes the lists afterwards.
#pragma once
#include "common.cuh"
#include "stages.cuh"

namespace pp {

template <int THREADS, int SLAB>
__device__ __forceinline__ void caf_prescan_body(const CafPreArgs &a, const pp_shape &s, const int bl,
                                                 float *stage, unsigned short *s_cells, uint64_t *bar,
                                                 int *warp_tot) {
  const int n = s.H * s.W;
  const float *p = a.caf + (size_t)bl * 9 * n;
  // caf_rest != NULL (host-mapped input): the eight geometry planes are read from the pinned host
  // buffer, only where a cell is active; the confidence plane was copied to the device
  const float *pr = a.caf_rest ? a.caf_rest + (size_t)bl * 9 * n : p;
  const float stride = (float)s.stride;
  float *ro = a.raw + (size_t)bl * 9 * a.cap;
  const bool filtered = a.use_min || a.use_max;
  uint32_t phase = 0;
  int cnt = 0;
  for (int slab = 0; slab < n; slab += SLAB) {
    const int m = min(n, slab + SLAB) - slab;
    const float *sp = stage_floats(p + slab, m, stage, bar, phase);
    const int n_act = block_active_cells_smem<THREADS>(sp, m, a.th, s_cells, warp_tot);   // caf_scored.py:42
    for (int base = 0; base < n_act; base += THREADS) {
      const int i = base + threadIdx.x;
      bool ok = i < n_act;
      float nine[9];
      if (ok) {
        const int c = slab + s_cells[i];
        nine[0] = sp[s_cells[i]];
#pragma unroll
        for (int r = 1; r < 9; r++)
          nine[r] = (a.skip_logb && (r == 3 || r == 7)) ? 0.0f : __ldg(pr + (size_t)r * n + c);
        if (filtered) {                                               // :47-55 (multi-scale only)
          const float dx = __fsub_rn(nine[1], nine[5]), dy = __fsub_rn(nine[2], nine[6]);
          const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
          if (a.use_min && !(d > a.min_dist)) ok = false;
          if (a.use_max && !(d < a.max_dist)) ok = false;
        }
      }
      int total = min(n_act - base, THREADS), off = i - base;
      if (filtered) off = block_flag_scan(ok, warp_tot, &total);
      if (ok) {
        const int k = cnt + off;
        if (k < a.cap) {
          ro[k] = nine[0];
#pragma unroll
          for (int r = 1; r < 9; r++) ro[(size_t)r * a.cap + k] = __fmul_rn(nine[r], stride);   // :58
        }
      }
      cnt += total;
    }
    __syncthreads();        // done with the staged slab before the next copy lands in it
  }
  if (threadIdx.x == 0) a.n_raw[bl] = min(cnt, a.cap);
}

}  // namespace pp
