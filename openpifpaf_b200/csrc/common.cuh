// common.cuh -- shared device helpers for libpifpaf_b200 (sm_100a only).
//
// Parity rule for this library: every parity-critical float expression is a
// sequence of single IEEE-754 binary32 operations in the reference's order.
// The whole library is compiled with -fmad=false (no FMA contraction), and the
// places that matter additionally spell out __fmul_rn/__fadd_rn/__fdiv_rn so the
// intent survives a change of flags.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pifpaf_b200.h"
#include "../../include/pifpaf_detmath.h"

#ifndef __CUDA_ARCH__
#define PP_ARCH_OK 1
#elif __CUDA_ARCH__ >= 1000
#define PP_ARCH_OK 1
#else
#error "libpifpaf_b200 is written for sm_100a (B200) only"
#endif

// Bounds / invariant checks in device code (build with PP_EXTRA_NVCC=-DPP_DEVICE_ASSERT: a violated
// check traps, the launch fails and every test that follows reports the CUDA error).  compute-sanitizer
// is closed on the B200 pool this library is developed on; these asserts over the whole GPU test suite
// are the substitute for its memcheck (profiles/r2_device_assert.md).  Compiled out otherwise.
#ifdef PP_DEVICE_ASSERT
#include <assert.h>
#define PP_DASSERT(cond) assert(cond)
#else
#define PP_DASSERT(cond) ((void)0)
#endif

namespace pp {

constexpr int kWarp = 32;

// ---- error plumbing -------------------------------------------------------
void set_cuda_error(cudaError_t e, const char *where);

#define PP_CUDA_CHECK(expr)                          \
  do {                                               \
    cudaError_t e__ = (expr);                        \
    if (e__ != cudaSuccess) {                        \
      pp::set_cuda_error(e__, #expr);                \
      return PP_ERR_CUDA;                            \
    }                                                \
  } while (0)

#define PP_LAUNCH_CHECK(name)                        \
  do {                                               \
    cudaError_t e__ = cudaPeekAtLastError();         \
    if (e__ != cudaSuccess) {                        \
      pp::set_cuda_error(e__, name);                 \
      return PP_ERR_CUDA;                            \
    }                                                \
  } while (0)

inline int hr_h(const pp_shape &s) { return (s.H - 1) * s.stride + 1; }
inline int hr_w(const pp_shape &s) { return (s.W - 1) * s.stride + 1; }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

int num_sms();

// ---- optional per-stage timing (pp_profile, api.cu) --------------------------
enum Stage { kStCompact = 0, kStSplat, kStCollect, kStSort, kStCaf, kStAssemble, kStNms };
void prof_begin_step(pp_profile *prof);
void prof_mark(pp_profile *prof, int stage, cudaStream_t st);   // start of `stage`
void debug_crumb(cudaStream_t st, int value);                      // no-op unless built with -DPP_BREADCRUMB
void prof_end_step(pp_profile *prof, cudaStream_t st);

// ---- exact float helpers ----------------------------------------------------
// clip() of functional.pyx:67-68: fmax(minv, fmin(maxv, v)); NaN -> maxv.
__device__ __forceinline__ float clipf(float v, float lo, float hi) {
  return fmaxf(lo, fminf(hi, v));
}

// approx_exp() of functional.pyx:57-64.  The reference evaluates
// (float)(1.0 + (double)x / 8.0); x/8 is exact and the double sum is either
// exact or far from a float rounding boundary, so one float add is identical.
__device__ __forceinline__ float approx_exp(float x) {
  if (x > 2.0f || x < -2.0f) return 0.0f;
  x = __fadd_rn(1.0f, __fmul_rn(x, 0.125f));
  x = __fmul_rn(x, x);
  x = __fmul_rn(x, x);
  x = __fmul_rn(x, x);
  return x;
}

// scalar_values() of functional.pyx:231-244 for one point.
__device__ __forceinline__ float scalar_value_at(const float *__restrict__ field, int fh, int fw,
                                                 float x, float y, float dflt) {
  float maxx = (float)fw - 1.0f, maxy = (float)fh - 1.0f;
  if (x < 0.0f || y < 0.0f || x > maxx || y > maxy) return dflt;
  return __ldg(field + (size_t)(int)y * fw + (int)x);
}

// ---- warp / block primitives ---------------------------------------------
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

// Exclusive prefix of a per-thread 0/1 flag over the whole CTA, in thread
// order.  `warp_tot` is shared scratch of (blockDim.x / 32) ints.  Returns the
// thread's offset; *total receives the CTA-wide count.  Contains two
// __syncthreads().
__device__ __forceinline__ int block_flag_scan(bool flag, int *warp_tot, int *total) {
  unsigned m = __ballot_sync(0xffffffffu, flag);
  int lane = lane_id(), w = warp_id();
  int in_warp = __popc(m & ((1u << lane) - 1u));
  if (lane == 0) warp_tot[w] = __popc(m);
  __syncthreads();
  int nw = blockDim.x >> 5;
  int base = 0, tot = 0;
  for (int i = 0; i < nw; i++) {
    int t = warp_tot[i];
    if (i < w) base += t;
    tot += t;
  }
  __syncthreads();
  *total = tot;
  return base + in_warp;
}

// Ordered list of the cells c in [c_lo, c_hi) with conf[c] > th (NaN fails), CTA-wide, in two
// sweeps: every warp owns a contiguous run of cells, counts its hits (four independent loads in
// flight per lane), the warp totals give each warp its base, and the second sweep (the plane is in
// L1 / L2 by then) writes the indices.  Two __syncthreads() in total instead of two per 256 cells.
// `list` receives c - c_lo (16 bits: a slab has at most 65536 cells) and must hold c_hi - c_lo
// entries; `warp_cnt` is shared scratch of THREADS / 32 ints.
template <int THREADS>
__device__ __forceinline__ int block_active_cells(const float *__restrict__ conf, int c_lo, int c_hi,
                                                  float th, unsigned short *list, int *warp_cnt) {
  constexpr int NW = THREADS / 32;
  constexpr int U = 4;
  const int lane = lane_id(), w = warp_id();
  const int n = c_hi - c_lo;
  const int seg = (((n + NW - 1) / NW) + 31) & ~31;          // cells per warp, a multiple of 32
  const int w0 = c_lo + w * seg, w1 = min(c_hi, w0 + seg);
  int cnt = 0;
  for (int base = w0; base < w1; base += 32 * U) {
    float v[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int c = base + 32 * u + lane;
      v[u] = c < w1 ? __ldg(conf + c) : th;
    }
#pragma unroll
    for (int u = 0; u < U; u++) cnt += __popc(__ballot_sync(0xffffffffu, v[u] > th));
  }
  if (lane == 0) warp_cnt[w] = cnt;
  __syncthreads();
  int at = 0, total = 0;
#pragma unroll
  for (int i = 0; i < NW; i++) {
    const int t = warp_cnt[i];
    if (i < w) at += t;
    total += t;
  }
  const unsigned below = (1u << lane) - 1u;
  if (cnt) {
    for (int base = w0; base < w1; base += 32 * U) {
      float v[U];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int c = base + 32 * u + lane;
        v[u] = c < w1 ? __ldg(conf + c) : th;
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        const bool hit = v[u] > th;
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (hit) list[at + __popc(m & below)] = (unsigned short)(base + 32 * u + lane - c_lo);
        at += __popc(m);
      }
    }
  }
  __syncthreads();
  return total;
}

// ---- TMA bulk staging of a linear run of floats into shared memory --------------------------------
// One elected thread arms an mbarrier with the byte count and issues ONE cp.async.bulk (global ->
// shared, no tensor map: a confidence plane is linear memory); everybody waits on the barrier.  A
// whole plane (26 KB at 641 px) is in flight at once instead of a few loads per thread per round
// trip, which is what bounded the compact / CafScored kernels.
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!done);
}

// Stage the floats [g, g + n) into `stage` (16-byte aligned, room for 4 n + 32 bytes).  The copy
// starts at the 16-byte boundary at or below g and ends at the one at or above g + n: callers pass
// planes that are followed by more planes of the same tensor and whose tensor base is 16-byte
// aligned, so the few extra bytes are inside the allocation.  CTA-collective: every thread calls it
// with the same arguments, `phase` is the per-thread parity of the barrier (starts at 0, flipped
// here).  The previous contents of `stage` must no longer be read by anyone (__syncthreads before).
__device__ __forceinline__ const float *stage_floats(const float *g, int n, float *stage, uint64_t *bar,
                                                     uint32_t &phase) {
  const uintptr_t a = (uintptr_t)g;
  const uint32_t mis = (uint32_t)(a & 15u);
  const uint32_t bytes = (mis + 4u * (uint32_t)n + 15u) & ~15u;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
                 ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(stage)), "l"(a - mis), "r"(bytes), "r"(smem_u32(bar)) : "memory");
  }
  mbar_wait(bar, phase);
  phase ^= 1u;
  return reinterpret_cast<const float *>(reinterpret_cast<const char *>(stage) + mis);
}

// block_active_cells on a plane staged in shared memory: `sp` = the n staged values.  Same result
// (ordered list of the cells with value > th, as 16-bit offsets) and the same two barriers.
template <int THREADS>
__device__ __forceinline__ int block_active_cells_smem(const float *sp, int n, float th, unsigned short *list,
                                                       int *warp_cnt) {
  constexpr int NW = THREADS / 32;
  const int lane = lane_id(), w = warp_id();
  const int seg = (((n + NW - 1) / NW) + 31) & ~31;          // cells per warp, a multiple of 32
  const int w0 = w * seg, w1 = min(n, w0 + seg);
  int cnt = 0;
  for (int base = w0; base < w1; base += 32) {
    const int c = base + lane;
    cnt += __popc(__ballot_sync(0xffffffffu, c < w1 && sp[c] > th));
  }
  if (lane == 0) warp_cnt[w] = cnt;
  __syncthreads();
  int at = 0, total = 0;
#pragma unroll
  for (int i = 0; i < NW; i++) {
    const int t = warp_cnt[i];
    if (i < w) at += t;
    total += t;
  }
  const unsigned below = (1u << lane) - 1u;
  if (cnt) {
    for (int base = w0; base < w1; base += 32) {
      const int c = base + lane;
      const bool hit = c < w1 && sp[c] > th;
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      PP_DASSERT(!hit || at + __popc(m & below) < n);
      if (hit) list[at + __popc(m & below)] = (unsigned short)c;
      at += __popc(m);
    }
  }
  __syncthreads();
  return total;
}

}  // namespace pp
