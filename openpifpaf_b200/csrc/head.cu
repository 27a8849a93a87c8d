// head.cu -- device epilogue of the fused head networks, producing the
// decoder's input fields.
//
// Replaces, for inference, the tail of CompositeFieldFused.forward
// (network/heads.py:406-448: PixelShuffle(2) + crop, sigmoid on the confidences,
// exp on the scales, channel split) and CifCafCollector.forward
// (network/heads.py:65-88: concatenation, index add, CAF channel reorder), i.e.
// about ten full passes of eager tensor ops, with one pass: every output float
// is produced from exactly one input float.
//
//   x    (B, C * 4^quad, h, w)  raw convolution output, C = n_fields * (1 + 4 * n_vectors)
//   out  (B, n_fields, 1 + 4 * n_vectors, H, W),  H = 2h-1, W = 2w-1 for quad = 1
//        channels [conf, (x_v, y_v, logb_v, scale_v) for every vector v]:
//        cif (n_vectors = 1) -> [c, x, y, b, s];  caf (2) -> [c, x1, y1, b1, s1, x2, y2, b2, s2]
#include "common.cuh"
#include "stages.cuh"

namespace pp {

constexpr int kHeadThreads = 256;
constexpr int kHeadUnroll = 8;

// One (image, field, channel) plane chunk.  KIND: 0 sigmoid, 1 + column, 2 + row, 3 copy, 4 exp.
// kHeadUnroll independent loads are in flight per thread (a one-load-per-thread version kept only
// ~8 KB per SM in flight: 1.35 TB/s); index math is 32-bit and row / col advance incrementally
// (the first version of this loop was issue-bound at 86 % with ~80 instructions per element).
template <bool QUAD, int KIND>
__device__ __forceinline__ void head_plane(const float *__restrict__ xi, float *__restrict__ o, int c,
                                           int h, int w, int W, int p0, int p_end) {
  const int step_row = kHeadThreads / W, step_col = kHeadThreads - step_row * W;
  const int hw = h * w;
  const int p = p0 + threadIdx.x;
  int row = p / W, col = p - row * W;
  float v[kHeadUnroll];
#pragma unroll
  for (int k = 0; k < kHeadUnroll; ++k) {
    int off;
    if (QUAD)          // PixelShuffle(2): out[c, 2i+di, 2j+dj] = in[4c + 2di + dj, i, j]
      off = (4 * c + 2 * (row & 1) + (col & 1)) * hw + (row >> 1) * w + (col >> 1);
    else
      off = c * hw + row * w + col;
    float r = 0.0f;
    if (p + k * kHeadThreads < p_end) r = __ldg(xi + off);
    if (KIND == 1) r = __fadd_rn(r, (float)col);
    if (KIND == 2) r = __fadd_rn(r, (float)row);
    v[k] = r;
    row += step_row; col += step_col;
    if (col >= W) { col -= W; ++row; }
  }
#pragma unroll
  for (int k = 0; k < kHeadUnroll; ++k) {
    float r = v[k];
    if (KIND == 0) r = __frcp_rn(__fadd_rn(1.0f, expf(-r)));
    if (KIND == 4) r = expf(r);
    if (p + k * kHeadThreads < p_end) __stcs(o + p + k * kHeadThreads, r);
  }
}

template <bool QUAD>
__global__ void __launch_bounds__(kHeadThreads)
head_epilogue_kernel(const float *__restrict__ x, float *__restrict__ out, int nf, int nv, int h, int w,
                     int H, int W) {
  const int nch = 1 + 4 * nv;
  const int plane = blockIdx.y;                 // (f, ch)
  const int b = blockIdx.z;
  const int f = plane / nch, ch = plane - f * nch;
  const int C = nf * nch;
  const float *xi = x + (size_t)b * C * (QUAD ? 4 : 1) * h * w;
  float *o = out + ((size_t)b * C + plane) * H * W;
  // the plane is cut into gridDim.x equal chunks (no short last CTA)
  const int HW = H * W;
  const int chunk = (HW + gridDim.x - 1) / gridDim.x;
  const int p0 = blockIdx.x * chunk, p_end = min(HW, p0 + chunk);
  // input channel (before the pixel shuffle) per output channel
  if (ch == 0) { head_plane<QUAD, 0>(xi, o, f, h, w, W, p0, p_end); return; }
  const int v = (ch - 1) >> 2, r = (ch - 1) & 3;
  if (r == 0) head_plane<QUAD, 1>(xi, o, nf + (f * nv + v) * 2, h, w, W, p0, p_end);
  else if (r == 1) head_plane<QUAD, 2>(xi, o, nf + (f * nv + v) * 2 + 1, h, w, W, p0, p_end);
  else if (r == 2) head_plane<QUAD, 3>(xi, o, nf + 2 * nv * nf + f * nv + v, h, w, W, p0, p_end);
  else head_plane<QUAD, 4>(xi, o, nf + 3 * nv * nf + f * nv + v, h, w, W, p0, p_end);
}

int head_epilogue_launch(const float *d_x, float *d_out, int B, int nf, int nv, int h, int w, int quad,
                         cudaStream_t st) {
  const int H = quad ? 2 * h - 1 : h, W = quad ? 2 * w - 1 : w;
  const int nch = 1 + 4 * nv;
  if (B <= 0 || nf <= 0 || nv <= 0 || H <= 0 || W <= 0 || nf * nch > 65535 || B > 65535) return PP_ERR_ARG;
  const int per_cta = kHeadThreads * kHeadUnroll;
  const int gx = (H * W + per_cta - 1) / per_cta;
  dim3 grid(gx, nf * nch, B);
  if ((long long)nf * nch * (quad ? 4 : 1) * h * w >= (1ll << 31)) return PP_ERR_ARG;   // 32-bit offsets per image
  if (quad) head_epilogue_kernel<true><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  else head_epilogue_kernel<false><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  PP_LAUNCH_CHECK("head_epilogue_kernel");
  return PP_OK;
}

}  // namespace pp
