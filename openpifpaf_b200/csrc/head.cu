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

// One band of rows of one (image, field, channel) plane.  KIND: 0 sigmoid, 1 + column, 2 + row,
// 3 copy, 4 exp.  The CTA is 32 columns x 8 rows of threads: a warp walks one output row, so stores are
// 128-byte runs and the pixel-shuffle gather reads two 64-byte runs (the even and the odd input
// channel).  All index arithmetic is per-thread constants plus adds: row / column parities never
// change for a thread (steps of 8 rows and 32 columns), so the input channel is fixed per thread.
// Eight loads are in flight per thread (2 rows x 4 column steps).
// History at batch 64 (caf, 574 MB moved): 1.37 TB/s one element per thread (latency-bound) ->
// 3.15 TB/s eight loads in flight, then issue-bound at 86 % with ~60 instructions per element
// (flat index, divisions, 64-bit offsets) -> this layout.
template <bool QUAD, int KIND>
__device__ __forceinline__ void head_band(const float *__restrict__ xi, float *__restrict__ o, int c,
                                          int h, int w, int W, int r0, int r1) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int hw = h * w;
  const int row0 = r0 + ty;
  // input offset of (row0, tx); rows advance by 8 (4 input rows), columns by 32 (16 input columns)
  int in0, in_row_step, in_col_step;
  if (QUAD) {   // PixelShuffle(2): out[c, 2i+di, 2j+dj] = in[4c + 2di + dj, i, j]
    in0 = (4 * c + 2 * (row0 & 1) + (tx & 1)) * hw + (row0 >> 1) * w + (tx >> 1);
    in_row_step = 4 * w; in_col_step = 16;
  } else {
    in0 = c * hw + row0 * w + tx;
    in_row_step = 8 * w; in_col_step = 32;
  }
  int out0 = row0 * W + tx;
  const float colf = (float)tx;
  for (int cb = 0; cb < W; cb += 128) {                 // 4 column steps per pass
    int in_r = in0 + (cb >> (QUAD ? 1 : 0)), out_r = out0 + cb;
    for (int row = row0; row < r1; row += 16) {         // 2 rows per pass
      float v[2][4];
#pragma unroll
      for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const bool ok = row + 8 * i < r1 && cb + tx + 32 * j < W;
          v[i][j] = ok ? __ldg(xi + in_r + i * in_row_step + j * in_col_step) : 0.0f;
        }
#pragma unroll
      for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          float r = v[i][j];
          if (KIND == 0) r = __frcp_rn(__fadd_rn(1.0f, expf(-r)));
          if (KIND == 1) r = __fadd_rn(r, colf + (float)(cb + 32 * j));
          if (KIND == 2) r = __fadd_rn(r, (float)(row + 8 * i));
          if (KIND == 4) r = expf(r);
          if (row + 8 * i < r1 && cb + tx + 32 * j < W) __stcs(o + out_r + i * 8 * W + 32 * j, r);
        }
      in_r += 2 * in_row_step;
      out_r += 16 * W;
    }
  }
}

constexpr int kHeadBandRows = 32;      // output rows per CTA (a multiple of 16)

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
  const int r0 = blockIdx.x * kHeadBandRows, r1 = min(H, r0 + kHeadBandRows);
  // input channel (before the pixel shuffle) per output channel
  if (ch == 0) { head_band<QUAD, 0>(xi, o, f, h, w, W, r0, r1); return; }
  const int v = (ch - 1) >> 2, r = (ch - 1) & 3;
  if (r == 0) head_band<QUAD, 1>(xi, o, nf + (f * nv + v) * 2, h, w, W, r0, r1);
  else if (r == 1) head_band<QUAD, 2>(xi, o, nf + (f * nv + v) * 2 + 1, h, w, W, r0, r1);
  else if (r == 2) head_band<QUAD, 3>(xi, o, nf + 2 * nv * nf + f * nv + v, h, w, W, r0, r1);
  else head_band<QUAD, 4>(xi, o, nf + 3 * nv * nf + f * nv + v, h, w, W, r0, r1);
}

int head_epilogue_launch(const float *d_x, float *d_out, int B, int nf, int nv, int h, int w, int quad,
                         cudaStream_t st) {
  const int H = quad ? 2 * h - 1 : h, W = quad ? 2 * w - 1 : w;
  const int nch = 1 + 4 * nv;
  if (B <= 0 || nf <= 0 || nv <= 0 || H <= 0 || W <= 0 || nf * nch > 65535 || B > 65535) return PP_ERR_ARG;
  if ((long long)H * W >= (1ll << 31)) return PP_ERR_ARG;
  const int gx = (H + kHeadBandRows - 1) / kHeadBandRows;
  dim3 grid(gx, nf * nch, B);
  if ((long long)nf * nch * (quad ? 4 : 1) * h * w >= (1ll << 31)) return PP_ERR_ARG;   // 32-bit offsets per image
  if (quad) head_epilogue_kernel<true><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  else head_epilogue_kernel<false><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  PP_LAUNCH_CHECK("head_epilogue_kernel");
  return PP_OK;
}

}  // namespace pp
