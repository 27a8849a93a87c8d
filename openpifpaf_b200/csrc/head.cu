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

// ---------------------------------------------------------------------------
// Round 2: the same pass with every global access made by the bulk-copy unit.  The input of one output
// plane is 4 adjacent channel planes (PixelShuffle: out[c, 2i+di, 2j+dj] = in[4c + 2di + dj, i, j]), i.e. one
// contiguous block, and so is the output plane: a CTA loads the rows it needs of the 4 channels with bulk
// copies (cp.async.bulk + mbarrier), interleaves and activates them out of shared memory into a tile laid out
// like the output rows, and sends the tile off with one bulk store.  No 164-byte row gathers from L2 (the
// round-1 kernel read 1.37 x the sectors it used) and no per-thread global address arithmetic.
constexpr int kHeadSmemBudget = 54 * 1024;     // 4 CTAs per SM

template <int KIND>
__device__ __forceinline__ float head_act(float r, int col, int row) {
  if (KIND == 0) r = __frcp_rn(__fadd_rn(1.0f, expf(-r)));
  if (KIND == 1) r = __fadd_rn(r, (float)col);
  if (KIND == 2) r = __fadd_rn(r, (float)row);
  if (KIND == 4) r = expf(r);
  return r;
}

template <bool QUAD, int KIND>
__device__ __forceinline__ void head_tile(const float *const (&in_q)[QUAD ? 4 : 1], float *out_s, int w, int W, int r0, int r1,
                                          int i0) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int row = r0 + ty; row < r1; row += kHeadThreads / 32) {
    const float *e = QUAD ? ((row & 1) ? in_q[2] : in_q[0]) + ((row >> 1) - i0) * w : in_q[0] + (row - r0) * w;
    const float *o = QUAD ? ((row & 1) ? in_q[3] : in_q[1]) + ((row >> 1) - i0) * w : e;
    float *dst = out_s + (row - r0) * W;
    for (int col = tx; col < W; col += 32) {
      const float r = QUAD ? ((col & 1) ? o[col >> 1] : e[col >> 1]) : e[col];
      dst[col] = head_act<KIND>(r, col, row);
    }
  }
}

template <bool QUAD>
__global__ void __launch_bounds__(kHeadThreads)
head_tma_kernel(const float *__restrict__ x, float *__restrict__ out, int nf, int nv, int h, int w, int H, int W,
                int band_rows, int in_stride) {
  extern __shared__ __align__(128) unsigned char head_smem[];
  __shared__ __align__(8) uint64_t bar;
  constexpr int NQ = QUAD ? 4 : 1;
  const int nch = 1 + 4 * nv;
  const int plane = blockIdx.y, b = blockIdx.z;
  const int f = plane / nch, ch = plane - f * nch;
  const int C = nf * nch, hw = h * w;
  const int v = (ch - 1) >> 2, rr = (ch - 1) & 3;
  // input channel (before the pixel shuffle) of this output channel (heads.py:65-88,406-448)
  const int c_in = ch == 0 ? f : (rr == 0 ? nf + (f * nv + v) * 2 : (rr == 1 ? nf + (f * nv + v) * 2 + 1
                   : (rr == 2 ? nf + 2 * nv * nf + f * nv + v : nf + 3 * nv * nf + f * nv + v)));
  const float *xi = x + (size_t)b * C * NQ * hw;
  float *o = out + ((size_t)b * C + plane) * H * W;
  const int r0 = blockIdx.x * band_rows, r1 = min(H, r0 + band_rows);
  const int i0 = QUAD ? r0 >> 1 : r0, i1 = QUAD ? (r1 - 1) >> 1 : r1 - 1;      // input rows [i0, i1]
  const int n_in = (i1 - i0 + 1) * w;
  float *in_s = reinterpret_cast<float *>(head_smem);
  float *out_tile = in_s + NQ * in_stride;
  const float *in_q[NQ];
  uint32_t bytes[NQ], total = 0;
#pragma unroll
  for (int q = 0; q < NQ; q++) {
    const uintptr_t a = (uintptr_t)(xi + (size_t)(NQ * c_in + q) * hw + (size_t)i0 * w);
    const uint32_t mis = (uint32_t)(a & 15u);
    bytes[q] = (mis + 4u * (uint32_t)n_in + 15u) & ~15u;
    total += bytes[q];
    in_q[q] = reinterpret_cast<const float *>(reinterpret_cast<const char *>(in_s + q * in_stride) + mis);
  }
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(total) : "memory");
#pragma unroll
    for (int q = 0; q < NQ; q++) {
      const uintptr_t a = (uintptr_t)(xi + (size_t)(NQ * c_in + q) * hw + (size_t)i0 * w) & ~(uintptr_t)15;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(in_s + q * in_stride)), "l"(a), "r"(bytes[q]), "r"(smem_u32(&bar)) : "memory");
    }
  }
  __syncthreads();                      // the barrier is initialised before anyone polls it
  mbar_wait(&bar, 0);
  // the tile is laid out like the output rows in memory, shifted so that 16-byte units line up
  float *og = o + (size_t)r0 * W;
  const int n_out = (r1 - r0) * W;
  const int lead = (int)(((uintptr_t)og & 15u) >> 2);
  float *out_s = out_tile + lead;
  if (ch == 0) head_tile<QUAD, 0>(in_q, out_s, w, W, r0, r1, i0);
  else if (rr == 0) head_tile<QUAD, 1>(in_q, out_s, w, W, r0, r1, i0);
  else if (rr == 1) head_tile<QUAD, 2>(in_q, out_s, w, W, r0, r1, i0);
  else if (rr == 2) head_tile<QUAD, 3>(in_q, out_s, w, W, r0, r1, i0);
  else head_tile<QUAD, 4>(in_q, out_s, w, W, r0, r1, i0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  const int head = lead ? min(4 - lead, n_out) : 0;
  const int mid = ((n_out - head) >> 2) << 2;
  const int t = threadIdx.x;
  if (t == 0) {
    if (mid > 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   :: "l"(og + head), "r"(smem_u32(out_s + head)), "r"((uint32_t)mid * 4u) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");     // shared memory outlives the read
    }
  } else if (t >= 32 && t < 36) {
    const int i = t - 32;
    if (i < head) og[i] = out_s[i];
  } else if (t >= 36 && t < 40) {
    const int i = head + mid + (t - 36);
    if (i < n_out) og[i] = out_s[i];
  }
}

// rows per band so that the input rows of NQ channels + the output tile fit the budget; 0 = not even two rows
static int head_band_rows(int h, int w, int H, int W, int quad, int *in_stride, size_t *smem) {
  const int nq = quad ? 4 : 1;
  for (int n_bands = 1; n_bands <= H; n_bands++) {
    const int rows = (H + n_bands - 1) / n_bands;
    const int in_rows = quad ? rows / 2 + 2 : rows;
    const int stride = (in_rows * w + 8 + 3) & ~3;
    const size_t bytes = ((size_t)nq * stride + (size_t)rows * W + 8) * sizeof(float);
    if (bytes <= (size_t)kHeadSmemBudget) {
      *in_stride = stride;
      *smem = bytes;
      return rows;
    }
    if (rows <= 2) break;
  }
  return 0;
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
  int in_stride = 0;
  size_t smem = 0;
  const int band_rows = head_band_rows(h, w, H, W, quad, &in_stride, &smem);
  if (band_rows > 0 && ((uintptr_t)d_x & 15u) == 0 && ((uintptr_t)d_out & 3u) == 0) {
    dim3 g2((H + band_rows - 1) / band_rows, nf * nch, B);
    if (quad) {
      PP_CUDA_CHECK(cudaFuncSetAttribute(head_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kHeadSmemBudget));
      head_tma_kernel<true><<<g2, kHeadThreads, smem, st>>>(d_x, d_out, nf, nv, h, w, H, W, band_rows, in_stride);
    } else {
      PP_CUDA_CHECK(cudaFuncSetAttribute(head_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kHeadSmemBudget));
      head_tma_kernel<false><<<g2, kHeadThreads, smem, st>>>(d_x, d_out, nf, nv, h, w, H, W, band_rows, in_stride);
    }
    PP_LAUNCH_CHECK("head_tma_kernel");
    return PP_OK;
  }
  if (quad) head_epilogue_kernel<true><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  else head_epilogue_kernel<false><<<grid, kHeadThreads, 0, st>>>(d_x, d_out, nf, nv, h, w, H, W);
  PP_LAUNCH_CHECK("head_epilogue_kernel");
  return PP_OK;
}

}  // namespace pp
