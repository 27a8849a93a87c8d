// shim.cu -- the reference's live functional.pyx primitives, one kernel each,
// with the reference's names and argument meaning (include/pifpaf_b200.h,
// "parity shim").  They exist for per-primitive known-answer tests against the
// compiled Cython module and for callers that still drive the decoder at the
// reference's granularity; the batch entry points do not go through them.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

// functional.pyx:105-141.  Sources are applied in index order; one CTA owns the
// whole field so the order of float adds per pixel is the reference's.  Each
// thread owns a set of pixels and walks the sources sequentially.
__global__ void shim_splat_kernel(float *field, int fh, int fw, const float *__restrict__ x,
                                  const float *__restrict__ y, const float *__restrict__ sigma,
                                  const float *__restrict__ v, int n, float truncate,
                                  float max_value) {
  const float truncate2 = __fmul_rn(truncate, truncate);
  const long long npx = (long long)fh * fw;
  for (long long pidx = (long long)blockIdx.x * blockDim.x + threadIdx.x; pidx < npx;
       pidx += (long long)gridDim.x * blockDim.x) {
    const int yy = (int)(pidx / fw), xx = (int)(pidx - (long long)yy * fw);
    float acc = field[pidx];
    for (int i = 0; i < n; i++) {
      const float cs = __ldg(sigma + i), cx = __ldg(x + i), cy = __ldg(y + i), cv = __ldg(v + i);
      const float cs2 = __fmul_rn(cs, cs);
      const float ts = __fmul_rn(truncate, cs);
      const int minx = (int)clipf(__fsub_rn(cx, ts), 0.0f, (float)(fw - 1));
      const int maxx = (int)clipf(__fadd_rn(__fadd_rn(cx, ts), 1.0f), (float)(minx + 1), (float)fw);
      const int miny = (int)clipf(__fsub_rn(cy, ts), 0.0f, (float)(fh - 1));
      const int maxy = (int)clipf(__fadd_rn(__fadd_rn(cy, ts), 1.0f), (float)(miny + 1), (float)fh);
      if (xx < minx || xx >= maxx || yy < miny || yy >= maxy) continue;
      const float dx = __fsub_rn((float)xx, cx), dy = __fsub_rn((float)yy, cy);
      const float dx2 = __fmul_rn(dx, dx), dy2 = __fmul_rn(dy, dy);
      const float d2 = __fadd_rn(dx2, dy2);
      if (d2 > __fmul_rn(truncate2, cs2)) continue;
      float vv;
      if (dx2 < 0.25f && dy2 < 0.25f) vv = cv;
      else vv = __fmul_rn(cv, approx_exp(__fdiv_rn(__fmul_rn(-0.5f, d2), cs2)));
      acc = __fadd_rn(acc, vv);
      acc = (acc < max_value) ? acc : max_value;
    }
    field[pidx] = acc;
  }
}

__global__ void shim_scalar_values_kernel(const float *__restrict__ field, int fh, int fw,
                                          const float *__restrict__ x, const float *__restrict__ y,
                                          int n, float dflt, float *__restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = scalar_value_at(field, fh, fw, x[i], y[i], dflt);
}

// functional.pyx:338-359: ordered compaction of the columns inside the box; one CTA.
__global__ void shim_caf_center_kernel(const float *__restrict__ f, int K, float x, float y,
                                       float sigma, float *__restrict__ out, int *__restrict__ count) {
  __shared__ int warp_tot[32];
  const float lox = __fsub_rn(x, sigma), hix = __fadd_rn(x, sigma);
  const float loy = __fsub_rn(y, sigma), hiy = __fadd_rn(y, sigma);
  int done = 0;
  for (int base = 0; base < K; base += blockDim.x) {
    int i = base + threadIdx.x;
    bool keep = false;
    if (i < K) {
      float cx = f[(size_t)K + i], cy = f[(size_t)2 * K + i];
      keep = !(cx < lox) && !(cx > hix) && !(cy < loy) && !(cy > hiy);
    }
    int tot;
    int off = block_flag_scan(keep, warp_tot, &tot);
    if (keep)
      for (int r = 0; r < 9; r++) out[(size_t)r * K + done + off] = f[(size_t)r * K + i];
    done += tot;
  }
  if (threadIdx.x == 0) *count = done;
}

__global__ void shim_occ_get_kernel(const uint8_t *__restrict__ field, int fh, int fw, float x,
                                    float y, float r, uint8_t *out) {
  float cx = clipf(__fdiv_rn(x, r), 0.0f, (float)(fw - 1));
  float cy = clipf(__fdiv_rn(y, r), 0.0f, (float)(fh - 1));
  *out = field[(size_t)(long long)cy * fw + (long long)cx];
}

__global__ void shim_occ_add_kernel(uint8_t *field, int fw, long long minx, long long maxx,
                                    long long miny, long long maxy, int value) {
  long long w = maxx - minx, n = w * (maxy - miny);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    long long yy = miny + i / w, xx = minx + i % w;
    field[(size_t)yy * fw + xx] = (uint8_t)(field[(size_t)yy * fw + xx] + value);
  }
}

}  // namespace pp

using namespace pp;

extern "C" {

int pp_scalar_square_add_gauss_with_max(float *d_field, int fh, int fw, const float *d_x,
                                        const float *d_y, const float *d_sigma, const float *d_v,
                                        int n, float truncate, float max_value, void *stream) {
  if (!d_field || fh <= 0 || fw <= 0 || n < 0 || (n > 0 && (!d_x || !d_y || !d_sigma || !d_v)))
    return PP_ERR_ARG;
  if (n == 0) return PP_OK;
  long long npx = (long long)fh * fw;
  int blocks = (int)((npx + 255) / 256);
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
  shim_splat_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_field, fh, fw, d_x, d_y, d_sigma,
                                                              d_v, n, truncate, max_value);
  PP_LAUNCH_CHECK("shim_splat_kernel");
  return PP_OK;
}

int pp_scalar_values(const float *d_field, int fh, int fw, const float *d_x, const float *d_y,
                     int n, float default_value, float *d_out, void *stream) {
  if (!d_field || fh <= 0 || fw <= 0 || n < 0 || (n > 0 && (!d_x || !d_y || !d_out)))
    return PP_ERR_ARG;
  if (n == 0) return PP_OK;
  shim_scalar_values_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      d_field, fh, fw, d_x, d_y, n, default_value, d_out);
  PP_LAUNCH_CHECK("shim_scalar_values_kernel");
  return PP_OK;
}

int pp_caf_center_s(const float *d_caf_field, int K, float x, float y, float sigma, float *d_out,
                    int32_t *d_count, void *stream) {
  if (K < 0 || !d_count || (K > 0 && (!d_caf_field || !d_out))) return PP_ERR_ARG;
  shim_caf_center_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(d_caf_field, K, x, y, sigma, d_out,
                                                              d_count);
  PP_LAUNCH_CHECK("shim_caf_center_kernel");
  return PP_OK;
}

int pp_scalar_nonzero_clipped_with_reduction(const uint8_t *d_field, int fh, int fw, float x,
                                             float y, float r, uint8_t *d_out, void *stream) {
  if (!d_field || !d_out || fh <= 0 || fw <= 0) return PP_ERR_ARG;
  shim_occ_get_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(d_field, fh, fw, x, y, r, d_out);
  PP_LAUNCH_CHECK("shim_occ_get_kernel");
  return PP_OK;
}

int pp_scalar_square_add_single(uint8_t *d_field, int fh, int fw, long long x, long long y,
                                long long sigma, int value, void *stream) {
  if (!d_field || fh <= 0 || fw <= 0) return PP_ERR_ARG;
  // decoder/utils.py:61-66
  long long minx = x - sigma > 0 ? x - sigma : 0, miny = y - sigma > 0 ? y - sigma : 0;
  long long t = x + sigma + 1 < fw ? x + sigma + 1 : fw;
  long long maxx = minx + 1 > t ? minx + 1 : t;
  t = y + sigma + 1 < fh ? y + sigma + 1 : fh;
  long long maxy = miny + 1 > t ? miny + 1 : t;
  if (maxx > fw) maxx = fw;
  if (maxy > fh) maxy = fh;
  if (maxx <= minx || maxy <= miny) return PP_OK;
  long long n = (maxx - minx) * (maxy - miny);
  int blocks = (int)((n + 255) / 256);
  if (blocks > 1024) blocks = 1024;
  shim_occ_add_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_field, fw, minx, maxx, miny, maxy,
                                                                value);
  PP_LAUNCH_CHECK("shim_occ_add_kernel");
  return PP_OK;
}

}  // extern "C"
