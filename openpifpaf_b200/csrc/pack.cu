// pack.cu -- gather of a batch's live result rows into one dense block (include/pifpaf_b200.h,
// "Packed results").  The reference returns Python lists of exactly the live annotations
// (decoder/generator/cifcaf.py:110-122, annotation.py:9-58); the fixed-capacity device arrays hold
// ann_cap rows per image, so copying them whole moves ~150x the live bytes.  One CTA per image.
#include "common.cuh"
#include "stages.cuh"

namespace pp {

int packed_record_words(int F) { return (4 + 9 * F + 1) & ~1; }
size_t packed_header_words(int B) { return ((size_t)4 + 3 * (size_t)B + 1) & ~(size_t)1; }
size_t packed_image_bytes(int F, int ann_cap) {
  return ((size_t)8 + (size_t)ann_cap * packed_record_words(F)) * sizeof(int32_t);
}

struct PackArgs {
  const float *ann_data, *ann_scales, *pre_data;
  const double *ann_score;
  const int *ann_source, *n_ann, *n_pre, *pre_order, *status;
  int32_t *out;
  int B, F, cap, rec_words;
  unsigned long long hdr_words;
};

__global__ void __launch_bounds__(128) pack_results_kernel(PackArgs a) {
  __shared__ int s_part[4];
  const int b = blockIdx.x, F = a.F, cap = a.cap;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // records of the images before this one
  int part = 0;
  for (int i = tid; i < b; i += blockDim.x) part += min(a.n_ann[i], cap);
  part = __reduce_add_sync(0xffffffffu, part);
  if (lane == 0) s_part[warp] = part;
  __syncthreads();
  const int first = s_part[0] + s_part[1] + s_part[2] + s_part[3];
  const int n = min(a.n_ann[b], cap);
  if (tid == 0) {
    a.out[4 + b] = n;
    a.out[4 + a.B + b] = a.status[b];
    a.out[4 + 2 * a.B + b] = a.n_pre ? a.n_pre[b] : 0;
    if (b == a.B - 1) {
      a.out[0] = a.B; a.out[1] = F; a.out[2] = first + n; a.out[3] = a.rec_words;
    }
  }
  const float *data = a.ann_data + (size_t)b * cap * F * 3;
  const float *scales = a.ann_scales + (size_t)b * cap * F;
  const float *pre = a.pre_data ? a.pre_data + (size_t)b * cap * F * 3 : nullptr;
  const int *order = a.pre_order ? a.pre_order + (size_t)b * cap * F * 2 : nullptr;
  for (int r = warp; r < n; r += blockDim.x >> 5) {          // one warp per record
    PP_DASSERT(first + r < a.B * cap);
    int32_t *rec = a.out + a.hdr_words + (size_t)(first + r) * a.rec_words;
    const int src = a.ann_source ? a.ann_source[(size_t)b * cap + r] : r;
    if (lane == 0) {
      const double sc = a.ann_score ? a.ann_score[(size_t)b * cap + r] : 0.0;
      *reinterpret_cast<double *>(rec) = sc;
      rec[2] = src;
      rec[3] = 0;
    }
    float *f = reinterpret_cast<float *>(rec + 4);
    for (int i = lane; i < 3 * F; i += 32) f[i] = data[(size_t)r * F * 3 + i];
    for (int i = lane; i < F; i += 32) f[3 * F + i] = scales[(size_t)r * F + i];
    const bool ok = src >= 0 && src < cap;
    for (int i = lane; i < 3 * F; i += 32) f[4 * F + i] = (pre && ok) ? pre[(size_t)src * F * 3 + i] : 0.0f;
    for (int i = lane; i < 2 * F; i += 32)
      rec[4 + 7 * F + i] = (order && ok) ? order[(size_t)src * F * 2 + i] : -1;
    if (lane == 0 && (9 * F & 1)) rec[4 + 9 * F] = 0;          // pad word
  }
}

int pack_launch(const pp_outputs &o, const pp_shape &s, const pp_caps &c, cudaStream_t st) {
  if (!o.packed || !o.ann_data || !o.ann_scales || !o.n_ann || !o.status) return PP_ERR_ARG;
  if (((uintptr_t)o.packed & 7u) != 0) return PP_ERR_ARG;
  PackArgs a;
  a.ann_data = o.ann_data; a.ann_scales = o.ann_scales; a.pre_data = o.pre_data;
  a.ann_score = o.ann_score; a.ann_source = o.ann_source; a.n_ann = o.n_ann; a.n_pre = o.n_pre;
  a.pre_order = o.pre_order; a.status = o.status;
  a.out = (int32_t *)o.packed;
  a.B = s.B; a.F = s.F; a.cap = c.ann_cap; a.rec_words = packed_record_words(s.F);
  a.hdr_words = packed_header_words(s.B);
  pack_results_kernel<<<s.B, 128, 0, st>>>(a);
  PP_LAUNCH_CHECK("pack_results_kernel");
  return PP_OK;
}

}  // namespace pp
