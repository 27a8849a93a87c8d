// api.cu -- extern "C" surface of libpifpaf_b200.so (see include/pifpaf_b200.h).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "stages.cuh"

namespace pp {

// ---- error text -------------------------------------------------------------
static thread_local char g_cuda_err[256] = "";

void set_cuda_error(cudaError_t e, const char *where) {
  snprintf(g_cuda_err, sizeof g_cuda_err, "%s: %s (%s)", where, cudaGetErrorName(e),
           cudaGetErrorString(e));
}

// SM count of the current device (cached per device ordinal)
int num_sms() {
  constexpr int kMaxDev = 64;
  static int cached[kMaxDev] = {0};
  int dev = 0, n = 148;
  if (cudaGetDevice(&dev) != cudaSuccess) return n;
  if (dev >= 0 && dev < kMaxDev && cached[dev] > 0) return cached[dev];
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  if (dev >= 0 && dev < kMaxDev) cached[dev] = n;
  return n;
}

}  // namespace pp

// ---- splat gate -------------------------------------------------------------------
struct pp_gate {
  static constexpr int kRing = 64;
  cudaEvent_t ev[kRing];
  unsigned long long issued;      // splats recorded so far
};

namespace pp {
// make `st` wait for the splat recorded last on this gate (any stream)
int gate_wait(pp_gate *g, cudaStream_t st) {
  if (!g || g->issued == 0) return PP_OK;
  PP_CUDA_CHECK(cudaStreamWaitEvent(st, g->ev[(g->issued - 1) % pp_gate::kRing], 0));
  return PP_OK;
}
// the splat just enqueued on `st` is the one the next decode waits for
int gate_signal(pp_gate *g, cudaStream_t st) {
  if (!g) return PP_OK;
  PP_CUDA_CHECK(cudaEventRecord(g->ev[g->issued % pp_gate::kRing], st));
  g->issued++;
  return PP_OK;
}
}  // namespace pp

extern "C" pp_gate *pp_gate_create(void) {
  pp_gate *g = (pp_gate *)calloc(1, sizeof(pp_gate));
  if (!g) return nullptr;
  for (int i = 0; i < pp_gate::kRing; i++)
    if (cudaEventCreateWithFlags(&g->ev[i], cudaEventDisableTiming) != cudaSuccess) {
      for (int j = 0; j < i; j++) cudaEventDestroy(g->ev[j]);
      free(g);
      return nullptr;
    }
  return g;
}

extern "C" void pp_gate_destroy(pp_gate *g) {
  if (!g) return;
  for (int i = 0; i < pp_gate::kRing; i++) cudaEventDestroy(g->ev[i]);
  free(g);
}

// ---- per-stage timer ----------------------------------------------------------
struct pp_profile {
  int max_steps;
  int step;        // completed steps
  int open;        // a step is being recorded
  cudaEvent_t *ev; // (max_steps, PP_N_STAGES + 1)
};

namespace pp {

void prof_begin_step(pp_profile *prof) {
  if (!prof) return;
  prof->open = prof->step < prof->max_steps;
}

#ifdef PP_BREADCRUMB
// Fault isolation (debug build): a one-thread kernel writes "stage about to start" into pinned host memory, one
// slot per stream; after a sticky CUDA error the host reads what every stream had reached (pp_debug_crumbs).
namespace {
__global__ void crumb_kernel(volatile int *slot, int value) { *slot = value; }
struct Crumbs { int *host = nullptr; int *dev = nullptr; cudaStream_t streams[64]; int n = 0; };
Crumbs g_crumbs;
}  // namespace
void debug_crumb(cudaStream_t st, int value) {
  Crumbs &c = g_crumbs;
  if (!c.host) {
    if (cudaHostAlloc((void **)&c.host, 64 * sizeof(int), cudaHostAllocMapped) != cudaSuccess) return;
    memset(c.host, 0, 64 * sizeof(int));
    cudaHostGetDevicePointer((void **)&c.dev, c.host, 0);
  }
  int k = 0;
  while (k < c.n && c.streams[k] != st) k++;
  if (k == c.n) { if (c.n == 64) return; c.streams[c.n++] = st; }
  crumb_kernel<<<1, 1, 0, st>>>(c.dev + k, value);
}
extern "C" int pp_debug_crumbs(int *out, int n) {
  for (int k = 0; k < g_crumbs.n && k < n; k++) out[k] = g_crumbs.host[k];
  return g_crumbs.n;
}
#else
void debug_crumb(cudaStream_t, int) {}
#endif

void prof_mark(pp_profile *prof, int stage, cudaStream_t st) {
  debug_crumb(st, 10 + stage);
  if (!prof || !prof->open) return;
  cudaEventRecord(prof->ev[(size_t)prof->step * (PP_N_STAGES + 1) + stage], st);
}

void prof_end_step(pp_profile *prof, cudaStream_t st) {
  debug_crumb(st, 99);
  if (!prof || !prof->open) return;
  cudaEventRecord(prof->ev[(size_t)prof->step * (PP_N_STAGES + 1) + PP_N_STAGES], st);
  prof->step++;
  prof->open = 0;
}

static bool shape_ok(const pp_shape *s) {
  return s && s->B > 0 && s->F > 0 && s->F <= PP_MAX_F && s->L >= 0 && s->L <= PP_MAX_L &&
         s->H > 0 && s->W > 0 && s->stride > 0;
}

static bool caps_ok(const pp_shape *s, const pp_caps *c) {
  return c && c->seed_cap > 0 && c->caf_cap > 0 && c->ann_cap > 0 && c->src_cap > 0 &&
         c->seed_cap < (1 << 24);
}

// ---- decode workspace layout ---------------------------------------------
struct DecodeWs {
  float *hr;                 // (B,F,Hh,Wh)      when the caller passes none
  float *seeds; int *n_seeds;
  float *fwd, *bwd; int *nfwd, *nbwd;
  float *bin_rows; int *bin_start;   // spatially binned copy of the lists for the assembly
  uint8_t *occ_cnt; int2 *occ_rec;
  void *seed_stage;
  float *pre_data, *pre_scales; int *pre_order; int *n_pre;
  float *cmp_data, *cmp_scales;   // force_complete: completed annotations (NMS input)
  void *stage;               // shared by the cifhr / seeds / nms stages (used one after another)
  size_t stage_bytes;
  size_t total;
};

static size_t max3(size_t a, size_t b, size_t c) { return a > b ? (a > c ? a : c) : (b > c ? b : c); }

static DecodeWs layout(void *base, const pp_shape &s, const pp_caps &c, bool need_hr) {
  DecodeWs w;
  char *p = (char *)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { char *q = p + off; off += align_up(bytes, 256); return (void *)q; };
  size_t B = s.B;
  w.hr = need_hr ? (float *)take(B * s.F * hr_h(s) * hr_w(s) * sizeof(float)) : nullptr;
  w.seeds = (float *)take(B * c.seed_cap * 5 * sizeof(float));
  w.n_seeds = (int *)take(B * sizeof(int));
  w.fwd = (float *)take(B * s.L * 9 * c.caf_cap * sizeof(float));
  w.bwd = (float *)take(B * s.L * 9 * c.caf_cap * sizeof(float));
  w.nfwd = (int *)take(B * s.L * sizeof(int));
  w.nbwd = (int *)take(B * s.L * sizeof(int));
  w.bin_rows = (float *)take(binned_rows_bytes(s, c));
  w.bin_start = (int *)take(binned_start_bytes(s));
  w.occ_cnt = (uint8_t *)take(B * c.seed_cap);
  w.occ_rec = (int2 *)take(B * c.seed_cap * sizeof(int2));
  w.pre_data = (float *)take(B * c.ann_cap * s.F * 3 * sizeof(float));
  w.pre_scales = (float *)take(B * c.ann_cap * s.F * sizeof(float));
  w.pre_order = (int *)take(B * c.ann_cap * s.F * 2 * sizeof(int));
  w.n_pre = (int *)take(B * sizeof(int));
  w.cmp_data = (float *)take(B * c.ann_cap * s.F * 3 * sizeof(float));
  w.cmp_scales = (float *)take(B * c.ann_cap * s.F * sizeof(float));
  // the seed stage reads the CifHr stage's compacted sources: its scratch is a region of its own
  w.stage_bytes = max3(cifhr_ws_bytes(s, c), (size_t)0, nms_ws_bytes(s, c));
  w.stage = take(w.stage_bytes);
  w.seed_stage = take(seeds_ws_bytes(s, c));
  w.total = off;
  return w;
}

}  // namespace pp

using namespace pp;

extern "C" {

const char *pp_error_string(int code) {
  switch (code) {
    case PP_OK: return "ok";
    case PP_ERR_ARG: return "invalid argument (shape, NULL pointer or unsupported option)";
    case PP_ERR_CAPACITY: return "a pp_caps capacity was exceeded";
    case PP_ERR_CUDA: return "CUDA runtime error (see pp_last_cuda_error)";
    case PP_ERR_WORKSPACE: return "workspace too small";
    case PP_ERR_STATIC_LIMIT: return "request exceeds the static scratch of a *_noalloc entry point";
    default: return "unknown error code";
  }
}

const char *pp_last_cuda_error(void) { return g_cuda_err; }

pp_profile *pp_profile_create(int max_steps) {
  if (max_steps <= 0) return nullptr;
  pp_profile *prof = (pp_profile *)calloc(1, sizeof(pp_profile));
  if (!prof) return nullptr;
  size_t n = (size_t)max_steps * (PP_N_STAGES + 1);
  prof->ev = (cudaEvent_t *)calloc(n, sizeof(cudaEvent_t));
  prof->max_steps = max_steps;
  for (size_t i = 0; prof->ev && i < n; i++)
    if (cudaEventCreate(&prof->ev[i]) != cudaSuccess) { prof->max_steps = 0; break; }
  if (!prof->ev || prof->max_steps == 0) { pp_profile_destroy(prof); return nullptr; }
  return prof;
}

void pp_profile_destroy(pp_profile *prof) {
  if (!prof) return;
  if (prof->ev) {
    size_t n = (size_t)(prof->max_steps > 0 ? prof->max_steps : 0) * (PP_N_STAGES + 1);
    for (size_t i = 0; i < n; i++)
      if (prof->ev[i]) cudaEventDestroy(prof->ev[i]);
    free(prof->ev);
  }
  free(prof);
}

void pp_profile_reset(pp_profile *prof) {
  if (prof) { prof->step = 0; prof->open = 0; }
}

int pp_profile_read(pp_profile *prof, float *h_ms, int32_t *n_steps) {
  if (!prof || !h_ms || !n_steps) return PP_ERR_ARG;
  *n_steps = prof->step;
  for (int s = 0; s < prof->step; s++) {
    cudaEvent_t *e = prof->ev + (size_t)s * (PP_N_STAGES + 1);
    PP_CUDA_CHECK(cudaEventSynchronize(e[PP_N_STAGES]));
    for (int k = 0; k < PP_N_STAGES; k++)
      PP_CUDA_CHECK(cudaEventElapsedTime(&h_ms[(size_t)s * PP_N_STAGES + k], e[k], e[k + 1]));
  }
  return PP_OK;
}

int pp_profile_read_abs(pp_profile *prof, float *h_ms, int32_t *n_steps) {
  if (!prof || !h_ms || !n_steps) return PP_ERR_ARG;
  *n_steps = prof->step;
  for (int s = 0; s < prof->step; s++) {
    cudaEvent_t *e = prof->ev + (size_t)s * (PP_N_STAGES + 1);
    PP_CUDA_CHECK(cudaEventSynchronize(e[PP_N_STAGES]));
    for (int k = 0; k <= PP_N_STAGES; k++)
      PP_CUDA_CHECK(cudaEventElapsedTime(&h_ms[(size_t)s * (PP_N_STAGES + 1) + k], prof->ev[0], e[k]));
  }
  return PP_OK;
}

const char *pp_stage_name(int stage) {
  static const char *names[PP_N_STAGES] = {"cif_compact", "cifhr_splat", "seeds_collect", "seeds_sort",
                                           "caf_scored", "assemble", "nms"};
  return (stage >= 0 && stage < PP_N_STAGES) ? names[stage] : "?";
}

/* Launches one single-scale pp_decode_batch enqueues with d_out->packed set (the bindings always set it):
 * 8 kernels (field scan = CIF compact + raw CAF candidates, strips kernel, CTA-per-strip kernel for wide
 * boxes [its work list is empty for ordinary pose fields], per-field seed sort, seed merge, CafScored scoring +
 * binning, assembly with the NMS as its tail, result gather) + 2 memsets (status, work cursors). */
int pp_decode_launch_count(const pp_params *p) {
  int n = PP_N_STAGES + 3;
  if (p && p->cifhr_warps < 0) n -= 1;  // the CTA-per-strip kernel alone
  if (p && p->force_complete) n += 4;   // second CafScored, its binning, complete_kernel, nms_kernel
  if (p && p->inverse) n += 1;          // annotations_inverse_kernel
  return n;
}
const char *pp_version(void) { return "pifpaf_b200 0.1.0 (sm_100a)"; }

void pp_default_params(pp_params *p) {
  if (!p) return;
  memset(p, 0, sizeof *p);
  p->nms_instance_threshold = 0.0;
  p->cif_threshold = 0.1f;
  p->seed_threshold = 0.2f;
  p->seed_score_scale = 1.0f;
  p->caf_threshold = 0.1f;
  p->cif_floor = 0.1f;
  p->cif_floor_compl = (float)(1.0 - 0.1);
  p->keypoint_threshold = 0.001f;
  p->nms_keypoint_threshold = 0.001f;
  p->nms_suppression = 0.0f;
  p->connection_method = 0;
  p->nms = 1;
  p->exp_mode = 1;
}

int pp_default_caps(const pp_shape *s, pp_caps *c) {
  // F up to PP_MAX_CATEGORIES so that the detection path can use the same defaults
  if (!s || !c || s->B <= 0 || s->F <= 0 || s->F > PP_MAX_CATEGORIES || s->H <= 0 || s->W <= 0 || s->stride <= 0)
    return PP_ERR_ARG;
  long long cells = (long long)s->H * s->W;
  long long seeds = cells * s->F;
  if (seeds >= (1 << 24)) seeds = (1 << 24) - 1;
  c->seed_cap = (int32_t)seeds;
  c->caf_cap = (int32_t)cells;
  c->ann_cap = 256;
  c->src_cap = (int32_t)cells;
  return PP_OK;
}

int pp_coco_skeleton(pp_skeleton *sk) {
  static const int coco[19][2] = {{16, 14}, {14, 12}, {17, 15}, {15, 13}, {12, 13}, {6, 12}, {7, 13},
                                  {6, 7},   {6, 8},   {7, 9},   {8, 10},  {9, 11},  {2, 3},  {1, 2},
                                  {1, 3},   {2, 4},   {3, 5},   {4, 6},   {5, 7}};
  if (!sk) return PP_ERR_ARG;
  memset(sk, 0, sizeof *sk);
  sk->L = 19;
  for (int l = 0; l < 19; l++) { sk->joints[l][0] = coco[l][0]; sk->joints[l][1] = coco[l][1]; }
  return PP_OK;
}

size_t pp_workspace_bytes(const pp_shape *s, const pp_caps *c) {
  if (!shape_ok(s) || !caps_ok(s, c)) return 0;
  return layout(nullptr, *s, *c, true).total;
}
size_t pp_workspace_bytes_ex(const pp_shape *s, const pp_caps *c, int32_t hr_external) {
  if (!shape_ok(s) || !caps_ok(s, c)) return 0;
  return layout(nullptr, *s, *c, hr_external == 0).total;
}

size_t pp_packed_bytes(const pp_shape *s, const pp_caps *c) {
  if (!s || !c || s->B <= 0 || s->F <= 0 || s->F > PP_MAX_F || c->ann_cap <= 0) return 0;
  return (size_t)s->B * packed_image_bytes(s->F, c->ann_cap);
}
int32_t pp_packed_record_words(int32_t F) { return F > 0 ? packed_record_words(F) : 0; }
size_t pp_packed_header_words(int32_t B) { return B > 0 ? packed_header_words(B) : 0; }
int pp_pack_results(const pp_outputs *d_out, const pp_shape *s, const pp_caps *c, void *stream) {
  if (!d_out || !s || !c || s->B <= 0 || s->F <= 0 || s->F > PP_MAX_F || c->ann_cap <= 0) return PP_ERR_ARG;
  return pack_launch(*d_out, *s, *c, (cudaStream_t)stream);
}

int32_t pp_sizeof(int32_t which) {
  switch (which) {
    case 0: return (int32_t)sizeof(pp_params);
    case 1: return (int32_t)sizeof(pp_image_meta);
    case 2: return (int32_t)sizeof(pp_outputs);
    case 3: return (int32_t)sizeof(pp_skeleton);
    case 4: return (int32_t)sizeof(pp_shape);
    case 5: return (int32_t)sizeof(pp_caps);
    case 6: return (int32_t)sizeof(pp_det_params);
    case 7: return (int32_t)sizeof(pp_det_outputs);
    default: return -1;
  }
}

int pp_annotations_inverse(float *d_ann_data, float *d_ann_scales, const int32_t *d_n_ann,
                           float *d_pre_data, const int32_t *d_n_pre, const pp_inverse *inv,
                           int32_t B, int32_t F, int32_t ann_cap, void *stream) {
  if (!d_ann_data || !d_ann_scales || !d_n_ann || !inv || (d_pre_data && !d_n_pre)) return PP_ERR_ARG;
  return annotations_inverse_launch(d_ann_data, d_ann_scales, d_n_ann, d_pre_data, d_n_pre, *inv, B, F,
                                    ann_cap, (cudaStream_t)stream);
}

int pp_head_epilogue(const float *d_x, float *d_out, int32_t B, int32_t n_fields, int32_t n_vectors,
                     int32_t h, int32_t w, int32_t quad, void *stream) {
  if (!d_x || !d_out) return PP_ERR_ARG;
  return head_epilogue_launch(d_x, d_out, B, n_fields, n_vectors, h, w, quad != 0, (cudaStream_t)stream);
}

int pp_cifhr_accumulate(const float *d_cif, float *d_hr, void *d_ws, size_t ws_bytes,
                        const pp_shape *s, const pp_caps *c, const pp_params *p,
                        int32_t *d_status, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !d_cif || !d_hr || !d_ws || !d_status) return PP_ERR_ARG;
  if (ws_bytes < cifhr_ws_bytes(*s, *c)) return PP_ERR_WORKSPACE;
  return cifhr_launch(d_cif, d_hr, d_ws, *s, *c, *p, d_status, (cudaStream_t)stream);
}

int pp_cifhr_accumulate_ex(const float *d_cif, const float *d_cif2, float *d_hr, int32_t merge, int32_t det,
                           void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_shape *s_hr,
                           const pp_caps *c, const pp_params *p, int32_t *d_status, void *stream) {
  if (!s || !(det ? (s->B > 0 && s->F > 0 && s->F <= PP_MAX_CATEGORIES && s->H > 0 && s->W > 0 && s->stride > 0)
                  : shape_ok(s)))
    return PP_ERR_ARG;
  if (!caps_ok(s, c) || !p || !d_cif || !d_hr || !d_ws || !d_status) return PP_ERR_ARG;
  pp_shape q = s_hr ? *s_hr : *s;
  q.B = s->B; q.F = s->F; q.L = s->L;
  if (q.H <= 0 || q.W <= 0 || q.stride <= 0) return PP_ERR_ARG;
  if (ws_bytes < cifhr_ws_bytes(q, *c)) return PP_ERR_WORKSPACE;
  return cifhr_launch_ex(d_cif, d_cif2, d_hr, d_ws, *s, q, *c, *p, p->cif_min_scale, d_status,
                         (cudaStream_t)stream, det != 0, merge != 0, nullptr, nullptr);
}

int pp_cif_seeds(const float *d_cif, const float *d_hr, float *d_seeds, int32_t *d_n_seeds,
                 void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                 const pp_params *p, int32_t *d_status, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !d_cif || !d_hr || !d_seeds || !d_n_seeds || !d_ws ||
      !d_status)
    return PP_ERR_ARG;
  if (ws_bytes < seeds_ws_bytes(*s, *c)) return PP_ERR_WORKSPACE;
  return seeds_launch(d_cif, d_hr, d_seeds, d_n_seeds, d_ws, *s, *c, *p, d_status,
                      (cudaStream_t)stream);
}

int pp_caf_scored(const float *d_caf, const float *d_hr, const pp_skeleton *sk, float score_th,
                  float *d_fwd, int32_t *d_nfwd, float *d_bwd, int32_t *d_nbwd,
                  const pp_shape *s, const pp_caps *c, const pp_params *p, int32_t *d_status,
                  void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !sk || !d_caf || !d_hr || !d_fwd || !d_nfwd ||
      !d_bwd || !d_nbwd || !d_status || sk->L != s->L)
    return PP_ERR_ARG;
  for (int l = 0; l < sk->L; l++)
    if (sk->joints[l][0] < 1 || sk->joints[l][1] < 1) return PP_ERR_ARG;
  return caf_scored_launch(d_caf, d_hr, *sk, score_th, d_fwd, d_nfwd, d_bwd, d_nbwd, *s, *c, *p,
                           d_status, (cudaStream_t)stream);
}

int pp_cifcaf_assemble(const float *d_seeds, const int32_t *d_n_seeds, const float *d_fwd,
                       const int32_t *d_nfwd, const float *d_bwd, const int32_t *d_nbwd,
                       const float *d_fwd2, const int32_t *d_nfwd2, const float *d_bwd2,
                       const int32_t *d_nbwd2, const pp_skeleton *sk, const pp_outputs *d_out,
                       void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                       const pp_params *p, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !sk || !d_out || !d_seeds || !d_n_seeds || !d_fwd ||
      !d_bwd || !d_nfwd || !d_nbwd || !d_ws || sk->L != s->L)
    return PP_ERR_ARG;
  if (!d_out->ann_data || !d_out->ann_scales || !d_out->n_ann || !d_out->status) return PP_ERR_ARG;
  if (p->force_complete && (!d_fwd2 || !d_nfwd2 || !d_bwd2 || !d_nbwd2)) return PP_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  // workspace: occ counters + pre buffers (when the caller keeps none) + nms scratch
  size_t B = s->B;
  size_t need_occ = align_up(B * c->seed_cap, 256);
  size_t need_rec = align_up(B * c->seed_cap * sizeof(int2), 256);
  size_t need_pd = align_up(B * c->ann_cap * s->F * 3 * sizeof(float), 256);
  size_t need_ps = align_up(B * c->ann_cap * s->F * sizeof(float), 256);
  size_t need_po = align_up(B * c->ann_cap * s->F * 2 * sizeof(int), 256);
  size_t need_np = align_up(B * sizeof(int), 256);
  size_t need_br = align_up(binned_rows_bytes(*s, *c), 256);
  size_t need_bs = align_up(binned_start_bytes(*s), 256);
  size_t need = need_occ + need_rec + 2 * need_pd + 2 * need_ps + need_po + need_np + need_br + need_bs +
                nms_ws_bytes(*s, *c);
  if (ws_bytes < need) return PP_ERR_WORKSPACE;
  char *q = (char *)d_ws;
  AsmArgs a = {};
  a.occ_cnt = (uint8_t *)q; q += need_occ;
  a.occ_rec = (int2 *)q; q += need_rec;
  a.pre_data = d_out->pre_data ? d_out->pre_data : (float *)q; q += need_pd;
  a.pre_scales = d_out->pre_scales ? d_out->pre_scales : (float *)q; q += need_ps;
  a.pre_order = d_out->pre_order ? d_out->pre_order : (int *)q; q += need_po;
  a.n_pre = d_out->n_pre ? d_out->n_pre : (int *)q; q += need_np;
  float *cmp_data = (float *)q; q += need_pd;
  float *cmp_scales = (float *)q; q += need_ps;
  float *bin_rows = (float *)q; q += need_br;
  int *bin_start = (int *)q; q += need_bs;
  a.seeds = d_seeds; a.n_seeds = d_n_seeds;
  int rcb = caf_bin_launch(d_fwd, d_nfwd, d_bwd, d_nbwd, bin_rows, bin_start, *s, *c, st);
  if (rcb) return rcb;
  a.rows = bin_rows; a.start = bin_start; a.geo = bin_geo(*s); a.nfwd = d_nfwd; a.nbwd = d_nbwd;
  a.status = d_out->status;
  a.seed_cap = c->seed_cap; a.caf_cap = c->caf_cap; a.ann_cap = c->ann_cap; a.L = s->L;
  a.hh = hr_h(*s); a.hw = hr_w(*s);
  if (p->initial) {
    if (!p->initial->data || !p->initial->scales || !p->initial->n || p->initial->cap <= 0) return PP_ERR_ARG;
    a.init_data = p->initial->data; a.init_scales = p->initial->scales;
    a.init_n = p->initial->n; a.init_cap = p->initial->cap;
  }
  if (!p->force_complete) {
    a.fuse_nms = 1;
    a.nms = nms_make_args(a.pre_data, a.pre_scales, a.n_pre, d_out->ann_data, d_out->ann_scales,
                          d_out->ann_score, d_out->ann_source, d_out->n_ann, q, *s, *c, *p);
  }
  int rc = assemble_launch(a, *sk, s->B, s->F, *p, st);
  if (rc) return rc;
  if (p->force_complete) {
    // the first lists' binned copy is no longer needed: reuse it for the completion lists
    if ((rc = caf_bin_launch(d_fwd2, d_nfwd2, d_bwd2, d_nbwd2, bin_rows, bin_start, *s, *c, st))) return rc;
    rc = complete_launch(a, bin_rows, bin_start, d_nfwd2, d_nbwd2, cmp_data, cmp_scales, *sk, s->B, s->F,
                         *p, st);
    if (rc) return rc;
    return nms_launch(cmp_data, cmp_scales, a.n_pre, d_out->ann_data, d_out->ann_scales,
                      d_out->ann_score, d_out->ann_source, d_out->n_ann, q, *s, *c, *p, st);
  }
  return PP_OK;      // the NMS ran inside the assembly kernel
}

int pp_keypoint_nms(float *d_ann_data, float *d_ann_scales, int32_t *d_n_ann, double *d_score,
                    int32_t *d_source, void *d_ws, size_t ws_bytes, const pp_shape *s,
                    const pp_caps *c, const pp_params *p, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !d_ann_data || !d_ann_scales || !d_n_ann || !d_ws)
    return PP_ERR_ARG;
  if (ws_bytes < nms_ws_bytes(*s, *c)) return PP_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  size_t bc = (size_t)s->B * c->ann_cap;
  size_t data_bytes = align_up(bc * s->F * 3 * sizeof(float), 256);
  char *stage = (char *)d_ws + nms_ws_staging_offset(*s, *c);
  float *in_data = (float *)stage;
  float *in_scales = (float *)(stage + data_bytes);
  int *n_in = (int *)(stage + 2 * data_bytes);
  PP_CUDA_CHECK(cudaMemcpyAsync(in_data, d_ann_data, bc * s->F * 3 * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  PP_CUDA_CHECK(cudaMemcpyAsync(in_scales, d_ann_scales, bc * s->F * sizeof(float),
                                cudaMemcpyDeviceToDevice, st));
  PP_CUDA_CHECK(cudaMemcpyAsync(n_in, d_n_ann, (size_t)s->B * sizeof(int), cudaMemcpyDeviceToDevice, st));
  return nms_launch(in_data, in_scales, n_in, d_ann_data, d_ann_scales, d_score, d_source, d_n_ann,
                    d_ws, *s, *c, *p, st);
}

// One decode over n_sets field sets (multi-scale FieldConfig, decoder/factory.py:153-180); the plain
// single-scale decode is the case n_sets = 1.
static int decode_core(const pp_field_set *sets, int n_sets, int hflip_pairs, const pp_skeleton *sk,
                       const pp_outputs *o, void *d_ws, size_t ws_bytes, const pp_shape *s,
                       const pp_caps *c, const pp_params *p, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !p || !sk || !o || !sets || n_sets < 1 || n_sets > PP_MAX_SETS ||
      !d_ws || sk->L != s->L)
    return PP_ERR_ARG;
  if (!o->ann_data || !o->ann_scales || !o->n_ann || !o->status) return PP_ERR_ARG;
  if (hflip_pairs && (n_sets & 1)) return PP_ERR_ARG;
  for (int k = 0; k < n_sets; k++)
    if (!sets[k].cif || !sets[k].caf || sets[k].H <= 0 || sets[k].W <= 0 || sets[k].stride <= 0)
      return PP_ERR_ARG;
  if (sets[0].H != s->H || sets[0].W != s->W || sets[0].stride != s->stride) return PP_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
#ifdef PP_DEBUG_SKIP
  // timing experiments only (scripts/stage_skip_probe.py): leave stages out; results are those of the last
  // full call on the same buffers.  1 seeds + sort, 2 CafScored + binning, 4 assembly, 8 NMS, 16 CifHr
  const int dbg_skip = getenv("PP_DEBUG_SKIP") ? atoi(getenv("PP_DEBUG_SKIP")) : 0;
#else
  const int dbg_skip = 0;
#endif
  const bool need_hr = o->hr == nullptr;
  DecodeWs w = layout(d_ws, *s, *c, need_hr);
  if (ws_bytes < w.total) return PP_ERR_WORKSPACE;
  float *hr = need_hr ? w.hr : o->hr;
  float *seeds = o->seeds ? o->seeds : w.seeds;
  int *n_seeds = o->n_seeds ? o->n_seeds : w.n_seeds;
  float *fwd = o->caf_fwd ? o->caf_fwd : w.fwd;
  float *bwd = o->caf_bwd ? o->caf_bwd : w.bwd;
  int *nfwd = o->n_caf_fwd ? o->n_caf_fwd : w.nfwd;
  int *nbwd = o->n_caf_bwd ? o->n_caf_bwd : w.nbwd;
  auto set_shape = [&](int k) {
    pp_shape q = *s;
    q.H = sets[k].H; q.W = sets[k].W; q.stride = sets[k].stride;
    return q;
  };

  PP_CUDA_CHECK(cudaMemsetAsync(o->status, 0, (size_t)s->B * sizeof(int), st));
  prof_begin_step(p->profile);
  int rc;
  // CifHr.fill (cif_hr.py:66-78): one map per group, merged with np.maximum; with 10 heads the
  // groups are the hflip pairs (i, i + 5), each accumulated with v / neighbors / 2
  const int n_groups = hflip_pairs ? n_sets / 2 : n_sets;
  // One field set with worst-case list capacity: the hr-independent half of CafScored rides in the CIF
  // compact launch (raw candidates into the forward lists) and one kernel after the splat finishes the
  // lists and bins them.  Otherwise (multi-scale, shrunken caf_cap) the stage kernels run one by one.
  const bool caf_split = n_sets == 1 && c->caf_cap >= s->H * s->W && !(dbg_skip & (2 | 16));
  CafPreArgs pre = {};
  if (caf_split) {
    const float stride = (float)s->stride;
    pre.caf = sets[0].caf; pre.caf_rest = sets[0].caf_rest; pre.raw = fwd; pre.n_raw = nfwd;
    pre.L = s->L; pre.cap = c->caf_cap; pre.th = p->caf_threshold;
    pre.min_dist = sets[0].caf_min_distance / stride; pre.use_min = sets[0].caf_min_distance != 0.0f;
    pre.max_dist = sets[0].caf_max_distance / stride; pre.use_max = sets[0].caf_max_distance > 0.0f;
    // host-mapped fields: every plane value read is a PCIe transaction; the two log-b rows only exist in
    // the CafScored lists a caller may ask for
    pre.skip_logb = sets[0].caf_rest != nullptr && !o->caf_fwd && !o->caf_bwd;
  }
  for (int g = 0; g < n_groups && !(dbg_skip & 16); g++) {
    const pp_shape q = set_shape(g);
    const float *second = hflip_pairs ? sets[g + n_groups].cif : nullptr;
    if (second && (sets[g + n_groups].H != q.H || sets[g + n_groups].W != q.W)) return PP_ERR_ARG;
    if ((rc = cifhr_launch_ex(sets[g].cif, second, hr, w.stage, q, *s, *c, *p, sets[g].cif_min_scale,
                              o->status, st, 0, g > 0, second ? nullptr : sets[g].cif_rest,
                              caf_split ? &pre : nullptr)))
      return rc;
  }
  // CifSeeds.fill (cif_seeds.py:56-65): every head on its own, one sorted list
  bool seeds_done = false;
  if (dbg_skip & 1) {
  } else if (n_sets == 1 && p->seed_threshold >= p->cif_threshold &&
             (p->cif_neighbors == 0.0f || p->cif_neighbors == 16.0f)) {
    // the seed candidates are a subset of the sources the CifHr stage compacted: skip the field scan
    if (seeds_segmented_ok(*s, *c)) {
      if ((rc = seeds_segmented_launch(cifhr_sources(w.stage, *s, *c), hr, seeds, n_seeds, w.seed_stage, *s, *c,
                                       *p, st, 0)))
        return rc;
      seeds_done = true;
    } else if ((rc = seeds_collect_from_sources_launch(cifhr_sources(w.stage, *s, *c), hr, w.seed_stage, *s, *c,
                                                       *p, st, 0))) {
      return rc;
    }
  } else {
    for (int k = 0; k < n_sets; k++)
      if ((rc = seeds_collect_launch(sets[k].cif, hr, w.seed_stage, set_shape(k), *s, *c, *p,
                                     sets[k].cif_min_scale, k == 0, st, 0)))
        return rc;
  }
  if (!(dbg_skip & 1) && !seeds_done &&
      (rc = seeds_sort_launch(seeds, n_seeds, w.seed_stage, s->B, *c, *p, o->status, st, 0)))
    return rc;
  // CafScored.fill (caf_scored.py:89-98): lists concatenated over the heads
  for (int k = 0; k < n_sets && !(dbg_skip & 2) && !caf_split; k++)
    if ((rc = caf_scored_launch_ex(sets[k].caf, hr, *sk, p->caf_threshold, fwd, nfwd, bwd, nbwd, set_shape(k),
                                   *s, *c, *p, sets[k].caf_min_distance, sets[k].caf_max_distance, k > 0,
                                   o->status, st, sets[k].caf_rest)))
      return rc;
  AsmArgs a = {};
  a.seeds = seeds; a.n_seeds = n_seeds;
  if (caf_split) {
    if ((rc = caf_score_bin_launch(hr, *sk, p->caf_threshold, fwd, nfwd, bwd, nbwd, w.bin_rows, w.bin_start, *s,
                                   *c, *p, o->status, st)))
      return rc;
  } else if (!(dbg_skip & 2) && (rc = caf_bin_launch(fwd, nfwd, bwd, nbwd, w.bin_rows, w.bin_start, *s, *c, st))) {
    return rc;
  }
  a.rows = w.bin_rows; a.start = w.bin_start; a.geo = bin_geo(*s); a.nfwd = nfwd; a.nbwd = nbwd;
  a.occ_cnt = w.occ_cnt; a.occ_rec = w.occ_rec;
  a.pre_data = o->pre_data ? o->pre_data : w.pre_data;
  a.pre_scales = o->pre_scales ? o->pre_scales : w.pre_scales;
  a.pre_order = o->pre_order ? o->pre_order : w.pre_order;
  a.n_pre = o->n_pre ? o->n_pre : w.n_pre;
  a.status = o->status;
  a.seed_cap = c->seed_cap; a.caf_cap = c->caf_cap; a.ann_cap = c->ann_cap; a.L = s->L;
  a.hh = hr_h(*s); a.hw = hr_w(*s);
  if (p->initial) {
    if (!p->initial->data || !p->initial->scales || !p->initial->n || p->initial->cap <= 0) return PP_ERR_ARG;
    a.init_data = p->initial->data; a.init_scales = p->initial->scales;
    a.init_n = p->initial->n; a.init_cap = p->initial->cap;
  }
  // without force_complete the keypoint NMS of an image runs as the tail of its assembly CTA
  const bool fuse_nms = !p->force_complete && !(dbg_skip & (4 | 8));
  if (fuse_nms) {
    a.fuse_nms = 1;
    a.nms = nms_make_args(a.pre_data, a.pre_scales, a.n_pre, o->ann_data, o->ann_scales, o->ann_score,
                          o->ann_source, o->n_ann, w.stage, *s, *c, *p);
  }
  if (!(dbg_skip & 4) && (rc = assemble_launch(a, *sk, s->B, s->F, *p, st))) return rc;
  const float *nms_data = a.pre_data, *nms_scales = a.pre_scales;
  if (p->force_complete) {                                           // cifcaf.py:110-115
    // second CafScored(score_th=0.0001) into the workspace lists (the first
    // ones are no longer needed, or live in the caller's buffers)
    for (int k = 0; k < n_sets; k++)
      if ((rc = caf_scored_launch_ex(sets[k].caf, hr, *sk, 0.0001f, w.fwd, w.nfwd, w.bwd, w.nbwd, set_shape(k),
                                     *s, *c, *p, sets[k].caf_min_distance, sets[k].caf_max_distance, k > 0,
                                     o->status, st, sets[k].caf_rest)))
        return rc;
    if ((rc = caf_bin_launch(w.fwd, w.nfwd, w.bwd, w.nbwd, w.bin_rows, w.bin_start, *s, *c, st)))
      return rc;
    if ((rc = complete_launch(a, w.bin_rows, w.bin_start, w.nfwd, w.nbwd, w.cmp_data, w.cmp_scales,
                              *sk, s->B, s->F, *p, st)))
      return rc;
    nms_data = w.cmp_data; nms_scales = w.cmp_scales;
  }
  if (fuse_nms) {
    prof_mark(p->profile, kStNms, st);      // (its time is inside the assembly stage)
    rc = PP_OK;
  } else {
    rc = (dbg_skip & 8) ? PP_OK
                        : nms_launch(nms_data, nms_scales, a.n_pre, o->ann_data, o->ann_scales, o->ann_score,
                                     o->ann_source, o->n_ann, w.stage, *s, *c, *p, st);
  }
  if (!rc && p->inverse)      // optional Preprocess.annotations_inverse on the results (inside the nms stage time)
    rc = annotations_inverse_launch(o->ann_data, o->ann_scales, o->n_ann, a.pre_data, a.n_pre, *p->inverse,
                                    s->B, s->F, c->ann_cap, st);
  debug_crumb(st, 50);
  if (!rc && o->packed) {      // one dense block of the live rows for the result copy
    pp_outputs po = *o;
    po.pre_data = a.pre_data; po.pre_scales = a.pre_scales; po.pre_order = a.pre_order; po.n_pre = a.n_pre;
    rc = pack_launch(po, *s, *c, st);
  }
  prof_end_step(p->profile, st);
  return rc;
}

int pp_decode_batch(const float *d_cif, const float *d_caf, const pp_skeleton *sk,
                    const pp_outputs *o, void *d_ws, size_t ws_bytes, const pp_shape *s,
                    const pp_caps *c, const pp_params *p, void *stream) {
  if (!s || !p) return PP_ERR_ARG;
  pp_field_set one = {d_cif, d_caf, s->H, s->W, s->stride, p->cif_min_scale, p->caf_min_distance,
                      p->caf_max_distance, nullptr, nullptr};
  return decode_core(&one, 1, 0, sk, o, d_ws, ws_bytes, s, c, p, stream);
}

// ---- host-mapped input ---------------------------------------------------------
// The decoder looks at every confidence value but at the other planes only where a cell is above
// the threshold (~1 % of the cells).  With the fields in pinned, device-mapped host memory the
// confidence planes are copied (strided DMA, 14 % of the bytes) and the kernels fetch the rest of
// the active cells over PCIe on demand.
static bool mapped_applies(const pp_params *p) {
  return p->host_mapped && !p->force_complete && p->seed_threshold >= p->cif_threshold &&
         (p->cif_neighbors == 0.0f || p->cif_neighbors == 16.0f);     // = the seeds come from the compacted sources
}

// device-visible alias of a pinned host buffer, or NULL
static const float *mapped_alias(const float *h) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, h) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (attr.type != cudaMemoryTypeHost || !attr.devicePointer) return nullptr;
  return (const float *)attr.devicePointer;
}

// copy the confidence plane of every (image, field) block of `planes` planes
static int copy_conf_planes(float *d, const float *h, size_t blocks, int planes, size_t cells,
                            cudaStream_t st) {
  const size_t pitch = (size_t)planes * cells * sizeof(float);
  PP_CUDA_CHECK(cudaMemcpy2DAsync(d, pitch, h, pitch, cells * sizeof(float), blocks,
                                  cudaMemcpyHostToDevice, st));
  return PP_OK;
}

static int decode_mapped(const float *d_cif, const float *d_caf, const float *cif_rest, const float *caf_rest,
                         const pp_skeleton *sk, const pp_outputs *o, void *d_ws, size_t ws_bytes,
                         const pp_shape *s, const pp_caps *c, const pp_params *p, void *stream) {
  pp_field_set one = {d_cif, d_caf, s->H, s->W, s->stride, p->cif_min_scale, p->caf_min_distance,
                      p->caf_max_distance, cif_rest, caf_rest};
  return decode_core(&one, 1, 0, sk, o, d_ws, ws_bytes, s, c, p, stream);
}

int pp_decode_batch_multiscale(const pp_field_set *sets, int32_t n_sets, int32_t hflip_pairs,
                               const pp_skeleton *sk, const pp_outputs *o, void *d_ws, size_t ws_bytes,
                               const pp_shape *s, const pp_caps *c, const pp_params *p, void *stream) {
  return decode_core(sets, n_sets, hflip_pairs, sk, o, d_ws, ws_bytes, s, c, p, stream);
}

// ---- CifDet ------------------------------------------------------------------
void pp_det_default_params(pp_det_params *p) {
  if (!p) return;
  memset(p, 0, sizeof *p);
  p->cif_threshold = 0.1f;
  p->seed_threshold = 0.2f;
  p->seed_score_scale = 1.0f;
  p->cif_min_scale = 0.0f;
  p->nms_suppression = 0.1f;
  p->nms_suppression_soft = 0.3f;
  p->nms_instance_threshold = 0.1f;
  p->nms_iou_threshold = 0.7f;
  p->nms_iou_threshold_soft = 0.5f;
  p->nms = 1;
}

static bool det_shape_ok(const pp_shape *s) {
  return s && s->B > 0 && s->F > 0 && s->F <= PP_MAX_CATEGORIES && s->H > 0 && s->W > 0 && s->stride > 0;
}

struct DetWs {
  float *hr; float *seeds; int *n_seeds;
  int *pre_field; float *pre_score; float *pre_box; int *n_pre;
  uint8_t *occ_cnt; int2 *occ_rec; float *work_score;
  void *stage, *seed_stage;
  size_t total;
};

static DetWs det_layout(void *base, const pp_shape &s, const pp_caps &c) {
  DetWs w;
  char *p = (char *)base;
  size_t off = 0;
  auto take = [&](size_t bytes) { void *r = p ? p + off : nullptr; off += align_up(bytes, 256); return r; };
  const size_t B = s.B, A = c.ann_cap;
  w.hr = (float *)take(B * s.F * hr_h(s) * hr_w(s) * sizeof(float));
  w.seeds = (float *)take(B * c.seed_cap * 6 * sizeof(float));
  w.n_seeds = (int *)take(B * sizeof(int));
  w.pre_field = (int *)take(B * A * sizeof(int));
  w.pre_score = (float *)take(B * A * sizeof(float));
  w.pre_box = (float *)take(B * A * 4 * sizeof(float));
  w.n_pre = (int *)take(B * sizeof(int));
  w.occ_cnt = (uint8_t *)take(B * c.seed_cap);
  w.occ_rec = (int2 *)take(B * c.seed_cap * sizeof(int2));
  w.work_score = (float *)take(B * A * sizeof(float));
  w.stage = take(cifhr_ws_bytes(s, c));
  w.seed_stage = take(seeds_ws_bytes(s, c, 1));
  w.total = off;
  return w;
}

size_t pp_cifdet_workspace_bytes(const pp_shape *s, const pp_caps *c) {
  if (!det_shape_ok(s) || !caps_ok(s, c)) return 0;
  return det_layout(nullptr, *s, *c).total;
}

int pp_cifdet_decode_batch(const float *d_field, const pp_det_outputs *o, void *d_ws, size_t ws_bytes,
                           const pp_shape *s, const pp_caps *c, const pp_det_params *p, void *stream) {
  if (!det_shape_ok(s) || !caps_ok(s, c) || !p || !o || !d_field || !d_ws || !o->det_field ||
      !o->det_score || !o->det_box || !o->n_det || !o->status)
    return PP_ERR_ARG;
  DetWs w = det_layout(d_ws, *s, *c);
  if (ws_bytes < w.total) return PP_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  pp_params q;
  pp_default_params(&q);
  q.cif_threshold = p->cif_threshold;
  q.seed_threshold = p->seed_threshold;
  q.seed_score_scale = p->seed_score_scale;
  q.cif_min_scale = p->cif_min_scale;
  PP_CUDA_CHECK(cudaMemsetAsync(o->status, 0, (size_t)s->B * sizeof(int), st));
  float *hr = o->hr ? o->hr : w.hr;
  int rc = cifhr_launch(d_field, hr, w.stage, *s, *c, q, o->status, st, 1);
  if (rc) return rc;
  float *seeds = o->seeds ? o->seeds : w.seeds;
  int *n_seeds = o->n_seeds ? o->n_seeds : w.n_seeds;
  if (q.seed_threshold >= q.cif_threshold && seeds_segmented_ok(*s, *c)) {
    if ((rc = seeds_segmented_launch(cifhr_sources(w.stage, *s, *c), hr, seeds, n_seeds, w.seed_stage, *s, *c, q,
                                     st, 1)))
      return rc;
  } else if (q.seed_threshold >= q.cif_threshold) {      // seeds from the compacted sources (see decode_core)
    if ((rc = seeds_collect_from_sources_launch(cifhr_sources(w.stage, *s, *c), hr, w.seed_stage, *s, *c, q,
                                                st, 1)))
      return rc;
    if ((rc = seeds_sort_launch(seeds, n_seeds, w.seed_stage, s->B, *c, q, o->status, st, 1))) return rc;
  } else if ((rc = seeds_launch(d_field, hr, seeds, n_seeds, w.seed_stage, *s, *c, q, o->status, st, 1))) {
    return rc;
  }
  DetArgs a;
  a.seeds = seeds; a.n_seeds = n_seeds;
  a.occ_cnt = w.occ_cnt; a.occ_rec = w.occ_rec;
  a.pre_field = o->pre_field ? o->pre_field : w.pre_field;
  a.pre_score = o->pre_score ? o->pre_score : w.pre_score;
  a.pre_box = o->pre_box ? o->pre_box : w.pre_box;
  a.n_pre = o->n_pre ? o->n_pre : w.n_pre;
  a.det_field = o->det_field; a.det_score = o->det_score; a.det_box = o->det_box; a.n_det = o->n_det;
  a.det_source = o->det_source;
  a.status = o->status;
  a.seed_cap = c->seed_cap; a.ann_cap = c->ann_cap; a.hh = hr_h(*s); a.hw = hr_w(*s); a.C = s->F;
  return det_assemble_launch(a, s->B, *p, w.work_score, st);
}

int pp_detections_inverse(float *d_box, const int32_t *d_n, const pp_inverse *inv, int32_t B,
                          int32_t ann_cap, void *stream) {
  if (!inv) return PP_ERR_ARG;
  return detections_inverse_launch(d_box, d_n, *inv, B, ann_cap, (cudaStream_t)stream);
}

size_t pp_detection_nms_workspace_bytes(int32_t B, int32_t cap) {
  return (B > 0 && cap > 0) ? det_nms_ws_bytes(B, cap) : 0;
}

int pp_detection_nms(const int32_t *d_field, const float *d_score, const float *d_box,
                     const int32_t *d_n, int32_t *d_out_field, float *d_out_score, float *d_out_box,
                     int32_t *d_out_source, int32_t *d_out_n, void *d_ws, size_t ws_bytes, int32_t B, int32_t cap,
                     const pp_det_params *p, void *stream) {
  if (!d_field || !d_score || !d_box || !d_n || !d_out_field || !d_out_score || !d_out_box || !d_out_n ||
      !d_ws || !p || B <= 0 || cap <= 0)
    return PP_ERR_ARG;
  if (ws_bytes < det_nms_ws_bytes(B, cap)) return PP_ERR_WORKSPACE;
  return det_nms_launch(d_field, d_score, d_box, d_n, d_out_field, d_out_score, d_out_box, d_out_source, d_out_n, d_ws,
                        B, cap, *p, (cudaStream_t)stream);
}

// ---- host entries ------------------------------------------------------------

struct OutField { size_t offset; size_t bytes_per_image; };

// byte size of one image's slice of every pp_outputs member, in declaration order
static void output_strides(const pp_shape &s, const pp_caps &c, size_t out[18]) {
  const size_t F = s.F, L = s.L, A = c.ann_cap;
  const size_t f = sizeof(float), i = sizeof(int32_t);
  out[0] = F * hr_h(s) * hr_w(s) * f;   // hr
  out[1] = (size_t)c.seed_cap * 5 * f;  // seeds
  out[2] = i;                           // n_seeds
  out[3] = L * 9 * c.caf_cap * f;       // caf_fwd
  out[4] = L * 9 * c.caf_cap * f;       // caf_bwd
  out[5] = L * i;                       // n_caf_fwd
  out[6] = L * i;                       // n_caf_bwd
  out[7] = A * F * 3 * f;               // ann_data
  out[8] = A * F * f;                   // ann_scales
  out[9] = A * sizeof(double);          // ann_score
  out[10] = A * i;                      // ann_source
  out[11] = i;                          // n_ann
  out[12] = A * F * 3 * f;              // pre_data
  out[13] = A * F * f;                  // pre_scales
  out[14] = A * F * 2 * i;              // pre_order
  out[15] = i;                          // n_pre
  out[16] = i;                          // status
  out[17] = packed_image_bytes(s.F, c.ann_cap);   // packed: block of a chunk starts at its first image's share
}

static_assert(sizeof(pp_outputs) == 18 * sizeof(void *), "pp_outputs member table out of date");
constexpr int kOutPacked = 17;

// the outputs of images [b0, ...) as their own pp_outputs
static pp_outputs slice_outputs(const pp_outputs &o, const size_t strides[18], size_t b0) {
  pp_outputs r = o;
  char **src = (char **)&r;
  for (int k = 0; k < 18; k++)
    if (src[k]) src[k] += b0 * strides[k];
  return r;
}

// every result array both sides have, except the packed block (its live part only: below)
static int copy_outputs_to_host(const pp_outputs &d, const pp_outputs &h, const size_t strides[18],
                                size_t b0, size_t nb, cudaStream_t st) {
  char *const *dp = (char *const *)&d;
  char *const *hp = (char *const *)&h;
  for (int k = 0; k < kOutPacked; k++)
    if (dp[k] && hp[k])
      PP_CUDA_CHECK(cudaMemcpyAsync(hp[k] + b0 * strides[k], dp[k] + b0 * strides[k], nb * strides[k],
                                    cudaMemcpyDeviceToHost, st));
  return PP_OK;
}

// bytes of a packed block of nb images that hold the header and the first `records` records
static size_t packed_prefix_bytes(size_t nb, int F, size_t records) {
  return (packed_header_words((int)nb) + records * (size_t)packed_record_words(F)) * sizeof(int32_t);
}

// status words of a finished packed block
static bool packed_block_overflowed(const void *h_block, size_t nb) {
  const int32_t *w = (const int32_t *)h_block;
  for (size_t b = 0; b < nb; b++)
    if ((w[4 + nb + b] & PP_STATUS_OVERFLOW_MASK) != 0) return true;
  return false;
}

int pp_decode_batch_host(const float *h_cif, const float *h_caf, float *d_cif, float *d_caf,
                         const pp_skeleton *sk, const pp_outputs *d_out, const pp_outputs *h_out,
                         void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                         const pp_params *p, void *stream) {
  if (!shape_ok(s) || !caps_ok(s, c) || !h_cif || !h_caf || !d_cif || !d_caf || !d_out || !h_out)
    return PP_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  size_t B = s->B, cells = (size_t)s->H * s->W;
  const float *m_cif = (p && mapped_applies(p)) ? mapped_alias(h_cif) : nullptr;
  const float *m_caf = m_cif ? mapped_alias(h_caf) : nullptr;
  int rc;
  if (m_cif && m_caf) {
    if ((rc = copy_conf_planes(d_cif, h_cif, B * s->F, 5, cells, st))) return rc;
    if ((rc = copy_conf_planes(d_caf, h_caf, B * s->L, 9, cells, st))) return rc;
    rc = decode_mapped(d_cif, d_caf, m_cif, m_caf, sk, d_out, d_ws, ws_bytes, s, c, p, stream);
  } else {
    PP_CUDA_CHECK(cudaMemcpyAsync(d_cif, h_cif, B * s->F * 5 * cells * sizeof(float),
                                  cudaMemcpyHostToDevice, st));
    PP_CUDA_CHECK(cudaMemcpyAsync(d_caf, h_caf, B * s->L * 9 * cells * sizeof(float),
                                  cudaMemcpyHostToDevice, st));
    rc = pp_decode_batch(d_cif, d_caf, sk, d_out, d_ws, ws_bytes, s, c, p, stream);
  }
  if (rc) return rc;
  size_t strides[18];
  output_strides(*s, *c, strides);
  if ((rc = copy_outputs_to_host(*d_out, *h_out, strides, 0, B, st))) return rc;
  const bool packed = d_out->packed && h_out->packed;
  if (packed)      // counts first ...
    PP_CUDA_CHECK(cudaMemcpyAsync(h_out->packed, d_out->packed, packed_prefix_bytes(B, s->F, 0),
                                  cudaMemcpyDeviceToHost, st));
  PP_CUDA_CHECK(cudaStreamSynchronize(st));
  if (packed) {    // ... then exactly the live rows
    const size_t total = (size_t)((const int32_t *)h_out->packed)[2];
    const size_t head = packed_prefix_bytes(B, s->F, 0);
    if (total > 0) {
      PP_CUDA_CHECK(cudaMemcpyAsync((char *)h_out->packed + head, (const char *)d_out->packed + head,
                                    packed_prefix_bytes(B, s->F, total) - head, cudaMemcpyDeviceToHost, st));
      PP_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    if (packed_block_overflowed(h_out->packed, B)) return PP_ERR_CAPACITY;
  }
  if (h_out->status)
    for (size_t b = 0; b < B; b++)
      if ((h_out->status[b] & PP_STATUS_OVERFLOW_MASK) != 0) return PP_ERR_CAPACITY;
  return PP_OK;
}

}  // extern "C"

struct pp_pipeline {
  int max_chunks, n_chunks, chunk_images, B;
  cudaStream_t s_in, s_out;
  cudaEvent_t ev_start;
  cudaEvent_t *ev_in, *ev_done, *ev_out;   // (max_chunks) each
  const int32_t *h_status;
  // packed result blocks (one per chunk): the copy-out moves header + rec_guess records per image; a chunk
  // with more live rows gets its remainder in pp_pipeline_wait, which also raises the guess
  char *h_packed;
  const char *d_packed;
  size_t packed_stride;
  int F, rec_guess;
};

extern "C" {

pp_pipeline *pp_pipeline_create(int max_chunks) {
  if (max_chunks <= 0) return nullptr;
  pp_pipeline *q = (pp_pipeline *)calloc(1, sizeof(pp_pipeline));
  if (!q) return nullptr;
  q->max_chunks = max_chunks;
  q->rec_guess = 6;
  q->ev_in = (cudaEvent_t *)calloc(3 * (size_t)max_chunks, sizeof(cudaEvent_t));
  bool ok = q->ev_in != nullptr;
  if (ok) {
    q->ev_done = q->ev_in + max_chunks;
    q->ev_out = q->ev_done + max_chunks;
    ok = cudaStreamCreateWithFlags(&q->s_in, cudaStreamNonBlocking) == cudaSuccess &&
         cudaStreamCreateWithFlags(&q->s_out, cudaStreamNonBlocking) == cudaSuccess &&
         cudaEventCreateWithFlags(&q->ev_start, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < 3 * max_chunks; i++)
      ok = cudaEventCreateWithFlags(&q->ev_in[i], cudaEventDisableTiming) == cudaSuccess;
  }
  if (!ok) { pp_pipeline_destroy(q); return nullptr; }
  return q;
}

void pp_pipeline_destroy(pp_pipeline *q) {
  if (!q) return;
  if (q->ev_in) {
    for (int i = 0; i < 3 * q->max_chunks; i++)
      if (q->ev_in[i]) cudaEventDestroy(q->ev_in[i]);
    free(q->ev_in);
  }
  if (q->ev_start) cudaEventDestroy(q->ev_start);
  if (q->s_in) cudaStreamDestroy(q->s_in);
  if (q->s_out) cudaStreamDestroy(q->s_out);
  free(q);
}

int pp_pipeline_chunks(const pp_pipeline *q) { return q ? q->n_chunks : 0; }

int pp_decode_batch_host_async(pp_pipeline *q, const float *h_cif, const float *h_caf, float *d_cif,
                               float *d_caf, const pp_skeleton *sk, const pp_outputs *d_out,
                               const pp_outputs *h_out, void *d_ws, size_t ws_bytes,
                               const pp_shape *s, const pp_caps *c, const pp_params *p,
                               int chunk_images, void *stream) {
  if (!q || !shape_ok(s) || !caps_ok(s, c) || !h_cif || !h_caf || !d_cif || !d_caf || !d_out ||
      !h_out || !p || !sk)
    return PP_ERR_ARG;
  if (chunk_images <= 0 || chunk_images > s->B) chunk_images = s->B;
  const int n_chunks = (s->B + chunk_images - 1) / chunk_images;
  if (n_chunks > q->max_chunks) return PP_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  // n_chunks is set once every chunk has been enqueued: after a failed submit pp_pipeline_wait returns
  // PP_ERR_ARG instead of waiting on events this call never recorded
  q->n_chunks = 0; q->chunk_images = chunk_images; q->B = s->B;
  q->h_status = h_out->status;
  const size_t cells = (size_t)s->H * s->W;
  const size_t cif_img = (size_t)s->F * 5 * cells, caf_img = (size_t)s->L * 9 * cells;
  size_t strides[18];
  output_strides(*s, *c, strides);
  const bool packed = d_out->packed && h_out->packed;
  q->h_packed = packed ? (char *)h_out->packed : nullptr;
  q->d_packed = packed ? (const char *)d_out->packed : nullptr;
  q->packed_stride = strides[kOutPacked];
  q->F = s->F;
  if (q->rec_guess > c->ann_cap) q->rec_guess = c->ann_cap;

  const float *m_cif = mapped_applies(p) ? mapped_alias(h_cif) : nullptr;
  const float *m_caf = m_cif ? mapped_alias(h_caf) : nullptr;
  const bool mapped = m_cif && m_caf;
  // the copy-in stream starts after whatever is already queued on `stream`
  PP_CUDA_CHECK(cudaEventRecord(q->ev_start, st));
  PP_CUDA_CHECK(cudaStreamWaitEvent(q->s_in, q->ev_start, 0));
  for (int k = 0; k < n_chunks; k++) {
    const size_t b0 = (size_t)k * chunk_images;
    const size_t nb = (size_t)s->B - b0 < (size_t)chunk_images ? (size_t)s->B - b0 : (size_t)chunk_images;
    if (mapped) {
      int rcc;
      if ((rcc = copy_conf_planes(d_cif + b0 * cif_img, h_cif + b0 * cif_img, nb * s->F, 5, cells, q->s_in)))
        return rcc;
      if ((rcc = copy_conf_planes(d_caf + b0 * caf_img, h_caf + b0 * caf_img, nb * s->L, 9, cells, q->s_in)))
        return rcc;
    } else {
      PP_CUDA_CHECK(cudaMemcpyAsync(d_cif + b0 * cif_img, h_cif + b0 * cif_img, nb * cif_img * sizeof(float),
                                    cudaMemcpyHostToDevice, q->s_in));
      PP_CUDA_CHECK(cudaMemcpyAsync(d_caf + b0 * caf_img, h_caf + b0 * caf_img, nb * caf_img * sizeof(float),
                                    cudaMemcpyHostToDevice, q->s_in));
    }
    PP_CUDA_CHECK(cudaEventRecord(q->ev_in[k], q->s_in));
  }
  for (int k = 0; k < n_chunks; k++) {
    const size_t b0 = (size_t)k * chunk_images;
    const size_t nb = (size_t)s->B - b0 < (size_t)chunk_images ? (size_t)s->B - b0 : (size_t)chunk_images;
    pp_shape sub = *s;
    sub.B = (int32_t)nb;
    const pp_outputs d_sub = slice_outputs(*d_out, strides, b0);
    PP_CUDA_CHECK(cudaStreamWaitEvent(st, q->ev_in[k], 0));
    // kernels of successive chunks are serial on `stream`, so they share one workspace
    pp_params p_sub = *p;
    pp_initial ini_sub;
    if (p->initial) {          // the chunk's slice of the initial annotations
      ini_sub = *p->initial;
      ini_sub.data += b0 * (size_t)ini_sub.cap * s->F * 3;
      ini_sub.scales += b0 * (size_t)ini_sub.cap * s->F;
      ini_sub.n += b0;
      p_sub.initial = &ini_sub;
    }
    pp_inverse inv_sub;
    if (p->inverse) {          // and of the per-image preprocessing records
      inv_sub = *p->inverse;
      inv_sub.meta += b0;
      p_sub.inverse = &inv_sub;
    }
    int rc = mapped
        ? decode_mapped(d_cif + b0 * cif_img, d_caf + b0 * caf_img, m_cif + b0 * cif_img, m_caf + b0 * caf_img,
                        sk, &d_sub, d_ws, ws_bytes, &sub, c, &p_sub, stream)
        : pp_decode_batch(d_cif + b0 * cif_img, d_caf + b0 * caf_img, sk, &d_sub, d_ws, ws_bytes, &sub, c,
                          &p_sub, stream);
    if (rc) return rc;
    PP_CUDA_CHECK(cudaEventRecord(q->ev_done[k], st));
    PP_CUDA_CHECK(cudaStreamWaitEvent(q->s_out, q->ev_done[k], 0));
    if ((rc = copy_outputs_to_host(*d_out, *h_out, strides, b0, nb, q->s_out))) return rc;
    if (packed)
      PP_CUDA_CHECK(cudaMemcpyAsync(q->h_packed + b0 * q->packed_stride, q->d_packed + b0 * q->packed_stride,
                                    packed_prefix_bytes(nb, s->F, nb * (size_t)q->rec_guess),
                                    cudaMemcpyDeviceToHost, q->s_out));
    PP_CUDA_CHECK(cudaEventRecord(q->ev_out[k], q->s_out));
  }
  // later work on `stream` is ordered after the last copy-out
  PP_CUDA_CHECK(cudaStreamWaitEvent(st, q->ev_out[n_chunks - 1], 0));
  q->n_chunks = n_chunks;
  return PP_OK;
}

int pp_pipeline_wait(pp_pipeline *q, int chunk) {
  if (!q || chunk >= q->n_chunks) return PP_ERR_ARG;
  const int k0 = chunk < 0 ? 0 : chunk, k1 = chunk < 0 ? q->n_chunks : chunk + 1;
  for (int k = k0; k < k1; k++) PP_CUDA_CHECK(cudaEventSynchronize(q->ev_out[k]));
  bool overflow = false;
  for (int k = k0; k < k1 && q->h_packed; k++) {
    const size_t b0 = (size_t)k * q->chunk_images;
    const size_t nb = (size_t)q->B - b0 < (size_t)q->chunk_images ? (size_t)q->B - b0 : (size_t)q->chunk_images;
    char *hb = q->h_packed + b0 * q->packed_stride;
    const size_t total = (size_t)((const int32_t *)hb)[2], have = nb * (size_t)q->rec_guess;
    if (total > have) {       // more live rows than guessed: fetch the rest now, guess higher from here on
      const size_t from = packed_prefix_bytes(nb, q->F, have), to = packed_prefix_bytes(nb, q->F, total);
      PP_CUDA_CHECK(cudaMemcpyAsync(hb + from, q->d_packed + b0 * q->packed_stride + from, to - from,
                                    cudaMemcpyDeviceToHost, q->s_out));
      PP_CUDA_CHECK(cudaStreamSynchronize(q->s_out));
      q->rec_guess = (int)((total + nb - 1) / nb) + 2;
    }
    overflow = overflow || packed_block_overflowed(hb, nb);
  }
  if (overflow) return PP_ERR_CAPACITY;
  if (q->h_status) {
    const int b0 = k0 * q->chunk_images;
    const int b1 = k1 * q->chunk_images < q->B ? k1 * q->chunk_images : q->B;
    for (int b = b0; b < b1; b++)
      if ((q->h_status[b] & PP_STATUS_OVERFLOW_MASK) != 0) return PP_ERR_CAPACITY;
  }
  return PP_OK;
}

}  // extern "C"
