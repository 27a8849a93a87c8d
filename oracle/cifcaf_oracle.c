/*
 * cifcaf_oracle.c -- CPU restatement of the reference CifCaf decoder.
 *
 * TEST INFRASTRUCTURE.  This file is the checker the CUDA path is compared
 * against; it is never the thing shipped or measured.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product (openpifpaf_b200/) must not.
 *
 * It restates, sequentially and in plain C, what
 *   /root/reference/openpifpaf/functional.pyx
 *   /root/reference/openpifpaf/decoder/{cif_hr,cif_seeds,caf_scored,occupancy,nms,utils}.py
 *   /root/reference/openpifpaf/decoder/generator/cifcaf.py
 *   /root/reference/openpifpaf/annotation.py
 * compute "as executed under numpy 2.3 (NEP 50)" -- every function cites the
 * lines it follows.  Arithmetic notes:
 *   * all array maths is binary32 with one rounding per operation
 *     (compile with -ffp-contract=off; x86-64 baseline has no FMA);
 *   * the two places where the reference leaves IEEE basic operations are
 *     numpy's SIMD float32 `exp` and libm `powf(sigma, 2)`
 *     (cifcaf.py:138-139).  exp_mode 0 uses libm expf/powf (closest to the
 *     reference), exp_mode 1 uses include/pifpaf_detmath.h and sigma*sigma,
 *     which the CUDA kernels reproduce bit for bit.
 *
 * Parity pin: the reference ships no golden vectors for this path
 * (SURVEY.md 8c).  This restatement is pinned against the reference itself
 * (oracle/_ref, staged by oracle/build_ref.py) by
 * tests/test_oracle_vs_reference.py and by the fixtures in tests/golden/
 * written by oracle/gen_golden.py.
 */
#define _POSIX_C_SOURCE 200809L
#include <math.h>
#include <pthread.h>
#include <unistd.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/pifpaf_detmath.h"

#define ORC_OK 0
#define ORC_ERR_CAPACITY (-1)
#define ORC_ERR_TIE (-2)   /* the reference would raise TypeError (cifcaf.py:261,285) */
#define ORC_ERR_ARG (-3)

typedef struct {
  double nms_instance_th;  /* nms.Keypoints.instance_threshold  nms.py:13 (compared in float64) */
  float cif_th;            /* CifHr.v_threshold            cif_hr.py:16  */
  float seed_th;           /* CifSeeds.threshold           cif_seeds.py:14 */
  float seed_score_scale;  /* CifSeeds.score_scale         cif_seeds.py:15 */
  float caf_th;            /* CafScored.default_score_th   caf_scored.py:14 */
  float cif_floor;         /* CafScored cif_floor          caf_scored.py:16 */
  float keypoint_th;       /* CifCaf.keypoint_threshold    cifcaf.py:33  */
  float nms_keypoint_th;   /* nms.Keypoints.keypoint_threshold  nms.py:14 */
  float nms_suppression;   /* nms.Keypoints.suppression    nms.py:12     */
  float cif_min_scale;     /* FieldConfig.cif_min_scales[0] */
  float caf_min_distance;  /* FieldConfig.caf_min_distances[0] */
  float caf_max_distance;  /* FieldConfig.caf_max_distances[0]; <= 0 means None */
  int connection_method;   /* 0 blend, 1 max               cifcaf.py:29  */
  int greedy;              /* cifcaf.py:32 */
  int force_complete;      /* cifcaf.py:31 */
  int do_nms;              /* nms is not None              cifcaf.py:117 */
  int exp_mode;            /* 0 libm, 1 deterministic      (see header)  */
  int stride;
  float cif_neighbors;     /* CifHr.neighbors             cif_hr.py:15 (0 = 16) */
} orc_config;

/* ------------------------------------------------------------------------- */
/* functional.pyx:57-68                                                       */

static float approx_exp(float x) {
  if (x > 2.0 || x < -2.0) return 0.0f;
  x = (float)(1.0 + (double)x / 8.0);
  x *= x;
  x *= x;
  x *= x;
  return x;
}

static float clipf(float v, float lo, float hi) {
  return (float)fmax((double)lo, fmin((double)hi, (double)v));
}

/* functional.pyx:105-141  scalar_square_add_gauss_with_max */
static void splat_gauss_with_max(float *field, long fh, long fw, const float *x, const float *y,
                                 const float *sigma, const float *v, long n, float truncate,
                                 float max_value) {
  float truncate2 = truncate * truncate;
  for (long i = 0; i < n; i++) {
    float csigma = sigma[i];
    float csigma2 = csigma * csigma;
    float cx = x[i], cy = y[i], cv = v[i];
    long minx = (long)clipf(cx - truncate * csigma, 0.0f, (float)(fw - 1));
    long maxx = (long)clipf(cx + truncate * csigma + 1, (float)(minx + 1), (float)fw);
    long miny = (long)clipf(cy - truncate * csigma, 0.0f, (float)(fh - 1));
    long maxy = (long)clipf(cy + truncate * csigma + 1, (float)(miny + 1), (float)fh);
    for (long xx = minx; xx < maxx; xx++) {
      float dxr = (float)xx - cx;
      float dx2 = dxr * dxr;
      for (long yy = miny; yy < maxy; yy++) {
        float dyr = (float)yy - cy;
        float dy2 = dyr * dyr;
        if (dx2 + dy2 > truncate2 * csigma2) continue;
        float vv;
        if (dx2 < 0.25 && dy2 < 0.25) {
          vv = cv;
        } else {
          float s = dx2 + dy2;
          float arg = (float)((-0.5 * (double)s) / (double)csigma2);
          vv = cv * approx_exp(arg);
        }
        float *cell = &field[yy * fw + xx];
        float t = *cell + vv;
        *cell = (t < max_value) ? t : max_value; /* Cython min(max_value, field) */
      }
    }
  }
}

/* functional.pyx:231-244  scalar_values, one element */
static float scalar_value_at(const float *field, long fh, long fw, float x, float y, float dflt) {
  float maxx = (float)fw - 1, maxy = (float)fh - 1;
  if (x < 0.0 || y < 0.0 || x > maxx || y > maxy) return dflt;
  return field[(long)y * fw + (long)x];
}

/* ------------------------------------------------------------------------- */
/* CifHr.fill_multiple / accumulate for one cif head -- cif_hr.py:26-65       */
/* cif: (F,5,H,W); hr: (F,Hh,Wh) written (zero-initialised here).             */

/* CifHr.accumulate of one head into ta (not zeroed here): cif_hr.py:26-41 */
/* Fields are independent (one map each, cif_hr.py:26-41): worker threads take whole fields from a
 * shared counter, so every map is still filled by ONE thread in the reference's source order and
 * nothing changes but the wall time (a raw random-init 641 px image: 91 s -> ~8 s on 16 cores). */
static int g_threads = 0;     /* 0 = one per online core */
void orc_set_threads(int n) { g_threads = n > 0 ? n : 0; }

typedef struct {
  const float *cif; int F, H, W; float stride, min_scale, cif_th, len_cifs, neighbors; long hh, hw; float *ta;
  int next, failed;
} cifhr_job;

static void *cifhr_worker(void *arg) {
  cifhr_job *j = (cifhr_job *)arg;
  long n = (long)j->H * j->W;
  float *buf = (float *)malloc(sizeof(float) * 4 * (size_t)n);
  if (!buf) { j->failed = 1; return NULL; }
  float *sx = buf, *sy = buf + n, *ss = buf + 2 * n, *sv = buf + 3 * n;
  const float stride = j->stride, min_scale = j->min_scale;
  for (;;) {
    int f = __sync_fetch_and_add(&j->next, 1);
    if (f >= j->F) break;
    const float *p = j->cif + (size_t)f * 5 * n;
    long k = 0;
    for (long c = 0; c < n; c++) {
      if (!(p[c] > j->cif_th)) continue;                         /* cif_hr.py:27 */
      if (min_scale != 0.0f && !(p[4 * n + c] > min_scale / stride)) continue; /* :28-29 */
      sx[k] = p[n + c] * stride;                                 /* :32 */
      sy[k] = p[2 * n + c] * stride;                             /* :33 */
      float sg = 0.5f * p[4 * n + c];
      sg = sg * stride;
      ss[k] = (sg != sg) ? sg : (1.0f > sg ? 1.0f : sg);         /* np.maximum :34 */
      sv[k] = p[c] / j->neighbors / j->len_cifs;                 /* :40 */
      k++;
    }
    splat_gauss_with_max(j->ta + (size_t)f * j->hh * j->hw, j->hh, j->hw, sx, sy, ss, sv, k, 1.0f, 1.0f);
  }
  free(buf);
  return NULL;
}

static int cifhr_accumulate(const float *cif, int F, int H, int W, int istride, float min_scale,
                            float cif_th, float len_cifs, float neighbors, long hh, long hw, float *ta) {
  cifhr_job job = {cif, F, H, W, (float)istride, min_scale, cif_th, len_cifs,
                   neighbors > 0.0f ? neighbors : 16.0f, hh, hw, ta, 0, 0};
  long nt = g_threads > 0 ? g_threads : sysconf(_SC_NPROCESSORS_ONLN);
  if (nt > F) nt = F;
  if (nt > 64) nt = 64;
  /* small maps: not worth a thread */
  if (nt <= 1 || (size_t)F * hh * hw < (size_t)1 << 21) {
    cifhr_worker(&job);
    return job.failed ? ORC_ERR_ARG : ORC_OK;
  }
  pthread_t th[64];
  int started = 0;
  for (long t = 0; t < nt - 1; t++)
    if (pthread_create(&th[started], NULL, cifhr_worker, &job) == 0) started++;
  cifhr_worker(&job);
  for (int t = 0; t < started; t++) pthread_join(th[t], NULL);
  return job.failed ? ORC_ERR_ARG : ORC_OK;
}

int orc_cifhr(const float *cif, int F, int H, int W, const orc_config *cfg, float *hr) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  memset(hr, 0, sizeof(float) * (size_t)F * hh * hw);
  return cifhr_accumulate(cif, F, H, W, cfg->stride, cfg->cif_min_scale, cfg->cif_th, 1.0f, cfg->cif_neighbors,
                          hh, hw, hr);
}

/* ------------------------------------------------------------------------- */
/* CifSeeds.fill_cif + get -- cif_seeds.py:23-54                              */
/* out: cap x 5 floats (v, f, x, y, s) sorted like sorted(seeds, reverse=True) */

static int seed_cmp_desc(const void *a, const void *b) {
  const float *p = (const float *)a, *q = (const float *)b;
  for (int i = 0; i < 5; i++) {
    if (p[i] > q[i]) return -1;
    if (p[i] < q[i]) return 1;
  }
  return 0;
}

/* CifSeeds.fill_cif of one head, appended at out[5 * k ...]; returns the new count (cif_seeds.py:23-50) */
static long seeds_append(const float *cif, int F, int H, int W, int istride, float min_scale,
                         const float *hr, long hh, long hw, const orc_config *cfg, float *out,
                         long cap, long k) {
  long n = (long)H * W;
  float stride = (float)istride;
  for (int f = 0; f < F; f++) {
    const float *p = cif + (size_t)f * 5 * n;
    for (long c = 0; c < n; c++) {
      float conf = p[c];
      if (!(conf > cfg->seed_th)) continue;                            /* :31 */
      if (min_scale != 0.0f && !(p[4 * n + c] > min_scale / stride)) continue;
      float x = p[n + c] * stride, y = p[2 * n + c] * stride;
      float hv = scalar_value_at(hr + (size_t)f * hh * hw, hh, hw, x, y, 0.0f); /* :37 */
      float a = 0.9f * hv, b = 0.1f * conf;
      float v = a + b;                                                 /* :38 */
      if (cfg->seed_score_scale != 1.0f) v = v * cfg->seed_score_scale; /* :41-42 */
      if (!(v > cfg->seed_th)) continue;                               /* :43 */
      if (k >= cap) return ORC_ERR_CAPACITY;
      float *o = out + 5 * k++;
      o[0] = v; o[1] = (float)f; o[2] = x; o[3] = y; o[4] = p[4 * n + c] * stride;
    }
  }
  return k;
}

long orc_seeds(const float *cif, int F, int H, int W, const float *hr, const orc_config *cfg,
               float *out, long cap) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  long k = seeds_append(cif, F, H, W, cfg->stride, cfg->cif_min_scale, hr, hh, hw, cfg, out, cap, 0);
  if (k < 0) return k;
  qsort(out, (size_t)k, 5 * sizeof(float), seed_cmp_desc);             /* :54 */
  return k;
}

/* ------------------------------------------------------------------------- */
/* CafScored.fill_caf -- caf_scored.py:32-87                                  */
/* caf: (L,9,H,W).  fwd/bwd: (L,9,cap) row-major per link, counts per link.   */

static int caf_scored_append(const float *caf, int L, int H, int W, int istride, float min_distance,
                             float max_distance, const float *hr, long hh, long hw, int F,
                             const int *skeleton /* 2L, 1-based */, const orc_config *cfg,
                             float score_th, float *fwd, int *nfwd, float *bwd, int *nbwd, long cap,
                             int append) {
  long n = (long)H * W;
  float stride = (float)istride;
  static const int brow[9] = {0, 5, 6, 7, 8, 1, 2, 3, 4};               /* :68 */
  for (int l = 0; l < L; l++) {
    const float *p = caf + (size_t)l * 9 * n;
    float *fo = fwd + (size_t)l * 9 * cap, *bo = bwd + (size_t)l * 9 * cap;
    int kf = append ? nfwd[l] : 0, kb = append ? nbwd[l] : 0;     /* np.concatenate :70,81 */
    int j1 = skeleton[2 * l] - 1, j2 = skeleton[2 * l + 1] - 1;
    for (long c = 0; c < n; c++) {
      float nine[9];
      if (!(p[c] > score_th)) continue;                                 /* :42 */
      for (int r = 0; r < 9; r++) nine[r] = p[(size_t)r * n + c];
      if (min_distance != 0.0f || max_distance > 0.0f) {
        float dx = nine[1] - nine[5], dy = nine[2] - nine[6];
        float d = sqrtf(dx * dx + dy * dy);                             /* :48,53 */
        if (min_distance != 0.0f && !(d > min_distance / stride)) continue;
        if (max_distance > 0.0f && !(d < max_distance / stride)) continue;
      }
      for (int r = 1; r < 9; r++) nine[r] = nine[r] * stride;           /* :58 */
      float score = nine[0];
      float sb = score, sf = score;
      if (cfg->cif_floor < 1.0f && j1 < F) {                            /* :62-64 */
        float h = scalar_value_at(hr + (size_t)j1 * hh * hw, hh, hw, nine[1], nine[2], 0.0f);
        float w = (1.0f - cfg->cif_floor) * h;
        w = cfg->cif_floor + w;
        sb = score * w;
      }
      if (cfg->cif_floor < 1.0f && j2 < F) {                            /* :73-75 */
        float h = scalar_value_at(hr + (size_t)j2 * hh * hw, hh, hw, nine[5], nine[6], 0.0f);
        float w = (1.0f - cfg->cif_floor) * h;
        w = cfg->cif_floor + w;
        sf = score * w;
      }
      if (sb > score_th) {                                              /* :67-70 */
        if (kb >= cap) return ORC_ERR_CAPACITY;
        for (int r = 0; r < 9; r++) bo[(size_t)r * cap + kb] = nine[brow[r]];
        bo[kb] = sb;
        kb++;
      }
      if (sf > score_th) {                                              /* :78-81 */
        if (kf >= cap) return ORC_ERR_CAPACITY;
        for (int r = 0; r < 9; r++) fo[(size_t)r * cap + kf] = nine[r];
        fo[kf] = sf;
        kf++;
      }
    }
    nfwd[l] = kf;
    nbwd[l] = kb;
  }
  return ORC_OK;
}

int orc_caf_scored(const float *caf, int L, int H, int W, const float *hr, int F,
                   const int *skeleton /* 2L, 1-based */, const orc_config *cfg, float score_th,
                   float *fwd, int *nfwd, float *bwd, int *nbwd, long cap) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  return caf_scored_append(caf, L, H, W, cfg->stride, cfg->caf_min_distance, cfg->caf_max_distance, hr,
                           hh, hw, F, skeleton, cfg, score_th, fwd, nfwd, bwd, nbwd, cap, 0);
}

/* ========================================================================= */
/* Assembly: cifcaf.py:57-65 (link tables), :124-217 (connection scoring),    */
/* :247-307 (_grow), :84-108 (seed loop), occupancy.py, nms.py                */

#define ORC_MAXF 64
#define ORC_MAXDEG 32
#define ORC_MAXEDGE 512   /* >= 4L: directed edges, queued once per _grow pass */

typedef struct {
  int F, L;
  int deg[ORC_MAXF];
  int nb_end[ORC_MAXF][ORC_MAXDEG];   /* by_source[start] keys in insertion order */
  int nb_link[ORC_MAXF][ORC_MAXDEG];
  int nb_fwd[ORC_MAXF][ORC_MAXDEG];
} orc_graph;

static int graph_init(orc_graph *g, int F, int L, const int *skeleton) {
  if (F > ORC_MAXF) return ORC_ERR_ARG;
  memset(g, 0, sizeof *g);
  g->F = F; g->L = L;
  if (4 * L > ORC_MAXEDGE) return ORC_ERR_ARG;
  for (int l = 0; l < L; l++) {                              /* cifcaf.py:62-65 */
    int a = skeleton[2 * l] - 1, b = skeleton[2 * l + 1] - 1;
    if (a < 0 || b < 0 || a >= F || b >= F) return ORC_ERR_ARG;
    for (int dir = 0; dir < 2; dir++) {
      int s = dir ? b : a, e = dir ? a : b, k;
      for (k = 0; k < g->deg[s]; k++) if (g->nb_end[s][k] == e) break;
      if (k == g->deg[s]) {
        if (k >= ORC_MAXDEG) return ORC_ERR_ARG;
        g->deg[s]++;
      }
      g->nb_end[s][k] = e;   /* dict assignment: value replaced, position kept */
      g->nb_link[s][k] = l;
      g->nb_fwd[s][k] = !dir;
    }
  }
  return ORC_OK;
}

typedef struct {
  const float *fwd, *bwd;  /* (L,9,cap) */
  const int *nfwd, *nbwd;
  long cap;
} orc_cafs;

typedef struct { float x, y, s, v; } xysv_t;

/* cifcaf.py:147-192 */
static xysv_t pick_target(const float *rows, long cap, const int *idx, const float *scores, int n,
                          int method) {
  xysv_t out = {0, 0, 0, 0};
  const float *xt = rows + 5 * cap, *yt = rows + 6 * cap, *st = rows + 8 * cap;
  if (method == 1) {                                          /* _target_with_maxscore */
    int m = 0;
    for (int i = 1; i < n; i++) if (scores[i] > scores[m]) m = i;
    out.x = xt[idx[m]]; out.y = yt[idx[m]]; out.s = st[idx[m]]; out.v = scores[m];
    return out;
  }
  if (n == 1) {                                               /* :164-170 */
    out.x = xt[idx[0]]; out.y = yt[idx[0]]; out.s = st[idx[0]]; out.v = scores[0] * 0.5f;
    return out;
  }
  /* np.argsort ascending, last two entries; ties resolved as a stable sort
   * would (larger index last) -- tie class T2 */
  int i1 = 0, i2 = -1;
  for (int i = 1; i < n; i++) {
    if (scores[i] >= scores[i1]) { i2 = i1; i1 = i; }
    else if (i2 < 0 || scores[i] >= scores[i2]) { i2 = i; }
  }
  float s1 = scores[i1], s2 = scores[i2];
  float x1 = xt[idx[i1]], y1 = yt[idx[i1]], t1 = st[idx[i1]];
  float x2 = xt[idx[i2]], y2 = yt[idx[i2]], t2 = st[idx[i2]];
  if (s2 < 0.01f || s2 < 0.5f * s1) {                         /* :178-179 */
    out.x = x1; out.y = y1; out.s = t1; out.v = s1 * 0.5f; return out;
  }
  float dx = x1 - x2, dy = y1 - y2;
  float dxx = dx * dx, dyy = dy * dy;
  float d = sqrtf(dxx + dyy);                                 /* :182 */
  if (d > t1 / 2.0f) {                                        /* :183-184 */
    out.x = x1; out.y = y1; out.s = t1; out.v = s1 * 0.5f; return out;
  }
  float den = s1 + s2;
  float a, b;
  a = s1 * x1; b = s2 * x2; out.x = (a + b) / den;            /* :186-191 */
  a = s1 * y1; b = s2 * y2; out.y = (a + b) / den;
  a = s1 * t1; b = s2 * t2; out.s = (a + b) / den;
  out.v = 0.5f * den;
  return out;
}

/* cifcaf.py:124-145 with functional.pyx:338-359 (caf_center_s) inlined */
static xysv_t grow_connection(float x, float y, float s, const float *rows, long cap, int K,
                              const orc_config *cfg, int *idx, float *scores) {
  xysv_t zero = {0, 0, 0, 0};
  float sigma_box = 2.0f * s;
  float lox = x - sigma_box, hix = x + sigma_box, loy = y - sigma_box, hiy = y + sigma_box;
  const float *cs = rows, *xs = rows + cap, *ys = rows + 2 * cap;
  int n = 0;
  for (int i = 0; i < K; i++) {
    if (xs[i] < lox) continue;
    if (xs[i] > hix) continue;
    if (ys[i] < loy) continue;
    if (ys[i] > hiy) continue;
    idx[n++] = i;
  }
  if (n == 0) return zero;
  float sigma = 0.5f * s;
  float sig2 = cfg->exp_mode == 0 ? powf(sigma, 2.0f) : sigma * sigma;
  for (int k = 0; k < n; k++) {
    int i = idx[k];
    float dx = x - xs[i], dy = y - ys[i];
    float dxx = dx * dx, dyy = dy * dy;
    float d = sqrtf(dxx + dyy);                               /* np.linalg.norm :134 */
    float d2 = d * d;
    float arg = (-0.5f * d2) / sig2;
    float e = cfg->exp_mode == 0 ? expf(arg) : pp_expf(arg);
    scores[k] = e * cs[i];                                    /* :139 */
  }
  return pick_target(rows, cap, idx, scores, n, cfg->connection_method);
}

typedef struct {
  int F;
  float data[ORC_MAXF][3];
  float js[ORC_MAXF];
  int n_dec; int dec[ORC_MAXF][2];
  int n_front; int front[ORC_MAXEDGE][2];
} orc_ann;

/* cifcaf.py:194-217 */
static xysv_t connection_value(const orc_ann *ann, const orc_graph *g, const orc_cafs *cf,
                               const orc_config *cfg, int start, int k, int reverse_match,
                               int *idx, float *scores) {
  xysv_t zero = {0, 0, 0, 0};
  int l = g->nb_link[start][k], fw = g->nb_fwd[start][k];
  const float *rf = (fw ? cf->fwd : cf->bwd) + (size_t)l * 9 * cf->cap;
  const float *rb = (fw ? cf->bwd : cf->fwd) + (size_t)l * 9 * cf->cap;
  int nf = fw ? cf->nfwd[l] : cf->nbwd[l], nb = fw ? cf->nbwd[l] : cf->nfwd[l];
  const float *xyv = ann->data[start];
  float ss = ann->js[start] > 0.0f ? ann->js[start] : 0.0f;          /* max(0.0, .) */
  xysv_t nw = grow_connection(xyv[0], xyv[1], ss, rf, cf->cap, nf, cfg, idx, scores);
  float prod = nw.v * xyv[2];
  float kp = sqrtf(prod);
  if (kp < cfg->keypoint_th) return zero;
  if (nw.v == 0.0f) return zero;
  float st = nw.s > 0.0f ? nw.s : 0.0f;
  if (reverse_match) {
    xysv_t rv = grow_connection(nw.x, nw.y, st, rb, cf->cap, nb, cfg, idx, scores);
    if (rv.s == 0.0f) return zero;
    float ax = fabsf(xyv[0] - rv.x), ay = fabsf(xyv[1] - rv.y);
    if (ax + ay > ss) return zero;
  }
  nw.v = kp;
  return nw;
}

typedef struct { float key; int evaluated; xysv_t t; int start, end, k; } fentry;

/* Python tuple order of (-score, xysv-or-None, start, end); *tie set when the
 * reference would compare a tuple with None */
static int fentry_less(const fentry *a, const fentry *b, int *tie) {
  if (a->key != b->key) return a->key < b->key;
  if (a->evaluated != b->evaluated) { *tie = 1; return a->evaluated < b->evaluated; }
  if (a->evaluated) {
    const float pa[4] = {a->t.x, a->t.y, a->t.s, a->t.v}, pb[4] = {b->t.x, b->t.y, b->t.s, b->t.v};
    for (int i = 0; i < 4; i++) if (pa[i] != pb[i]) return pa[i] < pb[i];
  }
  if (a->start != b->start) return a->start < b->start;
  return a->end < b->end;
}

typedef struct {
  fentry e[2 * ORC_MAXEDGE];
  int n;
  unsigned char queued[ORC_MAXF][ORC_MAXF];
} frontier_t;

static void add_to_frontier(frontier_t *fr, orc_ann *ann, const orc_graph *g, const float *cscale,
                            int start) {                                  /* cifcaf.py:251-263 */
  for (int k = 0; k < g->deg[start]; k++) {
    int end = g->nb_end[start][k];
    if (ann->data[end][2] > 0.0f) continue;
    if (fr->queued[start][end]) continue;
    float bound = sqrtf(ann->data[start][2]);
    if (cscale) bound = bound * cscale[g->nb_link[start][k]];
    fentry *e = &fr->e[fr->n++];
    e->key = -bound; e->evaluated = 0; e->start = start; e->end = end; e->k = k;
    fr->queued[start][end] = 1;
    ann->front[ann->n_front][0] = start; ann->front[ann->n_front][1] = end; ann->n_front++;
  }
}

static int frontier_pop(frontier_t *fr, fentry *out, int *tie) {
  if (fr->n == 0) return 0;
  int m = 0;
  for (int i = 1; i < fr->n; i++) if (fentry_less(&fr->e[i], &fr->e[m], tie)) m = i;
  *out = fr->e[m];
  fr->e[m] = fr->e[--fr->n];
  return 1;
}

/* cifcaf.py:247-307 */
static int grow(orc_ann *ann, const orc_graph *g, const orc_cafs *cf, const orc_config *cfg,
                const float *cscale, int reverse_match, frontier_t *fr, int *idx, float *scores) {
  int tie = 0;
  fr->n = 0;
  memset(fr->queued, 0, sizeof fr->queued);
  for (int j = 0; j < g->F; j++) {                                         /* :288-291 */
    if (ann->data[j][2] == 0.0f) continue;
    add_to_frontier(fr, ann, g, cscale, j);
  }
  for (;;) {
    fentry e;
    int have = 0;
    while (frontier_pop(fr, &e, &tie)) {                                   /* frontier_get :265-285 */
      if (e.evaluated) { have = 1; break; }
      if (ann->data[e.end][2] > 0.0f) continue;
      xysv_t nw = connection_value(ann, g, cf, cfg, e.start, e.k, reverse_match, idx, scores);
      if (nw.v == 0.0f) continue;
      float score = nw.v;
      e.t = nw; e.evaluated = 1;
      if (cfg->greedy) { e.key = -score; have = 1; break; }
      if (cscale) score = score * cscale[g->nb_link[e.start][e.k]];
      e.key = -score;
      fr->e[fr->n++] = e;
    }
    if (!have) break;
    if (ann->data[e.end][2] > 0.0f) continue;                              /* :297-298 */
    ann->data[e.end][0] = e.t.x; ann->data[e.end][1] = e.t.y; ann->data[e.end][2] = e.t.v;
    ann->js[e.end] = e.t.s;
    ann->dec[ann->n_dec][0] = e.start; ann->dec[ann->n_dec][1] = e.end; ann->n_dec++;
    add_to_frontier(fr, ann, g, cscale, e.end);
  }
  return tie ? ORC_ERR_TIE : ORC_OK;
}

/* ------------------------------------------------------------------------- */
/* Occupancy -- occupancy.py:10-47, utils.py:61-66, functional.pyx:281-286    */

typedef struct { unsigned char *m; long F, H, W; float min_scale_reduced; } orc_occ;

static int occ_init(orc_occ *o, long F, long shape_h, long shape_w) {
  o->F = F;
  o->min_scale_reduced = 2.0f;                          /* CifCaf: Occupancy(shape, 2, min_scale=4) cifcaf.py:84 */
  o->H = (long)((double)shape_h / 2.0);                 /* int(shape[1] / reduction) */
  o->W = (long)((double)shape_w / 2.0);
  if (o->H < 0 || o->W < 0) return ORC_ERR_ARG;
  o->m = (unsigned char *)calloc((size_t)(F * o->H * o->W) + 1, 1);
  return o->m ? ORC_OK : ORC_ERR_ARG;
}

static long long py_round_f(float v) {                  /* round(np.float32) -> int, half to even */
  double r = rint((double)v);
  if (r > 4e18) r = 4e18;
  if (r < -4e18) r = -4e18;
  return (long long)r;
}

static void occ_set(orc_occ *o, int f, float x, float y, float sigma) {   /* occupancy.py:31-39 */
  if (f >= o->F) return;
  long long xi = py_round_f(x / 2.0f), yi = py_round_f(y / 2.0f);
  float sh = sigma / 2.0f;
  long long si = py_round_f(sh > o->min_scale_reduced ? sh : o->min_scale_reduced); /* max(min_scale_reduced, sigma / r) */
  long long minx = xi - si > 0 ? xi - si : 0;
  long long miny = yi - si > 0 ? yi - si : 0;
  long long t = xi + si + 1 < o->W ? xi + si + 1 : o->W;
  long long maxx = minx + 1 > t ? minx + 1 : t;
  t = yi + si + 1 < o->H ? yi + si + 1 : o->H;
  long long maxy = miny + 1 > t ? miny + 1 : t;
  if (maxx > o->W) maxx = o->W;                          /* numpy slice clamps */
  if (maxy > o->H) maxy = o->H;
  unsigned char *p = o->m + (size_t)f * o->H * o->W;
  for (long long yy = miny; yy < maxy; yy++)
    for (long long xx = minx; xx < maxx; xx++) p[yy * o->W + xx] += 1;   /* uint8 wraps */
}

static int occ_get(const orc_occ *o, int f, float x, float y) {          /* occupancy.py:41-47 */
  if (f >= o->F) return 1;
  float cx = clipf(x / 2.0f, 0.0f, (float)(o->W - 1));
  float cy = clipf(y / 2.0f, 0.0f, (float)(o->H - 1));
  return o->m[(size_t)f * o->H * o->W + (long)cy * o->W + (long)cx] != 0;
}

/* ------------------------------------------------------------------------- */
/* Annotation.score -- annotation.py:24-28,60-71 (float64, numpy pairwise sum) */

static int cmp_float_asc(const void *a, const void *b) {
  float p = *(const float *)a, q = *(const float *)b;
  return (p > q) - (p < q);
}

static double ann_score(const orc_ann *a) {
  int F = a->F;
  float v[ORC_MAXF];
  double t[ORC_MAXF];
  for (int j = 0; j < F; j++) v[j] = a->data[j][2];
  qsort(v, (size_t)F, sizeof(float), cmp_float_asc);
  double wsum = (double)F + 2.0 * (F < 3 ? F : 3);       /* ones, first three = 3 */
  for (int j = 0; j < F; j++) {
    double w = (j < 3 ? 3.0 : 1.0) / wsum;
    t[j] = w * (double)v[F - 1 - j];
  }
  /* numpy pairwise_sum for n < 128 */
  if (F < 8) {
    double r = 0.0;
    for (int j = 0; j < F; j++) r += t[j];
    return r;
  }
  double r[8];
  int i;
  for (i = 0; i < 8; i++) r[i] = t[i];
  for (i = 8; i < F - (F % 8); i += 8)
    for (int j = 0; j < 8; j++) r[j] += t[i + j];
  double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  for (; i < F; i++) res += t[i];
  return res;
}

/* stable sort of annotation indices by descending score -- sorted(key=-score) */
static void sort_by_score(int *order, const double *score, int n) {
  for (int i = 1; i < n; i++) {
    int o = order[i], j = i - 1;
    while (j >= 0 && -score[order[j]] > -score[o]) { order[j + 1] = order[j]; j--; }
    order[j + 1] = o;
  }
}

/* nms.py:17-57.  anns/keep are rewritten in place; returns new count. */
static int nms_keypoints(orc_ann *anns, int n, const orc_config *cfg) {
  int F = n ? anns[0].F : 0;
  double *score = (double *)malloc(sizeof(double) * (size_t)(n + 1));
  int *order = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  orc_ann *tmp = (orc_ann *)malloc(sizeof(orc_ann) * (size_t)(n + 1));
  int m = 0;
  for (int a = 0; a < n; a++) {                                          /* :20-22 */
    for (int j = 0; j < F; j++)
      if (anns[a].data[j][2] < cfg->nms_keypoint_th)
        anns[a].data[j][0] = anns[a].data[j][1] = anns[a].data[j][2] = 0.0f;
    double s = ann_score(&anns[a]);
    if (s >= cfg->nms_instance_th) { tmp[m] = anns[a]; score[m] = s; order[m] = m; m++; }
  }
  if (m > 0) {
    float my = -INFINITY, mx = -INFINITY;                                 /* :27-31 */
    for (int a = 0; a < m; a++)
      for (int j = 0; j < F; j++) {
        if (tmp[a].data[j][1] > my) my = tmp[a].data[j][1];
        if (tmp[a].data[j][0] > mx) mx = tmp[a].data[j][0];
      }
    orc_occ occ;
    occ_init(&occ, F, (long)(my + 1.0f), (long)(mx + 1.0f));
    sort_by_score(order, score, m);                                       /* :33 */
    for (int oi = 0; oi < m; oi++) {
      orc_ann *an = &tmp[order[oi]];
      for (int f = 0; f < F; f++) {
        if (an->data[f][2] == 0.0f) continue;
        if (occ_get(&occ, f, an->data[f][0], an->data[f][1]))
          an->data[f][2] = an->data[f][2] * cfg->nms_suppression;
        else
          occ_set(&occ, f, an->data[f][0], an->data[f][1], an->js[f]);
      }
    }
    free(occ.m);
    int k = 0;                                                            /* :50-53 */
    for (int oi = 0; oi < m; oi++) {
      orc_ann *an = &tmp[order[oi]];
      for (int j = 0; j < F; j++)
        if (an->data[j][2] < cfg->nms_keypoint_th)
          an->data[j][0] = an->data[j][1] = an->data[j][2] = 0.0f;
      double s = ann_score(an);
      if (s >= cfg->nms_instance_th) { anns[k] = *an; score[k] = s; order[k] = k; k++; }
    }
    m = k;
    sort_by_score(order, score, m);
    for (int i = 0; i < m; i++) tmp[i] = anns[order[i]];
    for (int i = 0; i < m; i++) anns[i] = tmp[i];
  }
  free(score); free(order); free(tmp);
  return m;
}

/* ------------------------------------------------------------------------- */
/* _flood_fill -- cifcaf.py:309-331 (including the closure quirk: inside the   */
/* while loop the priority of newly queued entries is taken from the popped   */
/* entry's source xyv, not from the joint being expanded)                     */

typedef struct { double key; int end; float sx, sy, sv; float s; } ffentry;

static int ff_less(const ffentry *a, const ffentry *b) {
  if (a->key != b->key) return a->key < b->key;
  if (a->end != b->end) return a->end < b->end;
  if (a->sx != b->sx) return a->sx < b->sx;
  if (a->sy != b->sy) return a->sy < b->sy;
  if (a->sv != b->sv) return a->sv < b->sv;
  return a->s < b->s;
}

static void flood_fill(orc_ann *ann, const orc_graph *g) {
  int cap = g->F * ORC_MAXDEG * 2 + 8, n = 0;
  ffentry *q = (ffentry *)malloc(sizeof(ffentry) * (size_t)cap);
#define FF_ADD(start_, prio_)                                                   \
  for (int k_ = 0; k_ < g->deg[start_]; k_++) {                                 \
    int end_ = g->nb_end[start_][k_];                                           \
    if (ann->data[end_][2] > 0.0f) continue;                                    \
    if (n >= cap) { cap *= 2; q = (ffentry *)realloc(q, sizeof(ffentry) * (size_t)cap); } \
    ffentry *e_ = &q[n++];                                                      \
    e_->key = -(double)(prio_); e_->end = end_;                                 \
    e_->sx = ann->data[start_][0]; e_->sy = ann->data[start_][1];               \
    e_->sv = ann->data[start_][2]; e_->s = ann->js[start_];                     \
  }
  for (int j = 0; j < g->F; j++) {
    if (ann->data[j][2] == 0.0f) continue;
    FF_ADD(j, ann->data[j][2]);
  }
  while (n > 0) {
    int m = 0;
    for (int i = 1; i < n; i++) if (ff_less(&q[i], &q[m])) m = i;
    ffentry e = q[m];
    q[m] = q[--n];
    if (ann->data[e.end][2] > 0.0f) continue;
    ann->data[e.end][0] = e.sx; ann->data[e.end][1] = e.sy;
    ann->data[e.end][2] = 0.00001f;
    ann->js[e.end] = e.s;
    FF_ADD(e.end, e.sv);
  }
#undef FF_ADD
  free(q);
}

/* ------------------------------------------------------------------------- */
/* CifCaf.__call__ -- cifcaf.py:67-122; complete_annotations :333-351         */
/*                                                                           */
/* Outputs (caller-owned):                                                    */
/*   hr (F,Hh,Wh); seeds (seed_cap,5) + *n_seeds; fwd/bwd (L,9,caf_cap),      */
/*   nfwd/nbwd (L);                                                           */
/*   ann_data (ann_cap,F,3), ann_js (ann_cap,F), ann_ndec (ann_cap),          */
/*   ann_dec (ann_cap,F,2), ann_nfront (ann_cap), ann_front (ann_cap,4L,2)    */
/*   pre_* : the same, captured before force-complete / NMS (may be NULL).    */
/* Returns the number of annotations, or a negative ORC_ERR_*.  *tie_flag is   */
/* set when the reference would have raised TypeError on a heap tie.          */

static void export_anns(const orc_ann *anns, int n, int F, int L, float *data, float *js,
                        int *ndec, int *dec, int *nfront, int *front) {
  for (int a = 0; a < n; a++) {
    for (int j = 0; j < F; j++) {
      for (int c = 0; c < 3; c++) data[((size_t)a * F + j) * 3 + c] = anns[a].data[j][c];
      js[(size_t)a * F + j] = anns[a].js[j];
    }
    if (ndec) {
      ndec[a] = anns[a].n_dec;
      for (int i = 0; i < anns[a].n_dec && i < F; i++) {
        dec[((size_t)a * F + i) * 2] = anns[a].dec[i][0];
        dec[((size_t)a * F + i) * 2 + 1] = anns[a].dec[i][1];
      }
    }
    if (nfront) {
      nfront[a] = anns[a].n_front;
      for (int i = 0; i < anns[a].n_front && i < 4 * L; i++) {
        front[((size_t)a * 4 * L + i) * 2] = anns[a].front[i][0];
        front[((size_t)a * 4 * L + i) * 2 + 1] = anns[a].front[i][1];
      }
    }
  }
}

/* One field set of a multi-scale FieldConfig (field_config.py, decoder/factory.py:153-180). */
typedef struct {
  const float *cif, *caf;
  int H, W, stride;
  float cif_min_scale, caf_min_distance, caf_max_distance;   /* 0 = off / None */
} orc_set;

/* CifCaf.__call__ over n_sets field sets; set 0 defines the hr shape (cif_hr.py:44-51).
 * hflip_pairs: CifHr accumulates sets (k, k + n/2) together with len_cifs = 2 (cif_hr.py:66-72). */
int orc_decode_ms(const orc_set *sets, int n_sets, int hflip_pairs, int F, int L,
               const int *skeleton, const float *conf_scales, const orc_config *cfg,
               float *hr, float *seeds, long seed_cap, long *n_seeds,
               float *fwd, int *nfwd, float *bwd, int *nbwd, long caf_cap,
               int ann_cap, float *ann_data, float *ann_js, int *ann_ndec, int *ann_dec,
               int *ann_nfront, int *ann_front,
               int *pre_n, float *pre_data, float *pre_js, int *tie_flag,
               int init_n, const float *init_data, const float *init_js) {
  orc_graph g;
  int rc = graph_init(&g, F, L, skeleton);
  if (rc) return rc;
  long hh = (long)(sets[0].H - 1) * sets[0].stride + 1, hw = (long)(sets[0].W - 1) * sets[0].stride + 1;
  size_t hr_n = (size_t)F * hh * hw;
  int n_groups = hflip_pairs ? n_sets / 2 : n_sets;                             /* :80, cif_hr.py:66-78 */
  float *ta = n_groups > 1 ? (float *)malloc(sizeof(float) * hr_n) : NULL;
  for (int gi = 0; gi < n_groups; gi++) {
    float *dst = gi == 0 ? hr : ta;
    memset(dst, 0, sizeof(float) * hr_n);
    const orc_set *a = &sets[gi];
    float len_cifs = hflip_pairs ? 2.0f : 1.0f;
    if ((rc = cifhr_accumulate(a->cif, F, a->H, a->W, a->stride, a->cif_min_scale, cfg->cif_th, len_cifs,
                               cfg->cif_neighbors, hh, hw, dst))) return rc;
    if (hflip_pairs) {
      const orc_set *b = &sets[gi + n_groups];
      if ((rc = cifhr_accumulate(b->cif, F, b->H, b->W, a->stride, a->cif_min_scale, cfg->cif_th,
                                 len_cifs, cfg->cif_neighbors, hh, hw, dst))) return rc;
    }
    if (gi > 0)                                                                 /* np.maximum :62 */
      for (size_t i = 0; i < hr_n; i++) {
        float x = ta[i], y = hr[i];
        hr[i] = (x != x) ? x : ((y != y) ? y : (x > y ? x : y));
      }
  }
  free(ta);
  long ns = 0;
  for (int k = 0; k < n_sets; k++) {                                            /* :81, cif_seeds.py:56-65 */
    ns = seeds_append(sets[k].cif, F, sets[k].H, sets[k].W, sets[k].stride, sets[k].cif_min_scale, hr,
                      hh, hw, cfg, seeds, seed_cap, ns);
    if (ns < 0) return (int)ns;
  }
  qsort(seeds, (size_t)ns, 5 * sizeof(float), seed_cmp_desc);
  *n_seeds = ns;
  for (int k = 0; k < n_sets; k++) {                                            /* :82, caf_scored.py:89-98 */
    rc = caf_scored_append(sets[k].caf, L, sets[k].H, sets[k].W, sets[k].stride, sets[k].caf_min_distance,
                           sets[k].caf_max_distance, hr, hh, hw, F, skeleton, cfg, cfg->caf_th, fwd,
                           nfwd, bwd, nbwd, caf_cap, k > 0);
    if (rc) return rc;
  }
  orc_cafs cf = {fwd, bwd, nfwd, nbwd, caf_cap};

  orc_occ occ;
  if ((rc = occ_init(&occ, F, hh, hw))) return rc;                              /* :84 */
  orc_ann *anns = (orc_ann *)calloc((size_t)ann_cap + 1, sizeof(orc_ann));
  frontier_t *fr = (frontier_t *)malloc(sizeof(frontier_t));
  int *idx = (int *)malloc(sizeof(int) * (size_t)(caf_cap + 1));
  float *scores = (float *)malloc(sizeof(float) * (size_t)(caf_cap + 1));
  int n = 0, tie = 0;
  rc = ORC_OK;
  for (int ia = 0; ia < init_n && rc == ORC_OK; ia++) {                         /* :95-98 */
    if (n >= ann_cap) { rc = ORC_ERR_CAPACITY; break; }
    orc_ann *an = &anns[n];
    memset(an, 0, sizeof *an);
    an->F = F;
    for (int j = 0; j < F; j++) {
      for (int c = 0; c < 3; c++) an->data[j][c] = init_data[((size_t)ia * F + j) * 3 + c];
      an->js[j] = init_js[(size_t)ia * F + j];
    }
    if (grow(an, &g, &cf, cfg, conf_scales, 1, fr, idx, scores) == ORC_ERR_TIE) tie = 1;
    n++;
    for (int j = 0; j < F; j++) {
      if (an->data[j][2] == 0.0f) continue;
      occ_set(&occ, j, an->data[j][0], an->data[j][1], an->js[j]);
    }
  }
  for (long si = 0; si < ns && rc == ORC_OK; si++) {                            /* :100-108 */
    const float *sd = seeds + 5 * si;
    int f = (int)sd[1];
    if (occ_get(&occ, f, sd[2], sd[3])) continue;
    if (n >= ann_cap) { rc = ORC_ERR_CAPACITY; break; }
    orc_ann *an = &anns[n];
    memset(an, 0, sizeof *an);
    an->F = F;
    an->data[f][0] = sd[2]; an->data[f][1] = sd[3]; an->data[f][2] = sd[0];
    an->js[f] = sd[4];
    if (grow(an, &g, &cf, cfg, conf_scales, 1, fr, idx, scores) == ORC_ERR_TIE) tie = 1;
    n++;
    for (int j = 0; j < F; j++) {                                               /* :87-93 */
      if (an->data[j][2] == 0.0f) continue;
      occ_set(&occ, j, an->data[j][0], an->data[j][1], an->js[j]);
    }
  }
  free(occ.m);
  if (rc == ORC_OK && pre_n) {
    *pre_n = n;
    export_anns(anns, n, F, L, pre_data, pre_js, NULL, NULL, NULL, NULL);
  }
  if (rc == ORC_OK && cfg->force_complete) {                                    /* :333-351 */
    float *f2 = (float *)malloc(sizeof(float) * (size_t)L * 9 * caf_cap);
    float *b2 = (float *)malloc(sizeof(float) * (size_t)L * 9 * caf_cap);
    int *nf2 = (int *)malloc(sizeof(int) * (size_t)L), *nb2 = (int *)malloc(sizeof(int) * (size_t)L);
    for (int k = 0; k < n_sets && rc == ORC_OK; k++)
      rc = caf_scored_append(sets[k].caf, L, sets[k].H, sets[k].W, sets[k].stride,
                             sets[k].caf_min_distance, sets[k].caf_max_distance, hr, hh, hw, F, skeleton,
                             cfg, 0.0001f, f2, nf2, b2, nb2, caf_cap, k > 0);
    if (rc == ORC_OK) {
      orc_cafs c2 = {f2, b2, nf2, nb2, caf_cap};
      for (int a = 0; a < n; a++) {
        orc_ann *an = &anns[a];
        unsigned char unfilled[ORC_MAXF];
        int any_empty = 0;
        for (int j = 0; j < F; j++) unfilled[j] = an->data[j][2] == 0.0f;
        if (grow(an, &g, &c2, cfg, conf_scales, 0, fr, idx, scores) == ORC_ERR_TIE) tie = 1;
        for (int j = 0; j < F; j++) {
          if (unfilled[j] && an->data[j][2] > 0.0f)
            an->data[j][2] = 0.001f < an->data[j][2] ? 0.001f : an->data[j][2];
          if (an->data[j][2] == 0.0f) any_empty = 1;
        }
        if (any_empty) flood_fill(an, &g);
      }
    }
    free(f2); free(b2); free(nf2); free(nb2);
  }
  if (rc == ORC_OK && cfg->do_nms) n = nms_keypoints(anns, n, cfg);             /* :117-118 */
  if (rc == ORC_OK)
    export_anns(anns, n, F, L, ann_data, ann_js, ann_ndec, ann_dec, ann_nfront, ann_front);
  if (tie_flag) *tie_flag = tie;
  free(anns); free(fr); free(idx); free(scores);
  return rc == ORC_OK ? n : rc;
}

int orc_decode(const float *cif, const float *caf, int F, int L, int H, int W,
               const int *skeleton, const float *conf_scales, const orc_config *cfg,
               float *hr, float *seeds, long seed_cap, long *n_seeds,
               float *fwd, int *nfwd, float *bwd, int *nbwd, long caf_cap,
               int ann_cap, float *ann_data, float *ann_js, int *ann_ndec, int *ann_dec,
               int *ann_nfront, int *ann_front,
               int *pre_n, float *pre_data, float *pre_js, int *tie_flag,
               int init_n, const float *init_data, const float *init_js) {
  orc_set one = {cif, caf, H, W, cfg->stride, cfg->cif_min_scale, cfg->caf_min_distance,
                 cfg->caf_max_distance};
  return orc_decode_ms(&one, 1, 0, F, L, skeleton, conf_scales, cfg, hr, seeds, seed_cap, n_seeds, fwd,
                       nfwd, bwd, nbwd, caf_cap, ann_cap, ann_data, ann_js, ann_ndec, ann_dec, ann_nfront,
                       ann_front, pre_n, pre_data, pre_js, tie_flag, init_n, init_data, init_js);
}

/* ------------------------------------------------------------------------- */
/* Per-primitive entry points with the reference's own names, for known-answer */
/* tests against the compiled functional.pyx.                                 */

/* ------------------------------------------------------------------------- */
/* CifDet: generator/cifdet.py:26-52 with CifDetHr (cif_hr.py:84-100),      */
/* CifDetSeeds (cif_seeds.py:67-90) and nms.Detection (nms.py:60-102).                    */
/* field: (C,7,H,W) channels [v, x, y, _, w, h, _].                                       */

typedef struct {
  float cif_th;            /* CifHr.v_threshold        */
  float seed_th;           /* CifSeeds.threshold       */
  float seed_score_scale;  /* CifSeeds.score_scale     */
  float cif_min_scale;     /* FieldConfig.cif_min_scales[0] */
  float suppression;       /* nms.Detection.suppression       0.1 */
  float suppression_soft;  /* nms.Detection.suppression_soft  0.3 */
  float instance_th;       /* nms.Detection.instance_threshold 0.1 */
  float iou_th;            /* nms.Detection.iou_threshold     0.7 */
  float iou_th_soft;       /* nms.Detection.iou_threshold_soft 0.5 */
  int do_nms;
  int stride;
} orc_det_config;

int orc_cifdet_hr(const float *cif, int C_, int H, int W, const orc_det_config *cfg, float *hr) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  long n = (long)H * W;
  float *buf = (float *)malloc(sizeof(float) * 4 * (size_t)n);
  if (!buf) return ORC_ERR_ARG;
  float *sx = buf, *sy = buf + n, *ss = buf + 2 * n, *sv = buf + 3 * n;
  float stride = (float)cfg->stride;
  memset(hr, 0, sizeof(float) * (size_t)C_ * hh * hw);
  for (int f = 0; f < C_; f++) {
    const float *p = cif + (size_t)f * 7 * n;
    long k = 0;
    for (long c = 0; c < n; c++) {
      if (!(p[c] > cfg->cif_th)) continue;                          /* cif_hr.py:86 */
      if (cfg->cif_min_scale != 0.0f) {                             /* :87-89 */
        float lim = cfg->cif_min_scale / stride;
        if (!(p[4 * n + c] > lim) || !(p[5 * n + c] > lim)) continue;
      }
      float w = p[4 * n + c], h = p[5 * n + c];
      sx[k] = p[n + c] * stride;                                    /* :92 */
      sy[k] = p[2 * n + c] * stride;                                /* :93 */
      float m = (w != w) ? w : ((h != h) ? h : (w < h ? w : h));    /* np.minimum propagates NaN */
      float sg = 0.1f * m;
      sg = sg * stride;
      ss[k] = (sg != sg) ? sg : (1.0f > sg ? 1.0f : sg);            /* np.maximum :94 */
      sv[k] = p[c] / 16.0f / 1.0f;                                  /* :100 */
      k++;
    }
    splat_gauss_with_max(hr + (size_t)f * hh * hw, hh, hw, sx, sy, ss, sv, k, 1.0f, 1.0f);
  }
  free(buf);
  return ORC_OK;
}

static int det_seed_cmp_desc(const void *a, const void *b) {
  const float *p = (const float *)a, *q = (const float *)b;
  for (int i = 0; i < 6; i++) {
    if (p[i] > q[i]) return -1;
    if (p[i] < q[i]) return 1;
  }
  return 0;
}

/* out: cap x 6 floats (v, f, x, y, w, h) sorted like sorted(seeds, reverse=True) */
long orc_cifdet_seeds(const float *cif, int C_, int H, int W, const float *hr,
                      const orc_det_config *cfg, float *out, long cap) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  long n = (long)H * W, k = 0;
  float stride = (float)cfg->stride;
  for (int f = 0; f < C_; f++) {
    const float *p = cif + (size_t)f * 7 * n;
    for (long c = 0; c < n; c++) {
      float conf = p[c];
      if (!(conf > cfg->seed_th)) continue;                            /* cif_seeds.py:74 */
      if (cfg->cif_min_scale != 0.0f) {                                /* :75-77 */
        float lim = cfg->cif_min_scale / stride;
        if (!(p[4 * n + c] > lim) || !(p[5 * n + c] > lim)) continue;
      }
      float x = p[n + c] * stride, y = p[2 * n + c] * stride;
      float hv = scalar_value_at(hr + (size_t)f * hh * hw, hh, hw, x, y, 0.0f); /* :79 */
      float a = 0.9f * hv, b = 0.1f * conf;
      float v = a + b;                                                 /* :80 */
      if (cfg->seed_score_scale != 1.0f) v = v * cfg->seed_score_scale; /* :81-82 */
      if (!(v > cfg->seed_th)) continue;                               /* :83 */
      if (k >= cap) return ORC_ERR_CAPACITY;
      float *o = out + 6 * k++;
      o[0] = v; o[1] = (float)f; o[2] = x; o[3] = y;
      o[4] = p[4 * n + c] * stride; o[5] = p[5 * n + c] * stride;      /* :84 */
    }
  }
  qsort(out, (size_t)k, 6 * sizeof(float), det_seed_cmp_desc);
  return k;
}

/* nms.Detection.bbox_iou for one pair, float32 -- nms.py:67-77 */
static float det_iou(const float *a, const float *b) {
  float x1 = a[0] > b[0] ? a[0] : b[0];
  float y1 = a[1] > b[1] ? a[1] : b[1];
  float ax2 = a[0] + a[2], bx2 = b[0] + b[2];
  float ay2 = a[1] + a[3], by2 = b[1] + b[3];
  float x2 = ax2 < bx2 ? ax2 : bx2;
  float y2 = ay2 < by2 ? ay2 : by2;
  if (a[0] != a[0] || b[0] != b[0]) x1 = a[0] + b[0];        /* np.maximum / np.minimum propagate NaN */
  if (a[1] != a[1] || b[1] != b[1]) y1 = a[1] + b[1];
  if (ax2 != ax2 || bx2 != bx2) x2 = ax2 + bx2;
  if (ay2 != ay2 || by2 != by2) y2 = ay2 + by2;
  float dx = x2 - x1, dy = y2 - y1;
  float mx = (dx != dx) ? dx : (0.0f > dx ? 0.0f : dx);
  float my = (dy != dy) ? dy : (0.0f > dy ? 0.0f : dy);
  float inter = mx * my;
  float area_a = a[2] * a[3], area_b = b[2] * b[3];
  float den = area_a + area_b;
  den = den - inter;
  den = den + 1e-5f;
  return inter / den;
}

/* Stable sort of indices by descending score (sorted(key=-score)). */
static void det_sort(int *order, const float *score, int n) {
  for (int i = 1; i < n; i++) {
    int o = order[i], j = i;
    while (j > 0 && -score[order[j - 1]] > -score[o]) { order[j] = order[j - 1]; j--; }
    order[j] = o;
  }
}

/* Whole CifDet decode of one image.  pre_*: annotations before nms (seed order);
 * det_*: after nms.Detection.  Returns the number of final detections or < 0. */
int orc_cifdet_decode(const float *cif, int C_, int H, int W, const orc_det_config *cfg,
                      float *hr, float *seeds, long seed_cap, long *n_seeds_out,
                      int ann_cap, int *pre_field, float *pre_score, float *pre_box, int *n_pre_out,
                      int *det_field, float *det_score, float *det_box) {
  long hh = (long)(H - 1) * cfg->stride + 1, hw = (long)(W - 1) * cfg->stride + 1;
  int rc = orc_cifdet_hr(cif, C_, H, W, cfg, hr);
  if (rc) return rc;
  long ns = orc_cifdet_seeds(cif, C_, H, W, hr, cfg, seeds, seed_cap);
  if (ns < 0) return (int)ns;
  *n_seeds_out = ns;
  orc_occ occ;
  if ((rc = occ_init(&occ, C_, hh, hw))) return rc;
  occ.min_scale_reduced = 1.0f;                      /* Occupancy(shape, 2, min_scale=2.0) cifdet.py:36 */
  int n = 0;
  for (long i = 0; i < ns; i++) {                    /* cifdet.py:39-44 */
    const float *sd = seeds + 6 * i;
    int f = (int)sd[1];
    float x = sd[2], y = sd[3], w = sd[4], h = sd[5];
    if (occ_get(&occ, f, x, y)) continue;
    if (n >= ann_cap) { free(occ.m); return ORC_ERR_CAPACITY; }
    pre_field[n] = f; pre_score[n] = sd[0];
    pre_box[4 * n] = x - w / 2.0f; pre_box[4 * n + 1] = y - h / 2.0f;
    pre_box[4 * n + 2] = w; pre_box[4 * n + 3] = h;
    n++;
    float m = (h < w) ? h : w;                       /* Python min(w, h) */
    occ_set(&occ, f, x, y, 0.1f * m);
  }
  free(occ.m);
  *n_pre_out = n;
  if (!cfg->do_nms) {
    for (int i = 0; i < n; i++) {
      det_field[i] = pre_field[i]; det_score[i] = pre_score[i];
      memcpy(det_box + 4 * i, pre_box + 4 * i, 4 * sizeof(float));
    }
    return n;
  }
  /* nms.Detection.annotations -- nms.py:79-102 */
  int *order = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  float *score = (float *)malloc(sizeof(float) * (size_t)(n + 1));
  int m = 0;
  for (int i = 0; i < n; i++)
    if (pre_score[i] >= cfg->instance_th) order[m++] = i;            /* :82 */
  for (int i = 0; i < n; i++) score[i] = pre_score[i];
  det_sort(order, score, m);                                         /* :85 */
  for (int i = 1; i < m; i++) {                                      /* :88-96 */
    float max_iou = -1.0f;
    int any_nan = 0;
    for (int j = 0; j < i; j++) {
      if (!(score[order[j]] >= cfg->instance_th)) continue;
      float iou = det_iou(pre_box + 4 * order[i], pre_box + 4 * order[j]);
      if (iou != iou) any_nan = 1;
      if (iou > max_iou) max_iou = iou;
    }
    if (any_nan) continue;                                           /* np.max -> NaN, both tests false */
    if (max_iou > cfg->iou_th) score[order[i]] = score[order[i]] * cfg->suppression;
    else if (max_iou > cfg->iou_th_soft) score[order[i]] = score[order[i]] * cfg->suppression_soft;
  }
  int k = 0;
  for (int i = 0; i < m; i++)
    if (score[order[i]] >= cfg->instance_th) order[k++] = order[i];  /* :98 */
  det_sort(order, score, k);                                         /* :99 */
  for (int i = 0; i < k; i++) {
    det_field[i] = pre_field[order[i]]; det_score[i] = score[order[i]];
    memcpy(det_box + 4 * i, pre_box + 4 * order[i], 4 * sizeof(float));
  }
  free(order); free(score);
  return k;
}

int orc_det_config_size(void) { return (int)sizeof(orc_det_config); }

void orc_scalar_square_add_gauss_with_max(float *field, long fh, long fw, const float *x,
                                          const float *y, const float *sigma, const float *v,
                                          long n, float truncate, float max_value) {
  splat_gauss_with_max(field, fh, fw, x, y, sigma, v, n, truncate, max_value);
}

void orc_scalar_values(const float *field, long fh, long fw, const float *x, const float *y,
                       long n, float dflt, float *out) {
  for (long i = 0; i < n; i++) {
    float maxx = (float)fw - 1, maxy = (float)fh - 1;
    out[i] = dflt;
    if (x[i] < 0.0 || y[i] < 0.0 || x[i] > maxx || y[i] > maxy) continue;
    out[i] = field[(long)y[i] * fw + (long)x[i]];
  }
}

/* caf_center_s: returns the number of kept columns; out is (9, K) like in */
long orc_caf_center_s(const float *caf_field, long K, float x, float y, float sigma, float *out) {
  long n = 0;
  for (long i = 0; i < K; i++) {
    if (caf_field[1 * K + i] < x - sigma) continue;
    if (caf_field[1 * K + i] > x + sigma) continue;
    if (caf_field[2 * K + i] < y - sigma) continue;
    if (caf_field[2 * K + i] > y + sigma) continue;
    for (int r = 0; r < 9; r++) out[r * K + n] = caf_field[r * K + i];
    n++;
  }
  return n;
}

int orc_occupancy_get(const unsigned char *field, long fh, long fw, float x, float y, float r) {
  float cx = clipf(x / r, 0.0f, (float)(fw - 1));
  float cy = clipf(y / r, 0.0f, (float)(fh - 1));
  return field[(long)cy * fw + (long)cx];
}

void orc_occupancy_set(unsigned char *field, long fh, long fw, float x, float y, float sigma) {
  orc_occ o = {field, 1, fh, fw, 2.0f};
  occ_set(&o, 0, x, y, sigma);
}

double orc_annotation_score(const float *data /* (F,3) */, int F) {
  orc_ann a;
  a.F = F;
  for (int j = 0; j < F; j++) a.data[j][2] = data[3 * j + 2];
  return ann_score(&a);
}

float orc_detexp(float x) { return pp_expf(x); }
int orc_config_size(void) { return (int)sizeof(orc_config); }

