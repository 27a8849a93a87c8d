/*
 * pifpaf_b200.h -- C ABI of libpifpaf_b200.so, the sm_100a CUDA implementation
 * of the OpenPifPaf CifCaf decoder hot path.
 *
 * This library stands where the reference's only native module stands:
 *   /root/reference/openpifpaf/functional.pyx   (Cython, CPU, 4 live functions)
 * and additionally moves the numpy/Python stages that call it onto the device:
 *   decoder/cif_hr.py:26-65        -> pp_cifhr_accumulate
 *   decoder/cif_seeds.py:23-54     -> pp_cif_seeds
 *   decoder/caf_scored.py:32-87    -> pp_caf_scored
 *   decoder/generator/cifcaf.py:84-108,124-307 (+ occupancy.py) -> pp_cifcaf_assemble
 *   decoder/nms.py:17-57           -> pp_keypoint_nms
 *   decoder/generator/cifcaf.py:67-122 (whole __call__) -> pp_decode_batch[_host]
 *
 * Conventions
 *   - plain C: pointers, sizes, POD structs; no C++/torch types; nothing throws;
 *   - every entry returns PP_OK (0) or a negative PP_ERR_* code;
 *   - all buffers are CALLER-OWNED.  "d_" = device pointer, "h_" = host pointer;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *     device entries only enqueue work and never synchronise;
 *   - no global mutable state: re-entrant per stream / per workspace;
 *   - variable-length results use fixed capacities (pp_caps) + per-image counts;
 *     a capacity overflow is reported in the per-image `status` words
 *     (PP_STATUS_* bits), which the *_host entry turns into PP_ERR_CAPACITY.
 *
 * Tensor layouts (all float32, C-contiguous), matching what the reference's
 * CifCafCollector emits (network/heads.py:65-88) with a leading batch axis:
 *   cif  (B, F, 5, H, W)  channels [conf, x, y, logb, scale], x/y in field cells
 *   caf  (B, L, 9, H, W)  channels [conf, x1, y1, logb1, s1, x2, y2, logb2, s2]
 *   hr   (B, F, Hh, Wh)   Hh = (H-1)*stride+1, Wh = (W-1)*stride+1
 * cif / caf device pointers may point anywhere (4-byte aligned) inside an allocation whose base is
 * 16-byte aligned (cudaMalloc: 256): the kernels stage whole confidence planes with bulk copies
 * that round their byte range out to 16-byte boundaries.
 */
#ifndef PIFPAF_B200_H_
#define PIFPAF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PP_OK 0
#define PP_ERR_ARG (-1)       /* bad shape / NULL pointer / unsupported option */
#define PP_ERR_CAPACITY (-2)  /* a pp_caps capacity was exceeded               */
#define PP_ERR_CUDA (-3)      /* a CUDA runtime call failed (see pp_last_cuda_error) */
#define PP_ERR_WORKSPACE (-4) /* workspace too small                           */
#define PP_ERR_STATIC_LIMIT (-5) /* request exceeds the built-in scratch of a *_noalloc entry */

#define PP_MAX_F 32           /* keypoint types  */
#define PP_MAX_L 64           /* CAF links       */
#define PP_MAX_DEG 16         /* neighbours per joint */

/* per-image status bits written by the device entries */
#define PP_STATUS_SEED_OVERFLOW 1
#define PP_STATUS_CAF_OVERFLOW 2
#define PP_STATUS_ANN_OVERFLOW 4
#define PP_STATUS_SRC_OVERFLOW 8
#define PP_STATUS_OVERFLOW_MASK 15
/* Not an error (host entries do not fail on it): an image produced a connection score above 1, which
 * needs CAF confidences above 1 (raw, un-squashed head outputs).  The assembly evaluates frontier
 * edges eagerly; that equals the reference's lazy PriorityQueue only while evaluated scores stay at
 * or below their bound sqrt(v_start), i.e. while connection scores are <= 1 (cifcaf.py:258-285).
 * With this bit set the annotations are still deterministic, but their decoding order is not proven
 * to be the reference's. */
#define PP_STATUS_SCORE_ABOVE_ONE 16

typedef struct pp_shape {
  int32_t B, F, L, H, W, stride;
} pp_shape;

/* Capacities of the variable-length buffers, per image (0 = worst case). */
typedef struct pp_caps {
  int32_t seed_cap; /* seeds per image            (worst case F*H*W) */
  int32_t caf_cap;  /* candidates per link & direction (worst case H*W) */
  int32_t ann_cap;  /* annotations per image      (default 256)      */
  int32_t src_cap;  /* CifHr sources per field    (worst case H*W)   */
} pp_caps;

/* initial_annotations of CifCaf.__call__ (cifcaf.py:67-71,95-98): annotations that
 * are grown and marked occupied before the seeds are visited.  Device pointers. */
typedef struct pp_initial {
  const float *data;    /* (B, cap, F, 3) x, y, v */
  const float *scales;  /* (B, cap, F) joint_scales */
  const int32_t *n;     /* (B) annotations per image */
  int32_t cap;
} pp_initial;

/* Per-image preprocessing record for annotations_inverse (transforms/preprocess.py:35-82):
 * the decoded coordinates are mapped back to the original image.  The host fills this from
 * the reference's meta dict; trigonometry is done on the host (math.cos / math.sin of
 * -meta['rotation']['angle'] degrees, then rounded to float32 like numpy does for a Python
 * scalar against a float32 array). */
typedef struct pp_image_meta {
  double offset_x, offset_y;  /* meta['offset']                                  */
  double scale_x, scale_y;    /* meta['scale']                                   */
  double flip_w1;             /* meta['width_height'][0] - 1 (used when hflip)   */
  float rot_cos, rot_sin;     /* of -angle                                       */
  float rot_cx, rot_cy;       /* (rotation width - 1) / 2, (height - 1) / 2      */
  int32_t rotate;             /* angle != 0                                      */
  int32_t hflip;              /* meta['hflip']                                   */
} pp_image_meta;

typedef struct pp_inverse {
  const pp_image_meta *meta;    /* (B) device array                                            */
  const int32_t *hflip_target;  /* (F) device: row j moves to row hflip_target[j] when hflip;
                                   a permutation (meta['horizontal_swap']), or NULL = no swap */
} pp_inverse;

/* Decoder knobs; same meaning and defaults-by-configuration as the reference's
 * class attributes (decoder/factory.py:74-92). */
typedef struct pp_params {
  double nms_instance_threshold; /* nms.Keypoints.instance_threshold (float64 compare) */
  float cif_threshold;           /* CifHr.v_threshold          */
  float seed_threshold;          /* CifSeeds.threshold         */
  float seed_score_scale;        /* CifSeeds.score_scale       */
  float caf_threshold;           /* CafScored.default_score_th */
  float cif_floor;               /* CafScored cif_floor (0.1)  */
  float cif_floor_compl;         /* float32(1.0 - cif_floor), computed in double by the host */
  float keypoint_threshold;      /* CifCaf.keypoint_threshold  */
  float nms_keypoint_threshold;  /* nms.Keypoints.keypoint_threshold */
  float nms_suppression;         /* nms.Keypoints.suppression  */
  float cif_min_scale;           /* FieldConfig.cif_min_scales[i]   (0 = off)  */
  float caf_min_distance;        /* FieldConfig.caf_min_distances[i] (0 = off) */
  float caf_max_distance;        /* FieldConfig.caf_max_distances[i] (<=0 = None) */
  int32_t connection_method;     /* 0 = blend, 1 = max         */
  int32_t greedy;                /* CifCaf.greedy              */
  int32_t force_complete;        /* CifCaf.force_complete      */
  int32_t nms;                   /* 0: CifCaf(nms=None)        */
  int32_t exp_mode;              /* 1 = deterministic pp_expf (default), 0 = CUDA expf */
  uint32_t seed_mask;            /* FieldConfig.seed_mask as bits (0 = all fields seed, unless
                                    seed_mask_enabled); fields beyond PP_MAX_F cannot be masked */
  struct pp_profile *profile;    /* optional per-stage CUDA-event timer (NULL = off)    */
  const pp_initial *initial;     /* optional initial annotations (NULL = none)          */
  const pp_inverse *inverse;     /* optional annotations_inverse after the NMS (NULL = none) */
  int32_t host_mapped;           /* host entries only: copy just the confidence planes (1/5 of cif, 1/9
                                    of caf) and let the kernels read the other planes of the active
                                    cells straight from the pinned, device-mapped host buffers.  Used
                                    when seed_threshold >= cif_threshold and no force_complete;
                                    otherwise the full copy is made.  Same results.            */
  int32_t in_flight;             /* hint: batches the caller keeps in flight on other streams (0/1 =
                                    this one only).  > 1 makes the assembly kernel stage its lists in
                                    80 KB instead of a whole SM's shared memory, so that CTAs of the
                                    other batches' kernels can share its SMs.  Same results.     */
  /* tuning knobs (0 = the library's choice).  They change the schedule, never a result. */
  int32_t asm_heavy_seeds;       /* seeds per image above which the assembly indexes the seeds by
                                    (field, row band) for its occupancy updates (default 4096; < 0 = always) */
  int32_t asm_stage_kb;          /* KB of shared memory for the staged candidate lists (< 0 = none)      */
  int32_t cifhr_group;           /* strips per work item of the empty-strip producer (default 8)         */
  int32_t cifhr_ctas_per_sm;     /* persistent CTAs per SM of the strips kernel (default 1)              */
  int32_t cifhr_rshift;          /* log2 rows per strip + 1 (default 4 = 8 rows; clamped to what fits)   */
  float cif_neighbors;           /* CifHr.neighbors (cif_hr.py:15,40): v / neighbors; 0 = 16            */
  int32_t seed_mask_enabled;     /* 1: seed_mask applies even when it is all zero (no field seeds)      */
  int32_t cifhr_warps;           /* tuning: warps per CTA of the strips kernel, one strip each (default 3; 4 for
                                    images wider than 767 px; max 8); < 0: the round-1 CTA-per-strip kernel
                                    does every strip                                                      */
  int32_t cifhr_depth;           /* tuning: bulk stores the empty-strip producer keeps queued (default 8; > 16 = no limit) */
  struct pp_gate *splat_gate;    /* optional (NULL = off): decodes issued on different streams with the same
                                    gate run their CifHr splat kernels one after the other (see pp_gate) */
} pp_params;

/* A splat gate.  With several batches in flight on their own streams, the HBM-bound splat kernels of
 * different batches otherwise run side by side, finish together and leave the memory system idle while
 * all of them are in their latency-bound stages (a convoy).  Decodes that share a gate order ONLY their
 * splat kernels (event chain, submission order): each splat runs alone at full bandwidth, back to back
 * with the next, and the other stages of the other batches fill the SMs around it.  A gate belongs to
 * the host thread that submits the decodes; results never depend on it. */
typedef struct pp_gate pp_gate;
pp_gate *pp_gate_create(void);
void pp_gate_destroy(pp_gate *gate);

/* Skeleton: L links of 1-based joint ids (j1, j2), as in the reference's
 * COCO_PERSON_SKELETON; confidence_scales may be NULL (cifcaf.py:39,52). */
typedef struct pp_skeleton {
  int32_t L;
  int32_t joints[PP_MAX_L][2];
  int32_t has_confidence_scales;
  float confidence_scales[PP_MAX_L];
} pp_skeleton;

/* Result buffers of one batch (device pointers for pp_decode_batch, host
 * pointers for pp_decode_batch_host).  Any of the optional ones may be NULL. */
typedef struct pp_outputs {
  float *hr;            /* optional (B,F,Hh,Wh); if NULL the workspace copy is used */
  float *seeds;         /* optional (B,seed_cap,5): v, f, x, y, s sorted descending */
  int32_t *n_seeds;     /* optional (B)                                  */
  float *caf_fwd;       /* optional (B,L,9,caf_cap)                      */
  float *caf_bwd;       /* optional (B,L,9,caf_cap)                      */
  int32_t *n_caf_fwd;   /* optional (B,L)                                */
  int32_t *n_caf_bwd;   /* optional (B,L)                                */
  float *ann_data;      /* (B,ann_cap,F,3) x, y, v   -- final annotations, score order */
  float *ann_scales;    /* (B,ann_cap,F)   joint_scales                  */
  double *ann_score;    /* optional (B,ann_cap) Annotation.score()       */
  int32_t *ann_source;  /* optional (B,ann_cap) index into the pre-NMS list */
  int32_t *n_ann;       /* (B)                                           */
  float *pre_data;      /* optional (B,ann_cap,F,3) annotations before force-complete/NMS */
  float *pre_scales;    /* optional (B,ann_cap,F)                        */
  int32_t *pre_order;   /* optional (B,ann_cap,F,2) decoding_order (start,end), -1 padded */
  int32_t *n_pre;       /* optional (B)                                  */
  int32_t *status;      /* (B) PP_STATUS_* bits                          */
  void *packed;         /* optional, pp_packed_bytes(shape, caps) bytes, 256-byte aligned: the live
                           result rows of the batch gathered into one dense block (see "Packed
                           results" below) so that ONE small copy brings a batch's results to the host */
} pp_outputs;

/* ---- Packed results ---------------------------------------------------------------------------
 * The fixed-capacity result arrays above hold ann_cap rows per image of which a handful are live
 * (the reference returns a Python list of exactly the live annotations, cifcaf.py:110-122).  When
 * pp_outputs.packed is set, the decode entries finish with a gather kernel that writes, for the
 * images [0, B) of the call, one block of 32-bit words:
 *     word 0..3            B, F, total (= sum of n_ann), record_words
 *     word 4..4+B          n_ann[B]
 *     then                 status[B], n_pre[B], pad to an even word count
 *     then                 `total` records of record_words = pp_packed_record_words(F) words, image
 *                          after image, score order inside an image:
 *         [0..1] Annotation.score() as float64   [2] ann_source   [3] 0
 *         [4 ..)  ann_data (F,3)   ann_scales (F)   pre_data[ann_source] (F,3)   pre_order[ann_source] (F,2)
 * Host entries given h_out->packed copy header + live records only (pp_decode_batch_host: counts
 * first, then the rows; the pipelined entry: a guessed size in one copy, the rare remainder in
 * pp_pipeline_wait) and skip every other result array whose h_out member is NULL.  A chunk of the
 * pipelined entry starting at image b0 writes its block at packed + b0 * pp_packed_bytes(1 image). */
size_t pp_packed_bytes(const pp_shape *s, const pp_caps *c);   /* capacity for the worst case */
int32_t pp_packed_record_words(int32_t F);
size_t pp_packed_header_words(int32_t B);                      /* words before the first record */
/* The gather on its own (the decode entries call it when d_out->packed is set). */
int pp_pack_results(const pp_outputs *d_out, const pp_shape *s, const pp_caps *c, void *stream);

/* Per-stage device timing.  When pp_params.profile is set, pp_decode_batch
 * records one CUDA event before every stage and one after the last, on the
 * stream it launches on; nothing synchronises until pp_profile_read. */
#define PP_N_STAGES 7 /* cif_compact, cifhr_splat, seeds_collect, seeds_sort, caf_scored, assemble, nms */
typedef struct pp_profile pp_profile;
pp_profile *pp_profile_create(int max_steps);
void pp_profile_destroy(pp_profile *prof);
void pp_profile_reset(pp_profile *prof);
/* Synchronises the recorded events; h_ms (max_steps, PP_N_STAGES) <- ms per stage. */
int pp_profile_read(pp_profile *prof, float *h_ms, int32_t *n_steps);
/* The same events as absolute times: h_ms (max_steps, PP_N_STAGES + 1) <- ms from the first event of
 * step 0 to the start of every stage of every step and to the step's end (the steps may have run on
 * different streams: a timeline of batches in flight). */
int pp_profile_read_abs(pp_profile *prof, float *h_ms, int32_t *n_steps);
const char *pp_stage_name(int stage);
/* Kernel launches + memsets one pp_decode_batch enqueues (constant per call). */
int pp_decode_launch_count(const pp_params *p);

const char *pp_error_string(int code);
const char *pp_last_cuda_error(void);   /* thread-local text of the last CUDA failure */
const char *pp_version(void);

/* Fill defaults: params = SURVEY.md 8d decoder settings; caps = worst case. */
void pp_default_params(pp_params *p);
int pp_default_caps(const pp_shape *s, pp_caps *c);
int pp_coco_skeleton(pp_skeleton *sk);

/* Bytes of device scratch pp_decode_batch needs (includes hr when out->hr==NULL). */
size_t pp_workspace_bytes(const pp_shape *s, const pp_caps *c);
/* The same with the caller's own hr buffer in out->hr (hr_external != 0): the scratch without the hr map. */
size_t pp_workspace_bytes_ex(const pp_shape *s, const pp_caps *c, int32_t hr_external);

/* Device memory for hr maps (cif_hr.py:14-24 `accumulated`; B * F * hr_h * hr_w floats) that asks the driver
 * for compute data compression: cuMemCreate with CU_MEM_ALLOCATION_COMP_GENERIC, which cudaMalloc cannot
 * request.  The map is written once per decode, mostly with zeros, and B200 writes such data to HBM faster
 * when the allocation is compressible (measured with the strips kernel's store pattern: 7.5 TB/s against
 * 6.3 TB/s into a cudaMalloc buffer, scripts/micro/comp_probe.cu).  Pass the pointer as pp_outputs.hr.  On
 * the current device; *compressed (may be NULL) <- 1 when the allocation is compressible, 0 when the device
 * does not support it (the memory is then ordinary device memory).  Contents are identical either way.
 * pp_hr_free: no work may be in flight on the buffer.  Returns PP_OK or PP_ERR_CUDA / PP_ERR_ARG. */
int pp_hr_alloc(size_t bytes, void **d_ptr, int32_t *compressed);
int pp_hr_free(void *d_ptr);

/* ---- stage entry points (device pointers, enqueue only) ------------------ */

/* CifHr.fill for one cif head: d_hr <- min(1, sum of clipped Gaussians).
 * Replaces cif_hr.py:26-65 + functional.pyx:105-141.  Bit-exact. */
int pp_cifhr_accumulate(const float *d_cif, float *d_hr, void *d_ws, size_t ws_bytes,
                        const pp_shape *s, const pp_caps *c, const pp_params *p,
                        int32_t *d_status, void *stream);

/* CifHr.fill_multiple (cif_hr.py:42-64): up to two heads accumulated into ONE map with v / neighbors / n_heads
 * (d_cif2 = the second head of an hflip pair, or NULL), then either stored (merge = 0, the first call
 * of a CifHr object) or merged into the existing d_hr with np.maximum (merge = 1, every later scale). */
/* det != 0: CifDet fields (B,C,7,H,W), sigma = max(1, 0.1 min(w, h) stride) (CifDetHr, cif_hr.py:84-100).
 * s_hr: the shape (H, W, stride of the FIRST head filled) that defines the hr map, NULL = s. */
int pp_cifhr_accumulate_ex(const float *d_cif, const float *d_cif2, float *d_hr, int32_t merge, int32_t det,
                           void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_shape *s_hr,
                           const pp_caps *c, const pp_params *p, int32_t *d_status, void *stream);

/* CifSeeds.fill + get(): sorted seeds (v,f,x,y,s).  cif_seeds.py:23-54. */
int pp_cif_seeds(const float *d_cif, const float *d_hr, float *d_seeds, int32_t *d_n_seeds,
                 void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                 const pp_params *p, int32_t *d_status, void *stream);

/* CafScored.fill: forward/backward candidate lists.  caf_scored.py:32-87. */
int pp_caf_scored(const float *d_caf, const float *d_hr, const pp_skeleton *sk, float score_th,
                  float *d_fwd, int32_t *d_nfwd, float *d_bwd, int32_t *d_nbwd,
                  const pp_shape *s, const pp_caps *c, const pp_params *p, int32_t *d_status,
                  void *stream);

/* Seed loop + _grow + occupancy (+ force-complete) (+ keypoint NMS when p->nms).
 * cifcaf.py:84-122.  d_fwd2/d_bwd2 (score_th 1e-4 lists) only for force_complete. */
int pp_cifcaf_assemble(const float *d_seeds, const int32_t *d_n_seeds, const float *d_fwd,
                       const int32_t *d_nfwd, const float *d_bwd, const int32_t *d_nbwd,
                       const float *d_fwd2, const int32_t *d_nfwd2, const float *d_bwd2,
                       const int32_t *d_nbwd2, const pp_skeleton *sk, const pp_outputs *d_out,
                       void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                       const pp_params *p, void *stream);

/* nms.Keypoints.annotations on device buffers, in place.  nms.py:17-57. */
int pp_keypoint_nms(float *d_ann_data, float *d_ann_scales, int32_t *d_n_ann, double *d_score,
                    int32_t *d_source, void *d_ws, size_t ws_bytes, const pp_shape *s,
                    const pp_caps *c, const pp_params *p, void *stream);

/* pp_keypoint_nms without any device allocation by the caller or the library.  NO COUNTERPART IN
 * THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same inputs, outputs and results as pp_keypoint_nms; the scratch that entry takes as d_ws lives
 * in a static device array of PP_NMS_NOALLOC_BYTES inside the library (every GPU that loads the
 * library holds it until exit).  A request needing more (pp_keypoint_nms_noalloc_bytes(s, c) >
 * pp_keypoint_nms_noalloc_max_bytes(); with F = 17 that is B * ann_cap > 3577 annotation slots for B = 1)
 * returns PP_ERR_STATIC_LIMIT and enqueues nothing.  The call makes no cudaMalloc / cudaFree and
 * uses no allocating library routine; it only enqueues three device-to-device copies and one
 * kernel on `stream`.  All calls on one GPU share the array: calls that overlap on the GPU race --
 * keep them on one stream or order them with events.  Nothing in the array has to be cleared
 * between calls (every word read is written earlier in the same call). */
#define PP_NMS_NOALLOC_BYTES (4 * 1024 * 1024)
size_t pp_keypoint_nms_noalloc_max_bytes(void);
size_t pp_keypoint_nms_noalloc_bytes(const pp_shape *s, const pp_caps *c);   /* 0 = bad shape */
int pp_keypoint_nms_noalloc(float *d_ann_data, float *d_ann_scales, int32_t *d_n_ann, double *d_score,
                            int32_t *d_source, const pp_shape *s, const pp_caps *c,
                            const pp_params *p, void *stream);

/* Inference epilogue of a fused head network + field collection, one pass:
 * PixelShuffle(2) and crop (quad = 1), sigmoid on confidences, exp on scales,
 * absolute coordinates (regression + cell index) and the decoder's channel order.
 * Replaces the tail of CompositeFieldFused.forward (network/heads.py:406-448) and
 * CifCafCollector.forward (network/heads.py:65-88).
 *   d_x   (B, n_fields * (1 + 4 * n_vectors) * (quad ? 4 : 1), h, w) raw convolution output
 *   d_out (B, n_fields, 1 + 4 * n_vectors, H, W), H = quad ? 2h-1 : h, W likewise
 * n_vectors = 1 gives a cif field, 2 a caf field. */
int pp_head_epilogue(const float *d_x, float *d_out, int32_t B, int32_t n_fields, int32_t n_vectors,
                     int32_t h, int32_t w, int32_t quad, void *stream);

/* ---- multi-scale FieldConfig (decoder/factory.py:153-180, field_config.py) ---------------------
 * One field set per entry of FieldConfig.cif_indices / caf_indices: its own resolution, stride and
 * filters.  Set 0 defines the shape of the hr map (cif_hr.py:44-51). */
#define PP_MAX_SETS 16
typedef struct pp_field_set {
  const float *cif;        /* device (B,F,5,H,W) */
  const float *caf;        /* device (B,L,9,H,W) */
  int32_t H, W, stride;    /* FieldConfig.cif_strides[k] (= caf_strides[k]) */
  float cif_min_scale;     /* FieldConfig.cif_min_scales[k]                 */
  float caf_min_distance;  /* FieldConfig.caf_min_distances[k]              */
  float caf_max_distance;  /* FieldConfig.caf_max_distances[k]; 0 = None    */
  const float *cif_rest;   /* optional: where the non-confidence planes of cif / caf are read from  */
  const float *caf_rest;   /* (same layout; used by the host-mapped entries, NULL = cif / caf)      */
} pp_field_set;

/* pp_decode_batch over several field sets: CifHr maps merged with np.maximum (cif_hr.py:59-78;
 * hflip_pairs != 0: sets (k, k + n/2) are accumulated together with v / neighbors / 2, the
 * reference's 10-head case), seeds of all sets in one sorted list (cif_seeds.py:56-65), candidate
 * lists concatenated in set order (caf_scored.py:89-98).  `s` carries B, F, L and the H, W, stride
 * of set 0; capacities must cover the sums over the sets (seed_cap, caf_cap) and the largest
 * group (src_cap). */
int pp_decode_batch_multiscale(const pp_field_set *sets, int32_t n_sets, int32_t hflip_pairs,
                               const pp_skeleton *sk, const pp_outputs *d_out, void *d_ws,
                               size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                               const pp_params *p, void *stream);

/* ---- CifDet (detection decoder), generator/cifdet.py:17-52 ---------------------------------
 * Fields are (B, C, 7, H, W) float32 with channels [v, x, y, _, w, h, _] in field units, C
 * categories (<= PP_MAX_CATEGORIES).  pp_shape.F carries C, pp_shape.L is ignored. */
#define PP_MAX_CATEGORIES 256

typedef struct pp_det_params {
  float cif_threshold;            /* CifHr.v_threshold                 cif_hr.py:16   */
  float seed_threshold;           /* CifSeeds.threshold                cif_seeds.py:14 */
  float seed_score_scale;         /* CifSeeds.score_scale              cif_seeds.py:15 */
  float cif_min_scale;            /* FieldConfig.cif_min_scales[0]                     */
  float nms_suppression;          /* nms.Detection.suppression         nms.py:61 (0.1) */
  float nms_suppression_soft;     /* nms.Detection.suppression_soft    nms.py:62 (0.3) */
  float nms_instance_threshold;   /* nms.Detection.instance_threshold  nms.py:63 (0.1) */
  float nms_iou_threshold;        /* nms.Detection.iou_threshold       nms.py:64 (0.7) */
  float nms_iou_threshold_soft;   /* nms.Detection.iou_threshold_soft  nms.py:65 (0.5) */
  int32_t nms;                    /* 0: skip nms.Detection                            */
} pp_det_params;

/* Device result buffers of one batch of detections.  Optional members may be NULL. */
typedef struct pp_det_outputs {
  float *hr;            /* optional (B,C,Hh,Wh) CifDetHr.accumulated                   */
  float *seeds;         /* optional (B,seed_cap,6): v, f, x, y, w, h sorted descending */
  int32_t *n_seeds;     /* optional (B)                                                */
  int32_t *pre_field;   /* optional (B,ann_cap) detections before the NMS, seed order  */
  float *pre_score;     /* optional (B,ann_cap)                                        */
  float *pre_box;       /* optional (B,ann_cap,4) x, y, w, h; 16-byte aligned          */
  int32_t *n_pre;       /* optional (B)                                                */
  int32_t *det_field;   /* (B,ann_cap) AnnotationDet.field_i, score order              */
  float *det_score;     /* (B,ann_cap) AnnotationDet.score                             */
  float *det_box;       /* (B,ann_cap,4) AnnotationDet.bbox x, y, w, h; 16-byte aligned */
  int32_t *det_source;  /* optional (B,ann_cap) index into the pre list                */
  int32_t *n_det;       /* (B)                                                         */
  int32_t *status;      /* (B) PP_STATUS_* bits                                        */
} pp_det_outputs;

void pp_det_default_params(pp_det_params *p);
size_t pp_cifdet_workspace_bytes(const pp_shape *s, const pp_caps *c);
/* CifDet.__call__ for a batch: CifDetHr (cif_hr.py:84-100) -> CifDetSeeds (cif_seeds.py:67-90)
 * -> seed loop with Occupancy(shape, 2, min_scale=2.0) (cifdet.py:34-44) -> nms.Detection
 * (nms.py:60-102).  Asynchronous on `stream`. */
int pp_cifdet_decode_batch(const float *d_field, const pp_det_outputs *d_out, void *d_ws,
                           size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                           const pp_det_params *p, void *stream);

/* Preprocess.anndet_inverse (transforms/preprocess.py:84-94) on detection boxes, in place:
 * d_box (B,ann_cap,4) x, y, w, h.  Rotation of the corners + bounding box (utils.rotate_box),
 * offset, scale; like the reference, no hflip. */
int pp_detections_inverse(float *d_box, const int32_t *d_n, const pp_inverse *inv, int32_t B,
                          int32_t ann_cap, void *stream);

/* nms.Detection.annotations (nms.py:79-102) on its own: B lists of `cap` detections in any
 * order (d_box 16-byte aligned); results in score order.  Scratch: pp_detection_nms_workspace_bytes. */
size_t pp_detection_nms_workspace_bytes(int32_t B, int32_t cap);
int pp_detection_nms(const int32_t *d_field, const float *d_score, const float *d_box,
                     const int32_t *d_n, int32_t *d_out_field, float *d_out_score, float *d_out_box,
                     int32_t *d_out_source /* optional: input index per result */, int32_t *d_out_n, void *d_ws, size_t ws_bytes, int32_t B, int32_t cap,
                     const pp_det_params *p, void *stream);

/* sizeof of the ABI structs, for bindings to check their mirrors:
 * which = 0 pp_params, 1 pp_image_meta, 2 pp_outputs, 3 pp_skeleton, 4 pp_shape, 5 pp_caps,
 * 6 pp_det_params, 7 pp_det_outputs. */
int32_t pp_sizeof(int32_t which);

/* Preprocess.annotations_inverse (transforms/preprocess.py:35-82) for a whole batch, in place on
 * the decoder's result buffers: rotation (float32), offset and scale (float64, rounded to
 * float32 after each step, as numpy's in-place operators do), joint_scales / scale_x, hflip with
 * the horizontal keypoint swap.  d_pre_data (may be NULL) holds the as-grown rows that
 * decoding_order is rebuilt from; like the reference's decoding_order entries they get offset
 * and scale only. */
int pp_annotations_inverse(float *d_ann_data, float *d_ann_scales, const int32_t *d_n_ann,
                           float *d_pre_data, const int32_t *d_n_pre, const pp_inverse *inv,
                           int32_t B, int32_t F, int32_t ann_cap, void *stream);

/* ---- fused drivers ------------------------------------------------------- */

/* CifCaf.__call__ for a batch, everything device-resident (cifcaf.py:67-122). */
int pp_decode_batch(const float *d_cif, const float *d_caf, const pp_skeleton *sk,
                    const pp_outputs *d_out, void *d_ws, size_t ws_bytes, const pp_shape *s,
                    const pp_caps *c, const pp_params *p, void *stream);

/* Same, from HOST fields to HOST results: H2D copies, kernels, D2H copies and
 * one stream synchronise inside.  d_cif/d_caf/d_ws are caller-owned device
 * staging buffers; d_out are device result buffers, h_out their host mirrors
 * (only the non-NULL h_out members are copied back).  This is the call the
 * reference-side binding makes in place of Generator.batch's CPU decode
 * (decoder/generator/generator.py:84-101). */
int pp_decode_batch_host(const float *h_cif, const float *h_caf, float *d_cif, float *d_caf,
                         const pp_skeleton *sk, const pp_outputs *d_out, const pp_outputs *h_out,
                         void *d_ws, size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                         const pp_params *p, void *stream);

/* Pipelined variant of pp_decode_batch_host for pinned host buffers: the batch
 * is cut into chunks of `chunk_images`; the host->device copy of chunk c+1 and
 * the device->host copy of chunk c-1 (two internal streams of the pipeline
 * object) overlap the kernels of chunk c (on `stream`).  Only enqueues;
 * pp_pipeline_wait(pipe, c) blocks until chunk c's results are in h_out (so the
 * caller can consume chunk c while later chunks are in flight) and returns
 * PP_ERR_CAPACITY if one of its images overflowed; c = -1 waits for all. */
typedef struct pp_pipeline pp_pipeline;
pp_pipeline *pp_pipeline_create(int max_chunks);
void pp_pipeline_destroy(pp_pipeline *pipe);
int pp_decode_batch_host_async(pp_pipeline *pipe, const float *h_cif, const float *h_caf,
                               float *d_cif, float *d_caf, const pp_skeleton *sk,
                               const pp_outputs *d_out, const pp_outputs *h_out, void *d_ws,
                               size_t ws_bytes, const pp_shape *s, const pp_caps *c,
                               const pp_params *p, int chunk_images, void *stream);
int pp_pipeline_wait(pp_pipeline *pipe, int chunk);
int pp_pipeline_chunks(const pp_pipeline *pipe);

/* ---- parity shim: the reference's live functional.pyx primitives ---------
 * Same names / argument meaning as functional.pyx, on device buffers. */

/* functional.pyx:105-141; field (fh,fw) updated in place, n sources. */
int pp_scalar_square_add_gauss_with_max(float *d_field, int fh, int fw, const float *d_x,
                                        const float *d_y, const float *d_sigma, const float *d_v,
                                        int n, float truncate, float max_value, void *stream);
/* functional.pyx:231-244 */
int pp_scalar_values(const float *d_field, int fh, int fw, const float *d_x, const float *d_y,
                     int n, float default_value, float *d_out, void *stream);
/* functional.pyx:338-359; d_out (9,K) compacted in order, *d_count <- kept columns */
int pp_caf_center_s(const float *d_caf_field, int K, float x, float y, float sigma, float *d_out,
                    int32_t *d_count, void *stream);
/* functional.pyx:281-286; result (0/non-zero) written to *d_out */
int pp_scalar_nonzero_clipped_with_reduction(const uint8_t *d_field, int fh, int fw, float x,
                                             float y, float r, uint8_t *d_out, void *stream);
/* decoder/utils.py:61-66 (Occupancy.set's primitive); integer x, y, sigma */
int pp_scalar_square_add_single(uint8_t *d_field, int fh, int fw, long long x, long long y,
                                long long sigma, int value, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PIFPAF_B200_H_ */
