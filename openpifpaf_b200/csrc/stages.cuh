// stages.cuh -- internal launcher interface between api.cu and the stage files.
#pragma once
#include "common.cuh"

namespace pp {

// splat gate (api.cu)
int gate_wait(pp_gate *g, cudaStream_t st);
int gate_signal(pp_gate *g, cudaStream_t st);

size_t cifhr_ws_bytes(const pp_shape &s, const pp_caps &c);
// det = 1: CifDet fields (B,C,7,H,W) [v, x, y, _, w, h, _]; sigma / seed records follow
// CifDetHr / CifDetSeeds (cif_hr.py:84-100, cif_seeds.py:67-90); seeds are 6 floats (v, f, x, y, w, h)
int cifhr_launch(const float *d_cif, float *d_hr, void *d_ws, const pp_shape &s, const pp_caps &c,
                 const pp_params &p, int *d_status, cudaStream_t st, int det = 0);
// The compacted CIF sources of the last cifhr launch on this workspace (cifhr.cu), for the seed stage.
struct CifSrcView {
  const float4 *a;     // (B,F,src_cap) cx, cy, sigma^2, v / 16 / len_cifs
  const float2 *c;     // (B,F,src_cap) scale in pixels (CifDet: w, h in pixels)
  const int *n;        // (B,F)
  int src_cap;
};
CifSrcView cifhr_sources(void *d_ws, const pp_shape &s_hr, const pp_caps &c);
// Raw CAF candidate collection riding in the CIF compact launch (single field set; the part of
// CafScored.fill_caf that does not need the hr map: threshold, distance filters, stride scaling --
// caf_scored.py:42-58).  The lists are finished by caf_score_bin_launch once the hr map exists.
struct CafPreArgs {
  const float *caf;        // (B,L,9,H,W)
  const float *caf_rest;   // where planes 1..8 are read (host-mapped input), or NULL = caf
  float *raw;              // (B,L,9,cap): candidates above the threshold, rows 1..8 scaled by the stride, cell order
  int *n_raw;              // (B,L)
  int L, cap;
  float th, min_dist, max_dist;      // distances in field cells
  int use_min, use_max;
  int skip_logb;           // rows 3 and 7 (log-b, never used by the assembly) are written as 0 instead of
                           // being read: host-mapped input whose CafScored lists the caller does not keep
};
int cifhr_launch_ex(const float *d_cif, const float *d_cif2, float *d_hr, void *d_ws, const pp_shape &s,
                    const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_scale,
                    int *d_status, cudaStream_t st, int det, int merge, const float *cif_rest = nullptr,
                    const CafPreArgs *caf_pre = nullptr);
size_t seeds_ws_bytes(const pp_shape &s, const pp_caps &c, int det = 0);
// the two halves of seeds_launch, for multi-scale: collect per field set (reset = first set), sort once
int seeds_collect_launch(const float *d_cif, const float *d_hr, void *d_ws, const pp_shape &s,
                         const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_scale,
                         int reset, cudaStream_t st, int det);
// the same from the compacted CIF sources instead of the field (valid when seed_threshold >=
// cif_threshold, one field set, same min-scale filter): ~60 sources per plane instead of 6561 cells
int seeds_collect_from_sources_launch(const CifSrcView &src, const float *d_hr, void *d_ws, const pp_shape &s,
                                      const pp_caps &c, const pp_params &p, cudaStream_t st, int det);
// per-field sort + merge (seeds.cu): replaces collect-from-sources + sort when seeds_segmented_ok
bool seeds_segmented_ok(const pp_shape &s, const pp_caps &c);
int seeds_segmented_launch(const CifSrcView &src, const float *d_hr, float *d_seeds, int *d_n_seeds, void *d_ws,
                           const pp_shape &s, const pp_caps &c, const pp_params &p, cudaStream_t st, int det);
int seeds_sort_launch(float *d_seeds, int *d_n_seeds, void *d_ws, int B, const pp_caps &c,
                      const pp_params &p, int *d_status, cudaStream_t st, int det);
int seeds_launch(const float *d_cif, const float *d_hr, float *d_seeds, int *d_n_seeds, void *d_ws,
                 const pp_shape &s, const pp_caps &c, const pp_params &p, int *d_status,
                 cudaStream_t st, int det = 0);

// CifDet: seed loop with occupancy + nms.Detection (det.cu)
struct DetArgs {
  const float *seeds;      // (B, seed_cap, 6) sorted
  const int *n_seeds;
  uint8_t *occ_cnt;        // (B, seed_cap)
  int2 *occ_rec;           // (B, seed_cap)
  int *pre_field; float *pre_score; float *pre_box; int *n_pre;      // before the NMS, seed order
  int *det_field; float *det_score; float *det_box; int *n_det;      // after nms.Detection
  int *det_source;         // optional: index into the pre list
  int *status;
  int seed_cap, ann_cap, hh, hw, C;
};
size_t det_nms_ws_bytes(int B, int cap);
int det_nms_launch(const int *in_field, const float *in_score, const float *in_box, const int *n_in,
                   int *det_field, float *det_score, float *det_box, int *det_source, int *n_det,
                   void *d_ws, int B, int cap, const pp_det_params &p, cudaStream_t st);
int det_assemble_launch(const DetArgs &a, int B, const pp_det_params &p, float *work_score,
                        cudaStream_t st);

int caf_scored_launch(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                      float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                      const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st);
// s_hr: the shape that defines the hr map; append: continue the lists of an earlier field set
int caf_scored_launch_ex(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                         float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                         const pp_shape &s_hr, const pp_caps &c, const pp_params &p, float min_distance,
                         float max_distance, int append, int *d_status, cudaStream_t st,
                         const float *caf_rest = nullptr);
// Candidate lists re-ordered by spatial bin of their source end, for the assembly
// scans (caf_scored.cu: caf_bin_kernel).  Per (image, link, direction): seven rows
// (score, source x/y, target x/y/scale, original list index) of `cap` columns, and
// the start column of every bin (bins row-major, plus one last bin for NaN sources).
struct BinGeo {
  int bshift;      // bin edge = 1 << bshift hr pixels
  int nbx, nby;
  int nb;          // nbx * nby + 1
};
constexpr int kBinRows = 7;
constexpr int kMaxBins = 2048;
BinGeo bin_geo(const pp_shape &s);
size_t binned_rows_bytes(const pp_shape &s, const pp_caps &c);    // one (fwd + bwd) store
size_t binned_start_bytes(const pp_shape &s);
// CafScored.fill_caf from the raw candidates of CafPreArgs (in d_fwd / d_nfwd): hr gathers, both
// ordered compactions (the forward list in place), then the spatial binning of both lists.
int caf_score_bin_launch(const float *d_hr, const pp_skeleton &sk, float score_th, float *d_fwd, int *d_nfwd,
                         float *d_bwd, int *d_nbwd, float *d_rows, int *d_start, const pp_shape &s,
                         const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st);
int caf_bin_launch(const float *d_fwd, const int *d_nfwd, const float *d_bwd, const int *d_nbwd,
                   float *d_rows, int *d_start, const pp_shape &s, const pp_caps &c, cudaStream_t st);

struct NmsArgs {
  const float *in_data;    // (B,ann_cap,F,3)
  const float *in_scales;  // (B,ann_cap,F)
  const int *n_in;         // (B)
  float *work;             // (B,ann_cap,F,3) scratch copy of data
  double *score;           // (B,ann_cap) scratch
  int *order;              // (B,ann_cap) scratch
  int *order2;             // (B,ann_cap) scratch
  long long *boxes;        // (B,F,ann_cap,4) scratch
  float *out_data;         // (B,ann_cap,F,3)
  float *out_scales;       // (B,ann_cap,F)
  double *out_score;       // optional (B,ann_cap)
  int *out_source;         // optional (B,ann_cap)
  int *n_out;              // (B)
  int ann_cap, F;
  int do_nms;
  float kp_th, suppression;
  double inst_th;
};
// the argument block of one NMS over a batch; d_ws as sized by nms_ws_bytes
NmsArgs nms_make_args(const float *in_data, const float *in_scales, const int *n_in, float *out_data,
                      float *out_scales, double *out_score, int *out_source, int *n_out, void *d_ws,
                      const pp_shape &s, const pp_caps &c, const pp_params &p);

struct AsmArgs {
  const float *seeds;
  const int *n_seeds;
  const float *rows;        // (B, L, 2, kBinRows, caf_cap) binned candidate store
  const int *start;         // (B, L, 2, nb + 1)
  BinGeo geo;
  const int *nfwd, *nbwd;
  uint8_t *occ_cnt;         // (B, seed_cap) Occupancy.get() counters, one per seed
  int2 *occ_rec;            // (B, seed_cap) the seed's occupancy cell: x = cell x, y = cell y | field << 24
  float *pre_data;
  float *pre_scales;
  int *pre_order;
  int *n_pre;
  int *status;
  int seed_cap, caf_cap, ann_cap, L;
  int hh, hw;
  int stage_floats;   // set by assemble_launch
  int heavy_seeds;    // set by assemble_launch: above this many seeds the (field, band) seed index is used
  const float *init_data, *init_scales;   // initial annotations (optional)
  const int *init_n;
  int init_cap;
  int fuse_nms;       // run nms.Keypoints (or the nms=None copy) for the image right after its seed loop
  NmsArgs nms;
};
int assemble_launch(const AsmArgs &args, const pp_skeleton &sk, int B, int F, const pp_params &p,
                    cudaStream_t st);
int complete_launch(const AsmArgs &args, const float *rows2, const int *start2, const int *nfwd2,
                    const int *nbwd2, float *cmp_data, float *cmp_scales, const pp_skeleton &sk,
                    int B, int F, const pp_params &p, cudaStream_t st);
size_t nms_ws_bytes(const pp_shape &s, const pp_caps &c);
size_t nms_ws_staging_offset(const pp_shape &s, const pp_caps &c);
int nms_launch(const float *in_data, const float *in_scales, const int *n_in, float *out_data,
               float *out_scales, double *out_score, int *out_source, int *n_out, void *d_ws,
               const pp_shape &s, const pp_caps &c, const pp_params &p, cudaStream_t st);

int annotations_inverse_launch(float *ann_data, float *ann_scales, const int *n_ann, float *pre_data,
                               const int *n_pre, const pp_inverse &inv, int B, int F, int ann_cap,
                               cudaStream_t st);
int detections_inverse_launch(float *box, const int *n_det, const pp_inverse &inv, int B, int ann_cap,
                              cudaStream_t st);
// packed results (pack.cu)
int packed_record_words(int F);
size_t packed_header_words(int B);
size_t packed_image_bytes(int F, int ann_cap);      // capacity share of one image (block stride of a chunk)
int pack_launch(const pp_outputs &o, const pp_shape &s, const pp_caps &c, cudaStream_t st);
int head_epilogue_launch(const float *d_x, float *d_out, int B, int nf, int nv, int h, int w, int quad,
                         cudaStream_t st);

}  // namespace pp
