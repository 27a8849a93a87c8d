// stages.cuh -- internal launcher interface between api.cu and the stage files.
#pragma once
#include "common.cuh"

namespace pp {

size_t cifhr_ws_bytes(const pp_shape &s, const pp_caps &c);
int cifhr_launch(const float *d_cif, float *d_hr, void *d_ws, const pp_shape &s, const pp_caps &c,
                 const pp_params &p, int *d_status, cudaStream_t st);
size_t seeds_ws_bytes(const pp_shape &s, const pp_caps &c);
int seeds_launch(const float *d_cif, const float *d_hr, float *d_seeds, int *d_n_seeds, void *d_ws,
                 const pp_shape &s, const pp_caps &c, const pp_params &p, int *d_status,
                 cudaStream_t st);
int caf_scored_launch(const float *d_caf, const float *d_hr, const pp_skeleton &sk, float score_th,
                      float *d_fwd, int *d_nfwd, float *d_bwd, int *d_nbwd, const pp_shape &s,
                      const pp_caps &c, const pp_params &p, int *d_status, cudaStream_t st);
struct AsmArgs {
  const float *seeds;
  const int *n_seeds;
  const float *fwd, *bwd;
  const int *nfwd, *nbwd;
  uint8_t *occ_cnt;
  float *pre_data;
  float *pre_scales;
  int *pre_order;
  int *n_pre;
  int *status;
  int seed_cap, caf_cap, ann_cap, L;
  int hh, hw;
  int stage_floats;   // set by assemble_launch
  const float *init_data, *init_scales;   // initial annotations (optional)
  const int *init_n;
  int init_cap;
};
int assemble_launch(const AsmArgs &args, const pp_skeleton &sk, int B, int F, const pp_params &p,
                    cudaStream_t st);
int complete_launch(const AsmArgs &args, const float *fwd2, const float *bwd2, const int *nfwd2,
                    const int *nbwd2, float *cmp_data, float *cmp_scales, const pp_skeleton &sk,
                    int B, int F, const pp_params &p, cudaStream_t st);
size_t nms_ws_bytes(const pp_shape &s, const pp_caps &c);
size_t nms_ws_staging_offset(const pp_shape &s, const pp_caps &c);
int nms_launch(const float *in_data, const float *in_scales, const int *n_in, float *out_data,
               float *out_scales, double *out_score, int *out_source, int *n_out, void *d_ws,
               const pp_shape &s, const pp_caps &c, const pp_params &p, cudaStream_t st);


}  // namespace pp
