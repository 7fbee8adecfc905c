// Internal launchers of the non-GEMM kernels (crop.cu, tokens.cu, tokens2.cu, attention.cu, head.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "gemm.h"

namespace vipt {

int launch_crop(const uint8_t* const* frames, const int* frame_hw, const double* boxes, double factor, int S,
                int row_base, int rows_per_seq, void* apatch, int64_t group_stride, double* geom, int* err, int B,
                cudaStream_t st, int* slot_err = nullptr);      // slot_err (B) optional: sticky per-sequence flags
int crop_prepare();      // uploads the normalisation table; called outside stream capture (Engine::finalize, first launch_crop)
int launch_patchify(const float* img, int S, int row_base, int rows_per_seq, void* apatch, int64_t group_stride, int B,
                    cudaStream_t st);

int launch_ln_rows(const float* x_in, int n_in, const void* delta, const int* perm, int perm_stride, float* x_out, void* xn,
                   const float* gamma, const float* beta, int B, int n_out, int lens_t, cudaStream_t st);
int launch_prompt_a(const float* xa, const void* delta, int n_cur, const int* slot_of_pos, const float* xd_src, const float* wg_a,
                    const float* wg_d, const float* sc, float* xd, int B, int lens_t, int lens_full, cudaStream_t st);
int launch_fovea(const float* xd, float smooth, float* F, int B, int lens_t, int lens_full, cudaStream_t st);
int launch_prompt_b(const float* F, const float* w1, const float* b1, float* P, const float* base, const void* delta,
                    float* x, int n_cur, const int* slot_of_pos, const float* pos_embed, const float* gamma1,
                    const float* beta1, void* xn, int B, int lens_t, int lens_full, cudaStream_t st);
int launch_prompt_coop(float* x, const void* delta, int n_cur, const int* gi, float* F, float* X0, const float* wg_a,
                       const float* fold, const float* w1t, const float* b1, const float* gamma1, const float* beta1,
                       void* xn, int B, int lens_t, int lens_full, int num_sms, cudaStream_t st);
// prompt fold (prompt_mid.cu): part (6, B * n_cur, 20) row sums from the fc2 epilogue -> F (raster, in place), f16 (token rows,
// fp16, 64 columns of which 8 are written), rowstat (token rows) float2 (rstd, -mean * rstd) of norm1's input
int launch_prompt_mid(const float* part, int n_cur, const int* gi, float* F, const float* fold, const float* fold2, void* f16,
                      void* rowstat, int B, int lens_t, int lens_full, cudaStream_t st);
int launch_ce_topk(const float* ce_prob, int n_heads, int n_s, int n_keep, const int* gi_in, int* gi_out, int* perm,
                   int* removed, int* slot_of_pos, float* score_out, int lens_x, int B, cudaStream_t st);
int launch_final_norm(const float* x, const void* delta, int n_cur, const int* slot_of_pos, const float* gamma, const float* beta,
                      void* feat16, float* feat32, int B, int lens_t, int lens_full, cudaStream_t st);

int launch_assemble_rgbx(const void* bgr, const void* xbgr, int H, int W, void* out, cudaStream_t st);
int launch_assemble_rgbd_noalloc(const void* bgr, const void* depth16, int H, int W, int depth_clip, void* out,
                                 cudaStream_t st);      // scratch in a static __device__ array (frames.cu)
int launch_assemble_rgbd(const void* bgr, const void* depth16, int H, int W, int depth_clip, void* out, void* workspace,
                         cudaStream_t st);

struct AttnOp {
  CUtensorMap tm;
  int B, N, nkp;
};
int attention_plan(AttnOp* op, const void* qkv, int B, int N);
int attention_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride,
                     cudaStream_t st);     // one CTA per (sequence, head, tile) in latency mode, else attention4
int attention4_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride, int num_sms,
                      cudaStream_t st);    // 16 softmax warps, two threads per row, rotating K / V slots (attention4.cu)

int launch_head_final(const void* h4, const float* w5, const float* b5, const float* window, float* score_map,
                      float* size_map, float* offset_map, float* pred_boxes, double* state, const double* geom,
                      const int* frame_hw, double* out_boxes, float* out_scores, int B, cudaStream_t st);

}  // namespace vipt
