// Internal C++ interface of the tcgen05 GEMM (see gemm.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vipt_b200.h"

namespace vipt {

enum GemmEpilogue {
  EPI_BIAS_F16 = 0,    // out fp16 = acc + bias                      (qkv)
  EPI_GELU_F16 = 1,    // out fp16 = gelu_erf(acc + bias)            (mlp.fc1)
  EPI_RELU_F16 = 2,    // out fp16 = relu(acc + bias)                (head conv + folded BN)
  EPI_BIAS_F32 = 3,    // out fp32 = acc + bias                      (patch embed)
  EPI_RESID_F32 = 4,   // out fp32 = resid + acc + bias              (attn.proj, mlp.fc2; in place allowed)
  // LayerNorm fused into the GEMMs (2-CTA kernel only; gemm2.cu "LayerNorm fold"):
  EPI_RESID_STAT = 5,       // producer: x (fp32, in place) += acc + bias; x16 = fp16(x); per-row partial (sum, sum of
                            // squares) over 32 columns each                     (attn.proj, mlp.fc2)
  EPI_LNFOLD_F16 = 6,       // consumer: out fp16 = rstd * (acc - mean * colsum) + bias, W pre-scaled by the LayerNorm
                            // gain, bias = b + W beta                           (attn.qkv after norm1)
  EPI_LNFOLD_GELU_F16 = 7,  // consumer: the same followed by GELU              (mlp.fc1 after norm2)
  // Deep-prompt stage carried by the GEMMs (gemm2.cu "prompt fold"; vit_ce_prompt.py:268-310):
  EPI_RESID_PROMPT = 8,     // mlp.fc2: v = x + acc + bias in place, x16 = fp16(v), and per row and (tile, column half) the
                            // 20 sums {v, v^2, v.wa_0..7, v.u_0..7, v.b1, -} the prompt block and norm1 need
};

struct GemmParams {
  int M, N, K, G;
  void* out;
  int64_t ldc;               // elements
  int64_t out_group_stride;  // elements added per group g
  int64_t out_group_col;     // column offset added per group g
  const float* bias;         // (G, N) or null
  const float* resid;        // same layout as out (fp32) or null
  // implicit 3x3 convolution (pad 1) over a (B,16,16,C) fp16 feature map: A tiles are fetched by a 4-D TMA
  // map with shifted (x, y) coordinates (out-of-bounds = zero padding); k-block kb = tap * cb_per_tap + cb,
  // K order tap * cin + c; group g reads channels [g * conv_cin_stride, +cin).  0 = plain GEMM.
  int conv_cb_per_tap;
  int conv_cin_stride;
  // LayerNorm fold: statistics travel as (parts, M) float2 = (sum, sum of squares) over a part's columns
  float2* stat_out;          // EPI_RESID_STAT: parts = N / 32, part 2 g + h = the 16-column groups h and h + 2 of the 64 columns g
  const float2* stat_in;     // EPI_LNFOLD_*: partial sums of the K-wide rows of A
  int stat_parts;
  const float* colsum;       // EPI_LNFOLD_*: (N) sum over k of the fp16 weight row
  float ln_eps;
  // prompt fold, EPI_RESID_PROMPT: per-column constants (N, 20) floats = wa_0..7 (conv0_0 * gamma), u_0..7 (conv1x1), b1,
  // 3 pad; streamed with the x slabs
  const float* colw;
  float* part_out;           // EPI_RESID_PROMPT: (2 * N / 256, M, 20) partial sums
  // one extra k-block of 64 after the K / 64 regular ones, from the maps tmAx (M, 64) / tmBx (N, 64): a rank-<=64 update
  // A W^T + Ax Wx^T on the tensor pipe (prompt fold: Ax = the Fovea output f per row, Wx = W'U resp. U)
  int kext;
  // 2-CTA kernel: tiles >= tail_from (the last, partial round of the persistent loop) are cut into tail_split strips
  int tail_from, tail_split;
};

struct GemmOp {
  CUtensorMap tmA, tmB, tmC;   // tmC: output map of the 2-CTA kernel's TMA-store epilogue
  CUtensorMap tmB2;            // 2-CTA kernel: weight map with the box of a tail strip
  CUtensorMap tmAx, tmBx;      // 2-CTA kernel, kext: the extra k-block
  CUtensorMap tmC2;            // EPI_RESID_STAT: fp16 copy of the updated rows (tmC = the fp32 rows, loaded and stored)
  GemmParams p;
  int bn, epi;
};

int make_tmap_f16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box);
int make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box);
int gemm_plan(GemmOp* op, const void* A, int64_t lda, int64_t a_group_stride, const void* W, int64_t ldw,
              int64_t w_group_stride, int G, int M, int N, int K, int bn, int epi, void* out, int64_t ldc,
              int64_t out_group_stride, int64_t out_group_col, const float* bias, const float* resid);
// conv variants: `feat` (B*256, c_total) fp16; A tensor map is built over (c_total, 16, 16, B)
int gemm_plan_conv(GemmOp* op, const void* feat, int c_total, int cin, int cin_stride, const void* W, int G, int B, int N,
                   int bn, int epi, void* out, int64_t ldc, int64_t out_group_col, const float* bias);
int gemm2_plan_conv(GemmOp* op, const void* feat, int c_total, int cin, const void* W, int B, int N, int epi, void* out,
                    int64_t ldc, const float* bias);
int gemm_launch(const GemmOp& op, int num_sms, cudaStream_t st);   // dispatches 2-CTA plans (bn == 512) to gemm2
// 2-CTA (cta_group::2) kernel for large single-group GEMMs with N % 256 == 0 (gemm2.cu)
int gemm2_plan(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int epi,
               void* out, int64_t ldc, const float* bias, int num_clusters = 0);
int gemm2_launch(const GemmOp& op, int num_sms, cudaStream_t st);
// LayerNorm fold, producer: x (M, N) fp32 ld N updated in place, x16 (M, N) fp16, stat_out (N / 32, M)
int gemm2_plan_resid_stat(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x,
                          void* x16, const float* bias, float2* stat_out, int num_clusters = 0);
// prompt fold
int gemm2_plan_resid_prompt(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x,
                            void* x16, const float* bias, const float* colw20, float* part_out);
// extra k-block (any 2-CTA plan with 256-wide tiles): Ax (M, 64) fp16, Wx (N, 64) fp16, both row-major with ld 64
int gemm2_set_kext(GemmOp* op, const void* Ax, const void* Wx);
// LayerNorm fold, consumer: gemm2_plan(.., epi = EPI_LNFOLD_*, ..) followed by this; stat_parts = 0: stat_in (M) holds the
// final (rstd, -mean * rstd) of every row
void gemm2_set_lnfold(GemmOp* op, const float2* stat_in, int stat_parts, const float* colsum, float ln_eps);
int make_tmap_ex(CUtensorMap* out, int dtype_f32, int swizzle_bytes, const void* base, int rank, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box);

}  // namespace vipt
