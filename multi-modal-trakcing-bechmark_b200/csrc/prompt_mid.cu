// Deep-prompt stage of layers 1..11 at large batch, the part that is NOT a pass over the 768-wide rows
// (ViPT/lib/models/vipt/vit_ce_prompt.py:268-310; Prompt_block / Fovea :22-71).
//
// "Prompt fold": the 768-wide work of the stage lives in GEMM epilogues (gemm2.cu):
//   mlp.fc2 of block i-1 (EPI_RESID_PROMPT)   v = x + fc2(..) in place, x16 = fp16(v), and per row the sums
//                                              {v, v^2, v.wa_k, v.u_k, v.b1}      (wa = conv0_0 * gamma_a, u = conv1x1)
//   THIS KERNEL, one CTA per sequence          x0 = conv0_0(prompt_norms[i-1](v)) from the sums, Fovea softmax over the
//                                              positions of the sequence, x1 from the previous Fovea output by the rank-8
//                                              algebra of prompt_coop.cu, f = softmax * x0 + x1; then the LayerNorm
//                                              statistics of x = v + U f + b1 from the sums (no second pass over v)
//   attn.qkv of block i (LayerNorm fold + kext) norm1(x) Wqkv^T from x16 = fp16(v); the rank-8 term f (W'U)^T is one extra
//                                              k-block of the GEMM (A = f16 rows written here, 8 of 64 columns used)
//   attn.proj of block i (EPI_RESID_STAT + kext) x <- v + U f + b1 + proj(..), U f again as the extra k-block
// so the residual stream is read and written once per GEMM that changes it and never by a prompt kernel.
// Removed search positions (candidate elimination) enter the softmax normaliser as n_removed * exp(smooth * c_a - max),
// exactly as in prompt_coop.cu.
#include <cuda_fp16.h>

#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

namespace pm {
constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kParts = 6;              // (tile, column half) partial sums of a 768-wide row
constexpr int kPartFloats = 20;
constexpr float kLnEps = 1e-6f;
// fold (prompt_coop.cu layout) and fold2 (this kernel)
constexpr int kCA = 0, kA = 8, kA0 = 72, kCD = 80, kG = 88, kH = 152, kC0 = 160, kSmooth = 161, kFoldFloats = 168;
constexpr int kSwa = 0, kSu = 8, kSb = 16, kG9 = 24, kFold2Floats = 112;
}  // namespace pm

__global__ void __launch_bounds__(pm::kThreads, 1)
prompt_mid_kernel(const float* __restrict__ part, int M, int n_cur, const int* __restrict__ gi, float* __restrict__ F,
                  const float* __restrict__ fold, const float* __restrict__ fold2, __half* __restrict__ f16,
                  float2* __restrict__ rowstat, int lens_t, int lens_full) {
  using namespace pm;
  __shared__ float s_fold[kFoldFloats], s_fold2[kFold2Floats];
  __shared__ float s_red[2][kWarps * 8 * 2];
  __shared__ float s_ms[32];                     // [domain][channel][max, 1 / sum]
  __shared__ float4 s_sum[320 * 5];              // the 20 sums of every row of the sequence
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x;
  for (int i = tid; i < kFoldFloats; i += kThreads) s_fold[i] = fold[i];
  for (int i = tid; i < kFold2Floats; i += kThreads) s_fold2[i] = fold2[i];
  pdl_trigger();
  pdl_wait();
  __syncthreads();
  const float smooth = s_fold[kSmooth];
  const int lens_x = lens_full - lens_t;
  const int k = tid & 7, slot = tid >> 3;        // channel, row slot (64 per pass, 5 passes cover 320 rows)
  const size_t row_base = (size_t)b * n_cur;

  // The 6 partial sums of every row are added (fixed order) into shared memory first, one thread per (row, float4): 6
  // independent 16-byte loads each.  (Read per (row, channel) thread, 150 dependent scalar loads put the kernel at 18 us.)
  {
    const float4* p4 = reinterpret_cast<const float4*>(part);
    for (int idx = tid; idx < n_cur * 5; idx += kThreads) {
      const int r = idx / 5, j = idx - r * 5;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 t[kParts];
#pragma unroll
      for (int q = 0; q < kParts; ++q) t[q] = __ldcg(p4 + ((size_t)q * M + row_base + r) * 5 + j);
#pragma unroll
      for (int q = 0; q < kParts; ++q) { acc.x += t[q].x; acc.y += t[q].y; acc.z += t[q].z; acc.w += t[q].w; }
      s_sum[r * 5 + j] = acc;
    }
  }
  __syncthreads();
  // raster position and previous Fovea output of this thread's rows: in flight during the softmax statistics
  int posv[5];
  float4 fp0[5], fp1[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int r = slot + i * (kThreads / 8);
    posv[i] = 0;
    fp0[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    fp1[i] = fp0[i];
    if (r < n_cur) {
      posv[i] = r < lens_t ? r : lens_t + gi[b * lens_x + (r - lens_t)];
      fp0[i] = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + posv[i]) * 8);
      fp1[i] = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + posv[i]) * 8 + 4);
    }
  }
  float x0[5], sv[5], sq[5], dk[5], ev[5];
  float m[2] = {-INFINITY, -INFINITY};
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int r = slot + i * (kThreads / 8);
    x0[i] = 0.f; sv[i] = 0.f; sq[i] = 0.f; dk[i] = 0.f; ev[i] = 0.f;
    if (r < n_cur) {
      const float* pr = reinterpret_cast<const float*>(s_sum + r * 5);
      sv[i] = pr[0];
      sq[i] = pr[1];
      const float a = pr[2 + k];
      dk[i] = pr[10 + k];
      ev[i] = pr[18];
      const float mean = sv[i] * (1.0f / 768.0f);
      const float rstd = rsqrtf(fmaxf(fmaf(-mean, mean, sq[i] * (1.0f / 768.0f)), 0.f) + kLnEps);
      x0[i] = fmaf(rstd, fmaf(-mean, s_fold2[kSwa + k], a), s_fold[kCA + k]);
      const float l = x0[i] * smooth;
      if (r < lens_t) m[0] = fmaxf(m[0], l); else m[1] = fmaxf(m[1], l);
    }
  }
  // Fovea softmax statistics per (domain, channel): template rows and search rows are separate softmaxes
#pragma unroll
  for (int o = 8; o <= 16; o <<= 1) {
    m[0] = fmaxf(m[0], __shfl_xor_sync(0xffffffffu, m[0], o));
    m[1] = fmaxf(m[1], __shfl_xor_sync(0xffffffffu, m[1], o));
  }
  if (lane < 8) { s_red[0][(warp * 8 + lane) * 2] = m[0]; s_red[0][(warp * 8 + lane) * 2 + 1] = m[1]; }
  __syncthreads();
  const int n_removed_x = lens_x - (n_cur - lens_t);
  if (tid < 16) {
    const int dom = tid >> 3, kk = tid & 7;
    float Mx = -INFINITY;
    for (int w = 0; w < kWarps; ++w) Mx = fmaxf(Mx, s_red[0][(w * 8 + kk) * 2 + dom]);
    if (dom == 1 && n_removed_x > 0) Mx = fmaxf(Mx, smooth * s_fold[kCA + kk]);
    s_ms[(dom * 8 + kk) * 2] = Mx;
  }
  __syncthreads();
  float sm[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int r = slot + i * (kThreads / 8);
    if (r < n_cur) {
      if (r < lens_t) sm[0] += expf(x0[i] * smooth - s_ms[k * 2]);
      else sm[1] += expf(x0[i] * smooth - s_ms[(8 + k) * 2]);
    }
  }
#pragma unroll
  for (int o = 8; o <= 16; o <<= 1) {
    sm[0] += __shfl_xor_sync(0xffffffffu, sm[0], o);
    sm[1] += __shfl_xor_sync(0xffffffffu, sm[1], o);
  }
  if (lane < 8) { s_red[1][(warp * 8 + lane) * 2] = sm[0]; s_red[1][(warp * 8 + lane) * 2 + 1] = sm[1]; }
  __syncthreads();
  if (tid < 16) {
    const int dom = tid >> 3, kk = tid & 7;
    float S = 0.f;
    for (int w = 0; w < kWarps; ++w) S += s_red[1][(w * 8 + kk) * 2 + dom];
    if (dom == 1 && n_removed_x > 0) S += (float)n_removed_x * expf(smooth * s_fold[kCA + kk] - s_ms[(dom * 8 + kk) * 2]);
    s_ms[(dom * 8 + kk) * 2 + 1] = 1.0f / S;
  }
  __syncthreads();

  // per row: x1 from the previous Fovea output at the row's raster position, f, the statistics of x = v + U f + b1
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int r = slot + i * (kThreads / 8);
    const bool live = r < n_cur;
    float fk = 0.f, varp = 0.f, a = 0.f;
    int pos = 0;
    if (live) {
      pos = posv[i];
      const float4 f0 = fp0[i], f1 = fp1[i];
      const float f[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
      // var(f) = f^T G f + 2 h.f + c0 is the same for the 8 channels of a row: lane k contributes f_k (G_k.f + 2 h_k),
      // the 8 lanes of the row add up below (every lane of the warp takes part in the shuffles)
      float g = 0.f, g1 = 0.f;
      a = s_fold[kA0 + k];
#pragma unroll
      for (int qq = 0; qq < 8; qq += 2) {
        g += s_fold[kG + k * 8 + qq] * f[qq];
        g1 += s_fold[kG + k * 8 + qq + 1] * f[qq + 1];
      }
      const float4 fh = (k & 4) ? f1 : f0;             // f[k] without a dynamically indexed register array
      const float fsel = (k & 2) ? ((k & 1) ? fh.w : fh.z) : ((k & 1) ? fh.y : fh.x);
      varp = fsel * (g + g1 + 2.0f * s_fold[kH + k]);
#pragma unroll
      for (int p = 0; p < 8; ++p) a += s_fold[kA + k * 8 + p] * f[p];
    }
#pragma unroll
    for (int o = 1; o <= 4; o <<= 1) varp += __shfl_xor_sync(0xffffffffu, varp, o);
    if (live) {
      const int dom = r < lens_t ? 0 : 1;
      const float var = s_fold[kC0] + varp;
      const float x1 = a * rsqrtf(fmaxf(var, 0.f) + kLnEps) + s_fold[kCD + k];
      fk = expf(x0[i] * smooth - s_ms[(dom * 8 + k) * 2]) * s_ms[(dom * 8 + k) * 2 + 1] * x0[i] + x1;
    }
    __syncwarp();                                // the 8 lanes of a row have read F[pos] before anyone overwrites it
    if (live) {
      F[((size_t)b * lens_full + pos) * 8 + k] = fk;
      // extra k-block of qkv / proj: columns 0..7 = f, column 8 = 1 (qkv: times W'b1; proj: times 0), 9..63 stay zero
      f16[(row_base + r) * 64 + k] = __float2half_rn(fk);
      if (k == 0) f16[(row_base + r) * 64 + 8] = __float2half_rn(1.0f);
    }
    // sum over the 8 channels of the row (8 consecutive lanes): sum(x) and sum(x^2) of x = v + U f + b1
    //   sum x   = sum v + f.su + sb
    //   sum x^2 = sum v^2 + 2 (f.d + e) + [f 1] G9 [f 1]^T        d_k = v.u_k, e = v.b1, G9 = [U b1]^T [U b1]
    float fj[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) fj[j] = __shfl_sync(0xffffffffu, fk, (lane & ~7) + j);
    float gq = s_fold2[kG9 + k * 9 + 8] * 2.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) gq = fmaf(s_fold2[kG9 + k * 9 + j], fj[j], gq);
    float t1 = fk * s_fold2[kSu + k];            // -> f.su
    float t2 = fk * (2.0f * dk[i] + gq);         // -> 2 f.d + f^T G f + 2 f.g8
#pragma unroll
    for (int o = 1; o <= 4; o <<= 1) {
      t1 += __shfl_xor_sync(0xffffffffu, t1, o);
      t2 += __shfl_xor_sync(0xffffffffu, t2, o);
    }
    if (live && k == 0) {
      const float sx = sv[i] + t1 + s_fold2[kSb];
      const float sxx = sq[i] + t2 + 2.0f * ev[i] + s_fold2[kG9 + 8 * 9 + 8];
      const float mean = sx * (1.0f / 768.0f);
      const float rs = rsqrtf(fmaxf(fmaf(-mean, mean, sxx * (1.0f / 768.0f)), 0.f) + kLnEps);
      rowstat[row_base + r] = make_float2(rs, -mean * rs);
    }
  }
}

int launch_prompt_mid(const float* part, int n_cur, const int* gi, float* F, const float* fold, const float* fold2, void* f16,
                      void* rowstat, int B, int lens_t, int lens_full, cudaStream_t st) {
  if (!part || !gi || !F || !fold || !fold2 || !f16 || !rowstat || n_cur < lens_t || n_cur > lens_full ||
      lens_full > 5 * (pm::kThreads / 8) || lens_full > 320)
    return VIPT_ERR_ARG;
  return launch_pdl(prompt_mid_kernel, dim3(B), dim3(pm::kThreads), 0, st, part, B * n_cur, n_cur, gi, F, fold, fold2,
                    reinterpret_cast<__half*>(f16), reinterpret_cast<float2*>(rowstat), lens_t, lens_full) == cudaSuccess
             ? VIPT_OK
             : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
