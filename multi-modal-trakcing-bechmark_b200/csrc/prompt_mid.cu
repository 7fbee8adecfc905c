// Deep-prompt stage of layers 1..11 at large batch, the part that is NOT a pass over the 768-wide rows
// (ViPT/lib/models/vipt/vit_ce_prompt.py:268-310; Prompt_block / Fovea :22-71).
//
// "Prompt fold": the 768-wide work of the stage lives in GEMM epilogues (gemm2.cu):
//   mlp.fc2 of block i-1 (EPI_RESID_PROMPT)   v = x + fc2(..) in place, x16 = fp16(v), and per row the sums
//                                              {v, v^2, v.wa_k, v.u_k, v.b1}      (wa = conv0_0 * gamma_a, u = conv1x1)
//   THIS KERNEL, a pair of CTAs per sequence         x0 = conv0_0(prompt_norms[i-1](v)) from the sums, Fovea softmax over the
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
constexpr int kPass = 3;               // 64 row slots per pass: up to 192 rows per CTA (a CTA takes half of <= 320 rows)
constexpr int kMaxRows = kPass * (kThreads / 8);
constexpr int kParts = 6;              // (tile, column half) partial sums of a 768-wide row
constexpr float kLnEps = 1e-6f;
// fold (prompt_coop.cu layout) and fold2 (this kernel)
constexpr int kCA = 0, kA = 8, kA0 = 72, kCD = 80, kG = 88, kH = 152, kC0 = 160, kSmooth = 161, kFoldFloats = 168;
constexpr int kSwa = 0, kSu = 8, kSb = 16, kG9 = 24, kFold2Floats = 112;
}  // namespace pm

namespace {
__device__ __forceinline__ uint32_t pm_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void pm_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float pm_peer_f32(const float* local, uint32_t peer) {      // the peer CTA's copy of a shared variable
  uint32_t ra;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local)), "r"(peer));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(ra) : "memory");
  return v;
}
}  // namespace

// Two CTAs (one cluster) per sequence, each with half of the sequence's rows; the Fovea softmax statistics (per domain
// and channel: maximum, then sum of exponentials) are the only cross-row quantities and are exchanged through
// distributed shared memory.  (One CTA per sequence left 84 of 148 SMs idle and took 15 us per launch, 11 launches per
// step.)  thread = (row slot, channel k): the constants of channel k live in registers.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(pm::kThreads, 1)
prompt_mid_kernel(const float* __restrict__ part, int M, int n_cur, const int* __restrict__ gi, float* __restrict__ F,
                  const float* __restrict__ fold, const float* __restrict__ fold2, __half* __restrict__ f16,
                  float2* __restrict__ rowstat, int lens_t, int lens_full) {
  using namespace pm;
  __shared__ float s_fold[kFoldFloats], s_fold2[kFold2Floats];
  __shared__ float s_red[2][kWarps * 8 * 2];
  __shared__ float s_xmax[16], s_xsum[16];       // this CTA's maxima / sums per (domain, channel): read by the peer
  __shared__ float4 s_sum[kMaxRows * 5];         // the 20 sums of this CTA's rows
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = pm_ctarank();
  const int b = blockIdx.x >> 1;
  for (int i = tid; i < kFoldFloats; i += kThreads) s_fold[i] = fold[i];
  for (int i = tid; i < kFold2Floats; i += kThreads) s_fold2[i] = fold2[i];
  pdl_trigger();
  pdl_wait();
  __syncthreads();
  const int lens_x = lens_full - lens_t;
  const int k = tid & 7, slot = tid >> 3;        // channel, row slot
  const size_t row_base = (size_t)b * n_cur;
  const int half_rows = (n_cur + 1) >> 1;
  const int r_lo = (int)rank * half_rows, r_hi = min(n_cur, r_lo + half_rows);      // this CTA's rows [r_lo, r_hi)
  // constants of channel k
  const float smooth = s_fold[kSmooth], c_a = s_fold[kCA + k], c_d = s_fold[kCD + k], a0 = s_fold[kA0 + k];
  const float h2 = 2.0f * s_fold[kH + k], c0 = s_fold[kC0], swa = s_fold2[kSwa + k], su = s_fold2[kSu + k];
  float gk[8], ak[8], g9[9];
#pragma unroll
  for (int j = 0; j < 8; ++j) { gk[j] = s_fold[kG + k * 8 + j]; ak[j] = s_fold[kA + k * 8 + j]; }
#pragma unroll
  for (int j = 0; j < 9; ++j) g9[j] = s_fold2[kG9 + k * 9 + j];

  // The 6 partial sums of every row are added (fixed order) into shared memory first, one thread per (row, float4): 6
  // independent 16-byte loads each.  (Read per (row, channel) thread, 150 dependent scalar loads put the kernel at 18 us.)
  {
    const float4* p4 = reinterpret_cast<const float4*>(part);
    for (int idx = tid; idx < (r_hi - r_lo) * 5; idx += kThreads) {
      const int rl = idx / 5, j = idx - rl * 5;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      float4 t[kParts];
#pragma unroll
      for (int q = 0; q < kParts; ++q) t[q] = __ldcg(p4 + ((size_t)q * M + row_base + r_lo + rl) * 5 + j);
#pragma unroll
      for (int q = 0; q < kParts; ++q) { acc.x += t[q].x; acc.y += t[q].y; acc.z += t[q].z; acc.w += t[q].w; }
      s_sum[rl * 5 + j] = acc;
    }
  }
  // raster position and previous Fovea output of this thread's rows: in flight during the softmax statistics
  int posv[kPass];
  float4 fp0[kPass], fp1[kPass];
#pragma unroll
  for (int i = 0; i < kPass; ++i) {
    const int r = r_lo + slot + i * (kThreads / 8);
    posv[i] = 0;
    fp0[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    fp1[i] = fp0[i];
    if (r < r_hi) {
      posv[i] = r < lens_t ? r : lens_t + gi[b * lens_x + (r - lens_t)];
      fp0[i] = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + posv[i]) * 8);
      fp1[i] = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + posv[i]) * 8 + 4);
    }
  }
  __syncthreads();
  float x0[kPass], sv[kPass], sq[kPass], dk[kPass], ev[kPass];
  float m[2] = {-INFINITY, -INFINITY};
#pragma unroll
  for (int i = 0; i < kPass; ++i) {
    const int rl = slot + i * (kThreads / 8), r = r_lo + rl;
    x0[i] = 0.f; sv[i] = 0.f; sq[i] = 0.f; dk[i] = 0.f; ev[i] = 0.f;
    if (r < r_hi) {
      const float* pr = reinterpret_cast<const float*>(s_sum + rl * 5);
      sv[i] = pr[0];
      sq[i] = pr[1];
      const float a = pr[2 + k];
      dk[i] = pr[10 + k];
      ev[i] = pr[18];
      const float mean = sv[i] * (1.0f / 768.0f);
      const float rstd = rsqrtf(fmaxf(fmaf(-mean, mean, sq[i] * (1.0f / 768.0f)), 0.f) + kLnEps);
      x0[i] = fmaf(rstd, fmaf(-mean, swa, a), c_a);
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
  if (tid < 16) {
    const int dom = tid >> 3, kk = tid & 7;
    float Mx = -INFINITY;
    for (int w = 0; w < kWarps; ++w) Mx = fmaxf(Mx, s_red[0][(w * 8 + kk) * 2 + dom]);
    s_xmax[dom * 8 + kk] = Mx;
  }
  pm_cluster_sync();                             // (also a CTA barrier) both CTAs' maxima are written
  const int n_removed_x = lens_x - (n_cur - lens_t);
  float mx[2];
  mx[0] = fmaxf(s_xmax[k], pm_peer_f32(&s_xmax[k], rank ^ 1));
  mx[1] = fmaxf(s_xmax[8 + k], pm_peer_f32(&s_xmax[8 + k], rank ^ 1));
  if (n_removed_x > 0) mx[1] = fmaxf(mx[1], smooth * c_a);
  float sm[2] = {0.f, 0.f};
#pragma unroll
  for (int i = 0; i < kPass; ++i) {
    const int r = r_lo + slot + i * (kThreads / 8);
    if (r < r_hi) {
      if (r < lens_t) sm[0] += expf(x0[i] * smooth - mx[0]);
      else sm[1] += expf(x0[i] * smooth - mx[1]);
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
    s_xsum[dom * 8 + kk] = S;
  }
  pm_cluster_sync();
  float inv[2];
  {
    // rank 0's sum first, then rank 1's, then the removed positions: the same order in both CTAs
    const float l0 = s_xsum[k], p0 = pm_peer_f32(&s_xsum[k], rank ^ 1);
    const float l1 = s_xsum[8 + k], p1 = pm_peer_f32(&s_xsum[8 + k], rank ^ 1);
    const float S0 = rank == 0 ? l0 + p0 : p0 + l0;
    float S1 = rank == 0 ? l1 + p1 : p1 + l1;
    if (n_removed_x > 0) S1 += (float)n_removed_x * expf(smooth * c_a - mx[1]);
    inv[0] = 1.0f / S0;
    inv[1] = 1.0f / S1;
  }

  // per row: x1 from the previous Fovea output at the row's raster position, f, the statistics of x = v + U f + b1
#pragma unroll
  for (int i = 0; i < kPass; ++i) {
    const int r = r_lo + slot + i * (kThreads / 8);
    const bool live = r < r_hi;
    float fk = 0.f, varp = 0.f, a = a0;
    const int pos = posv[i];
    if (live) {
      const float4 f0 = fp0[i], f1 = fp1[i];
      // var(f) = f^T G f + 2 h.f + c0 is the same for the 8 channels of a row: lane k contributes f_k (G_k.f + 2 h_k),
      // the 8 lanes of the row add up below (every lane of the warp takes part in the shuffles)
      const float g = (gk[0] * f0.x + gk[2] * f0.z + gk[4] * f1.x + gk[6] * f1.z) +
                      (gk[1] * f0.y + gk[3] * f0.w + gk[5] * f1.y + gk[7] * f1.w);
      const float4 fh = (k & 4) ? f1 : f0;             // f[k] without a dynamically indexed register array
      const float fsel = (k & 2) ? ((k & 1) ? fh.w : fh.z) : ((k & 1) ? fh.y : fh.x);
      varp = fsel * (g + h2);
      a += ak[0] * f0.x + ak[1] * f0.y + ak[2] * f0.z + ak[3] * f0.w + ak[4] * f1.x + ak[5] * f1.y + ak[6] * f1.z + ak[7] * f1.w;
    }
#pragma unroll
    for (int o = 1; o <= 4; o <<= 1) varp += __shfl_xor_sync(0xffffffffu, varp, o);
    if (live) {
      const int dom = r < lens_t ? 0 : 1;
      const float x1 = a * rsqrtf(fmaxf(c0 + varp, 0.f) + kLnEps) + c_d;
      fk = expf(x0[i] * smooth - mx[dom]) * inv[dom] * x0[i] + x1;
    }
    __syncwarp();                                // (the 8 lanes of a row read F[pos] long before: registers)
    if (live) {
      F[((size_t)b * lens_full + pos) * 8 + k] = fk;
      // extra k-block of qkv / proj: columns 0..7 = f, column 8 = 1 (qkv: times W'b1; proj: times 0), 9..63 stay zero
      f16[(row_base + r) * 64 + k] = __float2half_rn(fk);
      if (k == 0) f16[(row_base + r) * 64 + 8] = __float2half_rn(1.0f);
    }
    // sum over the 8 channels of the row (8 consecutive lanes): sum(x) and sum(x^2) of x = v + U f + b1
    //   sum x   = sum v + f.su + sb
    //   sum x^2 = sum v^2 + 2 (f.d + e) + [f 1] G9 [f 1]^T        d_k = v.u_k, e = v.b1, G9 = [U b1]^T [U b1]
    float gq = g9[8] * 2.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) gq = fmaf(g9[j], __shfl_sync(0xffffffffu, fk, (lane & ~7) + j), gq);
    float t1 = fk * su;                          // -> f.su
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
  pm_cluster_sync();                             // the peer may still be reading this CTA's statistics
}

int launch_prompt_mid(const float* part, int n_cur, const int* gi, float* F, const float* fold, const float* fold2, void* f16,
                      void* rowstat, int B, int lens_t, int lens_full, cudaStream_t st) {
  if (!part || !gi || !F || !fold || !fold2 || !f16 || !rowstat || n_cur < lens_t || n_cur > lens_full ||
      (lens_full + 1) / 2 > pm::kMaxRows)
    return VIPT_ERR_ARG;
  return launch_pdl(prompt_mid_kernel, dim3(2 * B), dim3(pm::kThreads), 0, st, part, B * n_cur, n_cur, gi, F, fold, fold2,
                    reinterpret_cast<__half*>(f16), reinterpret_cast<float2*>(rowstat), lens_t, lens_full) == cudaSuccess
             ? VIPT_OK
             : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
