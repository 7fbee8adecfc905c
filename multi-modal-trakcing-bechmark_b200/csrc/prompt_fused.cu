// Deep-prompt stage of layers 1..11 (ViPT/lib/models/vipt/vit_ce_prompt.py:268-310) as ONE kernel:
//   x      <- x + delta                       (delta = fp16 output of the previous block's mlp.fc2)
//   x0     = conv0_0(prompt_norms[i-1](x))     per kept token          (Prompt_block, :63-65)
//   x1     = conv0_1(prompt_norms[i](P_prev))  per kept position       (:66)
//   F      = softmax_over_positions(smooth * x0) * x0 + x1             (Fovea, :37-47)
//   x      <- x + conv1x1(F)                   kept positions, current token order (:308-310)
//   xn     = blocks[i].norm1(x) in fp16        feeds the qkv GEMM      (attn_blocks.py:94)
//
// Each token row (768 fp32 + 768 fp16) is read from HBM exactly once and stays in registers across the
// Fovea softmax, whose normaliser spans all positions of a sequence: a thread-block CLUSTER of 8 CTAs
// owns one sequence (<= 40 rows per CTA, 16 warps, <= 3 rows per warp) and exchanges the per-channel softmax
// statistics of its rows through distributed shared memory.
//
// The prompt stream itself is never materialised.  P_prev = conv1x1_{i-1}(F_prev) has rank 8, so
//   conv0_1(LN(P_prev[t])) = (A f + a0) * rstd(f) + c_d,  var(f) = f^T G f + 2 h.f + c0,  f = F_prev[t]
// with 8x8 / 8-vectors folded once on the host in double (engine.cu); only F (8 floats per position)
// travels between layers.  Removed search positions are zero rows before prompt_norms[i-1]
// (:276-288), i.e. x0 = conv0_0(beta) = c_a there: they enter the softmax normaliser as
// n_removed * exp(smooth * c_a - max) and nothing else depends on them.
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

namespace pf {
constexpr int D = 768;
constexpr int kCluster = 8;
constexpr int kThreads = 512;          // 16 warps: enough warps per scheduler to hide the per-row dependency chains
constexpr int kWarps = kThreads / 32;
constexpr float kLnEps = 1e-6f;
// fold layout (floats), see PromptFoldHost in engine.cu
constexpr int kCA = 0, kA = 8, kA0 = 72, kCD = 80, kG = 88, kH = 152, kC0 = 160, kSmooth = 161, kFoldFloats = 168;
constexpr int kSmemBytes = 2 * 8 * D * 4 + kFoldFloats * 4 + 2 * 8 * 2 * 4 /*stat*/ + 2 * 8 * 2 * 4 /*M, invS*/ + 128 * 4 /*partials*/ +
                           kWarps * 3 * (8 + 8 + 1) * 4 /* x0, x1, pos */ + 64;
}  // namespace pf

__device__ __forceinline__ float pf_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ uint32_t pf_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ float pf_ld_dsmem(const float* local, uint32_t rank) {
  uint32_t a;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(smem_u32(local)), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory");
  return v;
}

template <int RPW>
__global__ void __cluster_dims__(pf::kCluster, 1, 1) __launch_bounds__(pf::kThreads, 1)
prompt_fused_kernel(float* __restrict__ x, const op_t* __restrict__ delta, int n_cur, const int* __restrict__ gi,
                    float* __restrict__ F, const float* __restrict__ wg_a, const float* __restrict__ fold,
                    const float* __restrict__ w1t, const float* __restrict__ b1, const float* __restrict__ gamma1,
                    const float* __restrict__ beta1, op_t* __restrict__ xn, int lens_t, int lens_full) {
  using namespace pf;
  extern __shared__ float4 pf_smem4[];
  float4* s_wa = pf_smem4;                       // [8][192] float4: conv0_0.weight * gamma_a
  float4* s_w1 = s_wa + 8 * 192;                 // [8][192] float4: conv1x1.weight transposed (k-major)
  float* s_fold = reinterpret_cast<float*>(s_w1 + 8 * 192);
  float* s_stat = s_fold + kFoldFloats;          // [2 domains][8][2]: local max, local sum of exp
  float* s_ms = s_stat + 32;                     // [2][8][2]: global max, 1 / global sum
  float* s_wst = s_ms + 32;                      // [4 warps][2][8][2] partial statistics
  float* s_x0 = s_wst + 128;                     // [kWarps * RPW][8]
  float* s_x1 = s_x0 + kWarps * 3 * 8;           // [kWarps * RPW][8]
  int* s_pos = reinterpret_cast<int*>(s_x1 + kWarps * 3 * 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = pf_cluster_rank();
  const int b = blockIdx.x / kCluster;
  const int rows_per_cta = (n_cur + kCluster - 1) / kCluster;
  const int lens_x = lens_full - lens_t;

  for (int i = tid; i < 8 * 192; i += kThreads) {
    s_wa[i] = reinterpret_cast<const float4*>(wg_a)[i];
    s_w1[i] = reinterpret_cast<const float4*>(w1t)[i];
  }
  if (tid < kFoldFloats) s_fold[tid] = fold[tid];

  // ---- phase 1: rows -> registers (x + delta), LayerNorm statistics, x0 per row
  float4 v[RPW][6];
  int slot[RPW];
#pragma unroll
  for (int i = 0; i < RPW; ++i) {
    const int local = warp + i * kWarps;
    const int r = (int)rank * rows_per_cta + local;
    slot[i] = (local < rows_per_cta && r < n_cur) ? r : -1;
    if (slot[i] >= 0) {
      const float4* xr = reinterpret_cast<const float4*>(x + ((size_t)b * n_cur + r) * D);
      const uint2* dr = reinterpret_cast<const uint2*>(delta + ((size_t)b * n_cur + r) * D);
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        v[i][j] = xr[lane + 32 * j];
        const uint2 u = dr[lane + 32 * j];
        const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
        const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
        v[i][j].x += d0.x; v[i][j].y += d0.y; v[i][j].z += d1.x; v[i][j].w += d1.y;
      }
      if (lane == 0) s_pos[warp * RPW + i] = r < lens_t ? r : lens_t + gi[b * lens_x + (r - lens_t)];
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) v[i][j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (lane == 0) s_pos[warp * RPW + i] = -1;
    }
  }
  __syncthreads();      // weights staged, s_pos visible
  const float smooth = s_fold[kSmooth];
  {
    float mean[RPW], q[RPW], dot[RPW][8];
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 6; ++j) s += v[i][j].x + v[i][j].y + v[i][j].z + v[i][j].w;
      mean[i] = pf_warp_sum(s) * (1.0f / D);
      q[i] = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) dot[i][k] = 0.f;
    }
    // weights outer, rows inner: one shared-memory float4 feeds RPW rows
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      float4 a[RPW];
#pragma unroll
      for (int i = 0; i < RPW; ++i) {
        a[i] = make_float4(v[i][j].x - mean[i], v[i][j].y - mean[i], v[i][j].z - mean[i], v[i][j].w - mean[i]);
        q[i] += a[i].x * a[i].x + a[i].y * a[i].y + a[i].z * a[i].z + a[i].w * a[i].w;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 w = s_wa[k * 192 + lane + 32 * j];
#pragma unroll
        for (int i = 0; i < RPW; ++i) dot[i][k] += w.x * a[i].x + w.y * a[i].y + w.z * a[i].z + w.w * a[i].w;
      }
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      const float rstd = rsqrtf(pf_warp_sum(q[i]) * (1.0f / D) + kLnEps);
      // transposed butterfly: 8 sums over 32 lanes in 9 shuffles; lane l ends with sum #(l >> 2)
      float a4[4], a2[2];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool hi = lane & 16;
        const float send = hi ? dot[i][k] : dot[i][k + 4];
        const float keep = hi ? dot[i][k + 4] : dot[i][k];
        a4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const bool hi = lane & 8;
        const float send = hi ? a4[k] : a4[k + 2];
        const float keep = hi ? a4[k + 2] : a4[k];
        a2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
      }
      float c;
      {
        const bool hi = lane & 4;
        const float send = hi ? a2[0] : a2[1];
        const float keep = hi ? a2[1] : a2[0];
        c = keep + __shfl_xor_sync(0xffffffffu, send, 4);
      }
      c += __shfl_xor_sync(0xffffffffu, c, 2);
      c += __shfl_xor_sync(0xffffffffu, c, 1);
      if ((lane & 3) == 0) s_x0[(warp * RPW + i) * 8 + (lane >> 2)] = c * rstd + s_fold[kCA + (lane >> 2)];
    }
  }
  __syncthreads();
  // stream d (rank-8 algebra on the previous layer's Fovea output): thread = (row, channel)
  for (int idx = tid; idx < kWarps * RPW * 8; idx += kThreads) {
    const int li = idx >> 3, k = idx & 7;
    const int pos = s_pos[li];
    if (pos < 0) continue;
    const float4 f0 = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + pos) * 8);
    const float4 f1 = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + pos) * 8 + 4);
    const float f[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
    float var = s_fold[kC0], a = s_fold[kA0 + k];
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      float g = 0.f;
#pragma unroll
      for (int qq = 0; qq < 8; ++qq) g += s_fold[kG + p * 8 + qq] * f[qq];
      var += f[p] * (g + 2.0f * s_fold[kH + p]);
      a += s_fold[kA + k * 8 + p] * f[p];
    }
    s_x1[idx] = a * rsqrtf(fmaxf(var, 0.f) + kLnEps) + s_fold[kCD + k];
  }
  // local softmax statistics: warps 0..3, thread = (channel k = lane & 7, row group g = 4 * warp + (lane >> 3));
  // partial (max, sum) per domain, combined over the 4 groups of a warp by shuffles, then over warps.
  if (warp < 4) {
    const int k = lane & 7, g = warp * 4 + (lane >> 3);
    float m[2] = {-INFINITY, -INFINITY}, sm[2] = {0.f, 0.f};
    for (int li = g; li < kWarps * RPW; li += 16) {
      const int pos = s_pos[li];
      if (pos >= 0) m[pos < lens_t ? 0 : 1] = fmaxf(m[pos < lens_t ? 0 : 1], s_x0[li * 8 + k] * smooth);
    }
    for (int li = g; li < kWarps * RPW; li += 16) {
      const int pos = s_pos[li];
      if (pos >= 0) sm[pos < lens_t ? 0 : 1] += expf(s_x0[li * 8 + k] * smooth - m[pos < lens_t ? 0 : 1]);
    }
#pragma unroll
    for (int o = 8; o <= 16; o <<= 1) {
#pragma unroll
      for (int d = 0; d < 2; ++d) {
        const float om = __shfl_xor_sync(0xffffffffu, m[d], o), os = __shfl_xor_sync(0xffffffffu, sm[d], o);
        const float M = fmaxf(m[d], om);
        sm[d] = (m[d] == -INFINITY ? 0.f : sm[d] * expf(m[d] - M)) + (om == -INFINITY ? 0.f : os * expf(om - M));
        m[d] = M;
      }
    }
    if (lane < 8) {
#pragma unroll
      for (int d = 0; d < 2; ++d) { s_wst[((warp * 2 + d) * 8 + k) * 2] = m[d]; s_wst[((warp * 2 + d) * 8 + k) * 2 + 1] = sm[d]; }
    }
  }
  __syncthreads();
  if (tid < 16) {
    const int dom = tid >> 3, k = tid & 7;
    float M = -INFINITY, S = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) M = fmaxf(M, s_wst[((w * 2 + dom) * 8 + k) * 2]);
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float mw = s_wst[((w * 2 + dom) * 8 + k) * 2];
      if (mw != -INFINITY) S += s_wst[((w * 2 + dom) * 8 + k) * 2 + 1] * expf(mw - M);
    }
    s_stat[tid * 2] = M;
    s_stat[tid * 2 + 1] = S;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (tid < 16) {
    const int dom = tid >> 3, k = tid & 7;
    float pm[kCluster], ps[kCluster];
    float M = -INFINITY;
#pragma unroll
    for (int c = 0; c < kCluster; ++c) {
      pm[c] = pf_ld_dsmem(&s_stat[tid * 2], c);
      ps[c] = pf_ld_dsmem(&s_stat[tid * 2 + 1], c);
      M = fmaxf(M, pm[c]);
    }
    const int n_removed = dom == 1 ? lens_x - (n_cur - lens_t) : 0;
    const float l_rem = smooth * s_fold[kCA + k];
    if (n_removed > 0) M = fmaxf(M, l_rem);
    float S = n_removed > 0 ? (float)n_removed * expf(l_rem - M) : 0.f;
#pragma unroll
    for (int c = 0; c < kCluster; ++c)
      if (ps[c] > 0.f) S += ps[c] * expf(pm[c] - M);
    s_ms[tid * 2] = M;
    s_ms[tid * 2 + 1] = 1.0f / S;
  }
  // peers may exit only after everyone has read their statistics: arrive now, wait at the end
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  __syncthreads();
  // Fovea output per (row, channel): F = softmax * x0 + x1; kept in s_x1, stored by position for the next layer
  for (int idx = tid; idx < kWarps * RPW * 8; idx += kThreads) {
    const int li = idx >> 3, k = idx & 7;
    const int pos = s_pos[li];
    if (pos < 0) continue;
    const int dom = pos < lens_t ? 0 : 1;
    const float x0 = s_x0[idx];
    const float fk = expf(x0 * smooth - s_ms[(dom * 8 + k) * 2]) * s_ms[(dom * 8 + k) * 2 + 1] * x0 + s_x1[idx];
    s_x1[idx] = fk;
    F[((size_t)b * lens_full + pos) * 8 + k] = fk;
  }
  __syncthreads();

  // ---- phase 2: up-projection (weights outer, rows inner), residual, block LayerNorm
  {
    float f[RPW][8], s[RPW];
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      const float4 f0 = *reinterpret_cast<const float4*>(&s_x1[(warp * RPW + i) * 8]);
      const float4 f1 = *reinterpret_cast<const float4*>(&s_x1[(warp * RPW + i) * 8 + 4]);
      f[i][0] = f0.x; f[i][1] = f0.y; f[i][2] = f0.z; f[i][3] = f0.w;
      f[i][4] = f1.x; f[i][5] = f1.y; f[i][6] = f1.z; f[i][7] = f1.w;
      s[i] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float4 bias = *reinterpret_cast<const float4*>(b1 + (lane + 32 * j) * 4);
      float4 p[RPW];
#pragma unroll
      for (int i = 0; i < RPW; ++i) p[i] = bias;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 w = s_w1[k * 192 + lane + 32 * j];
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
          p[i].x += w.x * f[i][k]; p[i].y += w.y * f[i][k]; p[i].z += w.z * f[i][k]; p[i].w += w.w * f[i][k];
        }
      }
#pragma unroll
      for (int i = 0; i < RPW; ++i) {
        v[i][j].x += p[i].x; v[i][j].y += p[i].y; v[i][j].z += p[i].z; v[i][j].w += p[i].w;
        s[i] += v[i][j].x + v[i][j].y + v[i][j].z + v[i][j].w;
      }
    }
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
      if (slot[i] < 0) continue;                   // warp-uniform
      const int r = slot[i];
      float4* xo = reinterpret_cast<float4*>(x + ((size_t)b * n_cur + r) * D);
#pragma unroll
      for (int j = 0; j < 6; ++j) xo[lane + 32 * j] = v[i][j];
      const float mean = pf_warp_sum(s[i]) * (1.0f / D);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const float a0 = v[i][j].x - mean, a1 = v[i][j].y - mean, a2 = v[i][j].z - mean, a3 = v[i][j].w - mean;
        q += a0 * a0 + a1 * a1 + a2 * a2 + a3 * a3;
      }
      const float rstd = rsqrtf(pf_warp_sum(q) * (1.0f / D) + kLnEps);
      uint2* no = reinterpret_cast<uint2*>(xn + ((size_t)b * n_cur + r) * D);
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const float4 g = *reinterpret_cast<const float4*>(gamma1 + (lane + 32 * j) * 4);
        const float4 be = *reinterpret_cast<const float4*>(beta1 + (lane + 32 * j) * 4);
        uint2 u;
        u.x = pack_op2((v[i][j].x - mean) * rstd * g.x + be.x, (v[i][j].y - mean) * rstd * g.y + be.y);
        u.y = pack_op2((v[i][j].z - mean) * rstd * g.z + be.z, (v[i][j].w - mean) * rstd * g.w + be.w);
        no[lane + 32 * j] = u;
      }
    }
  }
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int RPW>
static int launch_pf(float* x, const void* delta, int n_cur, const int* gi, float* F, const float* wg_a, const float* fold,
                     const float* w1t, const float* b1, const float* gamma1, const float* beta1, void* xn, int B,
                     int lens_t, int lens_full, cudaStream_t st) {
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(prompt_fused_kernel<RPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, pf::kSmemBytes) !=
        cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  prompt_fused_kernel<RPW><<<B * pf::kCluster, pf::kThreads, pf::kSmemBytes, st>>>(
      x, reinterpret_cast<const op_t*>(delta), n_cur, gi, F, wg_a, fold, w1t, b1, gamma1, beta1,
      reinterpret_cast<op_t*>(xn), lens_t, lens_full);
  return cudaGetLastError() == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

int launch_prompt_fused(float* x, const void* delta, int n_cur, const int* gi, float* F, const float* wg_a,
                        const float* fold, const float* w1t, const float* b1, const float* gamma1, const float* beta1,
                        void* xn, int B, int lens_t, int lens_full, cudaStream_t st) {
  if (!delta || n_cur < lens_t || n_cur > lens_full) return VIPT_ERR_ARG;
  const int rows_per_cta = (n_cur + pf::kCluster - 1) / pf::kCluster;
  const int rpw = (rows_per_cta + pf::kWarps - 1) / pf::kWarps;
  switch (rpw) {
    case 1: case 2:
      return launch_pf<2>(x, delta, n_cur, gi, F, wg_a, fold, w1t, b1, gamma1, beta1, xn, B, lens_t, lens_full, st);
    case 3:
      return launch_pf<3>(x, delta, n_cur, gi, F, wg_a, fold, w1t, b1, gamma1, beta1, xn, B, lens_t, lens_full, st);
  }
  return VIPT_ERR_ARG;
}

}  // namespace vipt
