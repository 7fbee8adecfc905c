// Prompt phase B (Fovea softmax + 8 -> 768 up-projection + residual + next LayerNorm), the
// candidate-elimination ranking, and the final norm + scatter.  See tokens.cu for the reference sites.
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

constexpr int D = 768;
constexpr float kLnEps = 1e-6f;
constexpr int kTileTok = 8;      // full-grid positions per CTA
constexpr int kBThreads = 192;   // 4 channels per thread

__device__ __forceinline__ float warp_sum2(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Fovea (vit_ce_prompt.py:33-47): per (sequence, domain) softmax over the domain's positions of
// x0 * smooth, per channel; f = softmax * x0 + x1.  Domains: template positions [0,64), search
// positions [64,320) -- separate Prompt_block calls (vit_ce_prompt.py:212-213, 301-302).  Removed
// search positions take part in the normaliser (their x0 is conv0_0(beta)).
// grid (B, 2), 256 threads: one position per thread.
__global__ void __launch_bounds__(256)
fovea_kernel(const float* __restrict__ xd, float smooth, float* __restrict__ F, int lens_t, int lens_full) {
  __shared__ float s_red[8][8];
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  const int b = blockIdx.x, dom = blockIdx.y, t = threadIdx.x;
  const int dom0 = dom == 0 ? 0 : lens_t;
  const int domL = dom == 0 ? lens_t : lens_full - lens_t;
  const int warp = t >> 5, lane = t & 31;
  const bool act = t < domL;
  float x0[8], x1[8], lg[8];
  if (act) {
    const float4* r = reinterpret_cast<const float4*>(xd + ((size_t)b * lens_full + dom0 + t) * 16);
    const float4 a = r[0], c = r[1], d = r[2], e = r[3];
    x0[0] = a.x; x0[1] = a.y; x0[2] = a.z; x0[3] = a.w; x0[4] = c.x; x0[5] = c.y; x0[6] = c.z; x0[7] = c.w;
    x1[0] = d.x; x1[1] = d.y; x1[2] = d.z; x1[3] = d.w; x1[4] = e.x; x1[5] = e.y; x1[6] = e.z; x1[7] = e.w;
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    lg[k] = act ? x0[k] * smooth : -INFINITY;
    float m = lg[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) s_red[warp][k] = m;
  }
  __syncthreads();
  float mx[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    float m = s_red[0][k];
#pragma unroll
    for (int w = 1; w < 8; ++w) m = fmaxf(m, s_red[w][k]);
    mx[k] = m;
  }
  __syncthreads();
  float e[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    e[k] = act ? expf(lg[k] - mx[k]) : 0.f;
    const float s = warp_sum2(e[k]);
    if (lane == 0) s_red[warp][k] = s;
  }
  __syncthreads();
  if (!act) return;
  float f[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += s_red[w][k];
    f[k] = (e[k] / s) * x0[k] + x1[k];
  }
  float4* o = reinterpret_cast<float4*>(F + ((size_t)b * lens_full + dom0 + t) * 8);
  o[0] = make_float4(f[0], f[1], f[2], f[3]);
  o[1] = make_float4(f[4], f[5], f[6], f[7]);
}

int launch_fovea(const float* xd, float smooth, float* F, int B, int lens_t, int lens_full, cudaStream_t st) {
  if (lens_t > 256 || lens_full - lens_t > 256) return VIPT_ERR_ARG;
  const cudaError_t e_ = launch_pdl(fovea_kernel, dim3(B, 2), 256, 0, st, xd, smooth, F, lens_t, lens_full);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

// Prompt phase B.  grid: B * (lens_full / 8); tile = 8 consecutive full-grid positions.
//   p          = conv1x1(f)                                          -> prompt stream P (all positions)
//   x[row]     = base[row] + p (+ pos_embed at layer 0)               -> kept positions only
//   xn[row]    = LayerNorm_norm1(x[row]) in fp16                      -> feeds the qkv GEMM
// All 8 residual rows are loaded up front (independent 3 KB rows in flight), then streamed out.
__global__ void __launch_bounds__(kBThreads, 3)
prompt_b_kernel(const float* __restrict__ F, const float* __restrict__ w1, const float* __restrict__ b1,
                float* __restrict__ P, const float* base, const op_t* __restrict__ delta, float* x,
                int n_cur,   // base may alias x (in-place); delta = fp16 mlp.fc2 output still to be added
                const int* __restrict__ slot_of_pos, const float* __restrict__ pos_embed,
                const float* __restrict__ gamma1, const float* __restrict__ beta1, op_t* __restrict__ xn, int lens_t,
                int lens_full) {
  __shared__ float s_f[kTileTok][8];
  __shared__ float s_part[kTileTok][6][2];
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  __shared__ int s_row[kTileTok];
  const int tid = threadIdx.x;
  const int tiles_per_b = lens_full / kTileTok;
  const int b = blockIdx.x / tiles_per_b;
  const int t0 = (blockIdx.x % tiles_per_b) * kTileTok;
  if (tid < kTileTok * 8) s_f[tid >> 3][tid & 7] = F[((size_t)b * lens_full + t0) * 8 + tid];
  if (tid >= 128 && tid < 128 + kTileTok) {
    const int tt = tid - 128, t = t0 + tt;
    int r = t;
    if (t >= lens_t && slot_of_pos) {
      const int s = slot_of_pos[b * (lens_full - lens_t) + (t - lens_t)];
      r = s < 0 ? -1 : lens_t + s;
    }
    s_row[tt] = r;
  }
  __syncthreads();

  const int c0 = tid * 4;
  float4 xv[kTileTok];
#pragma unroll
  for (int tt = 0; tt < kTileTok; ++tt) {
    const int r = s_row[tt];
    xv[tt] = r >= 0 ? *reinterpret_cast<const float4*>(base + ((size_t)b * n_cur + r) * D + c0)
                    : make_float4(0.f, 0.f, 0.f, 0.f);
    if (delta && r >= 0) {
      const uint2 u = *reinterpret_cast<const uint2*>(delta + ((size_t)b * n_cur + r) * D + c0);
      const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
      const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
      xv[tt].x += d0.x; xv[tt].y += d0.y; xv[tt].z += d1.x; xv[tt].w += d1.y;
    }
  }
  float w[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 lo = *reinterpret_cast<const float4*>(w1 + (c0 + i) * 8);
    const float4 hi = *reinterpret_cast<const float4*>(w1 + (c0 + i) * 8 + 4);
    w[i][0] = lo.x; w[i][1] = lo.y; w[i][2] = lo.z; w[i][3] = lo.w;
    w[i][4] = hi.x; w[i][5] = hi.y; w[i][6] = hi.z; w[i][7] = hi.w;
  }
  const float4 bias = *reinterpret_cast<const float4*>(b1 + c0);
  const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int tt = 0; tt < kTileTok; ++tt) {
    float p[4] = {bias.x, bias.y, bias.z, bias.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float f = s_f[tt][k];
#pragma unroll
      for (int i = 0; i < 4; ++i) p[i] += w[i][k] * f;
    }
    const int t = t0 + tt;
    if (P) *reinterpret_cast<float4*>(P + ((size_t)b * lens_full + t) * D + c0) = make_float4(p[0], p[1], p[2], p[3]);
    const int r = s_row[tt];
    float4 v = xv[tt];
    if (r >= 0) {
      v.x += p[0]; v.y += p[1]; v.z += p[2]; v.w += p[3];
      if (pos_embed) {
        const float4 pe = *reinterpret_cast<const float4*>(pos_embed + (size_t)t * D + c0);
        v.x += pe.x; v.y += pe.y; v.z += pe.z; v.w += pe.w;
      }
      *reinterpret_cast<float4*>(x + ((size_t)b * n_cur + r) * D + c0) = v;
    }
    xv[tt] = v;
    const float s = warp_sum2(v.x + v.y + v.z + v.w);
    if (lane == 0) s_part[tt][warp][0] = s;
  }
  __syncthreads();
  // two-pass variance (mean first), like torch's LayerNorm
  float mean[kTileTok];
#pragma unroll
  for (int tt = 0; tt < kTileTok; ++tt) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) s += s_part[tt][i][0];
    mean[tt] = s * (1.0f / D);
    const float a = xv[tt].x - mean[tt], bb = xv[tt].y - mean[tt], c = xv[tt].z - mean[tt], d = xv[tt].w - mean[tt];
    const float q = warp_sum2(a * a + bb * bb + c * c + d * d);
    if (lane == 0) s_part[tt][warp][1] = q;
  }
  __syncthreads();
  const float4 g = *reinterpret_cast<const float4*>(gamma1 + c0);
  const float4 be = *reinterpret_cast<const float4*>(beta1 + c0);
#pragma unroll
  for (int tt = 0; tt < kTileTok; ++tt) {
    const int r = s_row[tt];
    if (r < 0) continue;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 6; ++i) q += s_part[tt][i][1];
    const float rstd = rsqrtf(q * (1.0f / D) + kLnEps);
    uint2 u;
    u.x = pack_op2((xv[tt].x - mean[tt]) * rstd * g.x + be.x, (xv[tt].y - mean[tt]) * rstd * g.y + be.y);
    u.y = pack_op2((xv[tt].z - mean[tt]) * rstd * g.z + be.z, (xv[tt].w - mean[tt]) * rstd * g.w + be.w);
    *reinterpret_cast<uint2*>(xn + ((size_t)b * n_cur + r) * D + c0) = u;
  }
}

// Layer-0 form of phase B (input prompt, vit_ce_prompt.py:205-219 followed by blocks[0].norm1): every position is kept,
// there is no branch output to add and no prompt stream to store, so the kernel is a pure row stream -- one warp per
// token row, 12 independent 512-byte loads per warp in flight (E row + pos_embed row), conv1x1 staged k-major in shared
// memory (conflict-free 16-byte reads), LayerNorm by warp shuffles.  (The generic tile form above ran at 28 % occupancy
// with three block barriers per 8 rows: 74 us for 157 MB.)
__global__ void __launch_bounds__(256, 2)
prompt_b0_kernel(const float* __restrict__ F, const float* __restrict__ w1, const float* __restrict__ b1,
                 const float* __restrict__ base, float* __restrict__ x, const float* __restrict__ pos_embed,
                 const float* __restrict__ gamma1, const float* __restrict__ beta1, op_t* __restrict__ xn, int rows,
                 int lens_full) {
  __shared__ float4 s_wk[8 * 192];             // conv1x1.weight, k-major: [k][768]
  __shared__ float4 s_b[192], s_g[192], s_be[192];
  {
    float* swf = reinterpret_cast<float*>(s_wk);
    for (int i = threadIdx.x; i < D * 8; i += 256) swf[(i & 7) * D + (i >> 3)] = w1[i];      // (768, 8) -> [8][768]
    for (int i = threadIdx.x; i < 192; i += 256) {
      s_b[i] = reinterpret_cast<const float4*>(b1)[i];
      s_g[i] = reinterpret_cast<const float4*>(gamma1)[i];
      s_be[i] = reinterpret_cast<const float4*>(beta1)[i];
    }
  }
  __syncthreads();     // weights are static: staged before the upstream kernel has finished
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < rows; row += gridDim.x * 8) {
    const int t = row % lens_full;
    const float4* br = reinterpret_cast<const float4*>(base + (size_t)row * D);
    const float4* pr = reinterpret_cast<const float4*>(pos_embed + (size_t)t * D);
    float4 v[6], pe[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = __ldg(br + lane + 32 * j);
      pe[j] = __ldg(pr + lane + 32 * j);
    }
    const float4 f0 = __ldg(reinterpret_cast<const float4*>(F + (size_t)row * 8));
    const float4 f1 = __ldg(reinterpret_cast<const float4*>(F + (size_t)row * 8 + 4));
    const float f[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
#pragma unroll
    for (int j = 0; j < 6; ++j) { v[j].x += pe[j].x; v[j].y += pe[j].y; v[j].z += pe[j].z; v[j].w += pe[j].w; }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      float4 p = s_b[lane + 32 * j];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 w = s_wk[k * 192 + lane + 32 * j];
        p.x = fmaf(w.x, f[k], p.x); p.y = fmaf(w.y, f[k], p.y); p.z = fmaf(w.z, f[k], p.z); p.w = fmaf(w.w, f[k], p.w);
      }
      v[j].x += p.x; v[j].y += p.y; v[j].z += p.z; v[j].w += p.w;
      s += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    float4* xo = reinterpret_cast<float4*>(x + (size_t)row * D);
#pragma unroll
    for (int j = 0; j < 6; ++j) xo[lane + 32 * j] = v[j];
    const float mean = warp_sum2(s) * (1.0f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      q += a * a + b * b + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum2(q) * (1.0f / D) + kLnEps);
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float4 g = s_g[lane + 32 * j], be = s_be[lane + 32 * j];
      uint2 u;
      u.x = pack_op2((v[j].x - mean) * rstd * g.x + be.x, (v[j].y - mean) * rstd * g.y + be.y);
      u.y = pack_op2((v[j].z - mean) * rstd * g.z + be.z, (v[j].w - mean) * rstd * g.w + be.w);
      reinterpret_cast<uint2*>(xn + (size_t)row * D)[lane + 32 * j] = u;
    }
  }
}

int launch_prompt_b(const float* F, const float* w1, const float* b1, float* P, const float* base, const void* delta,
                    float* x, int n_cur, const int* slot_of_pos, const float* pos_embed, const float* gamma1,
                    const float* beta1, void* xn, int B, int lens_t, int lens_full, cudaStream_t st) {
  if (!P && !delta && !slot_of_pos && pos_embed && n_cur == lens_full && base != x) {      // layer 0: row-stream form
    const int rows = B * lens_full;
    int grid = (rows + 7) / 8;
    if (grid > 148 * 2) grid = 148 * 2;      // persistent: the 33 KB of weights are staged once per CTA
    return launch_pdl(prompt_b0_kernel, grid, 256, 0, st, F, w1, b1, base, x, pos_embed, gamma1, beta1, reinterpret_cast<op_t*>(xn),
                      rows, lens_full) == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
  }
  const cudaError_t e_ = launch_pdl(prompt_b_kernel, B * (lens_full / kTileTok), kBThreads, 0, st, F, w1, b1, P, base, reinterpret_cast<const op_t*>(delta), x, n_cur, slot_of_pos, pos_embed, gamma1, beta1, reinterpret_cast<op_t*>(xn), lens_t, lens_full);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

// ------------------------------------------------------------------------------------------------
// Candidate elimination ranking (attn_blocks.py:37-63).  One CTA per sequence.
//   score[j] = mean over heads of softmax prob (template token 27 -> search key j)
//   descending sort (ties: lower slot first), keep the first n_keep.
// Outputs: perm[r] = old slot of new slot r (r < n_keep); gi_out[r] = original raster position of new
// slot r; removed[r - n_keep] = raster positions of the dropped tokens in sorted order;
// slot_of_pos[pos] = new slot or -1.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ce_topk_kernel(const float* __restrict__ ce_prob, int n_heads, int n_s, int n_keep, const int* __restrict__ gi_in,
               int* __restrict__ gi_out, int* __restrict__ perm, int* __restrict__ removed,
               int* __restrict__ slot_of_pos, float* __restrict__ score_out, int lens_x) {
  __shared__ float s_score[256];
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  const int b = blockIdx.x, j = threadIdx.x;
  float sc = -INFINITY;
  if (j < n_s) {
    float s = 0.f;
    for (int h = 0; h < n_heads; ++h) s += ce_prob[((size_t)b * n_heads + h) * lens_x + j];
    sc = s / (float)n_heads;
    if (score_out) score_out[b * lens_x + j] = sc;
  }
  s_score[j] = sc;
  __syncthreads();
  if (j >= n_s) return;
  int rank = 0;
  for (int i = 0; i < n_s; ++i) {
    const float o = s_score[i];
    rank += (o > sc) || (o == sc && i < j);
  }
  const int pos = gi_in[b * lens_x + j];
  if (rank < n_keep) {
    perm[b * lens_x + rank] = j;
    gi_out[b * lens_x + rank] = pos;
    slot_of_pos[b * lens_x + pos] = rank;
  } else {
    removed[b * lens_x + (rank - n_keep)] = pos;
    slot_of_pos[b * lens_x + pos] = -1;
  }
}

int launch_ce_topk(const float* ce_prob, int n_heads, int n_s, int n_keep, const int* gi_in, int* gi_out, int* perm,
                   int* removed, int* slot_of_pos, float* score_out, int lens_x, int B, cudaStream_t st) {
  if (n_s > 256) return VIPT_ERR_ARG;
  const cudaError_t e_ = launch_pdl(ce_topk_kernel, B, 256, 0, st, ce_prob, n_heads, n_s, n_keep, gi_in, gi_out, perm, removed, slot_of_pos, score_out, lens_x);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

// ------------------------------------------------------------------------------------------------
// Final norm + recover (vit_ce_prompt.py:318-339): LN of the surviving rows, search rows scattered to
// raster order, removed positions EXACT zeros.  feat16 (B, 256, 768) feeds the head; feat32
// (B, 320, 768), optional, is the `backbone_feat` the model API returns.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
final_norm_kernel(const float* __restrict__ x, const op_t* __restrict__ delta, int n_cur,
                  const int* __restrict__ slot_of_pos,
                  const float* __restrict__ gamma, const float* __restrict__ beta, op_t* __restrict__ feat16,
                  float* __restrict__ feat32, int B, int lens_t, int lens_full) {
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= B * lens_full) return;
  const int b = row / lens_full, t = row % lens_full;
  int r = t;
  if (t >= lens_t && slot_of_pos) {
    const int s = slot_of_pos[b * (lens_full - lens_t) + (t - lens_t)];
    r = s < 0 ? -1 : lens_t + s;
  }
  if (t < lens_t && !feat32) return;
  float o[6][4];
  if (r >= 0) {
    const float* src = x + ((size_t)b * n_cur + r) * D;
    float4 v[6];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      v[j] = reinterpret_cast<const float4*>(src)[lane + 32 * j];
      if (delta) {
        const uint2 u = reinterpret_cast<const uint2*>(delta + ((size_t)b * n_cur + r) * D)[lane + 32 * j];
        const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
        const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
        v[j].x += d0.x; v[j].y += d0.y; v[j].z += d1.x; v[j].w += d1.y;
      }
      s += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    const float mean = warp_sum2(s) * (1.0f / D);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float a = v[j].x - mean, bb = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
      q += a * a + bb * bb + c * c + d * d;
    }
    const float rstd = rsqrtf(warp_sum2(q) * (1.0f / D) + kLnEps);
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const int c = (lane + 32 * j) * 4;
      const float4 g = *reinterpret_cast<const float4*>(gamma + c);
      const float4 be = *reinterpret_cast<const float4*>(beta + c);
      o[j][0] = (v[j].x - mean) * rstd * g.x + be.x; o[j][1] = (v[j].y - mean) * rstd * g.y + be.y;
      o[j][2] = (v[j].z - mean) * rstd * g.z + be.z; o[j][3] = (v[j].w - mean) * rstd * g.w + be.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 6; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const int c = (lane + 32 * j) * 4;
    if (feat32) *reinterpret_cast<float4*>(feat32 + (size_t)row * D + c) = make_float4(o[j][0], o[j][1], o[j][2], o[j][3]);
    if (t >= lens_t) {
      uint2 u;
      u.x = pack_op2(o[j][0], o[j][1]);
      u.y = pack_op2(o[j][2], o[j][3]);
      *reinterpret_cast<uint2*>(feat16 + ((size_t)b * (lens_full - lens_t) + (t - lens_t)) * D + c) = u;
    }
  }
}

int launch_final_norm(const float* x, const void* delta, int n_cur, const int* slot_of_pos, const float* gamma, const float* beta,
                      void* feat16, float* feat32, int B, int lens_t, int lens_full, cudaStream_t st) {
  const int rows = B * lens_full;
  const cudaError_t e_ = launch_pdl(final_norm_kernel, (rows + 7) / 8, 256, 0, st, x, reinterpret_cast<const op_t*>(delta), n_cur, slot_of_pos, gamma, beta, reinterpret_cast<op_t*>(feat16), feat32, B, lens_t, lens_full);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
