// HBM-bound token kernels of the ViPT backbone: LayerNorm (+ candidate-elimination gather), the two
// phases of the modality-complementary prompt block, the candidate-elimination ranking, and the final
// norm + scatter back to raster order.  One warp (or one CTA tile) per token row of 768 fp32 values.
//
// Reference sites (ViPT/lib/models/...):
//   prompt stage      vipt/vit_ce_prompt.py:205-219 (layer 0), :268-310 (layers 1..11)
//   Prompt_block/Fovea vipt/vit_ce_prompt.py:22-71
//   CE ranking        layers/attn_blocks.py:37-75
//   final recover     vipt/vit_ce_prompt.py:318-339
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

constexpr int D = 768;
constexpr float kLnEps = 1e-6f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// lane owns float4 chunks lane + 32*j, j = 0..5
__device__ __forceinline__ void load_row(const float* row, int lane, float4 (&v)[6]) {
#pragma unroll
  for (int j = 0; j < 6; ++j) v[j] = reinterpret_cast<const float4*>(row)[lane + 32 * j];
}
// fp16 branch output (attn.proj / mlp.fc2 GEMM result) of the same row: 4 halves per chunk
__device__ __forceinline__ float4 ld_delta4(const op_t* row, int chunk) {
  const uint2 u = reinterpret_cast<const uint2*>(row)[chunk];
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void add_delta_row(const op_t* drow, int lane, float4 (&v)[6]) {
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const float4 d = ld_delta4(drow, lane + 32 * j);
    v[j].x += d.x; v[j].y += d.y; v[j].z += d.z; v[j].w += d.w;
  }
}

__device__ __forceinline__ void row_stats(const float4 (&v)[6], float& mean, float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 6; ++j) s += v[j].x + v[j].y + v[j].z + v[j].w;
  mean = warp_sum(s) * (1.0f / D);
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
    q += a * a + b * b + c * c + d * d;
  }
  rstd = rsqrtf(warp_sum(q) * (1.0f / D) + kLnEps);
}

__device__ __forceinline__ void store_ln_f16(op_t* out, int lane, const float4 (&v)[6], float mean, float rstd,
                                             const float* __restrict__ gamma, const float* __restrict__ beta) {
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const int c = (lane + 32 * j) * 4;
    const float4 g = *reinterpret_cast<const float4*>(gamma + c);
    const float4 b = *reinterpret_cast<const float4*>(beta + c);
    uint2 u;
    u.x = pack_op2((v[j].x - mean) * rstd * g.x + b.x, (v[j].y - mean) * rstd * g.y + b.y);
    u.y = pack_op2((v[j].z - mean) * rstd * g.z + b.z, (v[j].w - mean) * rstd * g.w + b.w);
    *reinterpret_cast<uint2*>(out + c) = u;
  }
}

// ------------------------------------------------------------------------------------------------
// Residual add + LayerNorm:  x_out[row] = x_in[src] + delta[src];  xn[row] = LN(x_out[row]) in fp16.
// `delta` is the fp16 output of the preceding attn.proj / mlp.fc2 GEMM (the residual add of
// attn_blocks.py:95,103 happens here, coalesced, instead of in the GEMM epilogue).  `perm` gathers the
// search rows (candidate elimination: new slot r takes old slot perm[b][r]).  x_out may alias x_in
// when perm is null.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ln_rows_kernel(const float* x_in, int n_in, const op_t* __restrict__ delta, const int* __restrict__ perm, int perm_stride,
               float* x_out, op_t* __restrict__ xn, const float* __restrict__ gamma,
               const float* __restrict__ beta, int B, int n_out, int lens_t) {
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= B * n_out) return;
  const int b = row / n_out, r = row % n_out;
  int src = r;
  if (perm && r >= lens_t) src = lens_t + perm[b * perm_stride + (r - lens_t)];
  float4 v[6];
  load_row(x_in + ((size_t)b * n_in + src) * D, lane, v);
  if (delta) add_delta_row(delta + ((size_t)b * n_in + src) * D, lane, v);
  if (x_out) {
#pragma unroll
    for (int j = 0; j < 6; ++j) reinterpret_cast<float4*>(x_out + (size_t)row * D)[lane + 32 * j] = v[j];
  }
  float mean, rstd;
  row_stats(v, mean, rstd);
  store_ln_f16(xn + (size_t)row * D, lane, v, mean, rstd, gamma, beta);
}

int launch_ln_rows(const float* x_in, int n_in, const void* delta, const int* perm, int perm_stride, float* x_out, void* xn,
                   const float* gamma, const float* beta, int B, int n_out, int lens_t, cudaStream_t st) {
  const int rows = B * n_out;
  return launch_pdl(ln_rows_kernel, (rows + 7) / 8, 256, 0, st, x_in, n_in, reinterpret_cast<const op_t*>(delta), perm, perm_stride,
                    x_out, reinterpret_cast<op_t*>(xn), gamma, beta, B, n_out, lens_t) == cudaSuccess
             ? VIPT_OK
             : VIPT_ERR_LAUNCH;
}

// ------------------------------------------------------------------------------------------------
// Prompt phase A: per full-grid position, the two 768 -> 8 down-projections of the LayerNorm'ed
// streams, with LayerNorm folded into the projection:
//   conv0_k(LN(v)) = ((W.gamma) . v  -  mean * sum(W.gamma)) * rstd  +  (W . beta + bias)
// so each row is read exactly once.  A removed search position is an all-zero row (LN -> beta,
// vit_ce_prompt.py:279-288), which the same formula yields with v = 0.
//   stream a : current tokens x (gathered through slot_of_pos), norm = prompt_norms[i-1] (or [0])
//   stream d : previous prompt (or the aux-modality embedding at layer 0), norm = prompt_norms[i]
// Each warp handles kTok consecutive positions so a weight float4 read from shared memory is reused
// kTok times.  Output xd[(b*320 + t)*16 + k]: k<8 -> x0 (conv0_0), k>=8 -> x1 (conv0_1).
// ------------------------------------------------------------------------------------------------
constexpr int kTok = 4;

struct PromptAFold {            // device pointers, per layer
  const float* wg_a;            // [8][768]  conv0_0.weight * gamma_a
  const float* wg_d;            // [8][768]  conv0_1.weight * gamma_d
  const float* sc;              // [32]: s_a[8], c_a[8], s_d[8], c_d[8]
};

__global__ void __launch_bounds__(256, 2)
prompt_a_kernel(const float* __restrict__ xa, const op_t* __restrict__ delta, int n_cur,
                const int* __restrict__ slot_of_pos, const float* __restrict__ xd_src, PromptAFold w, float* __restrict__ xd, int B, int lens_t,
                int lens_full) {
  extern __shared__ float4 ws4[];    // [2][8][192] float4
  for (int i = threadIdx.x; i < 2 * 8 * 192; i += 256)
    ws4[i] = (i < 8 * 192) ? reinterpret_cast<const float4*>(w.wg_a)[i]
                           : reinterpret_cast<const float4*>(w.wg_d)[i - 8 * 192];
  __syncthreads();     // weights are static: staged before the upstream kernel has finished
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int groups_per_b = lens_full / kTok;
  const int n_tasks = B * groups_per_b * 2;          // task = (4 consecutive positions, stream)
  for (int task = blockIdx.x * 8 + (threadIdx.x >> 5); task < n_tasks; task += gridDim.x * 8) {
    const int stream = task & 1;
    const int wid = task >> 1;
    const int b = wid / groups_per_b;
    const int t0 = (wid % groups_per_b) * kTok;
    const float* rows[kTok];
    const op_t* drows[kTok];
#pragma unroll
    for (int tt = 0; tt < kTok; ++tt) {
      const int t = t0 + tt;
      drows[tt] = nullptr;
      if (stream == 0) {
        int r = t;
        if (t >= lens_t && slot_of_pos) {
          const int s = slot_of_pos[b * (lens_full - lens_t) + (t - lens_t)];
          r = s < 0 ? -1 : lens_t + s;
        }
        rows[tt] = r < 0 ? nullptr : xa + ((size_t)b * n_cur + r) * D;
        if (delta && r >= 0) drows[tt] = delta + ((size_t)b * n_cur + r) * D;
      } else {
        rows[tt] = xd_src + ((size_t)b * lens_full + t) * D;
      }
    }
    float acc[kTok][8], s1[kTok], s2[kTok];
#pragma unroll
    for (int tt = 0; tt < kTok; ++tt) {
      s1[tt] = 0.f; s2[tt] = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[tt][k] = 0.f;
    }
    const float4* wbase = ws4 + stream * 8 * 192;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      float4 v[3][kTok];                       // 12 independent 512-byte row segments in flight per warp
#pragma unroll
      for (int jj = 0; jj < 3; ++jj)
#pragma unroll
        for (int tt = 0; tt < kTok; ++tt)
        {
          v[jj][tt] = rows[tt] ? __ldg(reinterpret_cast<const float4*>(rows[tt]) + lane + 32 * (half * 3 + jj))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
          if (drows[tt]) {
            const float4 d = ld_delta4(drows[tt], lane + 32 * (half * 3 + jj));
            v[jj][tt].x += d.x; v[jj][tt].y += d.y; v[jj][tt].z += d.z; v[jj][tt].w += d.w;
          }
        }
#pragma unroll
      for (int jj = 0; jj < 3; ++jj) {
#pragma unroll
        for (int tt = 0; tt < kTok; ++tt) {
          const float4 a = v[jj][tt];
          s1[tt] += a.x + a.y + a.z + a.w;
          s2[tt] += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 wv = wbase[k * 192 + lane + 32 * (half * 3 + jj)];
#pragma unroll
          for (int tt = 0; tt < kTok; ++tt)
            acc[tt][k] += wv.x * v[jj][tt].x + wv.y * v[jj][tt].y + wv.z * v[jj][tt].z + wv.w * v[jj][tt].w;
        }
      }
    }
#pragma unroll
    for (int tt = 0; tt < kTok; ++tt) {
      const float mean = warp_sum(s1[tt]) * (1.0f / D);
      const float var = fmaxf(warp_sum(s2[tt]) * (1.0f / D) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + kLnEps);
      float mine = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float dot = warp_sum(acc[tt][k]);
        const float val = (dot - mean * w.sc[stream * 16 + k]) * rstd + w.sc[stream * 16 + 8 + k];
        if (lane == k) mine = val;
      }
      if (lane < 8) xd[((size_t)b * lens_full + t0 + tt) * 16 + stream * 8 + lane] = mine;
    }
  }
}

int launch_prompt_a(const float* xa, const void* delta, int n_cur, const int* slot_of_pos, const float* xd_src, const float* wg_a,
                    const float* wg_d, const float* sc, float* xd, int B, int lens_t, int lens_full, cudaStream_t st) {
  static bool attr = false;
  const int smem = 2 * 8 * 192 * 16;
  if (!attr) {
    if (cudaFuncSetAttribute(prompt_a_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  PromptAFold w{wg_a, wg_d, sc};
  const int tasks = B * (lens_full / kTok) * 2;
  int grid = (tasks + 7) / 8;
  if (grid > 148 * 2) grid = 148 * 2;      // persistent: the 48 KB weight tile is staged once per CTA
  const cudaError_t e_ = launch_pdl(prompt_a_kernel, grid, 256, smem, st, xa, reinterpret_cast<const op_t*>(delta), n_cur, slot_of_pos, xd_src, w, xd, B, lens_t, lens_full);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
