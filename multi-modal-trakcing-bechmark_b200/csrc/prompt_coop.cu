// Deep-prompt stage of layers 1..11 (ViPT/lib/models/vipt/vit_ce_prompt.py:268-310) as ONE cooperative,
// persistent kernel with a single grid-wide barrier:
//   x      <- x + delta                       (delta = fp16 output of the previous block's mlp.fc2)
//   x0     = conv0_0(prompt_norms[i-1](x))     per kept token          (Prompt_block, :63-65)
//   x1     = conv0_1(prompt_norms[i](P_prev))  per kept position       (:66)
//   F      = softmax_over_positions(smooth * x0) * x0 + x1             (Fovea, :37-47)
//   x      <- x + conv1x1(F)                   kept positions, current token order (:308-310)
//   xn     = blocks[i].norm1(x) in fp16        feeds the qkv GEMM      (attn_blocks.py:94)
//
// The prompt stream itself is never materialised.  P_prev = conv1x1_{i-1}(F_prev) has rank 8, so
//   conv0_1(LN(P_prev[t])) = (A f + a0) * rstd(f) + c_d,  var(f) = f^T G f + 2 h.f + c0,  f = F_prev[t]
// with 8x8 / 8-vectors folded once on the host in double (engine.cu); only F (8 floats per position)
// travels between layers.  Removed search positions are zero rows before prompt_norms[i-1]
// (:276-288), i.e. x0 = conv0_0(beta) = c_a there: they enter the softmax normaliser as
// n_removed * exp(smooth * c_a - max) and nothing else depends on them.
//
//   phase 1  every warp streams token rows (two at a time): v = x + delta, LayerNorm statistics and the
//            768 -> 8 down-projection x0 = conv0_0(prompt_norms[i-1](v)); x0 goes to a small global array
//   grid.sync()
//   phase 2  per CTA: Fovea softmax statistics of the (<= 3) sequences its rows belong to, from x0;
//            every warp re-reads its rows (L2 hits: they were read a few microseconds earlier by the same
//            SM's neighbours), forms F = softmax * x0 + x1 (x1 from the previous layer's F by the rank-8
//            algebra), adds the 8 -> 768 up-projection, writes x and xn = blocks[i].norm1(x) in fp16.
//
// Nothing is held across the barrier: 16 warps per SM stream independently, DRAM reads happen once (phase 2 hits
// L2).  (A round-1 variant with a cluster of 8 CTAs per sequence kept the rows in registers across a cluster barrier
// and exchanged the softmax statistics through distributed shared memory; it was limited to one CTA per SM and
// serialised load / math / barrier / math / store inside every CTA, measured slower, and is no longer shipped.)
//
// Rows reach the SM through a per-warp shared-memory slot filled by bulk async copies (cp.async.bulk +
// mbarrier): while a warp works on one row pair from registers, the copy of its next pair is already in
// flight, so ~150 KB per SM are outstanding all the time without spending registers on a second buffer
// (the register-only version spent 42 % of its time waiting for the loads of the pair it was about to use).
#include <cooperative_groups.h>

#include <algorithm>
#include <vector>

#include "kernels.h"
#include "sm100_ptx.cuh"

namespace cg = cooperative_groups;

#ifdef PC_TRACE      // timing experiments only: per-CTA phase timestamps (printed by launch_prompt_coop)
__device__ long long g_pc_trace[160 * 8];
#define PCT(ev)                                                                        \
  do {                                                                                 \
    if (threadIdx.x == 0) g_pc_trace[blockIdx.x * 8 + (ev)] = clock64();               \
  } while (0)
#else
#define PCT(ev)
#endif

namespace vipt {

namespace pc {
constexpr int D = 768;
constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxSeq = 4;             // sequences a CTA's row range may touch
constexpr float kLnEps = 1e-6f;
constexpr int kCA = 0, kA = 8, kA0 = 72, kCD = 80, kG = 88, kH = 152, kC0 = 160, kSmooth = 161, kFoldFloats = 168;
constexpr int kSlotBytes = 2 * D * 4 + 2 * D * 2;      // one row pair: two fp32 rows, then two fp16 branch rows
constexpr int kSmemBytes = 2 * 8 * D * 4 + 3 * D * 4 + kFoldFloats * 4 + kMaxSeq * 32 * 4 + 2 * kThreads * 2 * 4 +
                           kWarps * kSlotBytes + (kWarps + 1) * 8 + 64;
}  // namespace pc

// two independent butterflies interleaved: a 5-step shuffle chain is pure latency for a warp (this kernel runs 4 warps
// per scheduler; the reductions were its largest stall), two chains in flight cost the same time as one
__device__ __forceinline__ void pc_warp_sum2(float& a, float& b) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ta = __shfl_xor_sync(0xffffffffu, a, o), tb = __shfl_xor_sync(0xffffffffu, b, o);
    a += ta;
    b += tb;
  }
}

__device__ __forceinline__ float pc_warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// lane 0: start the bulk copies of rows [r0, r0 + nrows) (x fp32 and the branch output fp16) into this warp's slot
__device__ __forceinline__ void pc_prefetch(const float* __restrict__ x, const op_t* __restrict__ delta, size_t r0, int nrows,
                                            uint32_t slot, uint64_t* bar) {
  const uint32_t xb = nrows * pc::D * 4, db = nrows * pc::D * 2;
  fence_proxy_async_smem();                      // earlier generic reads of the slot are ordered before the async writes
  mbar_expect_tx(bar, xb + db);
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(slot), "l"(x + r0 * pc::D), "r"(xb), "r"(smem_u32(bar)) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(slot + 2 * pc::D * 4), "l"(delta + r0 * pc::D), "r"(db), "r"(smem_u32(bar)) : "memory");
}

// v = x + delta of row `i` (0/1) of the slot
__device__ __forceinline__ void pc_read_row(uint32_t slot, int i, int lane, float4 (&v)[6]) {
  const uint32_t xa = slot + i * pc::D * 4 + lane * 16, da = slot + 2 * pc::D * 4 + i * pc::D * 2 + lane * 8;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    v[j] = lds128f(xa + j * 512);
    uint32_t u0, u1;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(u0), "=r"(u1) : "r"(da + j * 256) : "memory");
    const float2 d0 = __half22float2(*reinterpret_cast<const __half2*>(&u0));
    const float2 d1 = __half22float2(*reinterpret_cast<const __half2*>(&u1));
    v[j].x += d0.x; v[j].y += d0.y; v[j].z += d1.x; v[j].w += d1.y;
  }
}

__global__ void __launch_bounds__(pc::kThreads, 1)
prompt_coop_kernel(float* __restrict__ x, const op_t* __restrict__ delta, int B, int n_cur, const int* __restrict__ gi,
                   float* __restrict__ F, float* __restrict__ X0, const float* __restrict__ wg_a,
                   const float* __restrict__ fold, const float* __restrict__ w1t, const float* __restrict__ b1,
                   const float* __restrict__ gamma1, const float* __restrict__ beta1, op_t* __restrict__ xn, int lens_t,
                   int lens_full) {
  using namespace pc;
  extern __shared__ float4 pc_smem4[];
  float4* s_wa = pc_smem4;                       // [8][192] float4: conv0_0.weight * gamma_a
  float4* s_w1 = s_wa + 8 * 192;                 // [8][192] float4: conv1x1.weight, k-major
  float4* s_b1 = s_w1 + 8 * 192;                 // [192] conv1x1.bias, then norm1 gamma, beta
  float4* s_g1 = s_b1 + 192;
  float4* s_be1 = s_g1 + 192;
  float* s_fold = reinterpret_cast<float*>(s_be1 + 192);
  float* s_ms = s_fold + kFoldFloats;            // [kMaxSeq][2 domains][8][2]: max, 1 / sum
  float* s_red = s_ms + kMaxSeq * 32;            // [2][kThreads][2] reduction scratch
  uint8_t* s_slots = reinterpret_cast<uint8_t*>(s_red + 2 * kThreads * 2);      // [kWarps][kSlotBytes], 16-byte aligned
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_slots + kWarps * kSlotBytes);
  uint64_t* wbar = s_bar + kWarps;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  PCT(0);
  const int lens_x = lens_full - lens_t;
  const int total = B * n_cur;
  int rpc = (total + (int)gridDim.x - 1) / (int)gridDim.x;
  rpc = (rpc + 1) & ~1;                          // even: rows are processed in pairs
  const int row_lo = min((int)blockIdx.x * rpc, total), row_hi = min(row_lo + rpc, total);

  const uint32_t slot = smem_u32(s_slots + warp * kSlotBytes);
  uint64_t* bar = s_bar + warp;
  uint32_t uses = 0;                             // completed phases of this warp's barrier
  if (lane == 0) mbar_init(bar, 1);
  fence_mbar_init();
  __syncwarp();
  // first row pair of phase 1: in flight while the weights are staged
  if (lane == 0 && row_lo + 2 * warp < row_hi)
    pc_prefetch(x, delta, (size_t)(row_lo + 2 * warp), min(2, row_hi - (row_lo + 2 * warp)), slot, bar);
  // weights: five bulk async copies on one CTA-level barrier (no LDG -> STS round trip through registers)
  if (tid == 0) {
    mbar_init(wbar, 1);
    fence_mbar_init();
    mbar_expect_tx(wbar, 2 * 8 * D * 4 + 3 * D * 4 + kFoldFloats * 4);
    auto bulk = [&](void* dst, const void* src, uint32_t bytes) {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(wbar)) : "memory");
    };
    bulk(s_wa, wg_a, 8 * D * 4);
    bulk(s_w1, w1t, 8 * D * 4);
    bulk(s_b1, b1, D * 4);
    bulk(s_g1, gamma1, D * 4);
    bulk(s_be1, beta1, D * 4);
    bulk(s_fold, fold, kFoldFloats * 4);
  }
  __syncthreads();               // barrier initialised before anybody waits on it
  pdl_trigger();                 // the next kernel's CTAs may start their prologue (sm100_ptx.cuh); this kernel itself is a
                                 // cooperative launch without the programmatic attribute, so it has nothing to wait for
  mbar_wait(wbar, 0);
  PCT(1);
  const float smooth = s_fold[kSmooth];

  // ------------------------------------------------------------------ phase 1: x0 per row
  for (int r0 = row_lo + 2 * warp; r0 < row_hi; r0 += 2 * kWarps) {
    const bool two = r0 + 1 < row_hi;
    float4 v[2][6];
    mbar_wait(bar, uses & 1); ++uses;
    pc_read_row(slot, 0, lane, v[0]);
    if (two) pc_read_row(slot, 1, lane, v[1]);
    else {
#pragma unroll
      for (int j = 0; j < 6; ++j) v[1][j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();
    if (lane == 0 && r0 + 2 * kWarps < row_hi)   // next pair of this warp: in flight during the math below
      pc_prefetch(x, delta, (size_t)(r0 + 2 * kWarps), min(2, row_hi - (r0 + 2 * kWarps)), slot, bar);
    // The 768 -> 8 projection is 2 x 6144 FMAs per row pair and lane-for-lane the bulk of this kernel's instructions:
    // packed fp32x2 FMAs on (x, y) / (z, w) column pairs (the float4 halves are register pairs already) halve them.
    float mean[2], q[2], dot[2][8];
    uint64_t q2[2], dot2[2][8], nm2[2];
    {
      float s[2] = {0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 6; ++j) s[i] += v[i][j].x + v[i][j].y + v[i][j].z + v[i][j].w;
      pc_warp_sum2(s[0], s[1]);
      mean[0] = s[0] * (1.0f / D);
      mean[1] = s[1] * (1.0f / D);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      nm2[i] = pk2(-mean[i], -mean[i]);
      q2[i] = 0ull;
#pragma unroll
      for (int k = 0; k < 8; ++k) dot2[i][k] = 0ull;
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      uint64_t a2[2][2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        a2[i][0] = add2(pk2(v[i][j].x, v[i][j].y), nm2[i]);
        a2[i][1] = add2(pk2(v[i][j].z, v[i][j].w), nm2[i]);
        q2[i] = fma2(a2[i][0], a2[i][0], q2[i]);
        q2[i] = fma2(a2[i][1], a2[i][1], q2[i]);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 w = s_wa[k * 192 + lane + 32 * j];
        const uint64_t w01 = pk2(w.x, w.y), w23 = pk2(w.z, w.w);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          dot2[i][k] = fma2(w01, a2[i][0], dot2[i][k]);
          dot2[i][k] = fma2(w23, a2[i][1], dot2[i][k]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      q[i] = hsum2(q2[i]);
#pragma unroll
      for (int k = 0; k < 8; ++k) dot[i][k] = hsum2(dot2[i][k]);
    }
    pc_warp_sum2(q[0], q[1]);
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float rstd = rsqrtf(q[i] * (1.0f / D) + kLnEps);
      // transposed butterfly: 8 sums over 32 lanes in 9 shuffles; lane l ends with sum #(l >> 2)
      float a4[4], a2[2];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool hi = lane & 16;
        a4[k] = (hi ? dot[i][k + 4] : dot[i][k]) + __shfl_xor_sync(0xffffffffu, hi ? dot[i][k] : dot[i][k + 4], 16);
      }
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        const bool hi = lane & 8;
        a2[k] = (hi ? a4[k + 2] : a4[k]) + __shfl_xor_sync(0xffffffffu, hi ? a4[k] : a4[k + 2], 8);
      }
      const bool hi4 = lane & 4;
      float c = (hi4 ? a2[1] : a2[0]) + __shfl_xor_sync(0xffffffffu, hi4 ? a2[0] : a2[1], 4);
      c += __shfl_xor_sync(0xffffffffu, c, 2);
      c += __shfl_xor_sync(0xffffffffu, c, 1);
      if ((lane & 3) == 0 && (i == 0 || two)) X0[(size_t)(r0 + i) * 8 + (lane >> 2)] = c * rstd + s_fold[kCA + (lane >> 2)];
    }
  }
  __syncthreads();
  PCT(2);
  __threadfence();
  cg::this_grid().sync();
  PCT(3);

  // ------------------------------------------------------------------ phase 2, sequence by sequence
  if (row_lo >= row_hi) return;
  const int b_lo = row_lo / n_cur, b_hi = (row_hi - 1) / n_cur;
  for (int b = b_lo; b <= b_hi; ++b) {
    const int lo = max(row_lo, b * n_cur), hi = min(row_hi, (b + 1) * n_cur);
    // first row pair of this warp in the segment: in flight during the statistics below
    if (lane == 0 && lo + 2 * warp < hi) pc_prefetch(x, delta, (size_t)(lo + 2 * warp), min(2, hi - (lo + 2 * warp)), slot, bar);
    // Fovea softmax statistics of sequence b (all its kept rows, wherever they were computed):
    // thread = (channel k, part); two block reductions (max, then sum of exp)
    {
      // thread = (channel k, part); lanes of a warp hold 4 parts x 8 channels: two shuffle steps fold the parts of a
      // warp, the 16 warp results go through shared memory, 16 threads finish (16 terms each, fixed order)
      const int k = tid & 7, part = tid >> 3;
      const float* x0b = X0 + (size_t)b * n_cur * 8;
      float l[5];                                // this thread's logits (n_cur <= 320: at most 5 rows per thread)
      float m[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int r = part + i * (kThreads / 8);
        l[i] = r < n_cur ? __ldcg(x0b + r * 8 + k) * smooth : -INFINITY;
        if (r < lens_t) m[0] = fmaxf(m[0], l[i]); else m[1] = fmaxf(m[1], l[i]);
      }
#pragma unroll
      for (int o = 8; o <= 16; o <<= 1) {
        m[0] = fmaxf(m[0], __shfl_xor_sync(0xffffffffu, m[0], o));
        m[1] = fmaxf(m[1], __shfl_xor_sync(0xffffffffu, m[1], o));
      }
      __syncthreads();                           // s_red / s_ms of the previous sequence are no longer read
      if (lane < 8) { s_red[(warp * 8 + lane) * 2] = m[0]; s_red[(warp * 8 + lane) * 2 + 1] = m[1]; }
      __syncthreads();
      if (tid < 16) {
        const int dom = tid >> 3, kk = tid & 7;
        float M = -INFINITY;
        for (int w = 0; w < kWarps; ++w) M = fmaxf(M, s_red[(w * 8 + kk) * 2 + dom]);
        const int n_removed = dom == 1 ? lens_x - (n_cur - lens_t) : 0;
        if (n_removed > 0) M = fmaxf(M, smooth * s_fold[kCA + kk]);
        s_ms[(dom * 8 + kk) * 2] = M;
      }
      __syncthreads();
      float sm[2] = {0.f, 0.f};
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int r = part + i * (kThreads / 8);
        if (r < n_cur) {                         // (no dynamically indexed register arrays: they live in local memory)
          if (r < lens_t) sm[0] += expf(l[i] - s_ms[k * 2]);
          else sm[1] += expf(l[i] - s_ms[(8 + k) * 2]);
        }
      }
#pragma unroll
      for (int o = 8; o <= 16; o <<= 1) {
        sm[0] += __shfl_xor_sync(0xffffffffu, sm[0], o);
        sm[1] += __shfl_xor_sync(0xffffffffu, sm[1], o);
      }
      if (lane < 8) { s_red[2 * kThreads + (warp * 8 + lane) * 2] = sm[0]; s_red[2 * kThreads + (warp * 8 + lane) * 2 + 1] = sm[1]; }
      __syncthreads();
      if (tid < 16) {
        const int dom = tid >> 3, kk = tid & 7;
        float S = 0.f;
        for (int w = 0; w < kWarps; ++w) S += s_red[2 * kThreads + (w * 8 + kk) * 2 + dom];
        const int n_removed = dom == 1 ? lens_x - (n_cur - lens_t) : 0;
        if (n_removed > 0) S += (float)n_removed * expf(smooth * s_fold[kCA + kk] - s_ms[(dom * 8 + kk) * 2]);
        s_ms[(dom * 8 + kk) * 2 + 1] = 1.0f / S;
      }
      __syncthreads();
    }
    PCT(4);
    for (int r0 = lo + 2 * warp; r0 < hi; r0 += 2 * kWarps) {
      const bool two = r0 + 1 < hi;
      float4 v[2][6];
      mbar_wait(bar, uses & 1); ++uses;
      pc_read_row(slot, 0, lane, v[0]);
      if (two) pc_read_row(slot, 1, lane, v[1]);
      else {
#pragma unroll
        for (int j = 0; j < 6; ++j) v[1][j] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      __syncwarp();
      if (lane == 0 && r0 + 2 * kWarps < hi)
        pc_prefetch(x, delta, (size_t)(r0 + 2 * kWarps), min(2, hi - (r0 + 2 * kWarps)), slot, bar);
      // Fovea output of the two rows: lanes 0..7 -> row 0, lanes 8..15 -> row 1 (channel = lane & 7)
      float fk = 0.f;
      {
        const int i = lane >> 3, k = lane & 7;
        if (lane < 16 && (i == 0 || two)) {
          const int r = r0 + i - b * n_cur;
          const int pos = r < lens_t ? r : lens_t + gi[b * lens_x + (r - lens_t)];
          const int dom = r < lens_t ? 0 : 1;
          const float4 f0 = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + pos) * 8);
          const float4 f1 = *reinterpret_cast<const float4*>(F + ((size_t)b * lens_full + pos) * 8 + 4);
          const float f[8] = {f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, f1.z, f1.w};
          float var = s_fold[kC0], a = s_fold[kA0 + k];
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            float g = 0.f, g1 = 0.f;          // two chains: this runs on 16 lanes per warp and is pure latency
#pragma unroll
            for (int qq = 0; qq < 8; qq += 2) {
              g += s_fold[kG + p * 8 + qq] * f[qq];
              g1 += s_fold[kG + p * 8 + qq + 1] * f[qq + 1];
            }
            g += g1;
            var += f[p] * (g + 2.0f * s_fold[kH + p]);
            a += s_fold[kA + k * 8 + p] * f[p];
          }
          const float x1 = a * rsqrtf(fmaxf(var, 0.f) + kLnEps) + s_fold[kCD + k];
          const float x0 = __ldcg(X0 + (size_t)(r0 + i) * 8 + k);
          fk = expf(x0 * smooth - s_ms[(dom * 8 + k) * 2]) * s_ms[(dom * 8 + k) * 2 + 1] * x0 + x1;
          __syncwarp(0x0000ffffu >> (two ? 0 : 8));
          F[((size_t)b * lens_full + pos) * 8 + k] = fk;       // all 8 lanes of a row have read F[pos] before this
        }
      }
      float f[2][8];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) f[i][k] = __shfl_sync(0xffffffffu, fk, i * 8 + k);
      uint64_t f2[2][8], s2[2] = {0ull, 0ull}, q2[2] = {0ull, 0ull};    // packed fp32x2 FMAs (see phase 1)
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) f2[i][k] = pk2(f[i][k], f[i][k]);
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const float4 bias = s_b1[lane + 32 * j];
        uint64_t p2[2][2];
#pragma unroll
        for (int i = 0; i < 2; ++i) { p2[i][0] = pk2(bias.x, bias.y); p2[i][1] = pk2(bias.z, bias.w); }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 w = s_w1[k * 192 + lane + 32 * j];
          const uint64_t w01 = pk2(w.x, w.y), w23 = pk2(w.z, w.w);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            p2[i][0] = fma2(w01, f2[i][k], p2[i][0]);
            p2[i][1] = fma2(w23, f2[i][k], p2[i][1]);
          }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint64_t v01 = add2(pk2(v[i][j].x, v[i][j].y), p2[i][0]), v23 = add2(pk2(v[i][j].z, v[i][j].w), p2[i][1]);
          upk2(v01, v[i][j].x, v[i][j].y);
          upk2(v23, v[i][j].z, v[i][j].w);
          s2[i] = add2(s2[i], add2(v01, v23));
          q2[i] = fma2(v01, v01, q2[i]);          // sum of squares in the same pass: one reduction round instead of two
          q2[i] = fma2(v23, v23, q2[i]);
        }
      }
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        if (i == 1 && !two) break;
        const size_t row = (size_t)r0 + i;
        float4* xo = reinterpret_cast<float4*>(x + row * D);
#pragma unroll
        for (int j = 0; j < 6; ++j) xo[lane + 32 * j] = v[i][j];
      }
      // LayerNorm statistics of both rows from (sum, sum of squares): four butterflies in flight together.  var =
      // E[v^2] - mean^2 in fp32 (|mean| << rms in the residual stream: relative error ~1e-6, far below the fp16 output)
      float st4[4] = {hsum2(s2[0]), hsum2(s2[1]), hsum2(q2[0]), hsum2(q2[1])};
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        float t[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) t[c] = __shfl_xor_sync(0xffffffffu, st4[c], o);
#pragma unroll
        for (int c = 0; c < 4; ++c) st4[c] += t[c];
      }
      float sc[2], sh[2];                        // xn = (v * sc + sh) * g + be
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float mean = st4[i] * (1.0f / D);
        const float var = fmaxf(st4[2 + i] * (1.0f / D) - mean * mean, 0.f);
        sc[i] = rsqrtf(var + kLnEps);
        sh[i] = -mean * sc[i];
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) {              // norm1 scale / shift read once for both rows
        const float4 g = s_g1[lane + 32 * j];
        const float4 be = s_be1[lane + 32 * j];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if (i == 1 && !two) break;
          uint2 u;
          u.x = pack_op2(fmaf(fmaf(v[i][j].x, sc[i], sh[i]), g.x, be.x), fmaf(fmaf(v[i][j].y, sc[i], sh[i]), g.y, be.y));
          u.y = pack_op2(fmaf(fmaf(v[i][j].z, sc[i], sh[i]), g.z, be.z), fmaf(fmaf(v[i][j].w, sc[i], sh[i]), g.w, be.w));
          reinterpret_cast<uint2*>(xn + ((size_t)r0 + i) * D)[lane + 32 * j] = u;
        }
      }
    }
  }
  __syncthreads();
  PCT(5);
}

int launch_prompt_coop(float* x, const void* delta, int n_cur, const int* gi, float* F, float* X0, const float* wg_a,
                       const float* fold, const float* w1t, const float* b1, const float* gamma1, const float* beta1,
                       void* xn, int B, int lens_t, int lens_full, int num_sms, cudaStream_t st) {
  static int ctas_per_sm = 0;
  if (!ctas_per_sm) {
    if (cudaFuncSetAttribute(prompt_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, pc::kSmemBytes) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, prompt_coop_kernel, pc::kThreads, pc::kSmemBytes) !=
            cudaSuccess || ctas_per_sm < 1)
      return VIPT_ERR_LAUNCH;
  }
  if (!delta || n_cur < lens_t || n_cur > lens_full || lens_full > 5 * (pc::kThreads / 8)) return VIPT_ERR_ARG;
  const int total = B * n_cur;
  int grid = ctas_per_sm * num_sms;
  const int want = (total + 2 * pc::kWarps - 1) / (2 * pc::kWarps);     // at least one row pair per warp
  if (grid > want) grid = want;
  const op_t* d = reinterpret_cast<const op_t*>(delta);
  op_t* xno = reinterpret_cast<op_t*>(xn);
  void* args[] = {&x, &d, &B, &n_cur, &gi, &F, &X0, &wg_a, &fold, &w1t, &b1, &gamma1, &beta1, &xno, &lens_t, &lens_full};
  if (cudaLaunchCooperativeKernel((const void*)prompt_coop_kernel, dim3(grid), dim3(pc::kThreads), args, pc::kSmemBytes, st) !=
      cudaSuccess)
    return VIPT_ERR_LAUNCH;
#ifdef PC_TRACE
  static int launches = 0;
  if (++launches == 25 || launches == 30) {
    cudaDeviceSynchronize();
    static long long h[160 * 8];
    cudaMemcpyFromSymbol(h, g_pc_trace, sizeof(h));
    // clock64 is per SM: only differences inside a CTA mean something
    auto stat = [&](int e1, int e0, const char* what) {
      std::vector<long long> d;
      for (int c = 0; c < grid; ++c) d.push_back(h[c * 8 + e1] - h[c * 8 + e0]);
      std::sort(d.begin(), d.end());
      fprintf(stderr, "  %-28s min %7lld  p50 %7lld  p90 %7lld  max %7lld clk\n", what, d[0], d[d.size() / 2], d[d.size() * 9 / 10], d.back());
    };
    fprintf(stderr, "prompt_coop n_cur %d grid %d\n", n_cur, grid);
    stat(1, 0, "prologue (weights staged)"); stat(2, 1, "phase 1"); stat(3, 2, "grid barrier"); stat(5, 3, "phase 2 (stats + rows)");
    stat(5, 0, "whole CTA");
  }
#endif
  return VIPT_OK;
}

}  // namespace vipt
