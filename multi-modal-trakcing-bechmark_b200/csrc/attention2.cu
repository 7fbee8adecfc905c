// Persistent, warp-specialised joint template+search attention (ViPT/lib/models/layers/attn.py:33-59)
// for N > 128 tokens.  One CTA per SM loops over (sequence, head) items; K and V of an item are loaded
// once and reused by all its 128-query tiles; the next item's K is prefetched into a second buffer.
//
//   warp 0        TMA producer   Q tile (16 KB), K (double-buffered), V
//   warp 1        MMA issuer     S_c = Q K_c^T  (two key chunks, c = 0/1, fp32 in TMEM columns [0, nkp))
//                                O_c = P_c V_c  (TMEM columns 320 / 384)
//   warps 2-5     softmax of chunk 0: two passes over TMEM (max, then exp2 -> fp16 P in swizzled smem)
//   warps 6-9     softmax of chunk 1
//   warps 10-13   epilogue: out = (e0 O_0 + e1 O_1) / (e0 s_0 + e1 s_1), e_c = exp(m_c - m); CE row
//
// Each chunk keeps its own running maximum m_c and sum s_c (no rescaling of TMEM accumulators); the two
// partial products are combined exactly in the epilogue.  While a softmax group works on tile t the
// MMA warp has already issued S of tile t+1 for the other chunk and P V of the previous one, so the
// tensor pipe, the MUFU pipe and the TMA loads overlap.  The (B,12,N,N) probability tensor of the
// reference never exists; on candidate-elimination layers the softmax row of template token 27 over
// the search keys is emitted in fp32 (attn_blocks.py:44-53).
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

namespace at2 {
constexpr int kHd = 64;
constexpr int kQTile = 128;
constexpr int kMaxKeys = 320;
constexpr int kThreads = 448;
constexpr int kOCol0 = 320, kOCol1 = 384;
constexpr int kBoxBytes = 64 * 128;              // one TMA box: 64 tokens x 64 fp16
enum { Q_FULL = 0, Q_EMPTY, K_FULL0, K_FULL1, K_EMPTY0, K_EMPTY1, V_FULL, V_EMPTY, S_FULL0, S_FULL1, S_EMPTY0, S_EMPTY1, P_FULL0, P_FULL1,
       O_FULL0, O_FULL1, O_EMPTY, ST_FULL0, ST_FULL1, NUM_BARS };
__host__ __device__ constexpr int kboxes(int nkp) { return (nkp + 63) / 64; }
__host__ __device__ constexpr int off_k(int) { return kQTile * 128; }
__host__ __device__ constexpr int off_v(int nkp) { return off_k(nkp) + 2 * kboxes(nkp) * kBoxBytes; }
__host__ __device__ constexpr int off_p(int nkp) { return off_v(nkp) + kboxes(nkp) * kBoxBytes; }
__host__ __device__ constexpr int off_stats(int nkp) { return off_p(nkp) + kboxes(nkp) * kQTile * 128; }
__host__ __device__ constexpr int smem_bytes(int nkp) {
  return 1024 + off_stats(nkp) + 2 * 2 * kQTile * 2 * 4 + 2 * kMaxKeys * 4 + NUM_BARS * 8 + 16;
}
}  // namespace at2

__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__global__ void __launch_bounds__(at2::kThreads, 1)
attention2_kernel(const __grid_constant__ CUtensorMap tm, op_t* __restrict__ out, float* __restrict__ ce_prob, int B,
                  int N, int nkp, int c0, int ce_row, int lens_t, int ce_stride, float scale_log2) {
  using namespace at2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + off_k(nkp);
  uint8_t* sV = smem + off_v(nkp);
  uint8_t* sP = smem + off_p(nkp);
  float* s_stats = reinterpret_cast<float*>(smem + off_stats(nkp));      // [2 parity][2 chunk][128][2]
  float* s_ce = s_stats + 2 * 2 * kQTile * 2;                            // [2][320]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_ce + 2 * kMaxKeys);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + NUM_BARS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nbox = kboxes(nkp);
  const int kbytes = nbox * kBoxBytes;
  const int T = (N + kQTile - 1) / kQTile;
  const int n_items = B * 12;
  const int w0 = c0, w1 = nkp - c0;              // chunk widths (multiples of 32)

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm);
    for (int i = 0; i < NUM_BARS; ++i) {
      const uint32_t cnt = (i == P_FULL0 || i == P_FULL1 || i == O_EMPTY || i == S_EMPTY0 || i == S_EMPTY1) ? 4u : (i == ST_FULL0 || i == ST_FULL1) ? 8u : 1u;
      mbar_init(&bars[i], cnt);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer
      int g = 0, n = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
        const int b = item / 12, h = item % 12;
        if (n == 0) {
          mbar_expect_tx(&bars[K_FULL0], kbytes);
          for (int i = 0; i < nbox; ++i) tma_load_3d(sK + i * kBoxBytes, &tm, &bars[K_FULL0], 768 + h * kHd, i * 64, b);
        }
        const int nxt = item + gridDim.x;
        if (nxt < n_items) {                      // prefetch the next item's K into the other buffer
          const int kb = (n + 1) & 1, j = (n + 1) >> 1;
          mbar_wait(&bars[K_EMPTY0 + kb], (j & 1) ^ 1);
          mbar_expect_tx(&bars[K_FULL0 + kb], kbytes);
          for (int i = 0; i < nbox; ++i)
            tma_load_3d(sK + kb * kbytes + i * kBoxBytes, &tm, &bars[K_FULL0 + kb], 768 + (nxt % 12) * kHd, i * 64, nxt / 12);
        }
        for (int t = 0; t < T; ++t, ++g) {
          mbar_wait(&bars[Q_EMPTY], (g & 1) ^ 1);
          mbar_expect_tx(&bars[Q_FULL], 2 * kBoxBytes);
          tma_load_3d(sQ, &tm, &bars[Q_FULL], h * kHd, t * kQTile, b);
          tma_load_3d(sQ + kBoxBytes, &tm, &bars[Q_FULL], h * kHd, t * kQTile + 64, b);
          if (t == 0) {
            mbar_wait(&bars[V_EMPTY], (n & 1) ^ 1);
            mbar_expect_tx(&bars[V_FULL], kbytes);
            for (int i = 0; i < nbox; ++i) tma_load_3d(sV + i * kBoxBytes, &tm, &bars[V_FULL], 1536 + h * kHd, i * 64, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------------------------------------ MMA issuer
      const uint32_t idesc_s0 = make_idesc(kQTile, w0), idesc_s1 = make_idesc(kQTile, w1);
      const uint32_t idesc_pv = make_idesc(kQTile, kHd, /*b_mn_major=*/true);
      const uint64_t qd = make_sdesc_sw128(smem_u32(sQ));
      auto issue_pv = [&](int u, int c, bool first_pv_of_tile, bool first_of_item, bool release_v, int n_of_u) {
        if (first_pv_of_tile) {
          mbar_wait(&bars[O_EMPTY], (u & 1) ^ 1);          // epilogue of tile u-1 has drained O_0 / O_1
          if (first_of_item) mbar_wait(&bars[V_FULL], n_of_u & 1);
          tc_fence_after();
        }
        const int j0 = (c == 0 ? 0 : c0) / 16, nj = (c == 0 ? w0 : w1) / 16;
        for (int jj = 0; jj < nj; ++jj) {
          const int J = j0 + jj;
          const uint64_t pd = make_sdesc_sw128(smem_u32(sP + (J >> 2) * (kQTile * 128) + (J & 3) * 32));
          const uint64_t vd = make_sdesc_sw128(smem_u32(sV + J * 16 * 128));
          umma_f16(tmem + (c == 0 ? kOCol0 : kOCol1), pd, vd, idesc_pv, jj != 0);
        }
        umma_commit(&bars[O_FULL0 + c]);
        if (release_v) umma_commit(&bars[V_EMPTY]);      // last P V of the item: V may be overwritten
      };
      int g = 0, n = 0;
      bool prev_first = false, prev_last = false;
      int prev_n = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
        const int kb = n & 1;
        mbar_wait(&bars[K_FULL0 + kb], (n >> 1) & 1);
        for (int t = 0; t < T; ++t, ++g) {
          mbar_wait(&bars[Q_FULL], g & 1);
          tc_fence_after();
          // Four events per tile, served in whatever order they fire: S buffer c drained -> issue S(g, c);
          // P(g-1, c) written -> issue P V of the previous tile.
          bool need_s[2] = {true, true}, need_pv[2] = {g > 0, g > 0};
          bool pv_started = false;
          while (need_s[0] | need_s[1] | need_pv[0] | need_pv[1]) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (need_s[c] && (g == 0 || mbar_try_wait(&bars[S_EMPTY0 + c], (g - 1) & 1))) {
                tc_fence_after();
                const uint64_t kd = make_sdesc_sw128(smem_u32(sK + kb * kbytes + (c == 0 ? 0 : c0) * 128));
#pragma unroll
                for (int k = 0; k < kHd / 16; ++k)
                  umma_f16(tmem + (c == 0 ? 0 : c0), qd + 2 * k, kd + 2 * k, c == 0 ? idesc_s0 : idesc_s1, k != 0);
                umma_commit(&bars[S_FULL0 + c]);
                need_s[c] = false;
                if (!need_s[c ^ 1]) {
                  umma_commit(&bars[Q_EMPTY]);
                  if (t == T - 1) umma_commit(&bars[K_EMPTY0 + kb]);
                }
              }
              if (need_pv[c] && mbar_try_wait(&bars[P_FULL0 + c], (g - 1) & 1)) {
                tc_fence_after();
                issue_pv(g - 1, c, !pv_started, prev_first, prev_last && !need_pv[c ^ 1], prev_n);
                pv_started = true;
                need_pv[c] = false;
              }
            }
          }
          prev_first = t == 0; prev_last = t == T - 1; prev_n = n;
        }
      }
      if (g > 0) {
        for (int c = 0; c < 2; ++c) {
          mbar_wait(&bars[P_FULL0 + c], (g - 1) & 1);
          tc_fence_after();
          issue_pv(g - 1, c, c == 0, prev_first, false, prev_n);
        }
      }
    }
  } else if (warp < 10) {
    // -------------------------------------------------------------- softmax groups
    const int c = (warp - 2) >> 2;
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const int key0 = c == 0 ? 0 : c0;
    const int ngroups = (c == 0 ? w0 : w1) / 32;
    const uint32_t t_row = tmem + ((uint32_t)(wr * 32) << 16);
    int g = 0, n = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
      for (int t = 0; t < T; ++t, ++g) {
        const int q0 = t * kQTile;
        mbar_wait(&bars[S_FULL0 + c], g & 1);
        tc_fence_after();
        if (q0 + wr * 32 < N) {                  // warp-uniform: at least one valid query row
          // TMEM loads are software-pipelined: the load of group i+1 is in flight while group i is processed
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // 4 independent chains (ILP)
          auto max_group = [&](const uint32_t (&v)[32], int col) {
            if (col + 32 <= N) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                  m4[i] = fmaxf(m4[i], fmaxf(__uint_as_float(v[j + 2 * i]), __uint_as_float(v[j + 2 * i + 1])));
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col + j < N) m4[j & 3] = fmaxf(m4[j & 3], __uint_as_float(v[j]));
            }
          };
          uint32_t va[32], vb[32];
          tmem_ld32(t_row + key0, va);
          for (int gi = 0; gi < ngroups; gi += 2) {
            tmem_ld_wait();
            if (gi + 1 < ngroups) tmem_ld32(t_row + key0 + (gi + 1) * 32, vb);
            max_group(va, key0 + gi * 32);
            if (gi + 1 < ngroups) {
              tmem_ld_wait();
              if (gi + 2 < ngroups) tmem_ld32(t_row + key0 + (gi + 2) * 32, va);
              max_group(vb, key0 + (gi + 1) * 32);
            }
          }
          tmem_ld32(t_row + key0, va);            // first group of pass 2, in flight during the wait below
          const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
          if (g > 0) mbar_wait(&bars[O_FULL0 + c], (g - 1) & 1);    // P V of the previous tile has consumed P_c
          const float msl = m * scale_log2;
          const bool ce_thread = ce_prob != nullptr && t == 0 && row == ce_row;
          float sum4[4] = {0.f, 0.f, 0.f, 0.f};
          auto exp_group = [&](const uint32_t (&v)[32], int col) {
            float p[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) p[j] = ex2_approx(fmaf(__uint_as_float(v[j]), scale_log2, -msl));
            if (col + 32 > N) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col + j >= N) p[j] = 0.f;
            }
#pragma unroll
            for (int j = 0; j < 32; ++j) sum4[j & 3] += p[j];
            if (ce_thread) {
#pragma unroll
              for (int j = 0; j < 32; ++j) s_ce[(n & 1) * kMaxKeys + col + j] = p[j];
            }
            uint8_t* prow = sP + (col >> 6) * (kQTile * 128) + row * 128;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 u;
              u.x = pack_op2(p[q * 8 + 0], p[q * 8 + 1]);
              u.y = pack_op2(p[q * 8 + 2], p[q * 8 + 3]);
              u.z = pack_op2(p[q * 8 + 4], p[q * 8 + 5]);
              u.w = pack_op2(p[q * 8 + 6], p[q * 8 + 7]);
              const int chunk = (((col >> 5) & 1) * 4 + q) ^ (row & 7);
              *reinterpret_cast<uint4*>(prow + chunk * 16) = u;
            }
          };
          auto s_drained = [&]() {               // S_c is in registers: the MMA warp may overwrite it
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[S_EMPTY0 + c]);
          };
          for (int gi = 0; gi < ngroups; gi += 2) {
            tmem_ld_wait();
            if (gi + 1 < ngroups) tmem_ld32(t_row + key0 + (gi + 1) * 32, vb); else s_drained();
            exp_group(va, key0 + gi * 32);
            if (gi + 1 < ngroups) {
              tmem_ld_wait();
              if (gi + 2 < ngroups) tmem_ld32(t_row + key0 + (gi + 2) * 32, va); else s_drained();
              exp_group(vb, key0 + (gi + 1) * 32);
            }
          }
          float* st = s_stats + (((g & 1) * 2 + c) * kQTile + row) * 2;
          st[0] = m;
          st[1] = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
        }
        else {
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[S_EMPTY0 + c]);
        }
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&bars[P_FULL0 + c]);
          mbar_arrive(&bars[ST_FULL0 + (g & 1)]);
        }
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue group
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const int et = threadIdx.x - 320;
    const uint32_t t_row = tmem + ((uint32_t)(wr * 32) << 16);
    int g = 0, n = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++n) {
      const int b = item / 12, h = item % 12;
      for (int t = 0; t < T; ++t, ++g) {
        const int q0 = t * kQTile;
        mbar_wait(&bars[ST_FULL0 + (g & 1)], (g >> 1) & 1);
        mbar_wait(&bars[O_FULL0], g & 1);
        mbar_wait(&bars[O_FULL1], g & 1);
        tc_fence_after();
        const float* st0 = s_stats + (((g & 1) * 2 + 0) * kQTile) * 2;
        const float* st1 = s_stats + (((g & 1) * 2 + 1) * kQTile) * 2;
        if (q0 + wr * 32 < N) {
          const float m0 = st0[row * 2], s0 = st0[row * 2 + 1], m1 = st1[row * 2], s1 = st1[row * 2 + 1];
          const float m = fmaxf(m0, m1);
          const float e0 = ex2_approx((m0 - m) * scale_log2), e1 = ex2_approx((m1 - m) * scale_log2);
          const float inv = 1.0f / (s0 * e0 + s1 * e1);
          const float a0 = e0 * inv, a1 = e1 * inv;
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t o0[32], o1[32];
            tmem_ld32(t_row + kOCol0 + hf * 32, o0);
            tmem_ld32(t_row + kOCol1 + hf * 32, o1);
            tmem_ld_wait();
            if (q0 + row < N) {
              op_t* o = out + ((size_t)b * N + q0 + row) * 768 + h * kHd + hf * 32;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  f[j] = __uint_as_float(o0[q * 8 + j]) * a0 + __uint_as_float(o1[q * 8 + j]) * a1;
                uint4 u;
                u.x = pack_op2(f[0], f[1]); u.y = pack_op2(f[2], f[3]);
                u.z = pack_op2(f[4], f[5]); u.w = pack_op2(f[6], f[7]);
                *reinterpret_cast<uint4*>(o + q * 8) = u;
              }
            }
          }
        }
        if (ce_prob != nullptr && t == 0) {      // normalised CE row, coalesced
          const float m0 = st0[ce_row * 2], s0 = st0[ce_row * 2 + 1], m1 = st1[ce_row * 2], s1 = st1[ce_row * 2 + 1];
          const float m = fmaxf(m0, m1);
          const float e0 = ex2_approx((m0 - m) * scale_log2), e1 = ex2_approx((m1 - m) * scale_log2);
          const float inv = 1.0f / (s0 * e0 + s1 * e1);
          float* ce_dst = ce_prob + ((size_t)b * 12 + h) * ce_stride;
          for (int col = lens_t + et; col < N; col += 128)
            ce_dst[col - lens_t] = s_ce[(n & 1) * kMaxKeys + col] * (col < c0 ? e0 : e1) * inv;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[O_EMPTY]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

int attention2_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride, int num_sms,
                      cudaStream_t st) {
  using namespace at2;
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attention2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(kMaxKeys)) !=
        cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  const int nkp = op.nkp;
  const int c0 = 32 * ((nkp + 63) / 64);
  if (nkp - c0 < 32 || c0 >= op.N || op.N > kMaxKeys) return VIPT_ERR_ARG;
  const int items = op.B * 12;
  const int grid = items < num_sms ? items : num_sms;
  const float scale_log2 = 0.125f * 1.4426950408889634f;
  attention2_kernel<<<grid, kThreads, smem_bytes(nkp), st>>>(op.tm, reinterpret_cast<op_t*>(out), ce_prob, op.B, op.N, nkp,
                                                             c0, ce_row, lens_t, ce_stride, scale_log2);
  return cudaGetLastError() == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
