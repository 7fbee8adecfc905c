// Persistent, warp-specialised joint template+search attention (ViPT/lib/models/layers/attn.py:33-59) for
// N > 128 tokens, third generation: the score MMAs run AHEAD of the softmax instead of in lock-step with it.
//
// One CTA per SM walks a contiguous range of (sequence, head, 128-query tile) units; K and V of a (sequence, head)
// item are loaded once per CTA that touches it and reused by its tiles.  The keys of a tile are cut into <= 3 chunks of <= 128 keys.  Chunks of all tiles and
// items form one stream k = 0, 1, 2, ...:
//
//   warp 0        TMA producer   Q tiles (double-buffered), K (double-buffered, next item prefetched), V
//   warp 1        score MMAs     S_k = Q K_k^T into TMEM buffer k % 3 (3 x 128 fp32 columns), up to three chunks
//                                ahead of the softmax
//   warp 14       value MMAs     O_k = P_k V_k into TMEM buffer k % 2 (2 x 64 columns); a second issuing thread,
//                                because one thread issuing ~30 small MMAs per tile was the bottleneck
//   warps 2-5     softmax of even chunks: max, exp2 -> fp16 P_k in swizzled smem, row sum
//   warps 6-9     softmax of odd chunks
//   warps 10-13   epilogue: drains O_k into registers and folds it into the running (m, l, acc) of the
//                 row (online softmax in registers -- TMEM accumulators are never rescaled); after the
//                 last chunk of a tile writes acc / l; CE layers also emit the fp32 softmax row of
//                 template token 27 over the search keys (attn_blocks.py:44-53)
//
// Why: in the lock-step version (attention2.cu) the chain  softmax(tile g) -> S_EMPTY -> Q(g+1) arrives ->
// S(g+1) MMA -> S_FULL -> softmax(g+1)  was serial; measured with clock64 the softmax warps spent 40-60 % of
// a tile waiting for S.  Here S_k only waits for the softmax of chunk k-3 and Q tiles are prefetched, so the
// softmax groups run back to back.
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

namespace at3 {
constexpr int kHd = 64;
constexpr int kQTile = 128;
constexpr int kMaxKeys = 320;
constexpr int kThreads = 480;
constexpr int kSBufCols = 128;                   // S buffers at TMEM columns 0 / 128 / 256
constexpr int kOCol = 384;                       // O buffers at 384 / 448
constexpr int kBoxBytes = 64 * 128;              // one TMA box: 64 tokens x 64 fp16
constexpr int kQBytes = kQTile * 128;
constexpr int kPBytes = 2 * kQTile * 128;        // one P buffer: 128 rows x 128 keys fp16
enum { Q_FULL = 0, Q_EMPTY = 2, K_FULL = 4, K_EMPTY = 6, V_FULL = 8, V_EMPTY = 9, S_FULL = 10, S_EMPTY = 13, P_FULL = 16,
       O_FULL = 18, O_EMPTY = 20, ST_FULL = 22, NUM_BARS = 26 };
__host__ __device__ constexpr int kboxes(int nkp) { return (nkp + 63) / 64; }
__host__ __device__ constexpr int off_k(int) { return 2 * kQBytes; }
__host__ __device__ constexpr int off_v(int nkp) { return off_k(nkp) + 2 * kboxes(nkp) * kBoxBytes; }
__host__ __device__ constexpr int off_p(int nkp) { return off_v(nkp) + kboxes(nkp) * kBoxBytes; }
__host__ __device__ constexpr int off_stats(int nkp) { return off_p(nkp) + 2 * kPBytes; }
constexpr int kStatsFloats = 4 * kQTile * 2;     // [4 chunks in flight][128 rows][m, sum]
constexpr int kCeFloats = 2 * kMaxKeys + 2 * 4 + 2 * 2;     // CE row probabilities, chunk maxima, final (m, 1/l)
__host__ __device__ constexpr int smem_bytes(int nkp) {
  return 1024 + off_stats(nkp) + (kStatsFloats + kCeFloats) * 4 + NUM_BARS * 8 + 16;
}
struct Chunks {
  int nc;
  int base[3];
  int w[3];
};
}  // namespace at3

namespace {

__device__ __forceinline__ float ex2f(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// tcgen05.mma with both smem descriptors given as (low word, shared high word): the issuing thread only
// adds to the 14-bit start-address field
__device__ __forceinline__ void umma_f16_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %5, 0;\n\t"
      "mov.b64 da, {%1, %3};\n\t"
      "mov.b64 db, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}\n"
      ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// position of chunk k of this CTA's stream
struct Pos {
  int n, t, j, g;
  __device__ __forceinline__ void advance(int NC, int T) {
    if (++j == NC) {
      j = 0; ++g;
      if (++t == T) { t = 0; ++n; }
    }
  }
};

}  // namespace

__global__ void __launch_bounds__(at3::kThreads, 1)
attention3_kernel(const __grid_constant__ CUtensorMap tm, op_t* __restrict__ out, float* __restrict__ ce_prob, int B,
                  int N, int nkp, const at3::Chunks ch, int ce_row, int lens_t, int ce_stride, float scale_log2) {
  using namespace at3;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + off_k(nkp);
  uint8_t* sV = smem + off_v(nkp);
  uint8_t* sP = smem + off_p(nkp);
  float* s_stats = reinterpret_cast<float*>(smem + off_stats(nkp));
  float* s_ce = s_stats + kStatsFloats;          // [2 items][320]
  float* s_cem = s_ce + 2 * kMaxKeys;            // [2 items][4] chunk maxima of the CE row
  float* s_cefin = s_cem + 2 * 4;                // [2 items][m, 1/l]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_cefin + 2 * 2);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + NUM_BARS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nbox = kboxes(nkp);
  const int kbytes = nbox * kBoxBytes;
  const int T = (N + kQTile - 1) / kQTile;
  const int NC = ch.nc;
  // chunk geometry by select, never by a dynamically indexed parameter array (that would go through local memory)
  const int cb1 = ch.base[1], cb2 = ch.base[2], cw0 = ch.w[0], cw1 = ch.w[1], cw2 = ch.w[2];
  auto chunk_base = [&](int j) { return j == 0 ? 0 : (j == 1 ? cb1 : cb2); };
  auto chunk_w = [&](int j) { return j == 0 ? cw0 : (j == 1 ? cw1 : cw2); };
  // Work is cut at (item, query tile) granularity: CTA c owns the contiguous unit range [u0, u1) of the item-major list of
  // B * 12 * T units (64 sequences at N = 320: 2304 units over 148 CTAs = 15.6, so every CTA gets 15 or 16 tiles; with whole
  // items it was 768 / 148 = 5.2, i.e. 18 tiles on some CTAs and 15 on the others).  A CTA whose range starts or ends inside
  // an item loads that item's K / V like any other and simply runs fewer of its tiles.
  const int n_units = B * 12 * T;
  const int u0 = (int)(((long long)blockIdx.x * n_units) / (int)gridDim.x);
  const int u1 = (int)(((long long)(blockIdx.x + 1) * n_units) / (int)gridDim.x);
  const int item0 = u0 / T, t0 = u0 - item0 * T;      // first item of this CTA and the tile it starts at
  const int n_local = u1 > u0 ? (u1 - 1) / T - item0 + 1 : 0;      // items touched
  const int per_item = T * NC;
  const int total = (u1 - u0) * NC;                   // chunks in this CTA's stream

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm);
    for (int i = 0; i < NUM_BARS; ++i) {
      const bool four = (i >= S_EMPTY && i < S_EMPTY + 3) || (i >= P_FULL && i < P_FULL + 2) ||
                        (i >= O_EMPTY && i < O_EMPTY + 2) || (i >= ST_FULL && i < ST_FULL + 4);
      mbar_init(&bars[i], four ? 4u : 1u);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0 && total > 0) {
      // ------------------------------------------------------------ TMA producer
      int g = 0;
      for (int n = 0; n < n_local; ++n) {
        const int item = item0 + n;
        const int b = item / 12, h = item % 12;
        const int ts = n == 0 ? t0 : 0;                  // tiles [ts, te) of this item belong to this CTA
        const int te = min(T, u1 - item * T);
        const int tv = te - ts > 1 ? ts + 1 : ts;        // V is requested after the second Q tile (needed later than both)
        if (n == 0) {
          mbar_expect_tx(&bars[K_FULL], kbytes);
          for (int i = 0; i < nbox; ++i) tma_load_3d(sK + i * kBoxBytes, &tm, &bars[K_FULL], 768 + h * kHd, i * 64, b);
        }
        for (int t = ts; t < te; ++t, ++g) {
          const int qb = g & 1;
          mbar_wait(&bars[Q_EMPTY + qb], ((g >> 1) & 1) ^ 1);
          mbar_expect_tx(&bars[Q_FULL + qb], 2 * kBoxBytes);
          tma_load_3d(sQ + qb * kQBytes, &tm, &bars[Q_FULL + qb], h * kHd, t * kQTile, b);
          tma_load_3d(sQ + qb * kQBytes + kBoxBytes, &tm, &bars[Q_FULL + qb], h * kHd, t * kQTile + 64, b);
          if (t == ts && n + 1 < n_local) {      // prefetch the next item's K into the other buffer
            const int nxt = item + 1, kb = (n + 1) & 1;
            mbar_wait(&bars[K_EMPTY + kb], (((n + 1) >> 1) & 1) ^ 1);
            mbar_expect_tx(&bars[K_FULL + kb], kbytes);
            for (int i = 0; i < nbox; ++i)
              tma_load_3d(sK + kb * kbytes + i * kBoxBytes, &tm, &bars[K_FULL + kb], 768 + (nxt % 12) * kHd, i * 64, nxt / 12);
          }
          if (t == tv) {
            mbar_wait(&bars[V_EMPTY], (n & 1) ^ 1);
            mbar_expect_tx(&bars[V_FULL], kbytes);
            for (int i = 0; i < nbox; ++i) tma_load_3d(sV + i * kBoxBytes, &tm, &bars[V_FULL], 1536 + h * kHd, i * 64, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    {
      // ------------------------------------------------------------ score MMAs: S_k = Q K_k^T, up to 3 chunks ahead.
      // The whole warp runs the loop (warp-uniform control flow and addresses, so descriptor math stays in
      // uniform registers); one elected lane issues the MMAs and commits.
      const uint32_t is0 = make_idesc(kQTile, cw0), is1 = make_idesc(kQTile, cw1 ? cw1 : cw0), is2 = make_idesc(kQTile, cw2 ? cw2 : cw0);
      const uint64_t d0 = make_sdesc_sw128(smem_u32(sQ));
      const uint32_t hi = (uint32_t)(d0 >> 32), q_lo = (uint32_t)d0;
      const uint32_t k_lo = (uint32_t)make_sdesc_sw128(smem_u32(sK));
      Pos ps = {0, t0, 0, 0};
      for (int ks = 0; ks < total; ++ks) {
        const int sb = ks % 3;
        if (ps.j == 0) {
          if (ps.t == 0 || ks == 0) mbar_wait(&bars[K_FULL + (ps.n & 1)], (ps.n >> 1) & 1);
          mbar_wait(&bars[Q_FULL + (ps.g & 1)], (ps.g >> 1) & 1);
        }
        if (ks >= 3) mbar_wait(&bars[S_EMPTY + sb], ((ks / 3) & 1) ^ 1);
        tc_fence_after();
        const uint32_t qa = q_lo + (uint32_t)(((ps.g & 1) * kQBytes) >> 4);
        const uint32_t ka = k_lo + (uint32_t)(((ps.n & 1) * kbytes + chunk_base(ps.j) * 128) >> 4);
        const uint32_t dcol = tmem + sb * kSBufCols, id = ps.j == 0 ? is0 : (ps.j == 1 ? is1 : is2);
        if (elect_one()) {
          umma_f16_lo(dcol, qa, ka, hi, id, 0);
          umma_f16_lo(dcol, qa + 2, ka + 2, hi, id, 1);
          umma_f16_lo(dcol, qa + 4, ka + 4, hi, id, 1);
          umma_f16_lo(dcol, qa + 6, ka + 6, hi, id, 1);
          umma_commit(&bars[S_FULL + sb]);
          if (ps.j == NC - 1) {
            umma_commit(&bars[Q_EMPTY + (ps.g & 1)]);
            if (ps.t == T - 1) umma_commit(&bars[K_EMPTY + (ps.n & 1)]);
          }
        }
        __syncwarp();
        ps.advance(NC, T);
      }
    }
  } else if (warp == 14) {
    {
      // ------------------------------------------------------------ value MMAs: O_k = P_k V_k (own issuing thread)
      const uint32_t idesc_pv = make_idesc(kQTile, kHd, /*b_mn_major=*/true);
      const uint64_t d0 = make_sdesc_sw128(smem_u32(sP));
      const uint32_t hi = (uint32_t)(d0 >> 32), p_lo = (uint32_t)d0;
      const uint32_t v_lo = (uint32_t)make_sdesc_sw128(smem_u32(sV));
      Pos pp = {0, t0, 0, 0};
      for (int kp = 0; kp < total; ++kp) {
        const int ob = kp & 1;
        if (pp.j == 0 && (pp.t == 0 || kp == 0)) mbar_wait(&bars[V_FULL], pp.n & 1);
        if (kp >= 2) mbar_wait(&bars[O_EMPTY + ob], ((kp >> 1) & 1) ^ 1);
        mbar_wait(&bars[P_FULL + ob], (kp >> 1) & 1);
        tc_fence_after();
        const int nj = chunk_w(pp.j) >> 4;
        const uint32_t pa = p_lo + (uint32_t)((ob * kPBytes) >> 4);
        const uint32_t va = v_lo + (uint32_t)((chunk_base(pp.j) * 128) >> 4);
        const uint32_t dcol = tmem + kOCol + ob * kHd;
        if (elect_one()) {
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            if (jj < nj)                            // A: 16 keys = 32 B inside the 128-B row, next 64-key block 16 KB on
              umma_f16_lo(dcol, pa + (uint32_t)(((jj >> 2) * (kQTile * 128) + (jj & 3) * 32) >> 4), va + jj * ((16 * 128) >> 4),
                          hi, idesc_pv, jj != 0);
          }
          umma_commit(&bars[O_FULL + ob]);
          if (pp.j == NC - 1 && pp.t == T - 1) umma_commit(&bars[V_EMPTY]);    // last P V of the item
        }
        __syncwarp();
        pp.advance(NC, T);
      }
    }
  } else if (warp < 10) {
    // -------------------------------------------------------------- softmax groups (even / odd chunks)
    const int grp = (warp - 2) >> 2;
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(wr * 32) << 16);
    const uint32_t pbuf = smem_u32(sP + grp * kPBytes) + row * 128;      // 32-bit shared address: STS, not generic ST
    for (int k = grp; k < total; k += 2) {
      const int kk = k + t0 * NC;                 // position in the item-aligned chunk stream
      const int n = kk / per_item, r = kk - n * per_item;
      const int t = r / NC, j = r - t * NC;
      const int q0 = t * kQTile;
      const int base = chunk_base(j);
      const int ngroups = chunk_w(j) >> 5;
      const int sb = k % 3;
      const uint32_t t_row = t_lane + sb * kSBufCols;
      mbar_wait(&bars[S_FULL + sb], (k / 3) & 1);
      tc_fence_after();
      if (q0 + wr * 32 < N) {                    // warp-uniform: at least one valid query row
        // TMEM loads are software-pipelined: the load of group i+1 is in flight while group i is processed
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // 4 independent chains (ILP)
        auto max_group = [&](const uint32_t (&v)[32], int col) {
          if (col + 32 <= N) {
#pragma unroll
            for (int jx = 0; jx < 32; jx += 8) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                m4[i] = fmaxf(m4[i], fmaxf(__uint_as_float(v[jx + 2 * i]), __uint_as_float(v[jx + 2 * i + 1])));
            }
          } else {
#pragma unroll
            for (int jx = 0; jx < 32; ++jx)
              if (col + jx < N) m4[jx & 3] = fmaxf(m4[jx & 3], __uint_as_float(v[jx]));
          }
        };
        uint32_t va[32], vb[32];
        tmem_ld32(t_row, va);
        for (int gi = 0; gi < ngroups; gi += 2) {
          tmem_ld_wait();
          if (gi + 1 < ngroups) tmem_ld32(t_row + (gi + 1) * 32, vb);
          max_group(va, base + gi * 32);
          if (gi + 1 < ngroups) {
            tmem_ld_wait();
            if (gi + 2 < ngroups) tmem_ld32(t_row + (gi + 2) * 32, va);
            max_group(vb, base + (gi + 1) * 32);
          }
        }
        tmem_ld32(t_row, va);                     // first group of pass 2, in flight during the wait below
        const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        if (k >= 2) mbar_wait(&bars[O_FULL + (k & 1)], ((k >> 1) & 1) ^ 1);    // P V of chunk k-2 has consumed this P buffer
        const float msl = m * scale_log2;
        const bool ce_thread = ce_prob != nullptr && t == 0 && row == ce_row;
        float sum4[4] = {0.f, 0.f, 0.f, 0.f};
        auto exp_group = [&](const uint32_t (&v)[32], int crel) {
          const int col = base + crel;
          float p[32];
#pragma unroll
          for (int jx = 0; jx < 32; ++jx) p[jx] = ex2f(fmaf(__uint_as_float(v[jx]), scale_log2, -msl));
          if (col + 32 > N) {
#pragma unroll
            for (int jx = 0; jx < 32; ++jx)
              if (col + jx >= N) p[jx] = 0.f;
          }
#pragma unroll
          for (int jx = 0; jx < 32; ++jx) sum4[jx & 3] += p[jx];
          if (ce_thread) {
#pragma unroll
            for (int jx = 0; jx < 32; ++jx) s_ce[(n & 1) * kMaxKeys + col + jx] = p[jx];
          }
          const uint32_t prow = pbuf + (crel >> 6) * (kQTile * 128);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int chunk = (((crel >> 5) & 1) * 4 + q) ^ (row & 7);
            sts128(prow + chunk * 16, pack_op2(p[q * 8 + 0], p[q * 8 + 1]), pack_op2(p[q * 8 + 2], p[q * 8 + 3]),
                   pack_op2(p[q * 8 + 4], p[q * 8 + 5]), pack_op2(p[q * 8 + 6], p[q * 8 + 7]));
          }
        };
        auto s_drained = [&]() {                 // S_k is in registers: the MMA warp may overwrite the buffer
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[S_EMPTY + sb]);
        };
        for (int gi = 0; gi < ngroups; gi += 2) {
          tmem_ld_wait();
          if (gi + 1 < ngroups) tmem_ld32(t_row + (gi + 1) * 32, vb); else s_drained();
          exp_group(va, gi * 32);
          if (gi + 1 < ngroups) {
            tmem_ld_wait();
            if (gi + 2 < ngroups) tmem_ld32(t_row + (gi + 2) * 32, va); else s_drained();
            exp_group(vb, (gi + 1) * 32);
          }
        }
        float* st = s_stats + ((k & 3) * kQTile + row) * 2;
        st[0] = m;
        st[1] = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
        if (ce_thread) s_cem[(n & 1) * 4 + j] = m;
      } else {
        // no valid row: still keep the barrier phases in step (the wait bounds how far this warp can run ahead)
        if (k >= 2) mbar_wait(&bars[O_FULL + (k & 1)], ((k >> 1) & 1) ^ 1);
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[S_EMPTY + sb]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&bars[ST_FULL + (k & 3)]);      // four deep: the epilogue may lag the softmax by up to 3 chunks
        mbar_arrive(&bars[P_FULL + (k & 1)]);
      }
    }
  } else {
    // -------------------------------------------------------------- epilogue group: online softmax in registers
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const int et = threadIdx.x - 320;
    const uint32_t t_lane = tmem + ((uint32_t)(wr * 32) << 16);
    float acc[kHd];
    float m_run = 0.f, l_run = 0.f;
    Pos p = {0, t0, 0, 0};
    for (int k = 0; k < total; ++k) {
      const int ob = k & 1;
      const int q0 = p.t * kQTile;
      const int item = item0 + p.n;
      const int b = item / 12, h = item % 12;
      mbar_wait(&bars[ST_FULL + (k & 3)], (k >> 2) & 1);
      mbar_wait(&bars[O_FULL + ob], (k >> 1) & 1);
      tc_fence_after();
      const bool valid = q0 + wr * 32 < N;
      if (valid) {
        const float* st = s_stats + ((k & 3) * kQTile + row) * 2;
        const float mc = st[0], sc = st[1];
        float a = 0.f, bb = 1.f;
        if (p.j == 0) {
          m_run = mc; l_run = sc;
        } else {
          const float mn = fmaxf(m_run, mc);
          a = ex2f((m_run - mn) * scale_log2);
          bb = ex2f((mc - mn) * scale_log2);
          l_run = l_run * a + sc * bb;
          m_run = mn;
        }
        // two halves through one 32-register buffer (acc[64] + 64 loaded values would spill)
        uint32_t o[32];
        tmem_ld32(t_lane + kOCol + ob * kHd, o);
        tmem_ld_wait();
        if (p.j == 0) {
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(o[i]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[i] = fmaf(acc[i], a, __uint_as_float(o[i]) * bb);
        }
        tmem_ld32(t_lane + kOCol + ob * kHd + 32, o);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[O_EMPTY + ob]);       // O_k is in registers
        if (p.j == 0) {
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[32 + i] = __uint_as_float(o[i]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[32 + i] = fmaf(acc[32 + i], a, __uint_as_float(o[i]) * bb);
        }
      } else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[O_EMPTY + ob]);
      }
      if (p.j == NC - 1) {
        if (valid && q0 + row < N) {
          const float inv = 1.0f / l_run;
          op_t* o = out + ((size_t)b * N + q0 + row) * 768 + h * kHd;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            uint4 u;
            u.x = pack_op2(acc[q * 8 + 0] * inv, acc[q * 8 + 1] * inv);
            u.y = pack_op2(acc[q * 8 + 2] * inv, acc[q * 8 + 3] * inv);
            u.z = pack_op2(acc[q * 8 + 4] * inv, acc[q * 8 + 5] * inv);
            u.w = pack_op2(acc[q * 8 + 6] * inv, acc[q * 8 + 7] * inv);
            *reinterpret_cast<uint4*>(o + q * 8) = u;
          }
        }
        if (ce_prob != nullptr && p.t == 0) {    // normalised CE row, coalesced
          float* fin = s_cefin + (p.n & 1) * 2;
          if (row == ce_row) { fin[0] = m_run; fin[1] = 1.0f / l_run; }
          asm volatile("bar.sync 1, 128;" ::: "memory");
          const float mf = fin[0], inv = fin[1];
          const float* cm = s_cem + (p.n & 1) * 4;
          const float e0 = ex2f((cm[0] - mf) * scale_log2);
          const float e1 = NC > 1 ? ex2f((cm[1] - mf) * scale_log2) : 0.f;
          const float e2 = NC > 2 ? ex2f((cm[2] - mf) * scale_log2) : 0.f;
          float* ce_dst = ce_prob + ((size_t)b * 12 + h) * ce_stride;
          for (int col = lens_t + et; col < N; col += 128) {
            const float e = (NC > 2 && col >= cb2) ? e2 : (NC > 1 && col >= cb1) ? e1 : e0;
            ce_dst[col - lens_t] = s_ce[(p.n & 1) * kMaxKeys + col] * e * inv;
          }
        }
      }
      p.advance(NC, T);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

int attention3_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride, int num_sms,
                      cudaStream_t st) {
  using namespace at3;
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attention3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(kMaxKeys)) !=
        cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  const int nkp = op.nkp;
  if (op.N > kMaxKeys || op.N <= kQTile || nkp % 32) return VIPT_ERR_ARG;
  // <= 3 chunks of <= 128 keys, widths multiples of 32 and as even as possible (the two softmax groups
  // alternate over the chunk stream)
  Chunks ch = {};
  ch.nc = (nkp + 127) / 128;
  const int units = nkp / 32;
  int basek = 0;
  for (int j = 0; j < 3; ++j) {
    const int u = j < ch.nc ? units / ch.nc + (j < units % ch.nc ? 1 : 0) : 0;
    ch.base[j] = basek;
    ch.w[j] = u * 32;
    basek += u * 32;
  }
  if (ch.base[ch.nc - 1] >= op.N) return VIPT_ERR_ARG;        // every chunk holds at least one real key
  const int items = op.B * 12;
  const int grid = items < num_sms ? items : num_sms;
  const float scale_log2 = 0.125f * 1.4426950408889634f;
  return launch_pdl(attention3_kernel, grid, kThreads, smem_bytes(nkp), st, op.tm, reinterpret_cast<op_t*>(out), ce_prob, op.B,
                    op.N, nkp, ch, ce_row, lens_t, ce_stride, scale_log2) == cudaSuccess
             ? VIPT_OK
             : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
