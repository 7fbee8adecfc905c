// Persistent, warp-specialised joint template+search attention (ViPT/lib/models/layers/attn.py:33-59) for N > 128
// tokens.  One CTA per SM walks a contiguous range of (sequence, head, 128-query tile) units; the keys of a tile are
// cut into <= 3 chunks of <= 128, and the chunks of all tiles form one stream k = 0, 1, 2, ...: S_k = Q K_k^T runs up
// to three chunks ahead of the softmax in TMEM; P_k goes through swizzled shared memory; O_k = P_k V_k is drained per
// chunk into an online softmax held in the epilogue warps' registers (TMEM accumulators are never rescaled).
// Built around what the round-2 measurements of its round-1 predecessor showed (profiles/r2_attention3_trace.txt,
// r2_attention3_ablation.txt, r2_ubench_softmax.txt):
//
//   * the softmax warps ran at half the rate the same loop reaches in isolation (6.7 vs 13.7 elements/clk/SM with
//     8 warps): 16 softmax warps now, TWO threads per query row (warps w and w+4 of a group own the same TMEM lane
//     quarter and split the chunk's 32-column blocks), the row maximum exchanged through shared memory and one
//     named barrier per chunk; 80 registers per thread instead of 128;
//   * with the softmax removed the kernel still took 59 of 90 us: the epilogue group spent 2 200 - 3 100 clk per tile
//     in the output store (every lane a different row: 32 cache lines per instruction).  Rows are now transposed
//     inside each group of 8 lanes (three butterfly exchanges), so one store instruction writes four whole 128-byte
//     rows;
//   * V was single-buffered (its load could only start after the last P V of the previous item, 14 % of the time in
//     V_FULL waits): K and V of consecutive items rotate through THREE slots (load i -> slot i % 3, K_n = load 2n,
//     V_n = load 2n+1), so V_n lands in the slot K_{n-1} left, three chunks before it is needed; K / V and Q have
//     their own producer warps, so a slot that is not free yet never delays a Q tile;
//   * no integer divisions in the per-chunk loops.
//
//   warps 0-7     softmax of even chunks (lane quarter = warp & 3, column half = (warp >> 2) & 1)
//   warps 8-15    softmax of odd chunks
//   warps 16-19   epilogue: online softmax in registers, output store, CE row (attn_blocks.py:44-53)
//   warp 20       TMA producer, Q tiles (double-buffered)
//   warp 23       TMA producer, K / V slots
//   warp 21       score MMAs    S_k into TMEM buffer k % 3
//   warp 22       value MMAs    O_k into TMEM buffer k % 2
// The producer / issuer warps have the HIGHEST warp ids on purpose: the warp arbiter of an SM sub-partition prefers
// higher ids, and these warps need a few dozen issue slots per chunk at exactly the moment their barrier opens --
// with the lowest ids they were served last behind four busy softmax warps (profiles/r2_attention4_trace.txt: P V MMAs
// issued 1 000 - 3 000 clk after their operands were ready).
#include "kernels.h"
#include "sm100_ptx.cuh"

#ifdef AT4_TRACE     // timing experiments only: event timestamps of CTA 0 (printed by attention4_launch)
__device__ long long g_at4_trace[20 * 64];
#define TRACE(ev, k)                                                                                    \
  do {                                                                                                  \
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && (k) < 64) g_at4_trace[(ev) * 64 + (k)] = clock64(); \
  } while (0)
#else
#define TRACE(ev, k)
#endif

namespace vipt {

namespace at4 {
constexpr int kHd = 64;
constexpr int kQTile = 128;
constexpr int kMaxKeys = 320;
constexpr int kThreads = 768;
constexpr int kSBufCols = 128;                   // S buffers at TMEM columns 0 / 128 / 256
constexpr int kOCol = 384;                       // O buffers at 384 / 448
constexpr int kBoxBytes = 64 * 128;              // one TMA box: 64 tokens x 64 fp16
constexpr int kQBytes = kQTile * 128;
constexpr int kPBytes = 2 * kQTile * 128;        // one P buffer: 128 rows x 128 keys fp16
constexpr int kRegsWg0 = 40, kRegsEpi = 120;      // setmaxnreg: 128 x (40 + 120) + 512 x 80 = 768 x 80 (the launch allocation)
constexpr int kAlignSlack = 512;                 // dynamic smem starts 1024-aligned in practice; checked at run time
enum { Q_FULL = 0, Q_EMPTY = 2, KV_FULL = 4, KV_EMPTY = 7, S_FULL = 10, S_EMPTY = 13, P_FULL = 16, O_FULL = 18,
       O_EMPTY = 20, ST_FULL = 22, NUM_BARS = 26 };
__host__ __device__ constexpr int kboxes(int nkp) { return (nkp + 63) / 64; }
__host__ __device__ constexpr int off_kv(int) { return 2 * kQBytes; }
__host__ __device__ constexpr int off_p(int nkp) { return off_kv(nkp) + 3 * kboxes(nkp) * kBoxBytes; }
__host__ __device__ constexpr int off_stats(int nkp) { return off_p(nkp) + 2 * kPBytes; }
constexpr int kStatsFloats = 4 * kQTile * 2;     // [4 chunks in flight][128 rows][m, sum]
constexpr int kHmaxFloats = 2 * 2 * 2 * kQTile;  // [group][parity of the group's iteration][column half][128 rows]
constexpr int kCeCols = 256;                     // CE row: search keys only
constexpr int kCeFloats = 2 * kCeCols + 2 * 4 + 2 * 2;     // CE row probabilities, chunk maxima, final (m, 1/l)
__host__ __device__ constexpr int smem_bytes(int nkp) {
  return kAlignSlack + off_stats(nkp) + (kStatsFloats + kHmaxFloats + kCeFloats) * 4 + NUM_BARS * 8 + 16;
}
struct Chunks {
  int nc;
  int base[3];
  int w[3];
};
}  // namespace at4

namespace {

__device__ __forceinline__ float ex2f4(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// tcgen05.mma with both smem descriptors given as (low word, shared high word)
__device__ __forceinline__ void umma4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
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

// Bounded mbarrier wait without the diagnostic printf of sm100_ptx.cuh's mbar_wait: a call site in every role loop costs
// registers around the call, and with 227 KB of shared memory carved out of the L1 every spill is an L2 round trip.
__device__ __forceinline__ void mbar_wait4(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 24)) __trap();
  }
}

template <int kRegs>
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs)); }
template <int kRegs>
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs)); }

// position of chunk k of this CTA's stream: local item n, query tile t, chunk j of the tile, tile counter g
struct Pos4 {
  int n, t, j, g;
  __device__ __forceinline__ void advance(int NC, int T) {
    if (++j == NC) {
      j = 0; ++g;
      if (++t == T) { t = 0; ++n; }
    }
  }
};

__device__ __forceinline__ uint4 shfl_xor_u4(uint4 v, int m) {
  v.x = __shfl_xor_sync(0xffffffffu, v.x, m); v.y = __shfl_xor_sync(0xffffffffu, v.y, m);
  v.z = __shfl_xor_sync(0xffffffffu, v.z, m); v.w = __shfl_xor_sync(0xffffffffu, v.w, m);
  return v;
}

}  // namespace

__global__ void __launch_bounds__(at4::kThreads, 1)
attention4_kernel(const __grid_constant__ CUtensorMap tm, op_t* __restrict__ out, float* __restrict__ ce_prob, int B,
                  int N, int nkp, const at4::Chunks ch, int ce_row, int lens_t, int ce_stride, float scale_log2) {
  using namespace at4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = smem + off_kv(nkp);
  uint8_t* sP = smem + off_p(nkp);
  float* s_stats = reinterpret_cast<float*>(smem + off_stats(nkp));
  float* s_hmax = s_stats + kStatsFloats;
  float* s_ce = s_hmax + kHmaxFloats;            // [2 items][256]
  float* s_cem = s_ce + 2 * kCeCols;             // [2 items][4] chunk maxima of the CE row
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
  // CTA c owns the contiguous unit range [u0, u1) of the item-major list of B * 12 * T (item, query tile) units.
  // Evaluated by every role for itself from an un-mergeable read of %ctaid (WORK()): computed once ahead of the role
  // dispatch these values become long-lived registers shared by all branches, which ptxas spills once the producer /
  // MMA warps drop to 40 registers -- and with 227 KB of shared memory carved out of the L1 a local-memory access in
  // a per-chunk loop is an L2 round trip.
#define WORK()                                                                              \
  int bid_, nblk_;                                                                          \
  asm volatile("mov.u32 %0, %%ctaid.x;" : "=r"(bid_));                                      \
  asm volatile("mov.u32 %0, %%nctaid.x;" : "=r"(nblk_));                                    \
  const int n_units = B * 12 * T;                                                           \
  const int u0 = (int)(((long long)bid_ * n_units) / nblk_);                                \
  const int u1 = (int)(((long long)(bid_ + 1) * n_units) / nblk_);                          \
  const int item0 = u0 / T, t0 = u0 - item0 * T;                                            \
  const int n_local = u1 > u0 ? (u1 - 1) / T - item0 + 1 : 0;                               \
  const int total = (u1 - u0) * NC;                                                         \
  const uint32_t tmem = *tmem_slot;                                                         \
  (void)n_local; (void)tmem; (void)item0; (void)t0

  for (int i = threadIdx.x; i < kStatsFloats; i += kThreads) s_stats[i] = 0.f;      // the softmax halves ADD their sums
  if (threadIdx.x == 0) {
    if (static_cast<int>(smem - smem_raw) > kAlignSlack) {
      printf("vipt: attention4 dynamic shared memory is not 512-byte aligned\n");
      __trap();
    }
    tma_prefetch_desc(&tm);
    for (int i = 0; i < NUM_BARS; ++i) {
      const bool eight = (i >= S_EMPTY && i < S_EMPTY + 3) || (i >= P_FULL && i < P_FULL + 2) || (i >= ST_FULL && i < ST_FULL + 4);
      const bool four = i >= O_EMPTY && i < O_EMPTY + 2;
      mbar_init(&bars[i], eight ? 8u : (four ? 4u : 1u));
    }
    fence_mbar_init();
  }
  if (warp == 21) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_trigger();
  pdl_wait();
  if (warp >= 20) {
    reg_dec<kRegsWg0>();
    if (warp == 20) {
      WORK();
      if (lane == 0 && total > 0) {
        // ------------------------------------------------------------ TMA producer: Q tiles
        int g = 0;
        for (int n = 0; n < n_local; ++n) {
          const int item = item0 + n;
          const int b = item / 12, h = item - b * 12;
          const int ts = n == 0 ? t0 : 0;                  // tiles [ts, te) of this item belong to this CTA
          const int te = min(T, u1 - item * T);
          for (int t = ts; t < te; ++t, ++g) {
            const int qb = g & 1;
            mbar_wait4(&bars[Q_EMPTY + qb], ((g >> 1) & 1) ^ 1);
            mbar_expect_tx(&bars[Q_FULL + qb], 2 * kBoxBytes);
            tma_load_3d(sQ + qb * kQBytes, &tm, &bars[Q_FULL + qb], h * kHd, t * kQTile, b);
            tma_load_3d(sQ + qb * kQBytes + kBoxBytes, &tm, &bars[Q_FULL + qb], h * kHd, t * kQTile + 64, b);
          }
        }
      }
    } else if (warp == 23) {
      WORK();
      if (lane == 0 && total > 0) {
        // ------------------------------------------------------------ TMA producer: K / V slots, load i -> slot i % 3
        for (int i = 0; i < 2 * n_local; ++i) {
          const int item = item0 + (i >> 1);
          const int b = item / 12, h = item - b * 12;
          const int slot = i % 3, use = i / 3;
          mbar_wait4(&bars[KV_EMPTY + slot], (use & 1) ^ 1);
          mbar_expect_tx(&bars[KV_FULL + slot], kbytes);
          const int col = ((i & 1) ? 1536 : 768) + h * kHd;
          for (int x = 0; x < nbox; ++x)
            tma_load_3d(sKV + slot * kbytes + x * kBoxBytes, &tm, &bars[KV_FULL + slot], col, x * 64, b);
        }
      }
    } else if (warp == 21) {
      // -------------------------------------------------------------- score MMAs: S_k = Q K_k^T, up to 3 chunks ahead.
      // The whole warp runs the loop (warp-uniform control flow, descriptor math in uniform registers); one elected
      // lane issues the MMAs and commits.
      const uint32_t is0 = make_idesc(kQTile, cw0), is1 = make_idesc(kQTile, cw1 ? cw1 : cw0), is2 = make_idesc(kQTile, cw2 ? cw2 : cw0);
      const uint64_t d0 = make_sdesc_sw128(smem_u32(sQ));
      const uint32_t hi = (uint32_t)(d0 >> 32), q_lo = (uint32_t)d0;
      const uint32_t k_lo = (uint32_t)make_sdesc_sw128(smem_u32(sKV));
      WORK();
      Pos4 ps = {0, t0, 0, 0};
      int kslot = 0, kuse = 0;                              // slot / use count of K_n: load 2n
      for (int ks = 0; ks < total; ++ks) {
        const int sb = ks % 3;
        if (ps.j == 0) {
          if (ps.t == 0 || ks == 0) {
            kslot = (2 * ps.n) % 3; kuse = (2 * ps.n) / 3;
            mbar_wait4(&bars[KV_FULL + kslot], kuse & 1);
          }
          mbar_wait4(&bars[Q_FULL + (ps.g & 1)], (ps.g >> 1) & 1);
        }
        if (ks >= 3) mbar_wait4(&bars[S_EMPTY + sb], ((ks / 3) & 1) ^ 1);
        TRACE(0, ks);
        tc_fence_after();
        const uint32_t qa = q_lo + (uint32_t)(((ps.g & 1) * kQBytes) >> 4);
        const uint32_t ka = k_lo + (uint32_t)((kslot * kbytes + chunk_base(ps.j) * 128) >> 4);
        const uint32_t dcol = tmem + sb * kSBufCols, id = ps.j == 0 ? is0 : (ps.j == 1 ? is1 : is2);
        if (elect_one()) {
          umma4(dcol, qa, ka, hi, id, 0);
          umma4(dcol, qa + 2, ka + 2, hi, id, 1);
          umma4(dcol, qa + 4, ka + 4, hi, id, 1);
          umma4(dcol, qa + 6, ka + 6, hi, id, 1);
          umma_commit(&bars[S_FULL + sb]);
          if (ps.j == NC - 1) {
            umma_commit(&bars[Q_EMPTY + (ps.g & 1)]);
            if (ps.t == T - 1) umma_commit(&bars[KV_EMPTY + kslot]);       // last S of the item: K slot free
          }
        }
        __syncwarp();
        ps.advance(NC, T);
      }
    } else {
      // -------------------------------------------------------------- value MMAs: O_k = P_k V_k (own issuing thread)
      const uint32_t idesc_pv = make_idesc(kQTile, kHd, /*b_mn_major=*/true);
      const uint64_t d0 = make_sdesc_sw128(smem_u32(sP));
      const uint32_t hi = (uint32_t)(d0 >> 32), p_lo = (uint32_t)d0;
      const uint32_t v_lo = (uint32_t)make_sdesc_sw128(smem_u32(sKV));
      WORK();
      Pos4 pp = {0, t0, 0, 0};
      int vslot = 0, vuse = 0;                              // slot / use count of V_n: load 2n + 1
      for (int kp = 0; kp < total; ++kp) {
        const int ob = kp & 1;
        TRACE(12, kp);
        if (pp.j == 0 && (pp.t == 0 || kp == 0)) {
          vslot = (2 * pp.n + 1) % 3; vuse = (2 * pp.n + 1) / 3;
          mbar_wait4(&bars[KV_FULL + vslot], vuse & 1);
        }
        TRACE(13, kp);
        if (kp >= 2) mbar_wait4(&bars[O_EMPTY + ob], ((kp >> 1) & 1) ^ 1);
        TRACE(14, kp);
        mbar_wait4(&bars[P_FULL + ob], (kp >> 1) & 1);
        TRACE(5, kp);
        tc_fence_after();
        const int nj = chunk_w(pp.j) >> 4;
        const uint32_t pa = p_lo + (uint32_t)((ob * kPBytes) >> 4);
        const uint32_t va = v_lo + (uint32_t)((vslot * kbytes + chunk_base(pp.j) * 128) >> 4);
        const uint32_t dcol = tmem + kOCol + ob * kHd;
        if (elect_one()) {
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            if (jj < nj)                            // A: 16 keys = 32 B inside the 128-B row, next 64-key block 16 KB on
              umma4(dcol, pa + (uint32_t)(((jj >> 2) * (kQTile * 128) + (jj & 3) * 32) >> 4), va + jj * ((16 * 128) >> 4), hi,
                    idesc_pv, jj != 0);
          }
          umma_commit(&bars[O_FULL + ob]);
          if (pp.j == NC - 1 && pp.t == T - 1) umma_commit(&bars[KV_EMPTY + vslot]);    // last P V of the item
        }
        __syncwarp();
        TRACE(15, kp);
        pp.advance(NC, T);
      }
    }
  } else if (warp < 16) {
    // ---------------------------------------------------------------- softmax groups (even / odd chunks), 8 warps each:
    // lane quarter = warp & 3; the two warps of a quarter split the chunk's 32-column blocks (half 0 the first
    // ceil(nb / 2), half 1 the rest) and exchange their maxima through s_hmax (double-buffered by iteration parity:
    // a warp can be at most one group barrier ahead of its partner)
    WORK();
    const int grp = warp >> 3;
    const int half = (warp >> 2) & 1;
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const uint32_t t_lane = tmem + ((uint32_t)(wr * 32) << 16);
    const uint32_t pbuf = smem_u32(sP + grp * kPBytes) + row * 128;      // 32-bit shared address: STS, not generic ST
    Pos4 p = {0, t0, 0, 0};
    if (grp == 1 && total > 1) p.advance(NC, T);
    for (int k = grp; k < total; k += 2) {
      const int q0 = p.t * kQTile;
      const int base = chunk_base(p.j);
      const int nb = chunk_w(p.j) >> 5;
      // (splitting the middle block of an odd block count by its 16-column halves between the two warps of a lane quarter
      // was measured: the extra code path cost more than the balance gained, 41.0 -> 44.1 us at 190 keys)
      const int nb0 = (nb + 1) >> 1;
      const int blk0 = half == 0 ? 0 : nb0, blk1 = half == 0 ? nb0 : nb;       // my blocks [blk0, blk1)
      const int sb = k % 3;
      const uint32_t t_row = t_lane + sb * kSBufCols;
      float* st = s_stats + ((k & 3) * kQTile + row) * 2;
      float* hm = s_hmax + ((grp * 2 + ((k >> 1) & 1)) * 2) * kQTile + row;     // [half 0], [half 1] at + kQTile
      const bool valid = q0 + wr * 32 < N;                  // warp-uniform: at least one real query row
      mbar_wait4(&bars[S_FULL + sb], (k / 3) & 1);
      if (wr == 2 && half == 0) TRACE(1, k);
      tc_fence_after();
      uint32_t v[32];
      if (valid) {
        float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // 4 independent chains
        for (int bi = blk0; bi < blk1; ++bi) {
          tmem_ld32(t_row + bi * 32, v);
          tmem_ld_wait();
          const int col = base + bi * 32;
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
        }
        hm[half * kQTile] = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
      }
      // every warp of the group, real rows or not: the partner's maximum is visible after this barrier
      if (grp == 0) asm volatile("bar.sync 1, 256;" ::: "memory");
      else asm volatile("bar.sync 2, 256;" ::: "memory");
      if (wr == 2 && half == 0) TRACE(2, k);
      if (valid) {
        const float m = fmaxf(hm[0], hm[kQTile]);
        if (k >= 2) mbar_wait4(&bars[O_FULL + (k & 1)], ((k >> 1) & 1) ^ 1);    // P V of chunk k-2 has consumed this P buffer
        if (wr == 2 && half == 0) TRACE(3, k);
        const float msl = m * scale_log2;
        const bool ce_thread = ce_prob != nullptr && p.t == 0 && row == ce_row;
        float sum4[4] = {0.f, 0.f, 0.f, 0.f};
        if (blk0 == blk1) {              // no block of my own (a 32-column chunk): S is not read by this warp
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars[S_EMPTY + sb]);
        }
        for (int bi = blk0; bi < blk1; ++bi) {
          tmem_ld32(t_row + bi * 32, v);
          tmem_ld_wait();
          if (bi + 1 == blk1) {          // S_k has been read for the last time by this warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars[S_EMPTY + sb]);
          }
          const int crel = bi * 32;
          const int col = base + crel;
          float pr[32];
#pragma unroll
          for (int jx = 0; jx < 32; ++jx) pr[jx] = ex2f4(fmaf(__uint_as_float(v[jx]), scale_log2, -msl));
          if (col + 32 > N) {
#pragma unroll
            for (int jx = 0; jx < 32; ++jx)
              if (col + jx >= N) pr[jx] = 0.f;
          }
#pragma unroll
          for (int jx = 0; jx < 32; ++jx) sum4[jx & 3] += pr[jx];
          if (ce_thread) {
            // one lane, its warp waits for it: a block lies either wholly among the template keys or wholly among the
            // search keys (lens_t % 32 == 0, checked at launch), so the 32 probabilities go out as 8 vector stores
            // instead of 32 predicated scalar ones (the CE lane's warp was the slowest of its group by ~1 800 clk per chunk)
            if (col >= lens_t && col + 32 - lens_t <= kCeCols) {
              const uint32_t dst = smem_u32(s_ce + (p.n & 1) * kCeCols + col - lens_t);
#pragma unroll
              for (int q = 0; q < 8; ++q)
                sts128(dst + q * 16, __float_as_uint(pr[q * 4 + 0]), __float_as_uint(pr[q * 4 + 1]), __float_as_uint(pr[q * 4 + 2]),
                       __float_as_uint(pr[q * 4 + 3]));
            }
          }
          const uint32_t prow = pbuf + (crel >> 6) * (kQTile * 128);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int c16 = (((crel >> 5) & 1) * 4 + q) ^ (row & 7);
            sts128(prow + c16 * 16, pack_op2(pr[q * 8 + 0], pr[q * 8 + 1]), pack_op2(pr[q * 8 + 2], pr[q * 8 + 3]),
                   pack_op2(pr[q * 8 + 4], pr[q * 8 + 5]), pack_op2(pr[q * 8 + 6], pr[q * 8 + 7]));
          }
        }
        // the statistics slot of chunk k - 4 was read (and its sum zeroed) by the epilogue before P V of chunk k - 2 was
        // issued, which the O_FULL wait above has seen complete.  Two addends into a zeroed slot: order-independent.
        const float sum = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
        asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(smem_u32(st + 1)), "f"(sum) : "memory");
        if (half == 0) st[0] = m;
        if (ce_thread && half == 0) s_cem[(p.n & 1) * 4 + p.j] = m;
      } else {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars[S_EMPTY + sb]);
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (wr == 2 && half == 0) TRACE(4, k);
      if (wr == 0 && half == 0) TRACE(17, k);      // the slowest warps of a group: lane quarter 0 (CE row; last valid quarter
      if (wr == 0 && half == 1) TRACE(18, k);      // of the third query tile)
      if (wr == 3 && half == 1) TRACE(19, k);
      if (lane == 0) {
        mbar_arrive(&bars[ST_FULL + (k & 3)]);      // four deep: the epilogue may lag the softmax by up to 3 chunks
        mbar_arrive(&bars[P_FULL + (k & 1)]);
      }
      p.advance(NC, T);
      p.advance(NC, T);
    }
  } else {
    // ---------------------------------------------------------------- epilogue group: online softmax in registers
    reg_inc<kRegsEpi>();
    WORK();
    const int wr = warp & 3;
    const int row = wr * 32 + lane;
    const int et = threadIdx.x - 512;
    const uint32_t t_lane = tmem + ((uint32_t)(wr * 32) << 16);
    float acc[kHd];
    float m_run = 0.f, l_run = 0.f;
    Pos4 p = {0, t0, 0, 0};
    for (int k = 0; k < total; ++k) {
      const int ob = k & 1;
      const int q0 = p.t * kQTile;
      mbar_wait4(&bars[ST_FULL + (k & 3)], (k >> 2) & 1);
      mbar_wait4(&bars[O_FULL + ob], (k >> 1) & 1);
      if (wr == 2) TRACE(6, k);
      tc_fence_after();
      const bool valid = q0 + wr * 32 < N;
      if (valid) {
        float* st = s_stats + ((k & 3) * kQTile + row) * 2;
        const float mc = st[0], sc = st[1];
        st[1] = 0.f;                                  // the slot's next users (chunk k + 4) add into it
        float a = 0.f, bb = 1.f;
        if (p.j == 0) {
          m_run = mc; l_run = sc;
        } else {
          const float mn = fmaxf(m_run, mc);
          a = ex2f4((m_run - mn) * scale_log2);
          bb = ex2f4((mc - mn) * scale_log2);
          l_run = l_run * a + sc * bb;
          m_run = mn;
        }
        // two halves through one 32-register buffer
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
        if (wr == 2) TRACE(7, k);
        if (wr == 0) TRACE(16, k);
        if (lane == 0) mbar_arrive(&bars[O_EMPTY + ob]);       // O_k is in registers, the statistics slot is read
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
      if (wr == 2) TRACE(9, k);
      if (p.j == NC - 1) {
        const int item = item0 + p.n;
        const int b = item / 12, h = item - b * 12;
        if (valid) {
          // 8 x 16-byte units per row.  Transposed inside each group of 8 lanes (butterfly over lane bits 2, 1, 0), so that
          // lane j of a group ends up with unit j of each of the group's 8 rows: store i then writes, per group, one whole
          // 128-byte row (4 rows per instruction instead of 32 sixteen-byte pieces of 32 different rows).
          const float inv = 1.0f / l_run;
          uint4 u[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            u[q].x = pack_op2(acc[q * 8 + 0] * inv, acc[q * 8 + 1] * inv);
            u[q].y = pack_op2(acc[q * 8 + 2] * inv, acc[q * 8 + 3] * inv);
            u[q].z = pack_op2(acc[q * 8 + 4] * inv, acc[q * 8 + 5] * inv);
            u[q].w = pack_op2(acc[q * 8 + 6] * inv, acc[q * 8 + 7] * inv);
          }
#pragma unroll
          for (int s = 4; s >= 1; s >>= 1) {
            const bool up = (lane & s) != 0;
#pragma unroll
            for (int x = 0; x < 8; ++x) {
              if (x & s) continue;
              const uint4 send = up ? u[x] : u[x | s];
              const uint4 recv = shfl_xor_u4(send, s);
              if (up) u[x] = recv; else u[x | s] = recv;
            }
          }
          if (wr == 2) TRACE(10, k);
          const int r0 = q0 + wr * 32 + (lane & ~7);
          op_t* o = out + ((size_t)b * N + r0) * 768 + h * kHd + (lane & 7) * 8;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (r0 + i < N) *reinterpret_cast<uint4*>(o + (size_t)i * 768) = u[i];
        }
        if (wr == 2) TRACE(11, k);
        if (ce_prob != nullptr && p.t == 0) {    // normalised CE row, coalesced
          float* fin = s_cefin + (p.n & 1) * 2;
          if (row == ce_row) { fin[0] = m_run; fin[1] = 1.0f / l_run; }
          asm volatile("bar.sync 3, 128;" ::: "memory");
          const float mf = fin[0], inv = fin[1];
          const float* cm = s_cem + (p.n & 1) * 4;
          const float e0 = ex2f4((cm[0] - mf) * scale_log2);
          const float e1 = NC > 1 ? ex2f4((cm[1] - mf) * scale_log2) : 0.f;
          const float e2 = NC > 2 ? ex2f4((cm[2] - mf) * scale_log2) : 0.f;
          float* ce_dst = ce_prob + ((size_t)b * 12 + h) * ce_stride;
          for (int col = lens_t + et; col < N; col += 128) {
            const float e = (NC > 2 && col >= cb2) ? e2 : (NC > 1 && col >= cb1) ? e1 : e0;
            ce_dst[col - lens_t] = s_ce[(p.n & 1) * kCeCols + col - lens_t] * e * inv;
          }
        }
      }
      if (wr == 2) TRACE(8, k);
      p.advance(NC, T);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 21) {
    tc_fence_after();
    tmem_dealloc<512>(*tmem_slot);
  }
#undef WORK
}

int attention4_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride, int num_sms,
                      cudaStream_t st) {
  using namespace at4;
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attention4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(kMaxKeys)) !=
        cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  const int nkp = op.nkp;
  if (op.N > kMaxKeys || op.N <= kQTile || nkp % 32) return VIPT_ERR_ARG;
  if (ce_prob != nullptr && (lens_t < kMaxKeys - kCeCols || ce_row >= lens_t || lens_t % 32 != 0)) return VIPT_ERR_ARG;
  // <= 3 chunks of <= 128 keys, widths multiples of 32 and as even as possible (the two softmax groups
  // alternate over the chunk stream)
  Chunks ch = {};
  // (as few chunks as possible: 192 keys as 2 + 2 + 2 blocks instead of 3 + 3 -- balanced halves, one more chunk per
  // tile -- was measured: 41.0 -> 47.1 us, i.e. ~1 100 clk of fixed cost per chunk: barrier, handshakes, pipeline bubbles)
  ch.nc = (nkp + 127) / 128;
  const int units = nkp / 32;
  int basek = 0;
  // An odd number of chunks per tile makes the two softmax groups swap chunks from tile to tile, so unequal widths are
  // fine there -- but the two warps of a lane quarter split a chunk by whole 32-column blocks, and a 3-block chunk costs
  // its group 2 block times.  320 keys: 4 + 4 + 2 blocks (5 block times per tile) instead of 4 + 3 + 3 (6).
  const bool by_pairs = (ch.nc & 1) && units % 2 == 0;
  const int un = by_pairs ? units / 2 : units;
  for (int j = 0; j < 3; ++j) {
    const int u = (j < ch.nc ? un / ch.nc + (j < un % ch.nc ? 1 : 0) : 0) * (by_pairs ? 2 : 1);
    ch.base[j] = basek;
    ch.w[j] = u * 32;
    basek += u * 32;
  }
  if (ch.base[ch.nc - 1] >= op.N) return VIPT_ERR_ARG;        // every chunk holds at least one real key
  const int items = op.B * 12;
  const int grid = items < num_sms ? items : num_sms;
  const float scale_log2 = 0.125f * 1.4426950408889634f;
  const cudaError_t e = launch_pdl(attention4_kernel, grid, kThreads, smem_bytes(nkp), st, op.tm, reinterpret_cast<op_t*>(out),
                                   ce_prob, op.B, op.N, nkp, ch, ce_row, lens_t, ce_stride, scale_log2);
#ifdef AT4_TRACE
  static int launches = 0;
  if (e == cudaSuccess && op.N == 320 && ++launches == 5) {
    cudaDeviceSynchronize();
    static long long h[20 * 64];
    cudaMemcpyFromSymbol(h, g_at4_trace, sizeof(h));
    const long long t0 = h[0];
    fprintf(stderr, "chunk  mmaS  smxS_FULL  smx_barrier  smx_Pfree  smx_P_FULL  mmaPV  epi_start  epi_O_EMPTY  epi_end  epi_fma  epi_transposed  epi_stored  pv_top pv_V pv_OEMPTY pv_issued epi0_O_EMPTY P_FULL_w0h0 P_FULL_w0h1 P_FULL_w3h1\n");
    for (int k = 0; k < 48; ++k) {
      fprintf(stderr, "%3d ", k);
      for (int ev = 0; ev < 20; ++ev) fprintf(stderr, " %8lld", h[ev * 64 + k] ? h[ev * 64 + k] - t0 : -1);
      fprintf(stderr, "\n");
    }
  }
#endif
  return e == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
