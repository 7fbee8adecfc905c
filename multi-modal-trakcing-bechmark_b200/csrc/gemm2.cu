// 2-CTA (cta_group::2) variant of the persistent tcgen05 GEMM for the large linear layers.
//
// A CTA pair (cluster of 2, one TPC) computes a 256 x 256 output tile with ONE tcgen05.mma.cta_group::2
// stream issued by the leader CTA: each CTA stages its own 128 rows of A and HALF of the weight tile
// (128 of the 256 W rows); the tensor cores of both SMs read both halves.  Per CTA and k-block that is
// 16 KB + 16 KB from L2 instead of 16 KB + 32 KB for the same 128 x 256 x 64 MACs -- the single-CTA
// kernel (gemm.cu) is bound by L2 -> SM operand traffic (~10 TB/s), not by the tensor pipe.
//
// Protocol (per smem stage s, per TMEM accumulator buffer a):
//   full[s]   (leader)   : leader producer arrives with expect_tx(2 x 32 KB); BOTH CTAs' TMA loads
//                          complete_tx on the leader's barrier (.cta_group::2 TMA form)
//   empty[s]  (each CTA) : tcgen05.commit multicast from the leader's MMA thread
//   tfull[a]  (each CTA) : tcgen05.commit multicast after the last k-block
//   tempty[a] (leader)   : 4 epilogue warps of each CTA arrive (peer: remote arrive through mapa)
//
// LayerNorm fold (north star (b): "GEMMs ... with LayerNorm, GELU and bias fused in"; reference sites
// ViPT/lib/models/layers/attn_blocks.py:94,103).  blocks[i].norm2 between attn.proj and mlp.fc1 (and, without deep
// prompts, norm1 between mlp.fc2 and the next attn.qkv) is not a pass over the residual stream any more:
//   producer  EPI_RESID_STAT   x <- x + acc + bias  in fp32, IN PLACE through one fp32 tensor map (the fourth warp
//             streams the x slabs of the tile in through a ring while the main loop runs; the epilogue adds, writes
//             the fp32 slab and an fp16 copy x16 back with TMA stores) and every thread -- it owns one row of the
//             tile -- keeps (sum, sum of squares) of its columns
//   consumer  EPI_LNFOLD_*     A = x16, W' = fp16(gamma * W):   LN(x) W^T + b = rstd (x16 W'^T - mean colsum(W')) + (b + W beta)
//             mean / rstd per row, two FMAs per element; the fourth warp prepares them one tile ahead
// The fp16 rounding of x instead of LN(x) gives the same operand error as long as |mean| is not large against the
// standard deviation of a row (measured on the synthetic weight sets: 2.9e-4 relative either way).
//
// Prompt fold (vit_ce_prompt.py:268-310; prompt_mid.cu has the whole picture): the deep-prompt stage of block i is
//   mlp.fc2 of block i-1, EPI_RESID_PROMPT: the residual epilogue also keeps, per row, the 17 dot products of
//             v = x + fc2(..) with conv0_0 * gamma (8), conv1x1 (8) and conv1x1.bias (the constants of a step's 32
//             columns travel with its x slab), so x0 = conv0_0(prompt_norms(v)) and the statistics of
//             x = v + U f + b1 need no pass over v
//   attn.qkv / attn.proj of block i: the rank-8 update U f enters as ONE EXTRA K-BLOCK (kext): Ax = f per row
//             (fp16, 8 of 64 columns used), Wx = W'U resp. U -- an eighth of a k-block of tensor work instead of 8 FMAs
//             per element in the epilogue (measured: +21 us per qkv launch in the epilogue form)
#include <stdlib.h>

#include "gemm.h"
#include "sm100_ptx.cuh"

namespace vipt {

#ifndef G2_ABL        // timing experiments only (scripts/ablate_gemm2.sh): 1 no output stores, 2 no prompt sums, 4 no x slab loads
#define G2_ABL 0
#endif

namespace g2 {
constexpr int BM = 128;            // rows per CTA (256 per pair)
constexpr int BK = 64;
constexpr int kSlabCols = 64;                        // epilogue slab: 128 rows x 64 fp16 = 16 KB, SW128
constexpr int kSlabBytes = BM * kSlabCols * 2;
constexpr int kABytes = BM * BK * 2;
// warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 epilogue feeder, 4..11 epilogue: 8 warps, 2 per TMEM lane group, 32 columns of a
// 64-column slab each.  -DG2_EPI16=1 builds the epilogues with real math per element (GELU, LayerNorm fold) with 16
// warps (4 per lane group, 16 columns each, 96 registers): measured and NOT kept -- fc1 + GELU 70.7 -> 72.7 us, fc1 with
// the LayerNorm fold 78.8 -> 83.0 us at 64 x 320 rows, step 3.80 -> 3.84 ms on the same box (twice the barrier
// participants and TMEM load instructions per slab for the same FMA work).
#ifndef G2_EPI16
#define G2_EPI16 0
#endif
__host__ __device__ constexpr int epi_warps(int epi) {
  return (G2_EPI16 && (epi == EPI_GELU_F16 || epi == EPI_LNFOLD_GELU_F16 || epi == EPI_LNFOLD_F16)) ? 16 : 8;
}
__host__ __device__ constexpr int threads_of(int epi) { return 128 + 32 * epi_warps(epi); }
constexpr int kRingMax = 3;            // residual epilogues: x slabs in flight towards the epilogue (barrier slots)
__host__ __device__ constexpr bool epi_resid(int epi) { return epi == EPI_RESID_STAT || epi == EPI_RESID_PROMPT; }
constexpr int kColwStep = 32 * 20 * 4;     // EPI_RESID_PROMPT: the 20 constants of a step's 32 columns travel with its x slab
// BN = columns per pair tile: 256, or 128 for N % 256 != 0.
// DEEPK (residual epilogues only): K >= 2048.  The main loop then needs its 5 stages (with 3 the fc2-sized GEMM ran at
// 51 % tensor activity, waiting for operands) and leaves the epilogue ~13 us per tile, so the epilogue gets the small
// configuration: 2 x slabs in flight, ONE output slab.  Shallow K (attn.proj): 3 stages, 3 slabs in, 2 out.
template <int BN, int EPI, int DEEPK>
struct Cfg {
  static constexpr bool kResid = epi_resid(EPI);
  static constexpr int kStages = kResid ? (DEEPK ? 5 : 3) : (BN == 256 ? 5 : 6);
  static constexpr int kBBytes = (BN / 2) * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kIn = DEEPK ? 2 : 3, kOut = DEEPK ? 1 : 2;
  // per ring slot; a multiple of 1024 so that every x slab keeps the 1024-byte alignment its 128-byte swizzle assumes
  static constexpr int kInBytes = kSlabBytes + (EPI == EPI_RESID_PROMPT ? (kColwStep + 1023) / 1024 * 1024 : 0);
  static constexpr int kOutBytes = kSlabBytes + kSlabBytes / 2;                               // fp32 slab + fp16 slab
  // plain epilogues: 4 staging slabs, 2 for fp16 outputs (double-buffered 64-column slabs; the LayerNorm-fold consumers
  // keep column sums and row statistics in the third), 4 for fp32 outputs (two 32-column slabs per step)
  static constexpr int kEpiBytes = kResid ? kIn * kInBytes + kOut * kOutBytes : 4 * kSlabBytes;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kEpiBytes + BN * 4 + (2 * kStages + 4 + 2 * kRingMax) * 8 + 16;
};
}  // namespace g2

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_cta(uint32_t local_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  // relaxed: the barrier only orders TMEM reads (completed by tcgen05.wait::ld + fence::before_thread_sync);
  // a release at cluster scope costs a MEMBAR.ALL per arrive
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load into this CTA's smem, completion signalled on a barrier that may live in the peer CTA
__device__ __forceinline__ void tma_load_3d_2cta(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0,
                                                 int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], "
      "[%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_2cta(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0,
                                                 int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void umma_f16_2cta(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2cta(uint64_t* bar) {     // arrives on the barrier of BOTH CTAs
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"((uint16_t)3)
      : "memory");
}

// TMEM -> registers: this warp's 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ void tmem_ldc(uint32_t taddr, uint32_t (&v)[32]) { tmem_ld32(taddr, v); }
__device__ __forceinline__ void tmem_ldc(uint32_t taddr, uint32_t (&v)[16]) { tmem_ld16(taddr, v); }

struct Gemm2Maps {
  CUtensorMap A, B, C, C2, B2, Ax, Bx;
};

template <int EPI, int BN, int DEEPK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(g2::threads_of(EPI), 1)
gemm2_kernel(const __grid_constant__ Gemm2Maps tm, GemmParams p) {
  using namespace g2;
  using C = Cfg<BN, EPI, DEEPK>;
  constexpr int kStages = C::kStages, kStageBytes = C::kStageBytes;
  constexpr bool kResid = epi_resid(EPI);
  constexpr bool kPrompt = EPI == EPI_RESID_PROMPT;          // residual + the sums of the prompt block
  constexpr bool kFold = EPI == EPI_LNFOLD_F16 || EPI == EPI_LNFOLD_GELU_F16;
  constexpr int kIn = C::kIn, kOut = C::kOut, kInBytes = C::kInBytes, kOutBytes = C::kOutBytes;
  constexpr int kEpiWarps = epi_warps(EPI), kEpiThreads = 32 * kEpiWarps;
  constexpr int CW = 64 / (kEpiWarps / 4);                   // columns of a 64-column slab per epilogue thread
  static_assert(kEpiWarps == 8 || (!kResid && EPI != EPI_BIAS_F32), "16 epilogue warps: fp16-output epilogues only");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint8_t* slab = smem + kStages * kStageBytes;            // epilogue staging (TMA store sources; residual: x slabs in)
  uint8_t* out_slab = slab + kIn * kInBytes;               // kResid: kOut x (fp32 slab, fp16 half slab)
  float* bias_s = reinterpret_cast<float*>(slab + C::kEpiBytes);
  float* cs_s = reinterpret_cast<float*>(slab + 2 * kSlabBytes);      // kFold: column sums (fp16 outputs use 2 slabs)
  float2* rowstat_s = reinterpret_cast<float2*>(slab + 2 * kSlabBytes + 4096);   // kFold: [2][128] (rstd, -mean * rstd)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bias_s + BN);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tfull_bar = empty_bar + kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* rfull_bar = tempty_bar + 2;                               // kResid: x slab ring; kFold: row statistics slots
  uint64_t* rempty_bar = rfull_bar + kRingMax;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rempty_bar + kRingMax);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int tiles_m = (p.M + 2 * BM - 1) / (2 * BM);
  const int tiles_n = p.N / BN;
  // Work units: the tiles of the complete rounds of the persistent loop, then the tiles of the last, partial round
  // cut into tail_split column strips (unit u >= tail_from: tile tail_from + (u - tail_from) / split, strip
  // (u - tail_from) % split).  Strips use the same k order, so results do not depend on the split.
  const int num_tiles = p.tail_from + (tiles_m * tiles_n - p.tail_from) * p.tail_split;
  const int num_kb = p.K / BK + p.kext;                  // kext: one more k-block from the maps Ax / Bx
  struct Unit { int tm, n0, width; };
  auto unit_of = [&](int u) -> Unit {
    int tile = u, strip = 0, width = BN;
    if (u >= p.tail_from) {
      const int v = u - p.tail_from;
      tile = p.tail_from + v / p.tail_split;
      strip = v % p.tail_split;
      width = BN / p.tail_split;
    }
    return Unit{tile / tiles_n, (tile % tiles_n) * BN + strip * width, width};
  };

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm.A);
    tma_prefetch_desc(&tm.B);
    tma_prefetch_desc(&tm.C);
    if (kResid) tma_prefetch_desc(&tm.C2);
    if (p.tail_split > 1) tma_prefetch_desc(&tm.B2);
    if (p.kext) { tma_prefetch_desc(&tm.Ax); tma_prefetch_desc(&tm.Bx); }
    for (int s = 0; s < kRingMax; ++s) {
      mbar_init(&rfull_bar[s], 1);
      mbar_init(&rempty_bar[s], kEpiWarps);
    }
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 2 * kEpiWarps);  // epilogue warps x 2 CTAs (leader's copy is the one used)
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();              // peer barriers are initialised before any remote arrive / multicast
  tc_fence_after();
  pdl_trigger();                   // programmatic dependent launch protocol, see sm100_ptx.cuh
  pdl_wait();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer (both CTAs): warp-uniform loop, one elected lane issues
    int s = 0;
    uint32_t ph = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      const Unit un = unit_of(t);
      const int tm_ = un.tm;
      const bool strip = un.width != BN;
      const uint32_t tx_bytes = 2 * (kABytes + (un.width / 2) * BK * 2);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = tiles + s * kStageBytes;
        uint8_t* b_dst = a_dst + kABytes;
        const uint32_t bar = mapa_cta(smem_u32(&full_bar[s]), 0);
        const bool ext = p.kext && kb == num_kb - 1;
        if (elect_one()) {
          // (the extra k-block always loads the full 128-row weight box: a strip's MMA reads its first rows only)
          if (leader) mbar_expect_tx(&full_bar[s], ext ? 2 * kStageBytes : tx_bytes);
          if (ext) {
            tma_load_3d_2cta(a_dst, &tm.Ax, bar, 0, tm_ * 2 * BM + (int)rank * BM, 0);
            tma_load_3d_2cta(b_dst, &tm.Bx, bar, 0, un.n0 + (int)rank * (un.width / 2), 0);
          } else {
            if (p.conv_cb_per_tap) {    // implicit 3x3 conv: the pair tile is one sequence's 16 x 16 map, 8 rows per CTA
              const int tap = kb / p.conv_cb_per_tap, cb = kb % p.conv_cb_per_tap;
              tma_load_4d_2cta(a_dst, &tm.A, bar, cb * BK, tap % 3 - 1, (int)rank * 8 + tap / 3 - 1, tm_);
            } else {
              tma_load_3d_2cta(a_dst, &tm.A, bar, kb * BK, tm_ * 2 * BM + (int)rank * BM, 0);
            }
            tma_load_3d_2cta(b_dst, strip ? &tm.B2 : &tm.B, bar, kb * BK, un.n0 + (int)rank * (un.width / 2), 0);
          }
        }
        __syncwarp();
        if (++s == kStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1 && leader) {
    // ------------------------------------------------ MMA issuer (leader CTA only).  The whole warp runs the loop so
    // that barriers, stage indices and descriptors stay warp-uniform (uniform registers feed UTCHMMA directly; a
    // single divergent thread made the compiler wrap every MMA in an R2UR broadcast loop); one elected lane issues.
    int s = 0;
    uint32_t ph = 0;
    int as = 0;
    uint32_t aph = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      const uint32_t idesc = make_idesc(2 * BM, t >= p.tail_from ? BN / p.tail_split : BN);
      mbar_wait(&tempty_bar[as], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(tiles + s * kStageBytes);
        const uint64_t adesc = make_sdesc_sw128(a_addr);
        const uint64_t bdesc = make_sdesc_sw128(a_addr + kABytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) umma_f16_2cta(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          umma_commit_2cta(&empty_bar[s]);
        }
        __syncwarp();
        if (++s == kStages) { s = 0; ph ^= 1; }
      }
      if (elect_one()) umma_commit_2cta(&tfull_bar[as]);
      __syncwarp();
      if (++as == 2) { as = 0; aph ^= 1; }
    }
  } else if (kFold && warp == 3) {
    // ------------------------------------------------ row statistics (both CTAs), one tile ahead of the epilogue: the
    // partial sums of this CTA's 128 rows -> (rstd, -mean * rstd) in shared memory.  (Read by the epilogue threads
    // themselves, the global-load latency sat on every tile's critical path: fc1 69.6 -> 94.2 us.)
    const float inv_k = 1.f / (float)(p.K);
    int it = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
      const int row0 = unit_of(t).tm * 2 * BM + (int)rank * BM;
      const int slot = it & 1;
      mbar_wait(&rempty_bar[slot], ((it >> 1) & 1) ^ 1);
      if (p.stat_parts == 0) {       // final statistics (prompt fold: prompt_mid_kernel wrote them)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int rr = row0 + i * 32 + lane;
          rowstat_s[slot * BM + i * 32 + lane] = rr < p.M ? __ldg(p.stat_in + rr) : make_float2(0.f, 0.f);
        }
      } else {
        // 4 rows per lane; the partial sums of a row are added in part order (the result does not depend on the
        // schedule), the loads go out in batches of 8 parts x 4 rows so that their latencies overlap
        float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
        for (int q0 = 0; q0 < p.stat_parts; q0 += 8) {
          float2 st[4][8];
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              const int rr = row0 + i * 32 + lane;
              st[i][u] = (q0 + u < p.stat_parts && rr < p.M) ? __ldg(p.stat_in + (size_t)(q0 + u) * p.M + rr) : make_float2(0.f, 0.f);
            }
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int u = 0; u < 8; ++u) {
              s1[i] += st[i][u].x;
              s2[i] += st[i][u].y;
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float mean = s1[i] * inv_k;
          const float rs = rsqrtf(fmaxf(fmaf(-mean, mean, s2[i] * inv_k), 0.f) + p.ln_eps);
          rowstat_s[slot * BM + i * 32 + lane] = make_float2(rs, -mean * rs);       // (rows >= M: never stored)
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&rfull_bar[slot]);
    }
  } else if (kResid && warp == 3) {
    // ------------------------------------------------ x slab producer (both CTAs): the fp32 rows this CTA's tile updates,
    // 32 columns at a time, in the order the epilogue consumes them (kPrompt: with the 20 constants of those columns);
    // runs ahead of the epilogue by the ring depth
    int rs = 0;
    uint32_t rph = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      const Unit un = unit_of(t);
      const int n0 = un.n0;
      const int row0 = un.tm * 2 * BM + (int)rank * BM;
      for (int j = 0; j < un.width / 32; ++j) {
        mbar_wait(&rempty_bar[rs], rph ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&rfull_bar[rs], ((G2_ABL & 4) ? 0 : kSlabBytes) + (kPrompt ? kColwStep : 0));
          if (!(G2_ABL & 4)) tma_load_3d(slab + rs * kInBytes, &tm.C, &rfull_bar[rs], n0 + j * 32, row0, 0);
          if (kPrompt)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(slab + rs * kInBytes + kSlabBytes)), "l"(p.colw + (size_t)(n0 + j * 32) * 20),
                           "r"(kColwStep), "r"(smem_u32(&rfull_bar[rs])) : "memory");
        }
        __syncwarp();
        if (++rs == kIn) { rs = 0; rph ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------ epilogue (both CTAs, own 128 rows)
    const int we = (warp - 4) & 3;           // TMEM lane group (warp % 4)
    const int hh = (warp - 4) >> 2;          // which CW-column part of every 64-column slab
    const int et = threadIdx.x - 128;
    const uint32_t bias_u32 = smem_u32(bias_s);
    int as = 0;
    uint32_t aph = 0;
    int rs = 0;                  // kResid: x slab ring position
    uint32_t rph = 0;
    int it = 0;                  // kFold: tile counter (row statistics slots)
    int ob = 0;                  // kResid: output slab, runs on across tiles
    float next_bias = 0.f, next_cs = 0.f;      // bias / column sum of column `et` of the next tile
    if (cluster_id < num_tiles) {
      const Unit nx = unit_of(cluster_id);
      if (et < nx.width) {
        next_bias = p.bias ? __ldg(p.bias + nx.n0 + et) : 0.f;
        if (kFold) next_cs = __ldg(p.colsum + nx.n0 + et);
      }
    }
    for (int t = cluster_id; t < num_tiles; t += num_clusters, ++it) {
      const Unit un = unit_of(t);
      const int tm_ = un.tm;
      const int n0 = un.n0;
      const int n_slabs = un.width / kSlabCols;
      const int row0 = tm_ * 2 * BM + (int)rank * BM;          // first row of this CTA's 128 x 256 tile
      const int r = we * 32 + lane;                            // row within the tile == TMEM lane
      // (bias_s of the previous tile was last read before that tile's final slab barrier.)  The values were fetched
      // while the previous tile was in the epilogue: read here, the global-load latency sat on every tile's critical
      // path (ncu: the staging stores were the top stall of the fc1 epilogue after the barrier waits).
      if (et < un.width) {
        bias_s[et] = next_bias;
        if (kFold) cs_s[et] = next_cs;
      }
      if (t + num_clusters < num_tiles) {
        const Unit nx = unit_of(t + num_clusters);
        if (et < nx.width) {
          next_bias = p.bias ? __ldg(p.bias + nx.n0 + et) : 0.f;
          if (kFold) next_cs = __ldg(p.colsum + nx.n0 + et);
        }
      }
      float ln_rs = 0.f, ln_nm = 0.f;                          // kFold: rstd and -mean * rstd of this thread's row
      if (kFold) {
        const int slot = it & 1;
        mbar_wait(&rfull_bar[slot], (it >> 1) & 1);
        const float2 st = rowstat_s[slot * BM + r];
        ln_rs = st.x;
        ln_nm = st.y;
        __syncwarp();
        if (lane == 0) mbar_arrive(&rempty_bar[slot]);
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
      mbar_wait(&tfull_bar[as], aph);
      tc_fence_after();
      if constexpr (kResid) {
        // x <- x + acc + bias, 32 columns per step: this warp set's threads take 16 of them (hh)
        const uint32_t taddr0 = tmem_base + as * BN + hh * 16 + ((uint32_t)(we * 32) << 16);
        float s1 = 0.f, s2 = 0.f;
        uint64_t pa[8];            // kPrompt: (v.wa_0, v.wa_1) .. (v.wa_6, v.wa_7), (v.u_0, v.u_1) .. (v.u_6, v.u_7)
        float pe = 0.f;            // kPrompt: v.b1
#pragma unroll
        for (int k = 0; k < 8; ++k) pa[k] = 0ull;
        uint32_t v[16];
        tmem_ld16(taddr0, v);
        const int n_steps = un.width / 32;
#pragma unroll 1
        for (int j = 0; j < n_steps; ++j) {
          mbar_wait(&rfull_bar[rs], rph);
          const uint32_t in_base = smem_u32(slab + rs * kInBytes);
          const uint32_t in_row = in_base + r * 128;
          const uint32_t brow = bias_u32 + (j * 32 + hh * 16) * 4;
          float f[16];
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 xq = lds128f(in_row + (((hh * 4 + q) ^ (r & 7)) * 16));
            const float4 bq = lds128f(brow + q * 16);
            f[q * 4 + 0] = (__uint_as_float(v[q * 4 + 0]) + bq.x) + xq.x;
            f[q * 4 + 1] = (__uint_as_float(v[q * 4 + 1]) + bq.y) + xq.y;
            f[q * 4 + 2] = (__uint_as_float(v[q * 4 + 2]) + bq.z) + xq.z;
            f[q * 4 + 3] = (__uint_as_float(v[q * 4 + 3]) + bq.w) + xq.w;
          }
          if (j + 1 < n_steps) tmem_ld16(taddr0 + (j + 1) * 32, v);
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            s1 += f[k];
            s2 = fmaf(f[k], f[k], s2);
          }
          if (kPrompt && !(G2_ABL & 2)) {         // 17 more row sums: packed FMAs on the (wa, u) pairs of the column, broadcast from smem
            const uint32_t wrow = in_base + kSlabBytes + hh * 16 * 80;
#pragma unroll
            for (int c = 0; c < 16; ++c) {
              const uint64_t ff = pk2(f[c], f[c]);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float4 w = lds128f(wrow + c * 80 + q * 16);
                pa[q * 2] = fma2(pk2(w.x, w.y), ff, pa[q * 2]);
                pa[q * 2 + 1] = fma2(pk2(w.z, w.w), ff, pa[q * 2 + 1]);
              }
              float b1c;
              asm volatile("ld.shared.f32 %0, [%1];" : "=f"(b1c) : "r"(wrow + c * 80 + 64) : "memory");
              pe = fmaf(b1c, f[c], pe);
            }
          } else if (j & 1) {    // a 64-column group is complete: part 2 * group + hh (this thread's 32 of its columns)
            if (p.stat_out && row0 + r < p.M)
              p.stat_out[(size_t)(((n0 >> 6) + (j >> 1)) * 2 + hh) * p.M + row0 + r] = make_float2(s1, s2);
            s1 = 0.f;
            s2 = 0.f;
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&rempty_bar[rs]);         // the x slab (and its constants) are consumed
          if (++rs == kIn) { rs = 0; rph ^= 1; }
          if (kOut == 1) {       // single output slab: its previous store must have drained before anybody writes it
            if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
          }
          const uint32_t o32 = smem_u32(out_slab + ob * kOutBytes) + r * 128;
#pragma unroll
          for (int q = 0; q < 4; ++q)
            sts128(o32 + (((hh * 4 + q) ^ (r & 7)) * 16), __float_as_uint(f[q * 4 + 0]), __float_as_uint(f[q * 4 + 1]),
                   __float_as_uint(f[q * 4 + 2]), __float_as_uint(f[q * 4 + 3]));
          // fp16 copy: 128 rows x 32 columns = 64-byte rows, 64-byte swizzle (16-byte chunk ^= (row / 2) % 4)
          const uint32_t o16 = smem_u32(out_slab + ob * kOutBytes + kSlabBytes) + r * 64;
#pragma unroll
          for (int q = 0; q < 2; ++q)
            sts128(o16 + (((hh * 2 + q) ^ ((r >> 1) & 3)) * 16), pack_op2(f[q * 8 + 0], f[q * 8 + 1]),
                   pack_op2(f[q * 8 + 2], f[q * 8 + 3]), pack_op2(f[q * 8 + 4], f[q * 8 + 5]), pack_op2(f[q * 8 + 6], f[q * 8 + 7]));
          fence_proxy_async_smem();
          // two output slabs: before the barrier the issuing thread makes sure the previous store has drained the slab
          // the next step writes
          if (kOut == 2 && et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          asm volatile("bar.sync 1, 256;" ::: "memory");
          if (et == 0 && !(G2_ABL & 1)) {
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(reinterpret_cast<uint64_t>(&tm.C)), "r"(smem_u32(out_slab + ob * kOutBytes)),
                           "r"(n0 + j * 32), "r"(row0), "r"(0)
                         : "memory");
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(reinterpret_cast<uint64_t>(&tm.C2)), "r"(smem_u32(out_slab + ob * kOutBytes + kSlabBytes)),
                           "r"(n0 + j * 32), "r"(row0), "r"(0)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          if (kOut == 2) ob ^= 1;
        }
        if (kPrompt && row0 + r < p.M) {       // 20 floats per row and (tile, half): s1, s2, v.wa_0..7, v.u_0..7, v.b1, 0
          float4* po = reinterpret_cast<float4*>(p.part_out + ((size_t)((n0 / BN) * 2 + hh) * p.M + row0 + r) * 20);
          float a[16];
#pragma unroll
          for (int k = 0; k < 8; ++k) upk2(pa[k], a[2 * k], a[2 * k + 1]);
          po[0] = make_float4(s1, s2, a[0], a[1]);
          po[1] = make_float4(a[2], a[3], a[4], a[5]);
          po[2] = make_float4(a[6], a[7], a[8], a[9]);
          po[3] = make_float4(a[10], a[11], a[12], a[13]);
          po[4] = make_float4(a[14], a[15], pe, 0.f);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(mapa_cta(smem_u32(&tempty_bar[as]), 0));
        if (++as == 2) { as = 0; aph ^= 1; }
        continue;
      }
      const uint32_t taddr0 = tmem_base + as * BN + hh * CW + ((uint32_t)(we * 32) << 16);
      uint32_t v[CW];
      tmem_ldc(taddr0, v);
#pragma unroll 1
      for (int sl = 0; sl < n_slabs; ++sl) {
        // fp16 outputs: one 64-column slab per step; fp32 outputs: each warp set fills its own 32-column slab
        uint8_t* buf = slab + (EPI == EPI_BIAS_F32 ? ((sl & 1) * 2 + hh) : (sl & 1)) * kSlabBytes;
        const uint32_t prow = smem_u32(buf) + r * 128;
        const uint32_t brow = bias_u32 + (sl * kSlabCols + hh * CW) * 4;
        float f[CW];
        tmem_ld_wait();
        if (kFold) {
          const uint32_t crow = smem_u32(cs_s) + (sl * kSlabCols + hh * CW) * 4;
#pragma unroll
          for (int j = 0; j < CW; j += 4) {
            const float4 bq = lds128f(brow + j * 4);
            const float4 cq = lds128f(crow + j * 4);
            f[j] = fmaf(ln_rs, __uint_as_float(v[j]), fmaf(ln_nm, cq.x, bq.x));
            f[j + 1] = fmaf(ln_rs, __uint_as_float(v[j + 1]), fmaf(ln_nm, cq.y, bq.y));
            f[j + 2] = fmaf(ln_rs, __uint_as_float(v[j + 2]), fmaf(ln_nm, cq.z, bq.z));
            f[j + 3] = fmaf(ln_rs, __uint_as_float(v[j + 3]), fmaf(ln_nm, cq.w, bq.w));
          }
        } else {
#pragma unroll
          for (int j = 0; j < CW; j += 4) {
            const float4 bq = lds128f(brow + j * 4);
            f[j] = __uint_as_float(v[j]) + bq.x; f[j + 1] = __uint_as_float(v[j + 1]) + bq.y;
            f[j + 2] = __uint_as_float(v[j + 2]) + bq.z; f[j + 3] = __uint_as_float(v[j + 3]) + bq.w;
          }
        }
        if (sl + 1 < n_slabs) tmem_ldc(taddr0 + (sl + 1) * kSlabCols, v);     // in flight during the math below
        if (EPI == EPI_GELU_F16 || EPI == EPI_LNFOLD_GELU_F16) {
#pragma unroll
          for (int j = 0; j < CW; j += 2) gelu_erf2x2(f[j], f[j + 1]);
        }
        if (EPI == EPI_RELU_F16) {
#pragma unroll
          for (int j = 0; j < CW; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        if (EPI == EPI_BIAS_F32) {
#pragma unroll
          for (int q = 0; q < CW / 4; ++q)                     // 128-byte rows of 32 fp32, 128-byte swizzle
            sts128(prow + ((q ^ (r & 7)) * 16), __float_as_uint(f[q * 4 + 0]), __float_as_uint(f[q * 4 + 1]),
                   __float_as_uint(f[q * 4 + 2]), __float_as_uint(f[q * 4 + 3]));
        } else {
#pragma unroll
          for (int q = 0; q < CW / 8; ++q) {
            const int chunk = (hh * (CW / 8) + q) ^ (r & 7);   // 128-byte swizzle, matches tmC
            sts128(prow + chunk * 16, pack_op2(f[q * 8 + 0], f[q * 8 + 1]), pack_op2(f[q * 8 + 2], f[q * 8 + 3]),
                   pack_op2(f[q * 8 + 4], f[q * 8 + 5]), pack_op2(f[q * 8 + 6], f[q * 8 + 7]));
          }
        }
        fence_proxy_async_smem();
        // one barrier per slab: before it, the issuing thread makes sure the PREVIOUS store has drained its
        // staging buffer, which is the buffer the next slab writes
        if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
        if (et == 0) {
          if (EPI == EPI_BIAS_F32) {
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2)
              asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                           ::"l"(reinterpret_cast<uint64_t>(&tm.C)), "r"(smem_u32(slab + ((sl & 1) * 2 + h2) * kSlabBytes)),
                             "r"(n0 + sl * kSlabCols + h2 * 32), "r"(row0), "r"(0)
                           : "memory");
          } else {
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(reinterpret_cast<uint64_t>(&tm.C)), "r"(smem_u32(buf)), "r"(n0 + sl * kSlabCols), "r"(row0),
                         "r"(0)
                         : "memory");
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_cta(smem_u32(&tempty_bar[as]), 0));
      if (++as == 2) { as = 0; aph ^= 1; }
    }
  }
  if (threadIdx.x == 128) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // staging smem drained
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();              // neither CTA may exit / free TMEM while the pair is still in flight
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
}

template <int EPI, int BN, int DEEPK>
static int launch2_t(const GemmOp& op, int num_sms, cudaStream_t st) {
  using C = g2::Cfg<BN, EPI, DEEPK>;
  static_assert(C::kSmemBytes <= 232448, "shared memory budget");
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(gemm2_kernel<EPI, BN, DEEPK>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes) != cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr_set = true;
  }
  const int tiles = op.p.tail_from + (((op.p.M + 255) / 256) * (op.p.N / BN) - op.p.tail_from) * op.p.tail_split;
  int grid = 2 * tiles;
  const int cap = num_sms & ~1;
  if (grid > cap) grid = cap;
  Gemm2Maps tm;
  tm.A = op.tmA; tm.B = op.tmB; tm.C = op.tmC; tm.C2 = op.tmC2; tm.B2 = op.tmB2; tm.Ax = op.tmAx; tm.Bx = op.tmBx;
  return launch_pdl(gemm2_kernel<EPI, BN, DEEPK>, grid, g2::threads_of(EPI), C::kSmemBytes, st, tm, op.p) == cudaSuccess ? VIPT_OK
                                                                                                                  : VIPT_ERR_LAUNCH;
}

int gemm2_launch(const GemmOp& op, int num_sms, cudaStream_t st) {
  const bool narrow = op.bn == 640;
  const bool deep = op.p.K >= 2048;
  switch (op.epi) {
    case EPI_BIAS_F16: return narrow ? launch2_t<EPI_BIAS_F16, 128, 0>(op, num_sms, st) : launch2_t<EPI_BIAS_F16, 256, 0>(op, num_sms, st);
    case EPI_GELU_F16: return narrow ? launch2_t<EPI_GELU_F16, 128, 0>(op, num_sms, st) : launch2_t<EPI_GELU_F16, 256, 0>(op, num_sms, st);
    case EPI_RELU_F16: return narrow ? launch2_t<EPI_RELU_F16, 128, 0>(op, num_sms, st) : launch2_t<EPI_RELU_F16, 256, 0>(op, num_sms, st);
    case EPI_BIAS_F32: return narrow ? launch2_t<EPI_BIAS_F32, 128, 0>(op, num_sms, st) : launch2_t<EPI_BIAS_F32, 256, 0>(op, num_sms, st);
    // LayerNorm / prompt fold: 256-wide pair tiles only
    case EPI_RESID_STAT:
      if (narrow) return VIPT_ERR_ARG;
      return deep ? launch2_t<EPI_RESID_STAT, 256, 1>(op, num_sms, st) : launch2_t<EPI_RESID_STAT, 256, 0>(op, num_sms, st);
    case EPI_RESID_PROMPT:
      if (narrow) return VIPT_ERR_ARG;
      return deep ? launch2_t<EPI_RESID_PROMPT, 256, 1>(op, num_sms, st) : launch2_t<EPI_RESID_PROMPT, 256, 0>(op, num_sms, st);
    case EPI_LNFOLD_F16: return narrow ? VIPT_ERR_ARG : launch2_t<EPI_LNFOLD_F16, 256, 0>(op, num_sms, st);
    case EPI_LNFOLD_GELU_F16: return narrow ? VIPT_ERR_ARG : launch2_t<EPI_LNFOLD_GELU_F16, 256, 0>(op, num_sms, st);
  }
  return VIPT_ERR_ARG;
}

// same operand conventions as gemm_plan, single group, N % 256 == 0; W box = 128 rows (half tile per CTA)
int gemm2_plan(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int epi,
               void* out, int64_t ldc, const float* bias, int num_clusters) {
  if (K % g2::BK != 0 || N % 128 != 0 || M <= 0 || epi == EPI_RESID_F32 || g2::epi_resid(epi)) return VIPT_ERR_ARG;
  if ((epi == EPI_LNFOLD_F16 || epi == EPI_LNFOLD_GELU_F16) && N % 256 != 0) return VIPT_ERR_ARG;
  // 128-wide pair tiles only when N is not a multiple of 256 (or forced, num_clusters < 0).  Choosing them for the
  // N = 768 layers to shorten the last wave (480 half-tiles = 3.5 rounds instead of 4) was measured and lost:
  // linear GEMMs 2.25 -> 2.32 ms per step, the narrow tile moves 1.5x the operand bytes per flop.
  const int bn = (N % 256 != 0 || num_clusters < 0) ? 128 : 256;
  const int osz = epi == EPI_BIAS_F32 ? 4 : 2;
  if ((ldc * osz) % 16 != 0 || (reinterpret_cast<uintptr_t>(out) & 15)) return VIPT_ERR_ARG;
  if ((lda * 2) % 16 != 0 || (ldw * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(A) & 15) ||
      (reinterpret_cast<uintptr_t>(W) & 15))
    return VIPT_ERR_ARG;
  uint64_t da[3] = {(uint64_t)K, (uint64_t)M, 1};
  uint64_t sa[2] = {(uint64_t)lda * 2, (uint64_t)M * lda * 2};
  uint32_t ba[3] = {g2::BK, g2::BM, 1};
  int r = make_tmap_f16(&op->tmA, A, 3, da, sa, ba);
  if (r) return r;
  uint64_t db[3] = {(uint64_t)K, (uint64_t)N, 1};
  uint64_t sb[2] = {(uint64_t)ldw * 2, (uint64_t)N * ldw * 2};
  uint32_t bb[3] = {g2::BK, (uint32_t)bn / 2, 1};
  r = make_tmap_f16(&op->tmB, W, 3, db, sb, bb);
  if (r) return r;
  uint64_t dc[3] = {(uint64_t)N, (uint64_t)M, 1};
  uint64_t sc[2] = {(uint64_t)ldc * osz, (uint64_t)M * ldc * osz};
  uint32_t bc[3] = {(uint32_t)(epi == EPI_BIAS_F32 ? 32 : g2::kSlabCols), g2::BM, 1};
  r = epi == EPI_BIAS_F32 ? make_tmap_f32(&op->tmC, out, 3, dc, sc, bc) : make_tmap_f16(&op->tmC, out, 3, dc, sc, bc);
  if (r) return r;
  op->tmC2 = op->tmC;
  op->tmB2 = op->tmB;
  op->tmAx = op->tmA;
  op->tmBx = op->tmB;
  // Tail of the persistent loop: T tiles over C clusters leave the last round partly empty (N = 768 at 64 x 320 rows:
  // 240 tiles = 3.24 rounds of 74, run as 4).  The tiles of that round are cut into 2 or 4 column strips when the strips
  // still fit ONE round; a strip costs more than its share of a tile (the A operand stream does not shrink: measured
  // 0.59 of a tile for a half, round 2), so the split must pay for itself.
  op->p.tail_from = ((M + 255) / 256) * (N / bn);
  op->p.tail_split = 1;
  if (num_clusters > 0 && bn == 256) {
    const int T = op->p.tail_from, rem = T % num_clusters;
    static const char* off = getenv("VIPT_NO_TAIL_SPLIT");
    if (!off && T > num_clusters && rem > 0) {
      int split = rem * 4 <= num_clusters ? 4 : rem * 2 <= num_clusters ? 2 : 1;
      if (split > 1) {
        uint32_t bb2[3] = {g2::BK, (uint32_t)(bn / split / 2), 1};
        r = make_tmap_f16(&op->tmB2, W, 3, db, sb, bb2);
        if (r) return r;
        op->p.tail_from = T - rem;
        op->p.tail_split = split;
      }
    }
  }
  op->p.M = M; op->p.N = N; op->p.K = K; op->p.G = 1;
  op->p.out = out; op->p.ldc = ldc; op->p.out_group_stride = 0; op->p.out_group_col = 0;
  op->p.bias = bias; op->p.resid = nullptr;
  op->p.conv_cb_per_tap = 0; op->p.conv_cin_stride = 0;
  op->p.stat_out = nullptr; op->p.stat_in = nullptr; op->p.stat_parts = 0; op->p.colsum = nullptr; op->p.ln_eps = 0.f;
  op->p.colw = nullptr; op->p.part_out = nullptr; op->p.kext = 0;
  op->bn = bn == 256 ? 512 : 640;      // marks a 2-CTA plan (256- / 128-wide pair tiles)
  op->epi = epi;
  return VIPT_OK;
}

int gemm2_plan_resid_stat(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x,
                          void* x16, const float* bias, float2* stat_out, int num_clusters) {
  if (N % 256 != 0 || !x || !x16 || !stat_out) return VIPT_ERR_ARG;
  // A / W maps and the fp32 map over x (box 32 columns x 128 rows, loaded and stored) come from the fp32-output plan
  int r = gemm2_plan(op, A, lda, W, ldw, M, N, K, EPI_BIAS_F32, x, N, bias, num_clusters);
  if (r) return r;
  if (reinterpret_cast<uintptr_t>(x16) & 15) return VIPT_ERR_ARG;
  uint64_t dc[3] = {(uint64_t)N, (uint64_t)M, 1};
  uint64_t sc[2] = {(uint64_t)N * 2, (uint64_t)M * N * 2};
  uint32_t bc[3] = {32, g2::BM, 1};
  r = make_tmap_ex(&op->tmC2, 0, 64, x16, 3, dc, sc, bc);
  if (r) return r;
  op->p.stat_out = stat_out;
  op->epi = EPI_RESID_STAT;
  return VIPT_OK;
}

int gemm2_plan_resid_prompt(GemmOp* op, const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x,
                            void* x16, const float* bias, const float* colw20, float* part_out) {
  if (!colw20 || !part_out || (reinterpret_cast<uintptr_t>(colw20) & 15) || (reinterpret_cast<uintptr_t>(part_out) & 15))
    return VIPT_ERR_ARG;
  // whole tiles only: the 20 sums of a row are kept per (tile, column half)
  int r = gemm2_plan_resid_stat(op, A, lda, W, ldw, M, N, K, x, x16, bias, reinterpret_cast<float2*>(part_out), 0);
  if (r) return r;
  op->p.stat_out = nullptr;
  op->p.colw = colw20;
  op->p.part_out = part_out;
  op->epi = EPI_RESID_PROMPT;
  return VIPT_OK;
}

int gemm2_set_kext(GemmOp* op, const void* Ax, const void* Wx) {
  if (op->bn != 512 || !Ax || !Wx || (reinterpret_cast<uintptr_t>(Ax) & 15) || (reinterpret_cast<uintptr_t>(Wx) & 15) ||
      op->p.conv_cb_per_tap)
    return VIPT_ERR_ARG;
  uint64_t da[3] = {64, (uint64_t)op->p.M, 1};
  uint64_t sa[2] = {128, (uint64_t)op->p.M * 128};
  uint32_t ba[3] = {g2::BK, g2::BM, 1};
  int r = make_tmap_f16(&op->tmAx, Ax, 3, da, sa, ba);
  if (r) return r;
  uint64_t db[3] = {64, (uint64_t)op->p.N, 1};
  uint64_t sb[2] = {128, (uint64_t)op->p.N * 128};
  uint32_t bb[3] = {g2::BK, 128, 1};
  r = make_tmap_f16(&op->tmBx, Wx, 3, db, sb, bb);
  if (r) return r;
  op->p.kext = 1;
  return VIPT_OK;
}

void gemm2_set_lnfold(GemmOp* op, const float2* stat_in, int stat_parts, const float* colsum, float ln_eps) {
  op->p.stat_in = stat_in;
  op->p.stat_parts = stat_parts;
  op->p.colsum = colsum;
  op->p.ln_eps = ln_eps;
}

int gemm2_plan_conv(GemmOp* op, const void* feat, int c_total, int cin, const void* W, int B, int N, int epi, void* out,
                    int64_t ldc, const float* bias) {
  if (cin % g2::BK != 0 || (c_total * 2) % 16 != 0) return VIPT_ERR_ARG;
  const int K = 9 * cin;
  // the plain plan builds valid W / C maps; its A map (a (M, K) view that does not exist) is replaced below
  int r = gemm2_plan(op, W, K, W, K, B * 256, N, K, epi, out, ldc, bias, 0);
  if (!r && op->bn != 512) return VIPT_ERR_ARG;
  if (r) return r;
  uint64_t d[4] = {(uint64_t)c_total, 16, 16, (uint64_t)B};
  uint64_t sb[3] = {(uint64_t)c_total * 2, (uint64_t)16 * c_total * 2, (uint64_t)256 * c_total * 2};
  uint32_t box[4] = {g2::BK, 16, 8, 1};
  r = make_tmap_f16(&op->tmA, feat, 4, d, sb, box);
  if (r) return r;
  op->p.conv_cb_per_tap = cin / g2::BK;
  op->p.conv_cin_stride = 0;
  return VIPT_OK;
}

}  // namespace vipt
