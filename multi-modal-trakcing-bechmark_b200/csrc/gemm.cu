// Persistent warp-specialised tcgen05 GEMM for sm_100a:  C[g] = epilogue(A[g] * W[g]^T + bias[g]).
//
//   A : (G, M, K) fp16, K contiguous          -> TMA, 128B-swizzled smem tiles 128 x 64
//   W : (G, N, K) fp16, K contiguous (nn.Linear / flattened conv weight layout) -> tiles BN x 64
//   accumulators: fp32 in TMEM, two buffers of BN columns so the epilogue of tile i overlaps the
//   main loop of tile i+1.
//   warp 0: TMA producer   warp 1: MMA issuer (one thread)   warp 2: TMEM allocator
//   warps 4-7: epilogue (tcgen05.ld -> bias / GELU / ReLU / residual -> global)
//
// Replaces the cuBLAS / cuDNN library calls of the reference's nn.Linear / Conv2d sites
// (ViPT/lib/models/layers/attn.py:37,53, timm Mlp fc1/fc2, patch_embed.py:28, head.py:8-21).
#include "gemm.h"

#include <stdio.h>

#include "sm100_ptx.cuh"

namespace vipt {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 256;
constexpr int kEpiThreads = 128;

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN >= 256) ? 4 : 6;
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kTmemCols = (2 * BN < 32) ? 32 : 2 * BN;
  // tiles (1024-aligned) + bias tile + barriers + tmem slot
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + BN * 4 + (2 * kStages + 4) * 8 + 16;
};

// GELU: the polynomial-erf form shared with the 2-CTA kernel (sm100_ptx.cuh), so that small and large batches
// produce identical activations.
__device__ __forceinline__ float gelu_erf(float x) {
  float a = x, b = 0.f;
  gelu_erf2x2(a, b);
  return a;
}

template <int BN, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  float* bias_s = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bias_s + BN);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tfull_bar = empty_bar + Cfg::kStages;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BM - 1) / BM;
  const int tiles_n = p.N / BN;
  const int num_tiles = p.G * tiles_m * tiles_n;
  const int num_kb = p.K / BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&tempty_bar[a], 4);   // one arrival per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_trigger();                   // programmatic dependent launch protocol, see sm100_ptx.cuh
  pdl_wait();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------ TMA producer
    int s = 0;
    uint32_t ph = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int tn = t % tiles_n;
      const int tm = (t / tiles_n) % tiles_m;
      const int g = t / (tiles_n * tiles_m);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint8_t* a_dst = tiles + s * Cfg::kStageBytes;
        uint8_t* b_dst = a_dst + Cfg::kABytes;
        if (elect_one()) {             // warp-uniform loop, one elected lane issues
          mbar_expect_tx(&full_bar[s], Cfg::kStageBytes);
          if (p.conv_cb_per_tap) {    // implicit 3x3 conv: 128 rows = 8 image rows x 16 columns of sequence tm / 2
            const int tap = kb / p.conv_cb_per_tap, cb = kb % p.conv_cb_per_tap;
            tma_load_4d(a_dst, &tmA, &full_bar[s], g * p.conv_cin_stride + cb * BK, tap % 3 - 1, (tm & 1) * 8 + tap / 3 - 1,
                        tm >> 1);
          } else {
            tma_load_3d(a_dst, &tmA, &full_bar[s], kb * BK, tm * BM, g);
          }
          tma_load_3d(b_dst, &tmB, &full_bar[s], kb * BK, tn * BN, g);
        }
        __syncwarp();
        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer: warp-uniform loop, one elected lane issues
    constexpr uint32_t idesc = make_idesc(BM, BN);
    int s = 0;
    uint32_t ph = 0;
    int as = 0;
    uint32_t aph = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      mbar_wait(&tempty_bar[as], aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(tiles + s * Cfg::kStageBytes);
        const uint64_t adesc = make_sdesc_sw128(a_addr);
        const uint64_t bdesc = make_sdesc_sw128(a_addr + Cfg::kABytes);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // +32 B per K=16 step inside the 128B swizzle atom (start-address field is in 16 B units)
            umma_f16(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[s]);
        }
        __syncwarp();
        if (++s == Cfg::kStages) { s = 0; ph ^= 1; }
      }
      if (elect_one()) umma_commit(&tfull_bar[as]);
      __syncwarp();
      if (++as == 2) { as = 0; aph ^= 1; }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------ epilogue
    const int we = warp - 4;                 // == warp % 4 : TMEM lane quarter of this warp
    const int et = threadIdx.x - 128;
    int as = 0;
    uint32_t aph = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int tn = t % tiles_n;
      const int tm = (t / tiles_n) % tiles_m;
      const int g = t / (tiles_n * tiles_m);
      const int n0 = tn * BN;
      // stage the bias tile (previous tile's readers are past the barrier below)
      asm volatile("bar.sync 1, 128;" ::: "memory");
      for (int i = et; i < BN; i += kEpiThreads) bias_s[i] = p.bias ? p.bias[(size_t)g * p.N + n0 + i] : 0.f;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(&tfull_bar[as], aph);
      tc_fence_after();
      const int row = tm * BM + we * 32 + lane;
      const bool row_ok = row < p.M;
      const size_t out_off = ((size_t)g * p.out_group_stride) + (size_t)row * p.ldc + (size_t)g * p.out_group_col + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem_base + as * BN + c * 32 + ((uint32_t)(we * 32) << 16), v);
        tmem_ld_wait();
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const float4 bq = lds128f(smem_u32(bias_s) + (c * 32 + j) * 4);
          f[j] = __uint_as_float(v[j]) + bq.x; f[j + 1] = __uint_as_float(v[j + 1]) + bq.y;
          f[j + 2] = __uint_as_float(v[j + 2]) + bq.z; f[j + 3] = __uint_as_float(v[j + 3]) + bq.w;
        }
        if (EPI == EPI_GELU_F16) {
#pragma unroll
          for (int j = 0; j < 32; j += 2) gelu_erf2x2(f[j], f[j + 1]);
        }
        if (EPI == EPI_RELU_F16) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        if (row_ok) {
          if (EPI == EPI_BIAS_F32 || EPI == EPI_RESID_F32) {
            float* o = reinterpret_cast<float*>(p.out) + out_off + c * 32;
            if (EPI == EPI_RESID_F32) {
              const float* r = p.resid + out_off + c * 32;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float4 rv = *reinterpret_cast<const float4*>(r + j);
                f[j] += rv.x; f[j + 1] += rv.y; f[j + 2] += rv.z; f[j + 3] += rv.w;
              }
            }
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<float4*>(o + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
          } else {
            op_t* o = reinterpret_cast<op_t*>(p.out) + out_off + c * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 u;
              u.x = pack_op2(f[j], f[j + 1]);
              u.y = pack_op2(f[j + 2], f[j + 3]);
              u.z = pack_op2(f[j + 4], f[j + 5]);
              u.w = pack_op2(f[j + 6], f[j + 7]);
              *reinterpret_cast<uint4*>(o + j) = u;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
      if (++as == 2) { as = 0; aph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static int make_tmap_dt(CUtensorMap* out, CUtensorMapDataType dt, const void* base, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B);
int make_tmap_ex(CUtensorMap* out, int dtype_f32, int swizzle_bytes, const void* base, int rank, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box) {
  const CUtensorMapSwizzle swz = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                 : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                 : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  return make_tmap_dt(out, dtype_f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, base, rank, dims,
                      strides_bytes, box, swz);
}
int make_tmap_f16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box) {
  return make_tmap_dt(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, base, rank, dims, strides_bytes, box);
}
int make_tmap_f32(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                  const uint32_t* box) {
  return make_tmap_dt(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, rank, dims, strides_bytes, box);
}
static int make_tmap_dt(CUtensorMap* out, CUtensorMapDataType dt, const void* base, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return VIPT_ERR_DRIVER;
  cuuint64_t d[5], st[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; b[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) st[i] = strides_bytes[i];
  CUresult r = enc(out, dt, (cuuint32_t)rank, const_cast<void*>(base), d, st, b, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "vipt: cuTensorMapEncodeTiled failed (%d) rank=%d dims=%llu,%llu box=%u,%u\n", (int)r, rank,
            (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
    return VIPT_ERR_DRIVER;
  }
  return VIPT_OK;
}

template <int BN, int EPI>
static int launch_t(const GemmOp& op, int num_sms, cudaStream_t st) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(gemm_kernel<BN, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes) !=
        cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr_set = true;
  }
  const int tiles = op.p.G * ((op.p.M + BM - 1) / BM) * (op.p.N / BN);
  const int grid = tiles < num_sms ? tiles : num_sms;
  return launch_pdl(gemm_kernel<BN, EPI>, grid, kThreads, Cfg::kSmemBytes, st, op.tmA, op.tmB, op.p) == cudaSuccess ? VIPT_OK
                                                                                                          : VIPT_ERR_LAUNCH;
}

template <int BN>
static int launch_bn(const GemmOp& op, int num_sms, cudaStream_t st) {
  switch (op.epi) {
    case EPI_BIAS_F16: return launch_t<BN, EPI_BIAS_F16>(op, num_sms, st);
    case EPI_GELU_F16: return launch_t<BN, EPI_GELU_F16>(op, num_sms, st);
    case EPI_RELU_F16: return launch_t<BN, EPI_RELU_F16>(op, num_sms, st);
    case EPI_BIAS_F32: return launch_t<BN, EPI_BIAS_F32>(op, num_sms, st);
    case EPI_RESID_F32: return launch_t<BN, EPI_RESID_F32>(op, num_sms, st);
  }
  return VIPT_ERR_ARG;
}

int gemm_launch(const GemmOp& op, int num_sms, cudaStream_t st) {
  switch (op.bn) {
    case 512:
    case 640: return gemm2_launch(op, num_sms, st);
    case 256: return launch_bn<256>(op, num_sms, st);
    case 128: return launch_bn<128>(op, num_sms, st);
    case 64: return launch_bn<64>(op, num_sms, st);
    case 32: return launch_bn<32>(op, num_sms, st);
  }
  return VIPT_ERR_ARG;
}

// weight map + parameters (everything but the A map)
static int plan_common(GemmOp* op, const void* W, int64_t ldw, int64_t w_group_stride, int G, int M, int N, int K, int bn,
                       int epi, void* out, int64_t ldc, int64_t out_group_stride, int64_t out_group_col, const float* bias,
                       const float* resid) {
  if (K % BK != 0 || N % bn != 0 || (bn != 256 && bn != 128 && bn != 64 && bn != 32) || M <= 0 || G <= 0 || epi > EPI_RESID_F32)
    return VIPT_ERR_ARG;
  if ((ldw * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(W) & 15)) return VIPT_ERR_ARG;
  uint64_t db[3] = {(uint64_t)K, (uint64_t)N, (uint64_t)G};
  uint64_t sb[2] = {(uint64_t)ldw * 2, (uint64_t)(G > 1 ? w_group_stride : (int64_t)N * ldw) * 2};
  uint32_t bb[3] = {BK, (uint32_t)bn, 1};
  int r = make_tmap_f16(&op->tmB, W, 3, db, sb, bb);
  if (r) return r;
  op->p.M = M; op->p.N = N; op->p.K = K; op->p.G = G;
  op->p.out = out; op->p.ldc = ldc; op->p.out_group_stride = out_group_stride; op->p.out_group_col = out_group_col;
  op->p.bias = bias; op->p.resid = resid;
  op->p.conv_cb_per_tap = 0; op->p.conv_cin_stride = 0;
  op->p.stat_out = nullptr; op->p.stat_in = nullptr; op->p.stat_parts = 0; op->p.colsum = nullptr; op->p.ln_eps = 0.f;
  op->p.tail_from = 0; op->p.tail_split = 1;
  op->p.colw = nullptr; op->p.part_out = nullptr; op->p.kext = 0;
  op->bn = bn; op->epi = epi;
  return VIPT_OK;
}

int gemm_plan(GemmOp* op, const void* A, int64_t lda, int64_t a_group_stride, const void* W, int64_t ldw,
              int64_t w_group_stride, int G, int M, int N, int K, int bn, int epi, void* out, int64_t ldc,
              int64_t out_group_stride, int64_t out_group_col, const float* bias, const float* resid) {
  if (M <= 0 || G <= 0 || (lda * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(A) & 15)) return VIPT_ERR_ARG;
  uint64_t da[3] = {(uint64_t)K, (uint64_t)M, (uint64_t)G};
  uint64_t sa[2] = {(uint64_t)lda * 2, (uint64_t)(G > 1 ? a_group_stride : (int64_t)M * lda) * 2};
  uint32_t ba[3] = {BK, BM, 1};
  int r = make_tmap_f16(&op->tmA, A, 3, da, sa, ba);
  if (r) return r;
  return plan_common(op, W, ldw, w_group_stride, G, M, N, K, bn, epi, out, ldc, out_group_stride, out_group_col, bias, resid);
}

static int make_conv_tmap(CUtensorMap* tm, const void* feat, int c_total, int B) {
  uint64_t d[4] = {(uint64_t)c_total, 16, 16, (uint64_t)B};
  uint64_t sb[3] = {(uint64_t)c_total * 2, (uint64_t)16 * c_total * 2, (uint64_t)256 * c_total * 2};
  uint32_t box[4] = {BK, 16, 8, 1};
  return make_tmap_f16(tm, feat, 4, d, sb, box);
}

int gemm_plan_conv(GemmOp* op, const void* feat, int c_total, int cin, int cin_stride, const void* W, int G, int B, int N,
                   int bn, int epi, void* out, int64_t ldc, int64_t out_group_col, const float* bias) {
  if (cin % BK != 0 || (c_total * 2) % 16 != 0 || (reinterpret_cast<uintptr_t>(feat) & 15)) return VIPT_ERR_ARG;
  const int K = 9 * cin, M = B * 256;
  int r = make_conv_tmap(&op->tmA, feat, c_total, B);
  if (r) return r;
  r = plan_common(op, W, K, (int64_t)N * K, G, M, N, K, bn, epi, out, ldc, 0, out_group_col, bias, nullptr);
  if (r) return r;
  op->p.conv_cb_per_tap = cin / BK;
  op->p.conv_cin_stride = cin_stride;
  return VIPT_OK;
}

}  // namespace vipt
