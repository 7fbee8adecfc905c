// Fused joint template+search attention for one (sequence, head, 128-query tile):
//   S = Q K^T (tcgen05, fp32 in TMEM) -> softmax in registers -> P (fp16, swizzled smem) -> O = P V
// and, on candidate-elimination layers, the softmax row of the CE template token (query 27) over the
// search keys is written out in fp32 for the ranking kernel.  The (B,12,N,N) probability tensor the
// reference materialises (ViPT/lib/models/layers/attn.py:40-52) never exists.
//
// q, k, v are read straight out of the qkv GEMM output (B*N, 2304) with TMA (3-D map: column, token,
// sequence), so tokens past N are zero-filled rather than belonging to the next sequence.
#include <stdlib.h>

#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

constexpr int kHd = 64;            // head dim
constexpr int kQTile = 128;
constexpr int kMaxKeys = 320;
constexpr int kAttThreads = 256;
// Shared memory (1024-aligned):  region A | V | softmax scratch | barriers
//   region A first holds Q (16 KB) and K (64-row boxes), and -- once S = Q K^T has completed -- is
//   overwritten by P (fp16, 64-key blocks of 128 x 128 B).  V follows.
// Tensor memory: S occupies columns [0, nkp); O reuses columns [0, 64) after P has been written, so
// N <= 256 needs 256 columns (two CTAs per SM), N = 320 needs 512.
constexpr int kSQ = 0;
constexpr int kSK = kQTile * 128;
__host__ __device__ constexpr int att_region_a(int nkp) {
  const int kb = (nkp + 63) / 64;
  const int qk = kQTile * 128 + kb * 64 * 128;
  const int p = kb * kQTile * 128;
  return qk > p ? qk : p;
}
__host__ __device__ constexpr int att_scratch_off(int nkp) { return att_region_a(nkp) + ((nkp + 63) / 64) * 64 * 128; }
__host__ __device__ constexpr int att_smem_bytes(int nkp) {
  return 1024 + att_scratch_off(nkp) + 2 * 2 * kQTile * 4 + kMaxKeys * 4 + 64;
}

template <uint32_t kTmemCols>
__global__ void __launch_bounds__(kAttThreads, kTmemCols <= 256 ? 2 : 1)
attention_kernel(const __grid_constant__ CUtensorMap tm, op_t* __restrict__ out, float* __restrict__ ce_prob,
                 int N, int nkp, int ce_row, int lens_t, int ce_stride, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int kSP = 0;                                          // P overwrites Q/K
  const int kSV = att_region_a(nkp);
  float* s_max = reinterpret_cast<float*>(smem + att_scratch_off(nkp));   // [2][128]
  float* s_sum = s_max + 2 * kQTile;                          // [2][128]
  float* s_ce = s_sum + 2 * kQTile;                           // [320] un-normalised CE row
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_ce + kMaxKeys);      // qk, v, s, o
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = qt * kQTile;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kboxes = (nkp + 63) / 64;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm);
    for (int i = 0; i < 4; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_trigger();
  pdl_wait();
  const uint32_t tmem = *tmem_slot;

  if (threadIdx.x == 0) {
    mbar_expect_tx(&bars[0], (2 + kboxes) * 64 * 128);
    tma_load_3d(smem + kSQ, &tm, &bars[0], h * kHd, q0, b);
    tma_load_3d(smem + kSQ + 64 * 128, &tm, &bars[0], h * kHd, q0 + 64, b);
    for (int i = 0; i < kboxes; ++i) tma_load_3d(smem + kSK + i * 64 * 128, &tm, &bars[0], 768 + h * kHd, i * 64, b);
    mbar_expect_tx(&bars[1], kboxes * 64 * 128);
    for (int i = 0; i < kboxes; ++i) tma_load_3d(smem + kSV + i * 64 * 128, &tm, &bars[1], 1536 + h * kHd, i * 64, b);
    // ---- S = Q K^T
    mbar_wait(&bars[0], 0);
    tc_fence_after();
    const uint64_t qd = make_sdesc_sw128(smem_u32(smem + kSQ));
    for (int n0 = 0; n0 < nkp; n0 += 256) {
      const int nn = min(256, nkp - n0);
      const uint32_t idesc = make_idesc(kQTile, nn);
      const uint64_t kd = make_sdesc_sw128(smem_u32(smem + kSK + n0 * 128));
#pragma unroll
      for (int k = 0; k < kHd / 16; ++k) umma_f16(tmem + n0, qd + 2 * k, kd + 2 * k, idesc, k != 0);
    }
    umma_commit(&bars[2]);
  }
  __syncwarp();

  // ---- softmax: thread = (row, column half)
  const int row = (warp & 3) * 32 + lane;
  const int hf = warp >> 2;
  const int chunks = nkp / 32;
  const int c_begin = hf == 0 ? 0 : (chunks + 1) / 2;
  const int c_end = hf == 0 ? (chunks + 1) / 2 : chunks;
  const uint32_t t_row = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  mbar_wait(&bars[2], 0);
  tc_fence_after();
  float m = -INFINITY;
  for (int c = c_begin; c < c_end; ++c) {
    uint32_t v[32];
    tmem_ld32(t_row + c * 32, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (c * 32 + j < N) m = fmaxf(m, __uint_as_float(v[j]));
  }
  s_max[hf * kQTile + row] = m;
  __syncthreads();
  m = fmaxf(s_max[row], s_max[kQTile + row]);
  const bool ce_cta = (ce_prob != nullptr) && (qt == 0);
  const bool ce_thread = ce_cta && (row == ce_row);
  float sum = 0.f;
  for (int c = c_begin; c < c_end; ++c) {
    uint32_t v[32];
    tmem_ld32(t_row + c * 32, v);
    tmem_ld_wait();
    float p[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      p[j] = (c * 32 + j < N) ? exp2f((__uint_as_float(v[j]) - m) * scale_log2) : 0.f;
      sum += p[j];
    }
    if (ce_thread) {
#pragma unroll
      for (int j = 0; j < 32; ++j) s_ce[c * 32 + j] = p[j];
    }
    // P[row][c*32 .. +32) -> swizzled K-major tile: block (c/2), 16-byte chunks (c%2)*4 .. +4
    const uint32_t prow = smem_u32(smem + kSP) + (c >> 1) * (kQTile * 128) + row * 128;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int chunk = ((c & 1) * 4 + q) ^ (row & 7);
      sts128(prow + chunk * 16, pack_op2(p[q * 8 + 0], p[q * 8 + 1]), pack_op2(p[q * 8 + 2], p[q * 8 + 3]),
             pack_op2(p[q * 8 + 4], p[q * 8 + 5]), pack_op2(p[q * 8 + 6], p[q * 8 + 7]));
    }
  }
  s_sum[hf * kQTile + row] = sum;
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  const float inv = 1.0f / (s_sum[row] + s_sum[kQTile + row]);
  if (ce_cta) {        // normalised CE row, written cooperatively (coalesced)
    const float ce_inv = 1.0f / (s_sum[ce_row] + s_sum[kQTile + ce_row]);
    float* ce_dst = ce_prob + ((size_t)b * gridDim.y + h) * ce_stride;
    for (int col = lens_t + (int)threadIdx.x; col < N; col += kAttThreads) ce_dst[col - lens_t] = s_ce[col] * ce_inv;
  }

  // ---- O = P V
  if (threadIdx.x == 0) {
    mbar_wait(&bars[1], 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc(kQTile, kHd, /*b_mn_major=*/true);
    for (int j = 0; j < nkp / 16; ++j) {
      const uint64_t pd = make_sdesc_sw128(smem_u32(smem + kSP + (j >> 2) * (kQTile * 128) + (j & 3) * 32));
      const uint64_t vd = make_sdesc_sw128(smem_u32(smem + kSV + j * 16 * 128));
      umma_f16(tmem, pd, vd, idesc, j != 0);      // O reuses S columns [0,64)
    }
    umma_commit(&bars[3]);
  }
  __syncwarp();
  mbar_wait(&bars[3], 0);
  tc_fence_after();
  {
    uint32_t v[32];
    tmem_ld32(t_row + hf * 32, v);
    tmem_ld_wait();
    if (q0 + row < N) {
      op_t* o = out + ((size_t)b * N + q0 + row) * 768 + h * kHd + hf * 32;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint4 u;
        u.x = pack_op2(__uint_as_float(v[q * 8 + 0]) * inv, __uint_as_float(v[q * 8 + 1]) * inv);
        u.y = pack_op2(__uint_as_float(v[q * 8 + 2]) * inv, __uint_as_float(v[q * 8 + 3]) * inv);
        u.z = pack_op2(__uint_as_float(v[q * 8 + 4]) * inv, __uint_as_float(v[q * 8 + 5]) * inv);
        u.w = pack_op2(__uint_as_float(v[q * 8 + 6]) * inv, __uint_as_float(v[q * 8 + 7]) * inv);
        *reinterpret_cast<uint4*>(o + q * 8) = u;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<kTmemCols>(tmem);
  }
}

int attention_plan(AttnOp* op, const void* qkv, int B, int N) {
  if (N > kMaxKeys || N < 1) return VIPT_ERR_ARG;
  uint64_t dims[3] = {2304, (uint64_t)N, (uint64_t)B};
  uint64_t strides[2] = {2304 * 2, (uint64_t)N * 2304 * 2};
  uint32_t box[3] = {64, 64, 1};
  int r = make_tmap_f16(&op->tm, qkv, 3, dims, strides, box);
  if (r) return r;
  op->B = B; op->N = N; op->nkp = ((N + 31) / 32) * 32;
  return VIPT_OK;
}

int attention_launch(const AttnOp& op, void* out, float* ce_prob, int ce_row, int lens_t, int ce_stride,
                     cudaStream_t st) {
  static bool attr = false;
  if (!attr) {
    if (cudaFuncSetAttribute(attention_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, att_smem_bytes(256)) !=
            cudaSuccess ||
        cudaFuncSetAttribute(attention_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             att_smem_bytes(kMaxKeys)) != cudaSuccess)
      return VIPT_ERR_LAUNCH;
    attr = true;
  }
  static int num_sms = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  // the persistent kernel needs >= one (sequence, head) item per SM to pay off; below that (latency mode) one CTA per
  // 128-query tile keeps more SMs busy
  if (op.N > kQTile && op.B * 12 >= num_sms) return attention4_launch(op, out, ce_prob, ce_row, lens_t, ce_stride, num_sms, st);
  dim3 grid((op.N + kQTile - 1) / kQTile, 12, op.B);
  const float scale_log2 = 0.125f * 1.4426950408889634f;
  const int smem = att_smem_bytes(op.nkp);
  const cudaError_t e =
      op.nkp <= 256 ? launch_pdl(attention_kernel<256>, grid, kAttThreads, smem, st, op.tm, reinterpret_cast<op_t*>(out), ce_prob,
                                 op.N, op.nkp, ce_row, lens_t, ce_stride, scale_log2)
                    : launch_pdl(attention_kernel<512>, grid, kAttThreads, smem, st, op.tm, reinterpret_cast<op_t*>(out), ce_prob,
                                 op.N, op.nkp, ce_row, lens_t, ce_stride, scale_log2);
  return e == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
