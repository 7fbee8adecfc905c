// sm_100a PTX wrappers: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld).
// Hand-written for this project; bit layouts follow the PTX ISA "tcgen05" matrix- and
// instruction-descriptor tables.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

namespace vipt {

// Operand element type of every tensor-core GEMM / attention operand.  fp16 (not bf16): same
// tcgen05 kind::f16 rate, 3 more mantissa bits -- needed for candidate-elimination parity
// (DESIGN.md "Precision").
typedef __half op_t;
constexpr uint32_t kOpFormat = 0;   // instruction-descriptor a/b format: 0 = F16, 1 = BF16

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// explicit shared-space vector accesses with 32-bit addresses (a generic pointer into dynamic smem compiles to
// ST.E / LD.E with 64-bit address arithmetic)
__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 lds128f(uint32_t saddr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr) : "memory");
  return v;
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug becomes a trap (reported as a launch failure) instead of a hung GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 24)) {
      printf("vipt: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

// ---------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2),
      "r"(c3)
      : "memory");
}
// generic-proxy smem writes -> visible to the async proxy (UMMA operand reads, TMA stores)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------- tcgen05 / TMEM
template <uint32_t kCols>
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst) {   // one full warp
  static_assert(kCols >= 32 && kCols <= 512 && (kCols & (kCols - 1)) == 0, "power of two in [32,512]");
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "n"(kCols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t kCols>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {      // same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives when all previously issued MMAs of this thread have completed
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// kind::f16 instruction descriptor: fp32 accumulate, K-major A; B K-major unless b_mn_major.
__host__ __device__ constexpr uint32_t make_idesc(uint32_t M, uint32_t N, bool b_mn_major = false) {
  return (1u << 4)                      // c_format = F32
         | (kOpFormat << 7)             // a_format
         | (kOpFormat << 10)            // b_format
         | (0u << 15)                   // a_major = K
         | ((b_mn_major ? 1u : 0u) << 16)
         | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// Shared-memory matrix descriptor, SWIZZLE_128B, 128-byte rows, 8-row groups 1024 B apart.
//   K-major operand : rows = M/N index, 64 fp16 of K per row (one swizzle atom wide)
//   MN-major operand: rows = K index, 64 fp16 of M/N per row
// Both use SBO = 1024 B; LBO is unused for a single atom in the leading dimension (encoded 1).
__device__ __forceinline__ uint64_t make_sdesc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);        // start address   [0,14)
  d |= static_cast<uint64_t>(1) << 16;                       // LBO (ignored)   [16,30)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;               // SBO = 1024 B    [32,46)
  d |= static_cast<uint64_t>(1) << 46;                       // descriptor version (sm_100)
  d |= static_cast<uint64_t>(2) << 61;                       // SWIZZLE_128B
  return d;
}

// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// two fp32 -> packed fp16x2, round-to-nearest, saturating to the finite range
__device__ __forceinline__ uint32_t pack_op2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ op_t to_op(float v) {
  uint32_t r = pack_op2(v, 0.f);
  return __ushort_as_half(static_cast<unsigned short>(r & 0xFFFF));
}

// ---------------------------------------------------------------- shared epilogue math
// GELU(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf(z) = z * P(z^2), |z| clamped to 3.2: a degree-8 polynomial
// evaluated entirely on the FMA pipe (max |erf error| 9e-5, below the fp16 resolution of the output).
// The MUFU-based form (rcp + ex2 per element) made the fc1 epilogue XU-bound: 875 TFLOP/s vs 1230 for qkv.
__device__ __forceinline__ uint64_t pk2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {   // packed fp32x2 FMA (sm_100)
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {   // packed fp32x2 add (sm_100)
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ void upk2(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
__device__ __forceinline__ float hsum2(uint64_t v) {
  float a, b;
  upk2(v, a, b);
  return a + b;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
// two elements at a time: x0 = lo, x1 = hi
__device__ __forceinline__ void gelu_erf2x2(float& x0, float& x1) {
  const float z0 = fminf(fmaxf(x0 * 0.70710678118654752f, -3.2f), 3.2f);
  const float z1 = fminf(fmaxf(x1 * 0.70710678118654752f, -3.2f), 3.2f);
  const uint64_t z = pk2(z0, z1);
  const uint64_t u = mul2(z, z);
  uint64_t p = pk2(3.263779069e-08f, 3.263779069e-08f);
  p = fma2(p, u, pk2(-1.673741013e-06f, -1.673741013e-06f));
  p = fma2(p, u, pk2(3.751051790e-05f, 3.751051790e-05f));
  p = fma2(p, u, pk2(-4.884092984e-04f, -4.884092984e-04f));
  p = fma2(p, u, pk2(4.166864145e-03f, 4.166864145e-03f));
  p = fma2(p, u, pk2(-2.504087559e-02f, -2.504087559e-02f));
  p = fma2(p, u, pk2(1.111956437e-01f, 1.111956437e-01f));
  p = fma2(p, u, pk2(-3.755427091e-01f, -3.755427091e-01f));
  p = fma2(p, u, pk2(1.128344418e+00f, 1.128344418e+00f));
  const uint64_t hx = mul2(pk2(x0, x1), pk2(0.5f, 0.5f));
  const uint64_t r = fma2(hx, mul2(p, z), hx);
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x0), "=f"(x1) : "l"(r));
}

// ---- programmatic dependent launch (PDL) -------------------------------------------------------------------
// In latency mode (engine.cu: pdl_select) every kernel of a tracking step is launched with
// cudaLaunchAttributeProgrammaticStreamSerialization (launch_pdl below), so inside the CUDA graph consecutive
// kernels are joined by programmatic edges: the CTAs
// of kernel n+1 become resident and run their prologue (barrier init, TMEM allocation, descriptor prefetch,
// weight staging) while kernel n drains its last tiles, and only then block in pdl_wait() until kernel n has
// completed and its writes are visible.  Protocol, identical in every kernel:
//   prologue that touches no global memory written by another kernel   (kernel parameters and weights are fine)
//   pdl_trigger();   // AFTER the prologue: a successor's CTA can only appear once every CTA of this grid owns its
//                    // shared memory and TMEM, so a successor can never take a resource this grid still needs
//   pdl_wait();      // all threads; before the first read of upstream data and before the first global write
// Both instructions are no-ops when the grid was launched without the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

bool pdl_enabled();      // engine.cu: set per engine by pdl_select()

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace vipt
