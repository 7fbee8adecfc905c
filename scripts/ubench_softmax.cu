// Micro-benchmarks behind the attention kernel's design (run on the GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/_bin/ubench_softmax scripts/ubench_softmax.cu
// Per SM, as a function of the number of warps W (4 / 8 / 16):
//   ldtm      tcgen05.ld 32x32b.x32 throughput (bytes / clk / SM)
//   mufu      ex2.approx throughput (ops / clk / SM)
//   pack      cvt.rn.satfinite.f16x2.f32 throughput (results / clk / SM)
//   softmax1  one-pass softmax block: ld 32 columns -> max -> ex2((s - m) * c) -> sum -> pack -> st.shared
//   softmax2  the same with a separate max pass over TMEM first (two TMEM reads per element)
//   softpoly  softmax1 with half of the exponentials evaluated by an FMA-pipe polynomial
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
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
__device__ __forceinline__ float ex2f(float x) {
  float r;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  asm volatile("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// 2^x on the FMA pipe: x <= 0; split into integer and fraction, degree-3 polynomial on [0,1), exponent by shift-add
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float fl = floorf(x);
  const float f = x - fl;
  float p = fmaf(f, 0.0555054f, 0.2402265f);
  p = fmaf(p, f, 0.6931472f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + ((int)fl << 23));
}

enum { M_LDTM = 0, M_MUFU, M_PACK, M_SOFT1, M_SOFT2, M_POLY, M_NUM };

template <int MODE>
__global__ void __launch_bounds__(512, 1) ubench(int iters, long long* cycles, float* sink) {
  extern __shared__ uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  const uint32_t t_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  const uint32_t srow = smem_u32(smem) + (threadIdx.x * 128);     // 128 B per thread
  float acc = 0.f;
  uint32_t v[32];
  __syncthreads();
  const long long t0 = clock64();
  if (MODE == M_LDTM) {
    for (int it = 0; it < iters; ++it) {
      tmem_ld32(t_lane + ((it * 32) & 511), v);
      if ((it & 3) == 3) { tmem_ld_wait(); acc += __uint_as_float(v[0]) + __uint_as_float(v[31]); }
    }
    tmem_ld_wait();
  } else if (MODE == M_MUFU) {
    float x[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = -0.01f * (lane + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; ++i) x[i] = ex2f(x[i]) - 1.5f;
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) acc += x[i];
  } else if (MODE == M_PACK) {
    float x[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) x[i] = 0.01f * (lane + i);
    uint32_t u = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        const uint32_t p = pack2(x[i], x[i + 1]);
        u ^= p;
        x[i] += 1.0f;
      }
    }
    acc = __uint_as_float(u);
  } else {
    const float c = 0.18f;
    for (int it = 0; it < iters; ++it) {
      const uint32_t ta = t_lane + ((it * 32) & 511);
      float m = -1e30f;
      if (MODE == M_SOFT2) {
        tmem_ld32(ta, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 2) m = fmaxf(m, fmaxf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
      }
      tmem_ld32(ta, v);
      tmem_ld_wait();
      if (MODE != M_SOFT2) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) m = fmaxf(m, fmaxf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
      }
      m = fminf(m, 1e30f);
      const float mc = m * c;
      float p[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float a = fmaf(__uint_as_float(v[i]), c, -mc);
        p[i] = (MODE == M_POLY && (i & 1)) ? ex2_poly(a) : ex2f(a);
      }
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int i = 0; i < 32; i += 2) { s0 += p[i]; s1 += p[i + 1]; }
      acc += s0 + s1;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(srow + (((((it & 1) << 2) | q) ^ (lane & 7)) * 16)),
                     "r"(pack2(p[q * 8 + 0], p[q * 8 + 1])), "r"(pack2(p[q * 8 + 2], p[q * 8 + 3])),
                     "r"(pack2(p[q * 8 + 4], p[q * 8 + 5])), "r"(pack2(p[q * 8 + 6], p[q * 8 + 7]))
                     : "memory");
      }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int MODE>
static void run(const char* name, int warps, int iters, double units_per_thread_iter, const char* unit) {
  const int grid = 148, threads = warps * 32;
  long long* cyc;
  float* sink;
  cudaMalloc(&cyc, grid * sizeof(long long));
  cudaMalloc(&sink, grid * 512 * sizeof(float));
  cudaFuncSetAttribute(ubench<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int rep = 0; rep < 2; ++rep) ubench<MODE><<<grid, threads, 65536>>>(iters, cyc, sink);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0;
  for (int i = 0; i < grid; ++i) mean += (double)h[i] / grid;
  const double units = (double)threads * iters * units_per_thread_iter;
  printf("%-9s warps %2d  cycles %9.0f  %8.2f %s / clk / SM   (%.1f clk per warp-iteration)\n", name, warps, mean, units / mean, unit,
         mean / iters);
  cudaFree(cyc);
  cudaFree(sink);
}

int main() {
  for (int w : {4, 8, 16}) run<M_LDTM>("ldtm", w, 4000, 128.0, "bytes");
  for (int w : {4, 8, 16}) run<M_MUFU>("mufu", w, 500, 32.0, "ex2");
  for (int w : {4, 8, 16}) run<M_PACK>("pack", w, 500, 16.0, "cvt.f16x2");
  for (int w : {4, 8, 16}) run<M_SOFT1>("softmax1", w, 1000, 32.0, "elements");
  for (int w : {4, 8, 16}) run<M_SOFT2>("softmax2", w, 1000, 32.0, "elements");
  for (int w : {4, 8, 16}) run<M_POLY>("softpoly", w, 1000, 32.0, "elements");
  return 0;
}
