// Fused crop + zero-pad + OpenCV-exact bilinear resize + normalise + patch-major (im2col) store.
//
// Replaces, in one HBM-bound kernel, the reference's host-side
//   sample_target          ViPT/lib/train/data/processing_utils.py:14-76   (cv2 crop/pad/resize)
//   PreprocessorMM.process ViPT/lib/test/tracker/data_utils.py:20-24       (/255, -mean, /std)
// and the im2col that Conv2d(k16,s16) (patch_embed.py:28) implies.  Output rows are tokens, columns
// are c*256 + ky*16 + kx, so the patch embedding is a plain GEMM against proj.weight.view(768,768).
//
// Integer arithmetic follows cv::resize INTER_LINEAR for 8-bit images bit-exactly (11-bit fixed-point
// coefficients, two-stage rounding; exact 2x downscale -> the generic INTER_AREA path that 6-channel
// images take: round-half-even of the 2x2 mean).  Geometry is evaluated in fp64 like the reference's
// Python floats (ceil, round-half-even = rint).
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

// PreprocessorMM.process maps a u8 value to ((v / 255) - mean) / std in fp32 (data_utils.py:22-23); the patch rows are
// fp16, so the whole map is a 3 x 256 table of fp16 values, computed once on the host with the same IEEE operations
// (two fp32 divisions per channel value made the kernel issue-bound: 550 instructions per output pixel).
static const float h_mean[3] = {0.485f, 0.456f, 0.406f};
static const float h_std[3] = {0.229f, 0.224f, 0.225f};
__device__ __align__(16) unsigned short g_norm_lut[3 * 256];

int crop_prepare() {
  static bool ready = false;
  if (ready) return VIPT_OK;
  unsigned short h[3 * 256];
  for (int c = 0; c < 3; ++c)
    for (int v = 0; v < 256; ++v) {
      volatile float q = (float)v / 255.0f;          // volatile: each step rounded to fp32, no contraction
      volatile float d = q - h_mean[c];
      volatile float r = d / h_std[c];
      h[c * 256 + v] = __half_as_ushort(__float2half_rn(r));
    }
  if (cudaMemcpyToSymbol(g_norm_lut, h, sizeof(h)) != cudaSuccess) return VIPT_ERR_LAUNCH;
  ready = true;
  return VIPT_OK;
}

struct Taps {
  int i0, i1;     // source indices in the padded crop (already clipped)
  int c0, c1;     // 11-bit coefficients
};

// horizontal taps: fraction zeroed at the borders; vertical: indices clipped, fraction kept
__device__ __forceinline__ Taps make_taps(int d, int src, double scale, bool horizontal) {
  float f = (float)__dsub_rn(__dmul_rn((double)d + 0.5, scale), 0.5);   // no FMA contraction: match the CPU
  int s = (int)floorf(f);
  f -= (float)s;
  Taps t;
  if (horizontal) {
    if (s < 0) { f = 0.f; s = 0; }
    if (s >= src - 1) { f = 0.f; s = src - 1; }
    t.i0 = s;
    t.i1 = min(s + 1, src - 1);
  } else {
    t.i0 = min(max(s, 0), src - 1);
    t.i1 = min(max(s + 1, 0), src - 1);
  }
  t.c0 = __float2int_rn((1.f - f) * 2048.f);
  t.c1 = __float2int_rn(f * 2048.f);
  return t;
}

// grid: (S*S/(8*256), B); one thread per 8 consecutive output pixels of a row, all 6 channels: the vertical taps, the
// row pointers and the patch-row address are shared by the 8 pixels and each channel leaves as one 16-byte store.
constexpr int kCropPix = 8;
__global__ void __launch_bounds__(256)
crop_kernel(const uint8_t* const* __restrict__ frames, const int* __restrict__ frame_hw,
            const double* __restrict__ boxes, double factor, int S, int row_base, int rows_per_seq,
            op_t* __restrict__ apatch, int64_t group_stride, double* __restrict__ geom, int* __restrict__ err,
            int* __restrict__ slot_err) {
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh); first kernel of a step: its
  pdl_wait();          // predecessor is the head of the previous step, which wrote the box state read here
  const int b = blockIdx.y;
  const int unit = blockIdx.x * 256 + threadIdx.x;   // over S * S / 8
  const int lgU = 31 - __clz(S) - 3;                 // S is 128 or 256: 16 or 32 units per row
  const int oy = unit >> lgU, ox0 = (unit & ((S >> 3) - 1)) << 3;
  __shared__ __align__(16) unsigned short s_lut[3 * 256];
  if (threadIdx.x < 96) reinterpret_cast<uint4*>(s_lut)[threadIdx.x] = reinterpret_cast<const uint4*>(g_norm_lut)[threadIdx.x];
  // the crop geometry of a sequence (fp64 sqrt, two fp64 divisions) is evaluated by one thread per CTA: evaluated
  // by every thread it made the kernel fp64-bound (87 us per 64-sequence step, 9 % of the HBM roofline)
  __shared__ double s_geo[2];
  __shared__ int s_gi[3];
  if (threadIdx.x == 0) {
    const double bx = boxes[b * 4 + 0], by = boxes[b * 4 + 1], bw = boxes[b * 4 + 2], bh = boxes[b * 4 + 3];
    const double c = ceil(sqrt(bw * bh) * factor);
    s_geo[0] = c;
    s_geo[1] = c >= 1.0 ? 1.0 / ((double)S / (double)(int)c) : 0.0;
    s_gi[0] = c >= 1.0 ? (int)c : 0;
    s_gi[1] = (int)rint(bx + 0.5 * bw - c * 0.5);
    s_gi[2] = (int)rint(by + 0.5 * bh - c * 0.5);
  }
  __syncthreads();
  const double csd = s_geo[0];
  if (!(csd >= 1.0)) {           // 'Too small bounding box.' (processing_utils.py:34)
    // the engine-wide flag is raised once per failing sequence (its per-slot flag is sticky until the next initialize),
    // so the sequences that keep tracking are not reported again on every later step
    if (unit == 0 && threadIdx.x == 0) {
      if (!slot_err) atomicExch(err, VIPT_ERR_SMALL_BOX);
      else if (slot_err[b] == 0) { slot_err[b] = VIPT_ERR_SMALL_BOX; atomicExch(err, VIPT_ERR_SMALL_BOX); }
    }
    return;
  }
  const int cs = s_gi[0];
  const int x1 = s_gi[1];
  const int y1 = s_gi[2];
  if (unit == 0 && geom) {
    geom[b * 4 + 0] = (double)x1; geom[b * 4 + 1] = (double)y1; geom[b * 4 + 2] = csd;
    geom[b * 4 + 3] = (double)S / csd;       // resize_factor (processing_utils.py:69)
  }
  const int H = frame_hw[b * 2 + 0], W = frame_hw[b * 2 + 1];
  const uint8_t* im = frames[b];
  // padded-crop pixel (py, px) is image pixel (y1+py, x1+px) when 0 <= Y <= H-2 and 0 <= X <= W-2
  // (the `+1` in x2_pad / y2_pad drops the last row / column whenever the crop overhangs), else 0.
  const bool even = (reinterpret_cast<uintptr_t>(im) & 1) == 0;      // a pixel is 6 bytes: 2-byte aligned when the frame is
  auto load6 = [&](const uint8_t* rowp, bool row_ok, int px, int (&v)[6]) {
    const int X = x1 + px;
    if (row_ok && X >= 0 && X <= W - 2) {
      const uint8_t* p = rowp + (size_t)X * 6;
      if (even) {                 // three 16-bit loads instead of six byte loads
        const unsigned short* q = reinterpret_cast<const unsigned short*>(p);
        const unsigned u0 = __ldg(q), u1 = __ldg(q + 1), u2 = __ldg(q + 2);
        v[0] = u0 & 255; v[1] = u0 >> 8; v[2] = u1 & 255; v[3] = u1 >> 8; v[4] = u2 & 255; v[5] = u2 >> 8;
      } else {
#pragma unroll
        for (int c = 0; c < 6; ++c) v[c] = p[c];
      }
    } else {
#pragma unroll
      for (int c = 0; c < 6; ++c) v[c] = 0;
    }
  };
  const bool area = cs == 2 * S;
  const double scale = s_geo[1];
  Taps ty;
  if (area) { ty.i0 = 2 * oy; ty.i1 = 2 * oy + 1; ty.c0 = ty.c1 = 0; }
  else ty = make_taps(oy, cs, scale, false);
  const int Y0 = y1 + ty.i0, Y1 = y1 + ty.i1;
  const bool ok0 = Y0 >= 0 && Y0 <= H - 2, ok1 = Y1 >= 0 && Y1 <= H - 2;
  const uint8_t* r0p = im + (size_t)(ok0 ? Y0 : 0) * W * 6;
  const uint8_t* r1p = im + (size_t)(ok1 ? Y1 : 0) * W * 6;
  uint32_t o[6][kCropPix / 2];
#pragma unroll
  for (int c = 0; c < 6; ++c)
#pragma unroll
    for (int q = 0; q < kCropPix / 2; ++q) o[c][q] = 0u;
#pragma unroll
  for (int px = 0; px < kCropPix; ++px) {
    const int ox = ox0 + px;
    int out[6];
    if (area) {
      int a[6], bb[6], c2[6], d[6];
      load6(r0p, ok0, 2 * ox, a); load6(r0p, ok0, 2 * ox + 1, bb); load6(r1p, ok1, 2 * ox, c2); load6(r1p, ok1, 2 * ox + 1, d);
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const int sum = a[c] + bb[c] + c2[c] + d[c];
        const int q = sum >> 2, r = sum & 3;
        out[c] = q + ((r == 3) || (r == 2 && (q & 1)));
      }
    } else {
      const Taps tx = make_taps(ox, cs, scale, true);
      int p00[6], p01[6], p10[6], p11[6];
      load6(r0p, ok0, tx.i0, p00); load6(r0p, ok0, tx.i1, p01); load6(r1p, ok1, tx.i0, p10); load6(r1p, ok1, tx.i1, p11);
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const int r0 = p00[c] * tx.c0 + p01[c] * tx.c1;
        const int r1 = p10[c] * tx.c0 + p11[c] * tx.c1;
        int v = (((ty.c0 * (r0 >> 4)) >> 16) + ((ty.c1 * (r1 >> 4)) >> 16) + 2) >> 2;
        out[c] = min(max(v, 0), 255);
      }
    }
#pragma unroll
    for (int c = 0; c < 6; ++c) o[c][px >> 1] |= (uint32_t)s_lut[(c % 3) * 256 + out[c]] << (16 * (px & 1));
  }
  const int tps = S / 16;
  const int token = (oy >> 4) * tps + (ox0 >> 4);
  const int col = (oy & 15) * 16 + (ox0 & 15);        // multiple of 8: 16-byte aligned in the fp16 patch row
  const size_t row = (size_t)b * rows_per_seq + row_base + token;
  unsigned short* dst = reinterpret_cast<unsigned short*>(apatch) + row * 768 + col;
#pragma unroll
  for (int c = 0; c < 6; ++c)
    *reinterpret_cast<uint4*>(dst + (size_t)(c / 3) * group_stride + (c % 3) * 256) = make_uint4(o[c][0], o[c][1], o[c][2], o[c][3]);
}

// fp32 NCHW (B,6,S,S) -> patch-major fp16 rows (the forward(template, search) entry).
__global__ void __launch_bounds__(256)
patchify_kernel(const float* __restrict__ img, int S, int row_base, int rows_per_seq, op_t* __restrict__ apatch,
                int64_t group_stride) {
  const int b = blockIdx.y;
  const int idx = blockIdx.x * 256 + threadIdx.x;   // over 6*S*S
  const int c = idx / (S * S);
  const int oy = (idx / S) % S, ox = idx % S;
  const float v = img[(size_t)b * 6 * S * S + idx];
  const int tps = S / 16;
  const int token = (oy >> 4) * tps + (ox >> 4);
  const int col = (oy & 15) * 16 + (ox & 15);
  const size_t row = (size_t)b * rows_per_seq + row_base + token;
  apatch[(size_t)(c / 3) * group_stride + row * 768 + (c % 3) * 256 + col] = to_op(v);
}

int launch_crop(const uint8_t* const* frames, const int* frame_hw, const double* boxes, double factor, int S,
                int row_base, int rows_per_seq, void* apatch, int64_t group_stride, double* geom, int* err, int B,
                cudaStream_t st, int* slot_err) {
  if ((S != 128 && S != 256) || crop_prepare()) return VIPT_ERR_ARG;
  if ((group_stride & 7) || (reinterpret_cast<uintptr_t>(apatch) & 15)) return VIPT_ERR_ARG;      // 16-byte stores
  dim3 grid(S * S / (kCropPix * 256), B);
  const cudaError_t e_ = launch_pdl(crop_kernel, grid, 256, 0, st, frames, frame_hw, boxes, factor, S, row_base, rows_per_seq,
                                    reinterpret_cast<op_t*>(apatch), group_stride, geom, err, slot_err);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

int launch_patchify(const float* img, int S, int row_base, int rows_per_seq, void* apatch, int64_t group_stride, int B,
                    cudaStream_t st) {
  dim3 grid(6 * S * S / 256, B);
  patchify_kernel<<<grid, 256, 0, st>>>(img, S, row_base, rows_per_seq, reinterpret_cast<op_t*>(apatch), group_stride);
  return cudaGetLastError() == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
