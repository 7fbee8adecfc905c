// Center head around the conv GEMMs (implicit 3x3 convolutions: gemm.cu / gemm2.cu conv mode): the fused
// tail (1x1 conv5 of the three towers, sigmoid/clamp, Hann window, argmax, box decode, map_box_back,
// clip_box, state update).  Reference: ViPT/lib/models/layers/head.py:130-201,
// ViPT/lib/test/tracker/vipt.py:78-88,112-118, ViPT/lib/utils/box_ops.py:97-106.
#include "kernels.h"
#include "sm100_ptx.cuh"

namespace vipt {

__device__ __forceinline__ float sigmoid_clamp(float v) {
  const float s = 1.0f / (1.0f + expf(-v));
  return fminf(fmaxf(s, 1e-4f), 1.0f - 1e-4f);
}

// block argmax with first-index tie-break (torch.max over the flattened map)
__device__ __forceinline__ void block_argmax(float v, int idx, float* s_v, int* s_i, float& best_v, int& best_i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { s_v[threadIdx.x >> 5] = v; s_i[threadIdx.x >> 5] = idx; }
  __syncthreads();
  best_v = s_v[0]; best_i = s_i[0];
  for (int w = 1; w < 8; ++w)
    if (s_v[w] > best_v || (s_v[w] == best_v && s_i[w] < best_i)) { best_v = s_v[w]; best_i = s_i[w]; }
}

// One CTA (256 threads = 16x16 positions) per sequence.
//   h4      : (B*256, 96) fp16 = conv4 outputs of the ctr | offset | size towers
//   w5 / b5 : conv5 weights (5, 32) rows: ctr, offset_x, offset_y, size_w, size_h ; biases (5)
//   maps    : score (B,256), size (B,2,256), offset (B,2,256) fp32; pred_boxes (B,4) un-windowed
//   tracking (state != null): windowed argmax -> box -> map_box_back -> clip_box -> state
__global__ void __launch_bounds__(256)
head_final_kernel(const op_t* __restrict__ h4, const float* __restrict__ w5, const float* __restrict__ b5,
                  const float* __restrict__ window, float* __restrict__ score_map, float* __restrict__ size_map,
                  float* __restrict__ offset_map, float* __restrict__ pred_boxes, double* __restrict__ state,
                  const double* __restrict__ geom, const int* __restrict__ frame_hw, double* __restrict__ out_boxes,
                  float* __restrict__ out_scores) {
  __shared__ float s_w5[5 * 32 + 5];
  pdl_trigger();       // programmatic dependent launch protocol (sm100_ptx.cuh)
  pdl_wait();
  __shared__ float s_v[8];
  __shared__ int s_i[8];
  __shared__ float s_size[2][256], s_off[2][256];
  const int b = blockIdx.x, pos = threadIdx.x;
  if (pos < 160) s_w5[pos] = w5[pos];
  if (pos < 5) s_w5[160 + pos] = b5[pos];
  __syncthreads();
  float hv[96];
  const uint4* src = reinterpret_cast<const uint4*>(h4 + ((size_t)b * 256 + pos) * 96);
#pragma unroll
  for (int i = 0; i < 12; ++i) {
    const uint4 u = src[i];
    const __half2* hp = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __half22float2(hp[j]);
      hv[i * 8 + j * 2] = f.x; hv[i * 8 + j * 2 + 1] = f.y;
    }
  }
  float o[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const int tower = (k == 0) ? 0 : (k < 3 ? 1 : 2);
    float a = s_w5[160 + k];
#pragma unroll
    for (int c = 0; c < 32; ++c) a += s_w5[k * 32 + c] * hv[tower * 32 + c];
    o[k] = a;
  }
  const float score = sigmoid_clamp(o[0]);
  const float sw = sigmoid_clamp(o[3]), sh = sigmoid_clamp(o[4]);
  score_map[b * 256 + pos] = score;
  offset_map[(b * 2 + 0) * 256 + pos] = o[1]; offset_map[(b * 2 + 1) * 256 + pos] = o[2];
  size_map[(b * 2 + 0) * 256 + pos] = sw; size_map[(b * 2 + 1) * 256 + pos] = sh;
  s_off[0][pos] = o[1]; s_off[1][pos] = o[2]; s_size[0][pos] = sw; s_size[1][pos] = sh;
  float bv; int bi;
  block_argmax(score, pos, s_v, s_i, bv, bi);           // forward()'s pred_boxes: un-windowed (head.py:136)
  if (pos == 0) {
    pred_boxes[b * 4 + 0] = ((float)(bi % 16) + s_off[0][bi]) / 16.0f;
    pred_boxes[b * 4 + 1] = ((float)(bi / 16) + s_off[1][bi]) / 16.0f;
    pred_boxes[b * 4 + 2] = s_size[0][bi]; pred_boxes[b * 4 + 3] = s_size[1][bi];
  }
  if (!state) return;
  const float resp = window[pos] * score;               // vipt.py:80
  block_argmax(resp, pos, s_v, s_i, bv, bi);
  if (pos == 0) {
    const float cxn = ((float)(bi % 16) + s_off[0][bi]) / 16.0f;
    const float cyn = ((float)(bi / 16) + s_off[1][bi]) / 16.0f;
    const double rf = geom[b * 4 + 3];
    const float rff = (float)rf;
    // pred_boxes.mean(0) * search_size / resize_factor in fp32, then .tolist() -> doubles (vipt.py:85-86)
    const double cx = (double)((cxn * 256.0f) / rff), cy = (double)((cyn * 256.0f) / rff);
    const double w = (double)((s_size[0][bi] * 256.0f) / rff), h = (double)((s_size[1][bi] * 256.0f) / rff);
    double* stt = state + b * 4;
    const double cx_prev = stt[0] + 0.5 * stt[2], cy_prev = stt[1] + 0.5 * stt[3];
    const double half_side = 0.5 * 256.0 / rf;
    const double cx_real = cx + (cx_prev - half_side), cy_real = cy + (cy_prev - half_side);
    double x1 = cx_real - 0.5 * w, y1 = cy_real - 0.5 * h;
    const double H = (double)frame_hw[b * 2 + 0], W = (double)frame_hw[b * 2 + 1], margin = 10.0;
    double x2 = x1 + w, y2 = y1 + h;                    // clip_box(box, H, W, margin=10)
    x1 = fmin(fmax(0.0, x1), W - margin); x2 = fmin(fmax(margin, x2), W);
    y1 = fmin(fmax(0.0, y1), H - margin); y2 = fmin(fmax(margin, y2), H);
    const double wn = fmax(margin, x2 - x1), hn = fmax(margin, y2 - y1);
    stt[0] = x1; stt[1] = y1; stt[2] = wn; stt[3] = hn;
    out_boxes[b * 4 + 0] = x1; out_boxes[b * 4 + 1] = y1; out_boxes[b * 4 + 2] = wn; out_boxes[b * 4 + 3] = hn;
    out_scores[b] = bv;
  }
}

int launch_head_final(const void* h4, const float* w5, const float* b5, const float* window, float* score_map,
                      float* size_map, float* offset_map, float* pred_boxes, double* state, const double* geom,
                      const int* frame_hw, double* out_boxes, float* out_scores, int B, cudaStream_t st) {
  const cudaError_t e_ = launch_pdl(head_final_kernel, B, 256, 0, st, reinterpret_cast<const op_t*>(h4), w5, b5, window, score_map, size_map, offset_map, pred_boxes, state, geom, frame_hw, out_boxes, out_scores);
  return e_ == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

}  // namespace vipt
