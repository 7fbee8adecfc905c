// Engine part 3: the HOST-frame path of the tracker API.  ViPTTrack.track(image) receives a host
// ndarray (ViPT/lib/test/tracker/vipt.py:64-69); the reference crops it on the CPU and uploads the
// 256x256x6 crop.  Here the crop runs on the device, so only the pixels the crop kernel can touch are
// uploaded: the image rectangle [x1, x1+crop_sz) x [y1, y1+crop_sz) clipped to the frame
// (processing_utils.py:32-47), into an engine-owned frame buffer of the full frame geometry (pixels
// outside the rectangle are stale and never read).
//
// Page-locked host frames (cudaHostAlloc / cudaHostRegister / torch pin_memory):
//   * ONE gather kernel per step reads the B rectangles straight from host memory (zero-copy, every byte once,
//     coalesced) and evaluates each rectangle ON THE DEVICE from the fp64 box state.  A step therefore does not
//     need the previous step's boxes on the host: up to kRoiSlots steps can be enqueued before their results
//     are collected (vipt_engine_collect), and the host never stalls the device.
//   * when a step is already in flight, the copy engine uploads -- on a side stream, next to that step -- the
//     rectangle of the last COLLECTED boxes widened by half a crop size, into the other of two frame sets; the
//     gather kernel then only reads what the exact rectangle needs beyond it (normally nothing, ~10 us).  An SM
//     kernel for this speculative part was tried and rejected: its resident CTAs delayed the persistent / cooperative
//     kernels of the running step (register file full) by ~0.5 ms.
// Pageable frames: one cudaMemcpy2DAsync per sequence, rectangle from a host mirror of the box state that collect /
// vipt_engine_results refresh (the whole frame when the mirror is stale).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "engine.h"

namespace vipt {

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      fprintf(stderr, "vipt: %s failed: %s (%s:%d)\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return VIPT_ERR_LAUNCH;                                                          \
    }                                                                                  \
  } while (0)

// Rectangle of the frame that the search crop of box state s can touch: sample_target's x1 / y1 / crop_sz
// (processing_utils.py:32-47, search factor 4.0, vipt.py:67) with one pixel of margin, widened by `grow` crop sizes
// on every side, clipped to the frame.  Empty when the crop kernel will report 'Too small bounding box.' or the
// crop lies outside the frame.
__host__ __device__ __forceinline__ void roi_rect(const double* s, int H, int W, double grow, int& x_lo, int& x_hi, int& y_lo, int& y_hi) {
  x_lo = x_hi = y_lo = y_hi = 0;
  const double csd = ceil(sqrt(s[2] * s[3]) * 4.0);
  if (!(csd >= 1.0) || csd > 1e8) return;
  const double x1 = rint(s[0] + 0.5 * s[2] - csd * 0.5), y1 = rint(s[1] + 0.5 * s[3] - csd * 0.5);
  const double m = 1.0 + floor(grow * csd);
  const double xl = fmax(0.0, x1 - m), xh = fmin((double)W, x1 + csd + m);
  const double yl = fmax(0.0, y1 - m), yh = fmin((double)H, y1 + csd + m);
  if (xh > xl && yh > yl) { x_lo = (int)xl; x_hi = (int)xh; y_lo = (int)yl; y_hi = (int)yh; }
}

// bytes [c0, c1) of row y (6 bytes per pixel): unaligned head, 16-byte vectors when src and dst share their
// alignment, tail
__device__ __forceinline__ void roi_copy_span(const uint8_t* src, uint8_t* dst, long long off, int width, bool vec, int lane) {
  if (width <= 0) return;
  const uint8_t* sp = src + off;
  uint8_t* t = dst + off;
  if (vec) {
    const int head = min(width, (int)((16 - (reinterpret_cast<uintptr_t>(t) & 15)) & 15));
    const int body = (width - head) >> 4;
    if (lane < head) t[lane] = sp[lane];
    const uint4* s4 = reinterpret_cast<const uint4*>(sp + head);
    uint4* t4 = reinterpret_cast<uint4*>(t + head);
    for (int i = lane; i < body; i += 32) t4[i] = s4[i];
    for (int i = head + body * 16 + lane; i < width; i += 32) t[i] = sp[i];
  } else {
    for (int i = lane; i < width; i += 32) t[i] = sp[i];
  }
}

// Zero-copy gather: one CTA column per sequence, warps take rows.  The exact rectangle is evaluated HERE from the
// device box state, so a step can be enqueued before the previous step's boxes have reached the host; only what
// `have` (a rectangle uploaded speculatively by the copy engine while the previous step ran) does not cover is read
// from host memory -- normally nothing.  Also publishes the frame-set pointers the crop kernel reads.
struct RoiD { const uint8_t* src; uint8_t* dst; int H, W; };
__global__ void __launch_bounds__(256) roi_gather_kernel(const __grid_constant__ Engine::RoiSrc srcs,
                                                         const __grid_constant__ Engine::RoiHave have, int use_have,
                                                         uint8_t* const* __restrict__ dst_tab, const int* __restrict__ hw,
                                                         const double* __restrict__ state, int* __restrict__ roi_bytes,
                                                         const uint8_t** ptrs_out) {
  const int b = blockIdx.y;
  const RoiD d = {srcs.p[b], dst_tab[b], hw[2 * b], hw[2 * b + 1]};
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const double s[4] = {state[4 * b], state[4 * b + 1], state[4 * b + 2], state[4 * b + 3]};
  int x_lo, x_hi, y_lo, y_hi;
  roi_rect(s, d.H, d.W, 0.0, x_lo, x_hi, y_lo, y_hi);
  int hx_lo = 0, hx_hi = 0, hy_lo = 0, hy_hi = 0;
  if (use_have) { hx_lo = have.r[b][0]; hx_hi = have.r[b][1]; hy_lo = have.r[b][2]; hy_hi = have.r[b][3]; }
  const bool covered = x_lo >= hx_lo && x_hi <= hx_hi && y_lo >= hy_lo && y_hi <= hy_hi;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ptrs_out[b] = d.dst;
    const int ix = max(0, min(x_hi, hx_hi) - max(x_lo, hx_lo)), iy = max(0, min(y_hi, hy_hi) - max(y_lo, hy_lo));
    roi_bytes[b] = (int)(((long long)(x_hi - x_lo) * (y_hi - y_lo) - (long long)ix * iy) * 6);
  }
  if (covered) return;
  const bool vec = ((reinterpret_cast<uintptr_t>(d.src) ^ reinterpret_cast<uintptr_t>(d.dst)) & 15) == 0;
  for (int y = y_lo + blockIdx.x * 8 + warp; y < y_hi; y += gridDim.x * 8) {
    const long long row = (long long)y * d.W;
    if (y < hy_lo || y >= hy_hi) {
      roi_copy_span(d.src, d.dst, (row + x_lo) * 6, (x_hi - x_lo) * 6, vec, lane);
    } else {      // the part of the row left and right of what is already there
      const int l_hi = min(x_hi, hx_lo), r_lo = max(x_lo, hx_hi);
      roi_copy_span(d.src, d.dst, (row + x_lo) * 6, (l_hi - x_lo) * 6, vec, lane);
      roi_copy_span(d.src, d.dst, (row + r_lo) * 6, (x_hi - r_lo) * 6, vec, lane);
    }
  }
}

int Engine::init_host(const void* const* host_frames, const int* hw_host, const double* boxes_host, cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  // (re)allocate the engine-owned frame buffers (two sets) when the geometry changes
  bool same = (int)own_frames.size() == 2 * B && (int)host_hw.size() == 2 * B;
  for (int i = 0; same && i < 2 * B; ++i) same = host_hw[i] == hw_host[i];
  CK(cudaStreamSynchronize(st));
  if (side_stream) CK(cudaStreamSynchronize(side_stream));
  if (!same) {
    for (void* p : own_frames) cudaFree(p);
    own_frames.assign(2 * B, nullptr);
    for (int i = 0; i < 2 * B; ++i) {
      const int b = i % B;
      const size_t bytes = (size_t)hw_host[2 * b] * hw_host[2 * b + 1] * 6;
      if (hw_host[2 * b] < 1 || hw_host[2 * b + 1] < 1) return VIPT_ERR_ARG;
      CK(cudaMalloc(&own_frames[i], bytes));
    }
    if (!own_ptrs_dev) {
      own_ptrs_dev = (const uint8_t**)dalloc(sizeof(void*) * B);
      own_tab_dev = (uint8_t**)dalloc(sizeof(void*) * 2 * B);
      if (!own_ptrs_dev || !own_tab_dev) return VIPT_ERR_LAUNCH;
    }
    CK(cudaMemcpyAsync(own_tab_dev, own_frames.data(), sizeof(void*) * 2 * B, cudaMemcpyHostToDevice, st));
    host_hw.assign(hw_host, hw_host + 2 * B);
  }
  CK(cudaMemcpyAsync(own_ptrs_dev, own_frames.data(), sizeof(void*) * B, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));     // own_frames.data() is host memory of this object: copy done before return
  ptrs_set = 0;
  host_step = 0;
  for (int b = 0; b < B; ++b)
    CK(cudaMemcpyAsync(own_frames[b], host_frames[b], (size_t)hw_host[2 * b] * hw_host[2 * b + 1] * 6,
                       cudaMemcpyHostToDevice, st));
  host_state.assign(boxes_host, boxes_host + 4 * B);
  spec_state = host_state;
  mirror_valid = true;
  res_pending = 0;
  return init_sequences((const void* const*)own_frames.data(), hw_host, boxes_host, st);
}

int Engine::set_state(const double* boxes_host, cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  host_state.assign(boxes_host, boxes_host + 4 * B);
  spec_state = host_state;
  CK(cudaMemcpyAsync(state, host_state.data(), sizeof(double) * 4 * B, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));
  mirror_valid = true;
  return VIPT_OK;
}

int Engine::track_host(const void* const* host_frames, bool use_graph, int64_t* bytes_copied, cudaStream_t st) {
  if (!finalized || own_frames.empty()) return VIPT_ERR_STATE;
  if (res_pending == kRoiSlots) return VIPT_ERR_STATE;        // collect() a step before enqueueing a fifth
  if (!res_host) {
    CK(cudaHostAlloc((void**)&res_host, res_slot_bytes() * kRoiSlots, cudaHostAllocDefault));
    roi_bytes_dev = (int*)dalloc(sizeof(int) * 4 * B);          // [2 frame sets][speculative, exact][B]
    if (!roi_bytes_dev) return VIPT_ERR_LAUNCH;
    CK(cudaMemsetAsync(roi_bytes_dev, 0, sizeof(int) * 4 * B, st));
  }
  int64_t total = 0;
  long long spec_total = 0;
  // zero-copy gather when every frame is page-locked and device-visible
  bool mapped = opt_zero_copy;
  std::vector<const uint8_t*> dev_src(B, nullptr);
  for (int b = 0; mapped && b < B; ++b) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host_frames[b]) != cudaSuccess || at.type != cudaMemoryTypeHost || !at.devicePointer) {
      cudaGetLastError();
      mapped = false;
    } else {
      dev_src[b] = (const uint8_t*)at.devicePointer;
    }
  }
  if (B > kRoiMaxB) mapped = false;
  const int set = mapped ? (int)(host_step & 1) : 0;
  if (mapped) {
    RoiSrc srcs;
    RoiHave have;
    for (int b = 0; b < B; ++b) srcs.p[b] = dev_src[b];
    uint8_t* const* dst_tab = own_tab_dev + (size_t)set * B;
    int* rb = roi_bytes_dev + (size_t)set * 2 * B;
    const bool speculate = res_pending > 0 && (int)spec_state.size() == 4 * B && opt_speculate;
    long long spec_bytes = 0;
    if (speculate) {
      // A step is in flight and its boxes are not known yet: while it runs, the copy engine uploads -- on a side
      // stream -- the rectangle of the last COLLECTED boxes widened by half a crop size on every side.  The gather
      // kernel below then fetches only what the exact rectangle needs beyond that.
      if (!side_stream) {
        CK(cudaStreamCreateWithFlags(&side_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&spec_ev, cudaEventDisableTiming));
      }
      if (set_free_ev[set]) CK(cudaStreamWaitEvent(side_stream, set_free_ev[set], 0));   // last read two steps ago
      for (int b = 0; b < B; ++b) {
        const int H = host_hw[2 * b], W = host_hw[2 * b + 1];
        int x_lo, x_hi, y_lo, y_hi;
        roi_rect(&spec_state[4 * b], H, W, 0.5, x_lo, x_hi, y_lo, y_hi);
        have.r[b][0] = x_lo; have.r[b][1] = x_hi; have.r[b][2] = y_lo; have.r[b][3] = y_hi;
        if (x_hi <= x_lo || y_hi <= y_lo) continue;
        const size_t off = ((size_t)y_lo * W + x_lo) * 6, width = (size_t)(x_hi - x_lo) * 6, rows = (size_t)(y_hi - y_lo);
        CK(cudaMemcpy2DAsync((uint8_t*)own_frames[(size_t)set * B + b] + off, (size_t)W * 6, (const uint8_t*)host_frames[b] + off,
                             (size_t)W * 6, width, rows, cudaMemcpyHostToDevice, side_stream));
        spec_bytes += (long long)(width * rows);
      }
      CK(cudaEventRecord(spec_ev, side_stream));
      CK(cudaStreamWaitEvent(st, spec_ev, 0));
    }
    roi_gather_kernel<<<dim3(speculate ? 2 : 16, B), 256, 0, st>>>(srcs, have, speculate ? 1 : 0, dst_tab, frame_hw, state, rb,
                                                                   own_ptrs_dev);
    CK(cudaGetLastError());
    spec_total = spec_bytes;
    ptrs_set = set;
    total = -1 - spec_total;                      // zero-copy part known after the step: vipt_engine_collect adds it
  } else {
    if (ptrs_set != 0) {
      CK(cudaMemcpyAsync(own_ptrs_dev, own_frames.data(), sizeof(void*) * B, cudaMemcpyHostToDevice, st));
      CK(cudaStreamSynchronize(st));
      ptrs_set = 0;
    }
    for (int b = 0; b < B; ++b) {
      const int H = host_hw[2 * b], W = host_hw[2 * b + 1];
      const uint8_t* src = (const uint8_t*)host_frames[b];
      uint8_t* dst = (uint8_t*)own_frames[b];
      int x_lo = 0, x_hi = W, y_lo = 0, y_hi = H;
      if (mirror_valid) {
        const double* s = &host_state[4 * b];
        const double csd = ceil(sqrt(s[2] * s[3]) * 4.0);             // search factor 4.0 (vipt.py:67)
        if (!(csd >= 1.0) || csd > 1e8) continue;                      // crop kernel reports 'Too small bounding box.'
        const double x1 = rint(s[0] + 0.5 * s[2] - csd * 0.5), y1 = rint(s[1] + 0.5 * s[3] - csd * 0.5);
        const double xl = fmax(0.0, x1 - 1.0), xh = fmin((double)W, x1 + csd + 1.0);
        const double yl = fmax(0.0, y1 - 1.0), yh = fmin((double)H, y1 + csd + 1.0);
        if (!(xh > xl) || !(yh > yl)) continue;                        // crop entirely outside the frame: all padding
        x_lo = (int)xl; x_hi = (int)xh; y_lo = (int)yl; y_hi = (int)yh;
      }
      const size_t off = ((size_t)y_lo * W + x_lo) * 6;
      const size_t width = (size_t)(x_hi - x_lo) * 6, rows = (size_t)(y_hi - y_lo);
      total += (int64_t)(width * rows);
      if (x_lo == 0 && x_hi == W)
        CK(cudaMemcpyAsync(dst + off, src + off, width * rows, cudaMemcpyHostToDevice, st));
      else
        CK(cudaMemcpy2DAsync(dst + off, (size_t)W * 6, src + off, (size_t)W * 6, width, rows, cudaMemcpyHostToDevice, st));
    }
  }
  if (bytes_copied) *bytes_copied = total;
  mirror_valid = false;          // the device state moves on; collect() / vipt_engine_results() refresh the mirror
  int r = track_enqueue(own_ptrs_dev, use_graph, st);
  if (r) return r;
  ++host_step;
  // this step's results -> pinned slot
  uint8_t* slot = res_host + (size_t)res_head * res_slot_bytes();
  CK(cudaMemcpyAsync(slot, out_boxes, sizeof(double) * 4 * B, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(slot + (size_t)B * 32, out_scores, sizeof(float) * B, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(slot + (size_t)B * 36, roi_bytes_dev + (size_t)set * 2 * B, sizeof(int) * 2 * B, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(slot + (size_t)B * 44, err, 4, cudaMemcpyDeviceToHost, st));
  if (!res_ev[res_head]) CK(cudaEventCreateWithFlags(&res_ev[res_head], cudaEventDisableTiming));
  CK(cudaEventRecord(res_ev[res_head], st));
  // frame set and byte counters of this step may be rewritten after this point (speculative uploads wait for it)
  if (!set_free_ev[set]) CK(cudaEventCreateWithFlags(&set_free_ev[set], cudaEventDisableTiming));
  CK(cudaEventRecord(set_free_ev[set], st));
  res_host_bytes[res_head] = total;
  res_head = (res_head + 1) % kRoiSlots;
  ++res_pending;
  return VIPT_OK;
}

// Results of the OLDEST step enqueued by track_host and not yet collected.
int Engine::collect(double* boxes_host, float* scores_host, int64_t* bytes_copied) {
  if (res_pending == 0 || !res_host) return VIPT_ERR_STATE;
  const int idx = (res_head - res_pending + kRoiSlots) % kRoiSlots;
  CK(cudaEventSynchronize(res_ev[idx]));
  const uint8_t* slot = res_host + (size_t)idx * res_slot_bytes();
  memcpy(boxes_host, slot, sizeof(double) * 4 * B);
  memcpy(scores_host, slot + (size_t)B * 32, sizeof(float) * B);
  long long bytes = res_host_bytes[idx];
  if (bytes < 0) {               // -1 - (bytes the copy engine uploaded speculatively); the gather kernel counted its own
    bytes = -1 - bytes;
    const int* rb = reinterpret_cast<const int*>(slot + (size_t)B * 36);
    for (int b = 0; b < B; ++b) bytes += rb[b];
  }
  if (bytes_copied) *bytes_copied = bytes;
  int herr = 0;
  memcpy(&herr, slot + (size_t)B * 44, 4);
  --res_pending;
  spec_state.assign(boxes_host, boxes_host + 4 * B);
  if (res_pending == 0) {          // the mirror of the box state is current again (pageable-frame path)
    host_state.assign(boxes_host, boxes_host + 4 * B);
    mirror_valid = true;
  }
  return herr;
}

}  // namespace vipt
