// Engine part 3: the HOST-frame path of the tracker API.  ViPTTrack.track(image) receives a host
// ndarray (ViPT/lib/test/tracker/vipt.py:64-69); the reference crops it on the CPU and uploads the
// 256x256x6 crop.  Here the crop runs on the device, so only the pixels the crop kernel can touch are
// uploaded: the image rectangle [x1, x1+crop_sz) x [y1, y1+crop_sz) clipped to the frame
// (processing_utils.py:32-47), one cudaMemcpy2DAsync per sequence into an engine-owned frame buffer
// of the full frame geometry (pixels outside the rectangle are stale and never read).  The rectangle
// is evaluated from a host mirror of the fp64 box state, which vipt_engine_results() refreshes; when
// the mirror is stale (steps enqueued without reading results) the whole frame is uploaded.
#include <math.h>
#include <stdio.h>

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

int Engine::init_host(const void* const* host_frames, const int* hw_host, const double* boxes_host, cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  // (re)allocate the engine-owned frame buffers when the geometry changes
  bool same = (int)own_frames.size() == B && (int)host_hw.size() == 2 * B;
  for (int i = 0; same && i < 2 * B; ++i) same = host_hw[i] == hw_host[i];
  if (!same) {
    CK(cudaStreamSynchronize(st));
    for (void* p : own_frames) cudaFree(p);
    own_frames.assign(B, nullptr);
    for (int b = 0; b < B; ++b) {
      const size_t bytes = (size_t)hw_host[2 * b] * hw_host[2 * b + 1] * 6;
      if (hw_host[2 * b] < 1 || hw_host[2 * b + 1] < 1) return VIPT_ERR_ARG;
      CK(cudaMalloc(&own_frames[b], bytes));
    }
    if (!own_ptrs_dev) {
      own_ptrs_dev = (const uint8_t**)dalloc(sizeof(void*) * B);
      if (!own_ptrs_dev) return VIPT_ERR_LAUNCH;
    }
    CK(cudaMemcpyAsync(own_ptrs_dev, own_frames.data(), sizeof(void*) * B, cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));     // own_frames.data() is host memory of this object: copy done before return
    host_hw.assign(hw_host, hw_host + 2 * B);
  }
  for (int b = 0; b < B; ++b)
    CK(cudaMemcpyAsync(own_frames[b], host_frames[b], (size_t)hw_host[2 * b] * hw_host[2 * b + 1] * 6,
                       cudaMemcpyHostToDevice, st));
  host_state.assign(boxes_host, boxes_host + 4 * B);
  mirror_valid = true;
  return init_sequences((const void* const*)own_frames.data(), hw_host, boxes_host, st);
}

int Engine::set_state(const double* boxes_host, cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  host_state.assign(boxes_host, boxes_host + 4 * B);
  CK(cudaMemcpyAsync(state, host_state.data(), sizeof(double) * 4 * B, cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));
  mirror_valid = true;
  return VIPT_OK;
}

int Engine::track_host(const void* const* host_frames, bool use_graph, int64_t* bytes_copied, cudaStream_t st) {
  if (!finalized || own_frames.empty()) return VIPT_ERR_STATE;
  int64_t total = 0;
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
    if (x_lo == 0 && x_hi == W)
      CK(cudaMemcpyAsync(dst + off, src + off, width * rows, cudaMemcpyHostToDevice, st));
    else
      CK(cudaMemcpy2DAsync(dst + off, (size_t)W * 6, src + off, (size_t)W * 6, width, rows, cudaMemcpyHostToDevice, st));
    total += (int64_t)(width * rows);
  }
  if (bytes_copied) *bytes_copied = total;
  mirror_valid = false;          // the device state moves on; vipt_engine_results() refreshes the mirror
  return track_enqueue(own_ptrs_dev, use_graph, st);
}

}  // namespace vipt
