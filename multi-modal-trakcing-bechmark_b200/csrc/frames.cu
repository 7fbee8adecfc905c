// Frame assembly, the step before the tracking path: get_x_frame / get_rgbd_frame
// (ViPT/lib/train/dataset/depth_utils.py:7-132) after image decode, on the device.
//   'rgbrgb'      (RGBT / RGBE, :54-56,128-130): BGR -> RGB on both images, interleaved to (H, W, 6)
//   'rgbcolormap' (RGBD, :19-24,48-52,119-123) : optional depth clip at min(3 * median, 10000), min-max
//                  normalisation to [0, 255] (cv2.normalize NORM_MINMAX on the 16-bit image), JET colour
//                  map (cv2.applyColorMap: BGR-ordered table), merged behind the RGB image
// The median and the extrema come from an exact 65536-bin histogram of the 16-bit depth image, so the
// result is bit-identical to the numpy / OpenCV sequence of the reference.
#include "kernels.h"

namespace vipt {

__constant__ uint8_t c_jet_bgr[256 * 3] = {
    128, 0, 0, 132, 0, 0, 136, 0, 0, 140, 0, 0, 144, 0, 0, 148, 0, 0, 152, 0, 0, 156, 0, 0,
    160, 0, 0, 164, 0, 0, 168, 0, 0, 172, 0, 0, 176, 0, 0, 180, 0, 0, 184, 0, 0, 188, 0, 0,
    192, 0, 0, 196, 0, 0, 200, 0, 0, 204, 0, 0, 208, 0, 0, 212, 0, 0, 216, 0, 0, 220, 0, 0,
    224, 0, 0, 228, 0, 0, 232, 0, 0, 236, 0, 0, 240, 0, 0, 244, 0, 0, 248, 0, 0, 252, 0, 0,
    255, 0, 0, 255, 4, 0, 255, 8, 0, 255, 12, 0, 255, 16, 0, 255, 20, 0, 255, 24, 0, 255, 28, 0,
    255, 32, 0, 255, 36, 0, 255, 40, 0, 255, 44, 0, 255, 48, 0, 255, 52, 0, 255, 56, 0, 255, 60, 0,
    255, 64, 0, 255, 68, 0, 255, 72, 0, 255, 76, 0, 255, 80, 0, 255, 84, 0, 255, 88, 0, 255, 92, 0,
    255, 96, 0, 255, 100, 0, 255, 104, 0, 255, 108, 0, 255, 112, 0, 255, 116, 0, 255, 120, 0, 255, 124, 0,
    255, 128, 0, 255, 132, 0, 255, 136, 0, 255, 140, 0, 255, 144, 0, 255, 148, 0, 255, 152, 0, 255, 156, 0,
    255, 160, 0, 255, 164, 0, 255, 168, 0, 255, 172, 0, 255, 176, 0, 255, 180, 0, 255, 184, 0, 255, 188, 0,
    255, 192, 0, 255, 196, 0, 255, 200, 0, 255, 204, 0, 255, 208, 0, 255, 212, 0, 255, 216, 0, 255, 220, 0,
    255, 224, 0, 255, 228, 0, 255, 232, 0, 255, 236, 0, 255, 240, 0, 255, 244, 0, 255, 248, 0, 255, 252, 0,
    254, 255, 2, 250, 255, 6, 246, 255, 10, 242, 255, 14, 238, 255, 18, 234, 255, 22, 230, 255, 26, 226, 255, 30,
    222, 255, 34, 218, 255, 38, 214, 255, 42, 210, 255, 46, 206, 255, 50, 202, 255, 54, 198, 255, 58, 194, 255, 62,
    190, 255, 66, 186, 255, 70, 182, 255, 74, 178, 255, 78, 174, 255, 82, 170, 255, 86, 166, 255, 90, 162, 255, 94,
    158, 255, 98, 154, 255, 102, 150, 255, 106, 146, 255, 110, 142, 255, 114, 138, 255, 118, 134, 255, 122, 130, 255, 126,
    126, 255, 130, 122, 255, 134, 118, 255, 138, 114, 255, 142, 110, 255, 146, 106, 255, 150, 102, 255, 154, 98, 255, 158,
    94, 255, 162, 90, 255, 166, 86, 255, 170, 82, 255, 174, 78, 255, 178, 74, 255, 182, 70, 255, 186, 66, 255, 190,
    62, 255, 194, 58, 255, 198, 54, 255, 202, 50, 255, 206, 46, 255, 210, 42, 255, 214, 38, 255, 218, 34, 255, 222,
    30, 255, 226, 26, 255, 230, 22, 255, 234, 18, 255, 238, 14, 255, 242, 10, 255, 246, 6, 255, 250, 1, 255, 254,
    0, 252, 255, 0, 248, 255, 0, 244, 255, 0, 240, 255, 0, 236, 255, 0, 232, 255, 0, 228, 255, 0, 224, 255,
    0, 220, 255, 0, 216, 255, 0, 212, 255, 0, 208, 255, 0, 204, 255, 0, 200, 255, 0, 196, 255, 0, 192, 255,
    0, 188, 255, 0, 184, 255, 0, 180, 255, 0, 176, 255, 0, 172, 255, 0, 168, 255, 0, 164, 255, 0, 160, 255,
    0, 156, 255, 0, 152, 255, 0, 148, 255, 0, 144, 255, 0, 140, 255, 0, 136, 255, 0, 132, 255, 0, 128, 255,
    0, 124, 255, 0, 120, 255, 0, 116, 255, 0, 112, 255, 0, 108, 255, 0, 104, 255, 0, 100, 255, 0, 96, 255,
    0, 92, 255, 0, 88, 255, 0, 84, 255, 0, 80, 255, 0, 76, 255, 0, 72, 255, 0, 68, 255, 0, 64, 255,
    0, 60, 255, 0, 56, 255, 0, 52, 255, 0, 48, 255, 0, 44, 255, 0, 40, 255, 0, 36, 255, 0, 32, 255,
    0, 28, 255, 0, 24, 255, 0, 20, 255, 0, 16, 255, 0, 12, 255, 0, 8, 255, 0, 4, 255, 0, 0, 255,
    0, 0, 252, 0, 0, 248, 0, 0, 244, 0, 0, 240, 0, 0, 236, 0, 0, 232, 0, 0, 228, 0, 0, 224,
    0, 0, 220, 0, 0, 216, 0, 0, 212, 0, 0, 208, 0, 0, 204, 0, 0, 200, 0, 0, 196, 0, 0, 192,
    0, 0, 188, 0, 0, 184, 0, 0, 180, 0, 0, 176, 0, 0, 172, 0, 0, 168, 0, 0, 164, 0, 0, 160,
    0, 0, 156, 0, 0, 152, 0, 0, 148, 0, 0, 144, 0, 0, 140, 0, 0, 136, 0, 0, 132, 0, 0, 128};

__global__ void __launch_bounds__(256)
rgbx_kernel(const uint8_t* __restrict__ bgr, const uint8_t* __restrict__ xbgr, uint8_t* __restrict__ out, int n) {
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    const uint8_t* a = bgr + (size_t)i * 3;
    const uint8_t* b = xbgr + (size_t)i * 3;
    uint8_t* o = out + (size_t)i * 6;
    o[0] = a[2]; o[1] = a[1]; o[2] = a[0];
    o[3] = b[2]; o[4] = b[1]; o[5] = b[0];
  }
}

__global__ void __launch_bounds__(256)
depth_hist_kernel(const uint16_t* __restrict__ dp, unsigned int* __restrict__ hist, int n) {
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) atomicAdd(&hist[dp[i]], 1u);
}

// one CTA of 1024 threads: order statistics (n-1)/2 and n/2, minimum, maximum -> clip value, scale, shift.
// params[0] = clip (as uint), params[1] = scale (float bits), params[2] = shift (float bits)
__global__ void __launch_bounds__(1024)
depth_params_kernel(const unsigned int* __restrict__ hist, int n, int depth_clip, unsigned int* __restrict__ params) {
  __shared__ unsigned int s_sum[1024];
  __shared__ int s_res[4];     // value of order stat k1, k2, min, max
  const int t = threadIdx.x;
  unsigned int loc = 0;
  for (int j = 0; j < 64; ++j) loc += hist[t * 64 + j];
  s_sum[t] = loc;
  if (t < 4) s_res[t] = t == 2 ? 65536 : -1;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {     // inclusive scan
    const unsigned int v = t >= o ? s_sum[t - o] : 0u;
    __syncthreads();
    s_sum[t] += v;
    __syncthreads();
  }
  const unsigned int before = s_sum[t] - loc;
  const unsigned int k1 = (unsigned int)(n - 1) / 2, k2 = (unsigned int)n / 2;
  if (loc) {
    unsigned int c = before;
    int first = -1, last = -1;
    for (int j = 0; j < 64; ++j) {
      const unsigned int h = hist[t * 64 + j];
      if (!h) continue;
      if (first < 0) first = t * 64 + j;
      last = t * 64 + j;
      if (k1 >= c && k1 < c + h) s_res[0] = t * 64 + j;
      if (k2 >= c && k2 < c + h) s_res[1] = t * 64 + j;
      c += h;
    }
    atomicMin(&s_res[2], first);
    atomicMax(&s_res[3], last);
  }
  __syncthreads();
  if (t == 0) {
    double smin = (double)s_res[2], smax = (double)s_res[3];
    unsigned int clip = 65535u;
    if (depth_clip) {
      const double median = 0.5 * ((double)s_res[0] + (double)s_res[1]);       // np.median
      const double max_depth = fmin(median * 3.0, 10000.0);
      clip = (unsigned int)max_depth;                                           // dp[dp > max_depth] = max_depth
      smin = fmin(smin, (double)clip);
      smax = fmin(smax, (double)clip);
    }
    // cv::normalize NORM_MINMAX, alpha 0, beta 255, then convertTo(16U) in float
    const double scale = 255.0 * (smax - smin > 2.220446049250313e-16 ? 1.0 / (smax - smin) : 0.0);
    const double shift = 0.0 - smin * scale;
    params[0] = clip;
    params[1] = __float_as_uint((float)scale);
    params[2] = __float_as_uint((float)shift);
  }
}

__global__ void __launch_bounds__(256)
rgbd_kernel(const uint8_t* __restrict__ bgr, const uint16_t* __restrict__ dp, const unsigned int* __restrict__ params,
            uint8_t* __restrict__ out, int n) {
  const unsigned int clip = params[0];
  const float a = __uint_as_float(params[1]), b = __uint_as_float(params[2]);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    unsigned int d = dp[i];
    d = d > clip ? clip : d;
    float v = rintf(__fmaf_rn((float)d, a, b));           // saturate_cast<ushort>(src * a + b): round half to even
    v = fminf(fmaxf(v, 0.f), 65535.f);
    const int idx = ((int)v) & 255;                       // np.asarray(dp, dtype=np.uint8)
    const uint8_t* c = bgr + (size_t)i * 3;
    uint8_t* o = out + (size_t)i * 6;
    o[0] = c[2]; o[1] = c[1]; o[2] = c[0];
    o[3] = c_jet_bgr[idx * 3]; o[4] = c_jet_bgr[idx * 3 + 1]; o[5] = c_jet_bgr[idx * 3 + 2];
  }
}

int launch_assemble_rgbx(const void* bgr, const void* xbgr, int H, int W, void* out, cudaStream_t st) {
  const int n = H * W;
  int grid = (n + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  rgbx_kernel<<<grid, 256, 0, st>>>((const uint8_t*)bgr, (const uint8_t*)xbgr, (uint8_t*)out, n);
  return cudaGetLastError() == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

int launch_assemble_rgbd(const void* bgr, const void* depth16, int H, int W, int depth_clip, void* out, void* workspace,
                         cudaStream_t st) {
  const int n = H * W;
  unsigned int* hist = (unsigned int*)workspace;            // 65536 bins + 4 parameters
  if (cudaMemsetAsync(hist, 0, 65536 * 4, st) != cudaSuccess) return VIPT_ERR_LAUNCH;
  int grid = (n + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  depth_hist_kernel<<<grid, 256, 0, st>>>((const uint16_t*)depth16, hist, n);
  depth_params_kernel<<<1, 1024, 0, st>>>(hist, n, depth_clip, hist + 65536);
  rgbd_kernel<<<grid, 256, 0, st>>>((const uint8_t*)bgr, (const uint16_t*)depth16, hist + 65536, (uint8_t*)out, n);
  return cudaGetLastError() == cudaSuccess ? VIPT_OK : VIPT_ERR_LAUNCH;
}

// ------------------------------------------------------------------------------------------------
// No-allocation variant of the 'rgbcolormap' assembly (vipt_op_assemble_rgbd_noalloc): the 65536-bin histogram + 4
// parameter words live in a file-scope device array instead of a caller-provided workspace, for callers that cannot
// call cudaMalloc (stream capture, tight serving loops, harnesses that forbid it).  262 160 bytes per GPU that loads
// the library, held until exit.  In host code `g_rgbd_scratch` is only a host-side stand-in for the symbol, never a
// device address: the address is asked for once per device, with that device current, and cached.
// Every call shares the array and it keeps the last call's contents, so the histogram is cleared at the start of
// each call; two calls that overlap on one GPU (different streams) race on it.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned int g_rgbd_scratch[65536 + 4];

int launch_assemble_rgbd_noalloc(const void* bgr, const void* depth16, int H, int W, int depth_clip, void* out,
                                 cudaStream_t st) {
  constexpr int kMaxDevices = 64;
  static unsigned int* scratch_of_device[kMaxDevices] = {};
  if (H < 1 || W < 1 || (long long)H * (long long)W > (long long)VIPT_NOALLOC_MAX_PIXELS) return VIPT_ERR_ARG;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return VIPT_ERR_LAUNCH;
  if (!scratch_of_device[dev]) {            // set-up, once per device, with that device current
    unsigned int* ptr = nullptr;
    if (cudaGetSymbolAddress((void**)&ptr, g_rgbd_scratch) != cudaSuccess || !ptr) return VIPT_ERR_LAUNCH;
    scratch_of_device[dev] = ptr;
  }
  return launch_assemble_rgbd(bgr, depth16, H, W, depth_clip, out, scratch_of_device[dev], st);
}

}  // namespace vipt
