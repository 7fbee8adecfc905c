// extern "C" entry points of libvipt_b200.so (declared in include/vipt_b200.h).
#include <stdio.h>
#include <string.h>

#include "gemm.h"

static thread_local char g_err[256] = "";
static void set_err(const char* m) { strncpy(g_err, m, sizeof(g_err) - 1); }

static int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

extern "C" {

const char* vipt_version(void) { return "vipt_b200 0.1 (sm_100a)"; }
const char* vipt_last_error(void) { return g_err; }

int vipt_op_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int bn, int epi,
                 void* out, int64_t ldc, const float* bias, const float* resid, void* stream) {
  vipt::GemmOp op;
  int r = vipt::gemm_plan(&op, A, lda, 0, W, ldw, 0, 1, M, N, K, bn, epi, out, ldc, 0, 0, bias, resid);
  if (r) { set_err("gemm_plan failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}

}  // extern "C"
