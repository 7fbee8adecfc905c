/* C ABI of libvipt_b200.so -- the drop-in boundary of the B200-native ViPT tracking forward.
 *
 * Plain pointers and sizes only; every device pointer is a CUDA device address, every stream a
 * cudaStream_t passed as void*.  All functions return VIPT_OK (0) or a negative error code; the
 * Python shims (vipt_b200.tracker / vipt_b200.model) turn non-zero codes into exceptions, mirroring
 * the reference, which reports errors as Python exceptions only.
 */
#ifndef VIPT_B200_H_
#define VIPT_B200_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VIPT_OK 0
#define VIPT_ERR_ARG -1      /* invalid argument / unsupported shape */
#define VIPT_ERR_LAUNCH -2   /* CUDA launch or runtime failure */
#define VIPT_ERR_DRIVER -3   /* driver entry point / tensor-map encode failure */
#define VIPT_ERR_STATE -4    /* weights not finalised, slot not initialised, ... */
#define VIPT_ERR_SMALL_BOX -5 /* 'Too small bounding box.' (processing_utils.py:34-35) */

const char* vipt_version(void);
const char* vipt_last_error(void);

/* ---- single-operator entry points (used by the per-kernel parity tests) ---- */

/* C = epi(A W^T + bias):  A (M,K) fp16 row-major lda, W (N,K) fp16 row-major ldw  (nn.Linear layout,
 * ViPT/lib/models/layers/attn.py:37,53).  epi: 0 bias->fp16, 1 bias+GELU(erf)->fp16, 2 bias+ReLU->fp16,
 * 3 bias->fp32, 4 resid+bias->fp32.  bn: N tile (256/128/64/32).  K % 64 == 0, N % bn == 0. */
int vipt_op_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int bn, int epi,
                 void* out, int64_t ldc, const float* bias, const float* resid, void* stream);

#ifdef __cplusplus
}
#endif
#endif
