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
 * 3 bias->fp32, 4 resid+bias->fp32.  bn: N tile (256/128/64/32), or 512 / 640 for the 2-CTA kernel with 256- / 128-wide pair tiles.  K % 64 == 0, N % tile == 0. */
int vipt_op_gemm(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int bn, int epi,
                 void* out, int64_t ldc, const float* bias, const float* resid, void* stream);

/* LayerNorm folded into a GEMM pair (north star (b); ViPT/lib/models/layers/attn_blocks.py:94,103: x = x + proj(..);
 * mlp(norm2(x))), 2-CTA kernel, N % 256 == 0.
 * Producer: x (M,N) fp32 <- x + A W^T + bias in place, x16 (M,N) fp16 = the updated rows, stat_out (N/32, M, 2)
 * fp32 = per-row (sum, sum of squares) of 32 columns each (part 2g + h: the 16-column groups h and h+2 of columns
 * [64g, 64g+64)). */
int vipt_op_gemm_resid_stat(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x, void* x16,
                            const float* bias, float* stat_out, void* stream);
/* Consumer: out fp16 = [gelu](rstd * (A W^T - mean * colsum) + bias) with mean / rstd of every row of A (width K) from
 * stat_in (stat_parts, M, 2); the caller passes W = fp16(gamma * W0), colsum = row sums of that fp16 W and
 * bias = b0 + W0 beta, which makes this LayerNorm(x) W0^T + b0.  stat_parts = 0: stat_in (M, 2) holds the final
 * (rstd, -mean * rstd) of every row.  Ax (M, 64) / Wx (N, 64) fp16 or both NULL: one extra k-block, acc += Ax Wx^T (the
 * rank-8 prompt update of ViPT/lib/models/vipt/vit_ce_prompt.py:308-310 on the tensor pipe). */
int vipt_op_gemm_lnfold(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int gelu, void* out,
                        int64_t ldc, const float* bias, const float* colsum, const float* stat_in, int stat_parts, float eps,
                        const void* Ax, const void* Wx, void* stream);
/* Prompt fold, producer (mlp.fc2 before a deep-prompt stage, vit_ce_prompt.py:268-310): x <- x + A W^T + bias in place,
 * x16 = fp16(x), and part_out (2*N/256, M, 20) fp32: per row and (256-column tile t, half h: the 16-column groups h, h+2, ..
 * of the tile) the sums {x, x^2, x.wa_0..7, x.u_0..7, x.b1, 0} over the part's columns; colw20 (N, 20) fp32 holds
 * wa_0..7, u_0..7, b1, 3 pad per column. */
int vipt_op_gemm_resid_prompt(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x, void* x16,
                              const float* bias, const float* colw20, float* part_out, void* stream);

/* Fused crop + zero pad + cv2-exact bilinear resize + /255, mean, std + patch-major store.  Replaces
 * sample_target (ViPT/lib/train/data/processing_utils.py:14-76) + PreprocessorMM.process
 * (ViPT/lib/test/tracker/data_utils.py:20-24).  frame_ptrs_dev: device array of B pointers to uint8
 * (H,W,6) frames; hw_dev (B,2) int H,W; boxes_dev (B,4) double x,y,w,h.  apatch: fp16 (2, B*rows_per_seq, 768),
 * token rows start at row_base, column = c*256 + ky*16 + kx.  geom_dev (B,4) double: x1, y1, crop_sz, resize_factor. */
int vipt_op_crop(const void* frame_ptrs_dev, const int* hw_dev, const double* boxes_dev, double factor, int S,
                 int row_base, int rows_per_seq, void* apatch, double* geom_dev, int* err_dev, int B, void* stream);

/* Fused joint attention over qkv (B*N, 2304) fp16 -> out (B*N, 768) fp16; when ce_prob != NULL also writes the
 * softmax row of template token 27 over the search keys, per head: ce_prob (B, 12, 256) fp32
 * (ViPT/lib/models/layers/attn.py:33-59, attn_blocks.py:44-53). */
int vipt_op_attention(const void* qkv, int B, int N, void* out, float* ce_prob, void* stream);

/* Candidate-elimination ranking (ViPT/lib/models/layers/attn_blocks.py:37-63): score[j] = mean over heads of
 * ce_prob (B, n_heads, 256)[.., j], j < n_s; descending order, exact ties keep the LOWER slot first (torch.sort is
 * not stable, so the reference leaves the order of exact ties unspecified).  gi_in/gi_out (B,256): raster position
 * of each slot before / after; perm (B,256): old slot of new slot r; removed (B,256): raster positions of the dropped
 * tokens in score order; slot_of_pos (B,256): new slot of a raster position or -1; score_out (B,256) or NULL. */
int vipt_op_ce_topk(const float* ce_prob, int n_heads, int n_s, int n_keep, const int* gi_in, int* gi_out, int* perm,
                    int* removed, int* slot_of_pos, float* score_out, int B, void* stream);

/* 3x3 convolution, padding 1, + bias + ReLU over a (B,16,16,c_total) fp16 feature map as an implicit GEMM (the
 * conv + folded BatchNorm + ReLU of ViPT/lib/models/layers/head.py:8-21).  Group g reads channels
 * [g*cin_stride, +cin) and writes columns [g*out_group_col, +N) of out (B*256, ldc) fp16; W (G, N, 9*cin) fp16 with
 * K order tap*cin + c, tap = ky*3 + kx; bias (G, N).  bn: N-tile of the single-CTA kernel (32..256), or 512 for the
 * 2-CTA kernel (G = 1, N % 256 == 0). */
int vipt_op_conv3x3(const void* feat, int c_total, int cin, int cin_stride, const void* W, int G, int B, int N, int bn,
                    void* out, int64_t ldc, int64_t out_group_col, const float* bias, void* stream);

/* Frame assembly after image decode (get_x_frame / get_rgbd_frame, ViPT/lib/train/dataset/depth_utils.py:7-132).
 * rgbx: dtype 'rgbrgb' -- two BGR uint8 (H,W,3) images -> RGB||RGB uint8 (H,W,6).
 * rgbd: dtype 'rgbcolormap' -- BGR uint8 (H,W,3) + uint16 depth (H,W): optional clip at min(3*median, 10000),
 * min-max normalisation to [0,255], JET colour map (BGR order, as cv2.applyColorMap returns it) -> (H,W,6).
 * workspace_dev: VIPT_ASSEMBLE_WORKSPACE_BYTES of device memory. */
#define VIPT_ASSEMBLE_WORKSPACE_BYTES ((65536 + 4) * 4)
int vipt_op_assemble_rgbx(const void* bgr_dev, const void* xbgr_dev, int H, int W, void* out_dev, void* stream);
int vipt_op_assemble_rgbd(const void* bgr_dev, const void* depth16_dev, int H, int W, int depth_clip, void* out_dev,
                          void* workspace_dev, void* stream);

/* No-allocation variant of vipt_op_assemble_rgbd.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same inputs, outputs and results as vipt_op_assemble_rgbd, but the
 * histogram scratch ((65536 + 4) * 4 = 262 160 bytes) is a file-scope `static __device__` array of the library: every
 * GPU that loads the library holds it until exit; nothing is allocated or freed by the call (no cudaMalloc / cudaFree,
 * no allocating library routine).  Limits: H * W <= VIPT_NOALLOC_MAX_PIXELS, otherwise VIPT_ERR_ARG with a message in
 * vipt_last_error().  Every call on a GPU shares the one array (cleared at the start of each call): calls that overlap
 * on the same GPU (different streams, no ordering) race; serialise them or use the workspace variant. */
#define VIPT_NOALLOC_MAX_PIXELS (4096 * 4096)
int vipt_op_assemble_rgbd_noalloc(const void* bgr_dev, const void* depth16_dev, int H, int W, int depth_clip,
                                  void* out_dev, void* stream);

/* ---- engine: the drop-in for ViPTTrack.initialize/track (ViPT/lib/test/tracker/vipt.py:41-110) and for
 * build_viptrack(...).forward(template, search, ce_template_mask) (ViPT/lib/models/vipt/ostrack_prompt.py:39-58),
 * batched over `batch` independent sequences ---- */
int vipt_engine_create(void** out_engine);
int vipt_engine_destroy(void* engine);
/* one call per state_dict entry (fp32, host memory), names exactly as in the reference checkpoint['net']
 * (vipt.py:21); then finalize packs them (fp16 GEMM operands, BN folded into the head convs, LN folded into
 * the prompt down-projections) for `batch` concurrent sequences; deep = 1 for vipt_deep, 0 for vipt_shaw. */
int vipt_engine_set_tensor(void* engine, const char* name, const float* host_data, int64_t numel);
int vipt_engine_finalize(void* engine, int batch, int deep);
/* model-level forward on device tensors: template (B,6,128,128) fp32, search (B,6,256,256) fp32.  Outputs are read
 * with vipt_engine_buffer: score_map, size_map, offset_map, pred_boxes, removed, feat32 (if want_feat32).
 * debug_stop > 0 stops after that stage (parity tests only). */
int vipt_engine_forward(void* engine, const float* template_dev, const float* search_dev, int want_feat32,
                        int debug_stop, void* stream);
/* initialize(): frame pointers (host array of B device pointers, uint8 HWC 6 channels), hw (B,2), init boxes (B,4). */
int vipt_engine_init(void* engine, const void* const* frame_dev_ptrs_host, const int* hw_host,
                     const double* boxes_host, void* stream);
/* track(): enqueue one step for all B sequences (crop at the current state -> backbone -> head -> state update).
 * frame_ptrs_dev: DEVICE array of B frame pointers.  No host synchronisation. use_graph: replay a CUDA graph. */
int vipt_engine_track_enqueue(void* engine, const void* frame_ptrs_dev, int use_graph, void* stream);
/* copies the latest boxes (B,4) [x,y,w,h] and best scores (B) to host memory and synchronises the stream. */
int vipt_engine_results(void* engine, double* boxes_host, float* scores_host, void* stream);
/* Per-sequence error flags (B ints: 0 or VIPT_ERR_SMALL_BOX), sticky until the next init.  A sequence whose box
 * collapses ('Too small bounding box.', processing_utils.py:34) keeps its last crop and no longer disturbs the others:
 * init / results / collect report VIPT_ERR_SMALL_BOX ONCE per newly failing sequence; this call tells which ones and
 * lowers the engine-wide flag, so the caller can drop them and keep tracking the rest (the reference loses only the
 * offending sequence's process, RGBT_workspace/test_rgbt_mgpus.py:66-117). */
int vipt_engine_slot_errors(void* engine, int* slot_err_host, void* stream);
/* HOST-frame variants (the reference hands ViPTTrack host ndarrays, vipt.py:41,64): frames are B host pointers to
 * uint8 (H,W,6) images.  init_host uploads the whole frames into engine-owned device buffers; track_host brings in
 * only the rectangle the search crop can read (processing_utils.py:32-47) and enqueues one step.  Page-locked
 * frames (cudaHostAlloc / cudaHostRegister / torch pin_memory): one gather kernel reads all B rectangles straight
 * from host memory, the rectangle evaluated on the device from the box state, so up to 4 steps may be enqueued
 * before their results are collected; bytes_copied is then -1 at enqueue and reported by collect.  Pageable frames:
 * one 2-D copy per sequence, rectangle from the host mirror of the state (whole frame when the mirror is stale).
 * LIFETIME: the gather kernel and the speculative uploads read the frames asynchronously, after track_host has
 * returned.  Every frame passed to vipt_engine_track_host must stay valid and unmodified until THAT step has been
 * collected (vipt_engine_collect) or the stream has been synchronised through vipt_engine_results.
 * collect returns the boxes / scores of the OLDEST uncollected track_host step (VIPT_ERR_STATE when there is none).
 * set_state overrides the box state (teacher-forced evaluation). */
int vipt_engine_init_host(void* engine, const void* const* host_frames, const int* hw_host, const double* boxes_host,
                          void* stream);
int vipt_engine_track_host(void* engine, const void* const* host_frames, int use_graph, int64_t* bytes_copied,
                           void* stream);
int vipt_engine_collect(void* engine, double* boxes_host, float* scores_host, int64_t* bytes_copied);
int vipt_engine_set_state(void* engine, const double* boxes_host, void* stream);
/* one un-graphed step with a CUDA-event pair around every kernel launch; sums (ms) and launch counts per
 * kernel class: 0 linear GEMMs, 1 head conv GEMMs, 2 attention, 3 prompt blocks, 4 LayerNorm/CE/final,
 * 5 crop, 6 (unused: the head convolutions are implicit GEMMs), 7 head tail.  Used by bench.py for the roofline line. */
int vipt_engine_profile_step(void* engine, const void* frame_ptrs_dev, float* ms_by_class_host, int* n_by_class_host,
                             void* stream);
int64_t vipt_engine_info(void* engine, const char* what);
int vipt_engine_buffer(void* engine, const char* name, void** dev_ptr, int64_t* bytes);

#ifdef __cplusplus
}
#endif
#endif
