// extern "C" entry points of libvipt_b200.so (declared in include/vipt_b200.h).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "engine.h"

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
  int r = (bn == 512 || bn == 640) ? vipt::gemm2_plan(&op, A, lda, W, ldw, M, N, K, epi, out, ldc, bias, bn == 640 ? -1 : num_sms() / 2)
                    : vipt::gemm_plan(&op, A, lda, 0, W, ldw, 0, 1, M, N, K, bn, epi, out, ldc, 0, 0, bias, resid);
  if (r) { set_err("gemm_plan failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}

int vipt_op_gemm_resid_stat(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x, void* x16,
                            const float* bias, float* stat_out, void* stream) {
  vipt::GemmOp op;
  int r = vipt::gemm2_plan_resid_stat(&op, A, lda, W, ldw, M, N, K, x, x16, bias, (float2*)stat_out, num_sms() / 2);
  if (r) { set_err("gemm2_plan_resid_stat failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}
int vipt_op_gemm_resid_prompt(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, float* x, void* x16,
                              const float* bias, const float* colw20, float* part_out, void* stream) {
  vipt::GemmOp op;
  int r = vipt::gemm2_plan_resid_prompt(&op, A, lda, W, ldw, M, N, K, x, x16, bias, colw20, part_out);
  if (r) { set_err("gemm2_plan_resid_prompt failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}
int vipt_op_gemm_lnfold(const void* A, int64_t lda, const void* W, int64_t ldw, int M, int N, int K, int gelu, void* out,
                        int64_t ldc, const float* bias, const float* colsum, const float* stat_in, int stat_parts, float eps,
                        const void* Ax, const void* Wx, void* stream) {
  if (!colsum || !stat_in || stat_parts < 0 || (Ax == nullptr) != (Wx == nullptr)) return VIPT_ERR_ARG;
  vipt::GemmOp op;
  int r = vipt::gemm2_plan(&op, A, lda, W, ldw, M, N, K, gelu ? vipt::EPI_LNFOLD_GELU_F16 : vipt::EPI_LNFOLD_F16, out, ldc, bias,
                             num_sms() / 2);
  if (r) { set_err("gemm2_plan (LayerNorm fold) failed"); return r; }
  vipt::gemm2_set_lnfold(&op, (const float2*)stat_in, stat_parts, colsum, eps);
  if (Ax && (r = vipt::gemm2_set_kext(&op, Ax, Wx))) { set_err("gemm2_set_kext failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}

int vipt_op_crop(const void* frame_ptrs_dev, const int* hw_dev, const double* boxes_dev, double factor, int S,
                 int row_base, int rows_per_seq, void* apatch, double* geom_dev, int* err_dev, int B, void* stream) {
  return vipt::launch_crop((const uint8_t* const*)frame_ptrs_dev, hw_dev, boxes_dev, factor, S, row_base, rows_per_seq,
                           apatch, (int64_t)B * rows_per_seq * 768, geom_dev, err_dev, B, (cudaStream_t)stream);
}

int vipt_op_attention(const void* qkv, int B, int N, void* out, float* ce_prob, void* stream) {
  vipt::AttnOp op;
  int r = vipt::attention_plan(&op, qkv, B, N);
  if (r) { set_err("attention_plan failed"); return r; }
  return vipt::attention_launch(op, out, ce_prob, vipt::kCeToken, vipt::kLensT, vipt::kLensX, (cudaStream_t)stream);
}

int vipt_op_ce_topk(const float* ce_prob, int n_heads, int n_s, int n_keep, const int* gi_in, int* gi_out, int* perm,
                    int* removed, int* slot_of_pos, float* score_out, int B, void* stream) {
  if (!ce_prob || !gi_in || !gi_out || !perm || !removed || !slot_of_pos || n_keep < 1 || n_keep > n_s) return VIPT_ERR_ARG;
  return vipt::launch_ce_topk(ce_prob, n_heads, n_s, n_keep, gi_in, gi_out, perm, removed, slot_of_pos, score_out,
                              vipt::kLensX, B, (cudaStream_t)stream);
}

int vipt_op_conv3x3(const void* feat, int c_total, int cin, int cin_stride, const void* W, int G, int B, int N, int bn,
                    void* out, int64_t ldc, int64_t out_group_col, const float* bias, void* stream) {
  vipt::GemmOp op;
  int r = bn == 512 ? vipt::gemm2_plan_conv(&op, feat, c_total, cin, W, B, N, vipt::EPI_RELU_F16, out, ldc, bias)
                    : vipt::gemm_plan_conv(&op, feat, c_total, cin, cin_stride, W, G, B, N, bn, vipt::EPI_RELU_F16, out, ldc,
                                           out_group_col, bias);
  if (r) { set_err("conv plan failed"); return r; }
  r = vipt::gemm_launch(op, num_sms(), (cudaStream_t)stream);
  if (r) set_err(cudaGetErrorString(cudaGetLastError()));
  return r;
}

int vipt_op_assemble_rgbx(const void* bgr_dev, const void* xbgr_dev, int H, int W, void* out_dev, void* stream) {
  if (!bgr_dev || !xbgr_dev || !out_dev || H < 1 || W < 1) return VIPT_ERR_ARG;
  return vipt::launch_assemble_rgbx(bgr_dev, xbgr_dev, H, W, out_dev, (cudaStream_t)stream);
}
int vipt_op_assemble_rgbd(const void* bgr_dev, const void* depth16_dev, int H, int W, int depth_clip, void* out_dev,
                          void* workspace_dev, void* stream) {
  if (!bgr_dev || !depth16_dev || !out_dev || !workspace_dev || H < 1 || W < 1) return VIPT_ERR_ARG;
  return vipt::launch_assemble_rgbd(bgr_dev, depth16_dev, H, W, depth_clip, out_dev, workspace_dev, (cudaStream_t)stream);
}

int vipt_op_assemble_rgbd_noalloc(const void* bgr_dev, const void* depth16_dev, int H, int W, int depth_clip,
                                  void* out_dev, void* stream) {
  if (!bgr_dev || !depth16_dev || !out_dev || H < 1 || W < 1) { set_err("vipt_op_assemble_rgbd_noalloc: null pointer or empty image"); return VIPT_ERR_ARG; }
  if ((long long)H * (long long)W > (long long)VIPT_NOALLOC_MAX_PIXELS) {
    set_err("vipt_op_assemble_rgbd_noalloc: H * W exceeds VIPT_NOALLOC_MAX_PIXELS (16 777 216); use vipt_op_assemble_rgbd "
            "with a caller-provided workspace");
    return VIPT_ERR_ARG;
  }
  int r = vipt::launch_assemble_rgbd_noalloc(bgr_dev, depth16_dev, H, W, depth_clip, out_dev, (cudaStream_t)stream);
  if (r) set_err("vipt_op_assemble_rgbd_noalloc: cudaGetSymbolAddress or launch failed");
  return r;
}

// ---------------------------------------------------------------- engine
int vipt_engine_create(void** out) {
  if (!out) return VIPT_ERR_ARG;
  vipt::Engine* e = new vipt::Engine();
  e->opt_zero_copy = getenv("VIPT_NO_ZEROCOPY") == nullptr;
  e->opt_speculate = getenv("VIPT_NO_SPECULATION") == nullptr;
  e->opt_lnfold = getenv("VIPT_NO_LNFOLD") == nullptr;
  e->opt_promptfold = getenv("VIPT_NO_PROMPT_FOLD") == nullptr;
  e->opt_promptfold_ce = getenv("VIPT_PROMPT_FOLD_NO_CE") == nullptr;
  *out = e;
  return VIPT_OK;
}
int vipt_engine_destroy(void* e) {
  delete (vipt::Engine*)e;
  return VIPT_OK;
}
int vipt_engine_set_tensor(void* e, const char* name, const float* host_data, int64_t numel) {
  if (!e || !name || !host_data || numel <= 0) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->set_tensor(name, host_data, numel);
}
int vipt_engine_finalize(void* e, int batch, int deep) {
  if (!e || batch < 1) return VIPT_ERR_ARG;
  int r = ((vipt::Engine*)e)->finalize(batch, deep != 0);
  if (r) set_err("finalize failed (missing / mis-sized state_dict entry or CUDA error; see stderr)");
  return r;
}
int vipt_engine_forward(void* e, const float* template_dev, const float* search_dev, int want_feat32, int debug_stop,
                        void* stream) {
  if (!e || !template_dev || !search_dev) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->forward_images(template_dev, search_dev, want_feat32 != 0, debug_stop, (cudaStream_t)stream);
}
int vipt_engine_init(void* e, const void* const* frame_dev_ptrs_host, const int* hw_host, const double* boxes_host,
                     void* stream) {
  if (!e || !frame_dev_ptrs_host || !hw_host || !boxes_host) return VIPT_ERR_ARG;
  int r = ((vipt::Engine*)e)->init_sequences(frame_dev_ptrs_host, hw_host, boxes_host, (cudaStream_t)stream);
  if (r == VIPT_ERR_SMALL_BOX) set_err("Too small bounding box.");
  return r;
}
int vipt_engine_track_enqueue(void* e, const void* frame_ptrs_dev, int use_graph, void* stream) {
  if (!e || !frame_ptrs_dev) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->track_enqueue(frame_ptrs_dev, use_graph != 0, (cudaStream_t)stream);
}
int vipt_engine_profile_step(void* e, const void* frame_ptrs_dev, float* ms_by_class_host, int* n_by_class_host,
                             void* stream) {
  if (!e || !frame_ptrs_dev || !ms_by_class_host || !n_by_class_host) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->profile_step(frame_ptrs_dev, ms_by_class_host, n_by_class_host, (cudaStream_t)stream);
}
int vipt_engine_results(void* ev, double* boxes_host, float* scores_host, void* stream) {
  vipt::Engine* e = (vipt::Engine*)ev;
  if (!e || !boxes_host || !scores_host) return VIPT_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  int herr = 0;
  if (cudaMemcpyAsync(boxes_host, e->out_boxes, sizeof(double) * 4 * e->B, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaMemcpyAsync(scores_host, e->out_scores, sizeof(float) * e->B, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaMemcpyAsync(&herr, e->err, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaStreamSynchronize(st) != cudaSuccess) {
    set_err(cudaGetErrorString(cudaGetLastError()));
    return VIPT_ERR_LAUNCH;
  }
  if (herr == VIPT_ERR_SMALL_BOX) set_err("Too small bounding box.");
  e->host_state.assign(boxes_host, boxes_host + 4 * e->B);     // mirror for the ROI uploads (engine3.cu)
  e->spec_state = e->host_state;
  e->mirror_valid = true;
  e->res_pending = 0;            // the latest state has been read: earlier uncollected steps are dropped
  return herr;
}
int vipt_engine_slot_errors(void* ev, int* slot_err_host, void* stream) {
  vipt::Engine* e = (vipt::Engine*)ev;
  if (!e || !slot_err_host || !e->finalized) return VIPT_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemcpyAsync(slot_err_host, e->err + 1, sizeof(int) * e->B, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaMemsetAsync(e->err, 0, sizeof(int), st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
    set_err(cudaGetErrorString(cudaGetLastError()));
    return VIPT_ERR_LAUNCH;
  }
  return VIPT_OK;
}
int vipt_engine_init_host(void* e, const void* const* host_frames, const int* hw_host, const double* boxes_host,
                          void* stream) {
  if (!e || !host_frames || !hw_host || !boxes_host) return VIPT_ERR_ARG;
  int r = ((vipt::Engine*)e)->init_host(host_frames, hw_host, boxes_host, (cudaStream_t)stream);
  if (r == VIPT_ERR_SMALL_BOX) set_err("Too small bounding box.");
  return r;
}
int vipt_engine_track_host(void* e, const void* const* host_frames, int use_graph, int64_t* bytes_copied, void* stream) {
  if (!e || !host_frames) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->track_host(host_frames, use_graph != 0, bytes_copied, (cudaStream_t)stream);
}
int vipt_engine_collect(void* e, double* boxes_host, float* scores_host, int64_t* bytes_copied) {
  if (!e || !boxes_host || !scores_host) return VIPT_ERR_ARG;
  int r = ((vipt::Engine*)e)->collect(boxes_host, scores_host, bytes_copied);
  if (r == VIPT_ERR_SMALL_BOX) set_err("Too small bounding box.");
  else if (r == VIPT_ERR_STATE) set_err("no enqueued host-frame step to collect");
  return r;
}
int vipt_engine_set_state(void* e, const double* boxes_host, void* stream) {
  if (!e || !boxes_host) return VIPT_ERR_ARG;
  return ((vipt::Engine*)e)->set_state(boxes_host, (cudaStream_t)stream);
}
int64_t vipt_engine_info(void* ev, const char* what) {
  vipt::Engine* e = (vipt::Engine*)ev;
  if (!e || !what) return -1;
  if (!strcmp(what, "launches_per_step")) return e->launches_per_step;
  if (!strcmp(what, "batch")) return e->B;
  if (!strcmp(what, "promptfold_layers")) { int n = 0; for (int i = 0; i < vipt::kDepth; ++i) n += e->PW[i].fold_on; return n; }
  if (!strcmp(what, "lnfold_layers")) { int n = 0; for (int i = 0; i < vipt::kDepth; ++i) n += e->L[i].fuse_ln2 + e->L[i].fuse_ln1; return n; }
  if (!strcmp(what, "final_tokens")) return e->n_tok[vipt::kDepth];
  if (!strncmp(what, "n_tok_", 6)) { int i = atoi(what + 6); return (i >= 0 && i <= vipt::kDepth) ? e->n_tok[i] : -1; }
  return -1;
}
int vipt_engine_buffer(void* ev, const char* name, void** ptr, int64_t* bytes) {
  vipt::Engine* e = (vipt::Engine*)ev;
  if (!e || !name || !ptr || !bytes || !e->finalized) return VIPT_ERR_ARG;
  const int64_t R = (int64_t)e->B * vipt::kLensFull, RH = (int64_t)e->B * vipt::kLensX;
  struct Ent { const char* n; void* p; int64_t b; } tab[] = {
      {"score_map", e->score_map, RH * 4}, {"size_map", e->size_map, 2 * RH * 4}, {"offset_map", e->offset_map, 2 * RH * 4},
      {"pred_boxes", e->pred_boxes, e->B * 16}, {"removed", e->removed, 3 * RH * 4}, {"ce_score", e->ce_score, 3 * RH * 4},
      {"feat32", e->feat32, R * 768 * 4}, {"feat16", e->feat16, RH * 768 * 2}, {"x", e->x_final ? e->x_final : e->xA, R * 768 * 4},
      {"xA", e->xA, R * 768 * 4}, {"xB", e->xB, R * 768 * 4}, {"xn", e->xn, R * 768 * 2}, {"dbuf", e->dbuf, R * 768 * 2},
      {"XD", e->XD, R * 16 * 4}, {"FV", e->FV, R * 8 * 4}, {"E", e->E, 2 * R * 768 * 4}, {"qkv", e->qkv, R * 2304 * 2}, {"att", e->att, R * 768 * 2},
      {"hbuf", e->hbuf, R * 3072 * 2}, {"gi", e->gi_final ? e->gi_final : e->giA, RH * 4}, {"slot_of_pos", e->slot_of_pos, RH * 4},
      {"h1", e->h1, RH * 768 * 2}, {"h4", e->h4, RH * 96 * 2}, {"state", e->state, e->B * 32}, {"out_boxes", e->out_boxes, e->B * 32},
      {"out_scores", e->out_scores, e->B * 4}, {"geom_x", e->geom_x, e->B * 32}, {"geom_z", e->geom_z, e->B * 32},
      {"apatch", e->apatch, 2 * R * 768 * 2}, {"ce_prob", e->ce_prob, (int64_t)e->B * 12 * 256 * 4}, {"perm", e->perm, RH * 4},
  };
  for (auto& t : tab)
    if (!strcmp(t.n, name)) { *ptr = t.p; *bytes = t.b; return VIPT_OK; }
  set_err("unknown buffer name");
  return VIPT_ERR_ARG;
}

}  // extern "C"
