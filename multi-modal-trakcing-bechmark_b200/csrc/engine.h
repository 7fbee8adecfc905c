// Device-resident batched ViPT tracker engine: packed weights, activation buffers sized for a fixed
// number of concurrent sequences, pre-encoded TMA descriptors, per-sequence state (boxes in fp64).
#pragma once
#include <cuda_fp16.h>

#include <map>
#include <string>
#include <vector>

#include "kernels.h"

namespace vipt {

constexpr int kDepth = 12;
constexpr int kLensT = 64;      // template tokens (128/16)^2
constexpr int kLensX = 256;     // search tokens (256/16)^2
constexpr int kLensFull = 320;
constexpr int kEmbed = 768;
constexpr int kMlp = 3072;
constexpr int kHeads = 12;
constexpr int kCeToken = 27;    // CTR_POINT on the 8x8 template grid (ce_utils.py:22-24)

struct LayerW {
  __half *w_qkv, *w_proj, *w_fc1, *w_fc2;
  float *b_qkv, *b_proj, *b_fc1, *b_fc2;
  float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
};
struct PromptW {
  float *wg_a, *wg_d, *sc, *w1, *b1;
  float *fold, *w1t;     // layers >= 1: folded rank-8 algebra + k-major conv1x1 of the fused kernel
  float smooth;
};

enum KernelClass { K_GEMM = 0, K_GEMM_HEAD, K_ATTN, K_PROMPT, K_NORM, K_CROP, K_IM2COL, K_HEAD_FINAL, K_NUM_CLASSES };
struct ProfRec { int cls; cudaEvent_t a, b; };

struct Engine {
  int B = 0;               // concurrent sequences (fixed)
  bool deep = true;
  bool finalized = false;
  int num_sms = 148;
  std::map<std::string, std::vector<float>> raw;   // state_dict as uploaded, fp32
  std::vector<void*> allocs;

  // packed weights
  __half* w_pe = nullptr; float* b_pe = nullptr;   // (2,768,768), (2,768)
  float* pos_embed = nullptr;                       // (320,768) z then x
  LayerW L[kDepth];
  PromptW PW[kDepth];
  float *norm_g = nullptr, *norm_b = nullptr;
  __half *w_c1, *w_c2, *w_c3, *w_c4;               // head convs (BN folded), K order = tap*cin + c
  float *b_c1, *b_c2, *b_c3, *b_c4, *w5, *b5, *window;

  // CE schedule
  int n_tok[kDepth + 1];     // tokens entering layer i (attention N); n_tok[12] = final
  int ce_keep[kDepth];       // 0 if not a CE layer, else number of kept search tokens

  // activations
  __half *apatch, *xn, *dbuf, *qkv, *att, *hbuf, *feat16, *h1, *h2, *h3, *h4;
  float *E, *xA, *xB, *XD, *FV, *ce_prob, *ce_score, *feat32;
  int *giA, *giB, *perm, *removed, *slot_of_pos, *ident, *err;
  // outputs / state
  float *score_map, *size_map, *offset_map, *pred_boxes, *out_scores;
  double *state, *geom_z, *geom_x, *out_boxes;
  int* frame_hw;                      // (B,2) H,W
  const uint8_t** frame_ptrs;         // (B) device array of frame pointers

  // plans (TMA descriptors encoded once)
  GemmOp g_pe, g_qkv[kDepth], g_proj[kDepth], g_fc1[kDepth], g_fc2[kDepth];
  GemmOp g_c1, g_c2, g_c3, g_c4;
  AttnOp a_op[kDepth];
  int n_mlp[kDepth];         // tokens in layer i's MLP (after CE)

  float* x_final = nullptr;   // residual stream / index list after the last block (debug, forward outputs)
  int* gi_final = nullptr;

  // host-frame path (engine3.cu): engine-owned frame buffers + host mirror of the box state
  std::vector<void*> own_frames;
  const uint8_t** own_ptrs_dev = nullptr;
  std::vector<int> host_hw;
  std::vector<double> host_state;
  bool mirror_valid = false;
  int init_host(const void* const* host_frames, const int* hw_host, const double* boxes_host, cudaStream_t st);
  int track_host(const void* const* host_frames, bool use_graph, int64_t* bytes_copied, cudaStream_t st);
  int set_state(const double* boxes_host, cudaStream_t st);

  // CUDA graph of one tracking step
  cudaGraphExec_t graph_exec = nullptr;
  long launches_per_step = 0;

  std::vector<ProfRec>* prof = nullptr;
  void prof_begin(int cls, cudaStream_t st);
  void prof_end(cudaStream_t st);
  int profile_step(const void* frame_ptrs_dev, float* ms_by_class, int* n_by_class, cudaStream_t st);

  ~Engine();
  void* dalloc(size_t bytes);
  int set_tensor(const char* name, const float* data, int64_t numel);
  int finalize(int batch, bool deep_prompt);
  int finalize_buffers();
  int forward_images(const float* template_dev, const float* search_dev, bool want_feat32, int debug_stop,
                     cudaStream_t st);
  int run_backbone_head(bool tracking, bool want_feat32, int debug_stop, cudaStream_t st);
  int init_sequences(const void* const* frames_host_ptrs, const int* hw_host, const double* boxes_host, cudaStream_t st);
  int track_enqueue(const void* frame_ptrs_dev, bool use_graph, cudaStream_t st);
};

}  // namespace vipt
