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
  // LayerNorm fold (gemm2.cu): when fuse_ln2 / fuse_ln1 is set, w_fc1 / w_qkv hold fp16(gamma * W), b_* hold
  // b + W beta and cs_* the column sums of the fp16 weight rows
  float *cs_fc1 = nullptr, *cs_qkv = nullptr;
  bool fuse_ln2 = false;     // norm2 folded into attn.proj (residual + statistics) and mlp.fc1
  bool fuse_ln1 = false;     // shallow prompt only: norm1 folded into the previous block's mlp.fc2 and attn.qkv
};
struct PromptW {
  float *wg_a, *wg_d, *sc, *w1, *b1;
  float *fold, *w1t;     // layers >= 1: folded rank-8 algebra + k-major conv1x1 of the fused kernel
  float smooth;
  // prompt fold (gemm2.cu / prompt_mid.cu), layers >= 1 at large batch: per-column constants of the fc2 and qkv epilogues,
  // the sums / Gram matrix the statistics of x = v + U f + b1 need, and attn.proj's bias + conv1x1.bias
  bool fold_on = false;
  float *colw20 = nullptr, *fold2 = nullptr, *b_proj_r8 = nullptr, *b_qkv_r8 = nullptr, *cs_qkv_r8 = nullptr;
  __half *wx_qkv = nullptr, *wx_proj = nullptr;      // extra k-block weights: W'U (2304, 64) and U (768, 64), 8 columns used
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
  float* part = nullptr;                     // prompt fold: (6, B * 320, 20) row sums
  __half* F16 = nullptr;                     // prompt fold: Fovea output per token row, fp16 (B * 320, 64), 8 columns used
  float2* rowstat = nullptr;                 // prompt fold: (B * 320) (rstd, -mean * rstd) of norm1's input
  float2* lnstat = nullptr;           // LayerNorm fold: (24, B * 320) partial (sum, sum of squares) per row
  int *giA, *giB, *perm, *removed, *slot_of_pos, *ident, *err;
  // outputs / state
  float *score_map, *size_map, *offset_map, *pred_boxes, *out_scores;
  double *state, *geom_z, *geom_x, *out_boxes;
  int* frame_hw;                      // (B,2) H,W
  const uint8_t** frame_ptrs;         // (B) device array of frame pointers

  // plans (TMA descriptors encoded once)
  GemmOp g_pe, g_pe2[2], g_qkv[kDepth], g_proj[kDepth], g_fc1[kDepth], g_fc2[kDepth];
  GemmOp g_c1, g_c2, g_c3, g_c4;
  AttnOp a_op[kDepth];
  bool pe_pair = false;      // patch embed on the 2-CTA kernel (one launch per modality)
  int n_mlp[kDepth];         // tokens in layer i's MLP (after CE)

  float* x_final = nullptr;   // residual stream / index list after the last block (debug, forward outputs)
  int* gi_final = nullptr;

  // host-frame path (engine3.cu): engine-owned frame buffers + host mirror of the box state
  std::vector<void*> own_frames;           // [2][B]: two sets, a step uses set (step & 1)
  const uint8_t** own_ptrs_dev = nullptr;  // the set the NEXT enqueued step reads (written in stream order)
  int ptrs_set = 0;                        // set own_ptrs_dev holds after the work enqueued so far
  long host_step = 0;
  cudaStream_t side_stream = nullptr;      // speculative gathers run here, next to the previous step
  cudaEvent_t set_free_ev[2] = {};         // recorded after the last step that read frame set i
  cudaEvent_t spec_ev = nullptr;
  std::vector<int> host_hw;
  std::vector<double> host_state;
  bool mirror_valid = false;
  // host-frame path switches, read ONCE from the environment when the engine is created:
  // VIPT_NO_ZEROCOPY=1 uploads page-locked frames with per-sequence 2-D copies instead of the gather kernel,
  // VIPT_NO_SPECULATION=1 disables the copy-engine upload next to the running step.  Both covered by tests.
  bool opt_zero_copy = true, opt_speculate = true;
  // VIPT_NO_LNFOLD=1 keeps every LayerNorm a separate pass (ln_rows_kernel); read once at engine creation, covered by a test
  bool opt_lnfold = true;
  // VIPT_NO_PROMPT_FOLD=1 keeps the deep-prompt stage in prompt_coop_kernel at every layer; same rules
  bool opt_promptfold = true;
  bool opt_promptfold_ce = true;           // VIPT_PROMPT_FOLD_NO_CE=1: blocks 3 / 6 / 9 keep prompt_coop_kernel (measurement switch)
  void plan_schedule();                    // CE schedule + which LayerNorms are folded (needs B, deep, num_sms)
  // zero-copy gather: the device-visible addresses of the B host frames travel as a kernel parameter
  static constexpr int kRoiSlots = 4;      // steps that may be enqueued before their results are collected
  static constexpr int kRoiMaxB = 128;
  struct RoiSrc {
    const uint8_t* p[kRoiMaxB];
  };
  struct RoiHave {                         // rectangle already present in the frame set: x_lo, x_hi, y_lo, y_hi
    int r[kRoiMaxB][4];
  };
  std::vector<double> spec_state;          // boxes of the last collected step (speculative uploads)
  uint8_t** own_tab_dev = nullptr;         // [2 sets][B] frame buffers, device copy of own_frames
  int* roi_bytes_dev = nullptr;            // [2][B] bytes gathered per sequence in the last step: speculative, exact
  // per-step results of the host-frame path: pinned ring, one slot per enqueued step (vipt_engine_collect)
  uint8_t* res_host = nullptr;             // [kRoiSlots][boxes B*32 | scores B*4 | roi bytes 2*B*4 | err 4 | pad]
  cudaEvent_t res_ev[kRoiSlots] = {};
  long long res_host_bytes[kRoiSlots] = {};   // >= 0: bytes copied, known on the host at enqueue (pageable path)
  int res_head = 0, res_pending = 0;
  size_t res_slot_bytes() const { return (size_t)B * 44 + 16; }
  int collect(double* boxes_host, float* scores_host, int64_t* bytes_copied);
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

void pdl_select(int batch, int num_sms);     // engine.cu: programmatic dependent launch on/off for the next launches

}  // namespace vipt
