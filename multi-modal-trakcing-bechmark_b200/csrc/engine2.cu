// Engine part 2: buffers, plans, per-step orchestration.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

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
#define RET(x)          \
  do {                  \
    int r_ = (x);       \
    if (r_) return r_;  \
  } while (0)

void Engine::prof_begin(int cls, cudaStream_t st) {
  if (!prof) return;
  ProfRec r;
  r.cls = cls;
  cudaEventCreate(&r.a);
  cudaEventCreate(&r.b);
  cudaEventRecord(r.a, st);
  prof->push_back(r);
}
void Engine::prof_end(cudaStream_t st) {
  if (!prof) return;
  cudaEventRecord(prof->back().b, st);
}

int Engine::finalize_buffers() {
  const size_t R = (size_t)B * kLensFull;
  const size_t RH = (size_t)B * kLensX;
#define AL(ptr, type, count)                          \
  do {                                                \
    ptr = (type*)dalloc(sizeof(type) * (count));      \
    if (!ptr) return VIPT_ERR_LAUNCH;                 \
  } while (0)
  AL(apatch, __half, 2 * R * 768);
  AL(E, float, 2 * R * 768);
  AL(xA, float, R * 768); AL(xB, float, R * 768);
  AL(XD, float, R * 16); AL(FV, float, R * 8); AL(lnstat, float2, 24 * R); AL(part, float, 6 * R * 20); AL(F16, __half, R * 64); AL(rowstat, float2, R);
  AL(xn, __half, R * 768); AL(dbuf, __half, R * 768); AL(qkv, __half, R * 2304); AL(att, __half, R * 768); AL(hbuf, __half, R * 3072);
  AL(ce_prob, float, (size_t)B * kHeads * kLensX); AL(ce_score, float, 3 * RH);
  AL(feat16, __half, RH * 768); AL(feat32, float, R * 768);
  AL(h1, __half, RH * 768); AL(h2, __half, RH * 384); AL(h3, __half, RH * 192); AL(h4, __half, RH * 96);
  AL(giA, int, RH); AL(giB, int, RH); AL(perm, int, RH); AL(removed, int, 3 * RH); AL(slot_of_pos, int, RH);
  AL(ident, int, RH); AL(err, int, 1 + B);      // [0] engine-wide flag, [1 + b] sticky per-sequence flags
  AL(score_map, float, RH); AL(size_map, float, 2 * RH); AL(offset_map, float, 2 * RH);
  AL(pred_boxes, float, B * 4); AL(out_scores, float, B);
  AL(state, double, B * 4); AL(geom_z, double, B * 4); AL(geom_x, double, B * 4); AL(out_boxes, double, B * 4);
  AL(frame_hw, int, B * 2);
  frame_ptrs = (const uint8_t**)dalloc(sizeof(void*) * B);
  if (!frame_ptrs) return VIPT_ERR_LAUNCH;
#undef AL
  {
    std::vector<int> id(RH);
    for (size_t i = 0; i < RH; ++i) id[i] = (int)(i % kLensX);
    CK(cudaMemcpy(ident, id.data(), RH * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(err, 0, 4 * (1 + B)));
    RET(crop_prepare());
    CK(cudaMemset(removed, 0, 3 * RH * 4));
    CK(cudaMemset(F16, 0, R * 64 * 2));      // columns 8..63 of the extra k-block stay zero
  }
  // ---- GEMM / attention plans
  const int M0 = B * kLensFull;
  // patch embed of both modalities (fp32 out): two launches of the 2-CTA kernel at large batch, else one grouped launch
  pe_pair = ((M0 + 255) / 256) * 3 >= num_sms / 2;
  if (pe_pair) {
    for (int g = 0; g < 2; ++g)
      RET(gemm2_plan(&g_pe2[g], apatch + (size_t)g * M0 * 768, 768, w_pe + (size_t)g * 768 * 768, 768, M0, 768, 768, EPI_BIAS_F32,
                     E + (size_t)g * M0 * 768, 768, b_pe + g * 768, num_sms / 2));
  } else {
    RET(gemm_plan(&g_pe, apatch, 768, (int64_t)M0 * 768, w_pe, 768, 768 * 768, 2, M0, 768, 768, 256, EPI_BIAS_F32, E, 768,
                  (int64_t)M0 * 768, 0, b_pe, nullptr));
  }
  // Large batches: 2-CTA 256 x 256 tiles.  Small batches (latency mode): the pair kernel would fill only a few
  // SMs, so the single-CTA kernel with narrow N tiles spreads the weight stream over the whole chip.
  auto plan_lin = [&](GemmOp* op, const void* A, int64_t lda, const void* W, int M, int N, int K, int epi, void* out,
                      const float* bias) -> int {
    const int pair_tiles = ((M + 255) / 256) * (N / 256);
    if (pair_tiles >= num_sms / 2) return gemm2_plan(op, A, lda, W, K, M, N, K, epi, out, N, bias, num_sms / 2);
    const int mt = (M + 127) / 128;
    int bn = 128;
    {
      // These GEMMs are bound by the operand stream into each SM ((128 + bn) x 64 fp16 per k-block), not by the MMAs:
      // pick the tile width that minimises  rounds x (128 + bn), ties to the wider tile (less total L2 traffic).
      // E.g. one sequence (3 row tiles): qkv 108 x bn64 tiles in one round instead of 216 x bn32 in two.  Measured
      // against the previous rule (narrowest tile that still gives >= #SMs tiles): step 0.848 -> 0.828 ms at one
      // sequence, 1.058 -> 0.979 at 4, 1.297 -> 1.181 at 8, 1.628 -> 1.574 at 16.
      int best = 1 << 30;
      for (int c = 128; c >= 32; c >>= 1) {
        const int tiles = mt * (N / c), rounds = (tiles + num_sms - 1) / num_sms, cost = rounds * (128 + c);
        if (cost < best) { best = cost; bn = c; }
      }
    }
    return gemm_plan(op, A, lda, 0, W, K, 0, 1, M, N, K, bn, epi, out, N, 0, 0, bias, nullptr);
  };
  float* x_cur = xA;            // the residual stream ping-pongs between xA and xB at every candidate elimination
  float* x_oth = xB;
  for (int i = 0; i < kDepth; ++i) {
    const int Ma = B * n_tok[i], Mm = B * n_mlp[i];
    // LayerNorm fold (gemm2.cu): the producer GEMM adds its branch to x in place and leaves x16 (in xn) + row statistics,
    // the consumer GEMM applies mean / rstd in its epilogue
    if (PW[i].fold_on) {
      // prompt fold (prompt_mid.cu): x16 = fp16(v), the rank-8 prompt update enters in the epilogue
      RET(gemm2_plan(&g_qkv[i], xn, 768, L[i].w_qkv, 768, Ma, 2304, 768, EPI_LNFOLD_F16, qkv, 2304, PW[i].b_qkv_r8, num_sms / 2));
      gemm2_set_lnfold(&g_qkv[i], rowstat, 0, PW[i].cs_qkv_r8, 1e-6f);      // final statistics from prompt_mid_kernel
      RET(gemm2_set_kext(&g_qkv[i], F16, PW[i].wx_qkv));
    } else if (L[i].fuse_ln1) {
      RET(gemm2_plan(&g_qkv[i], xn, 768, L[i].w_qkv, 768, Ma, 2304, 768, EPI_LNFOLD_F16, qkv, 2304, L[i].b_qkv, num_sms / 2));
      gemm2_set_lnfold(&g_qkv[i], lnstat, 24, L[i].cs_qkv, 1e-6f);
    } else {
      RET(plan_lin(&g_qkv[i], xn, 768, L[i].w_qkv, Ma, 2304, 768, EPI_BIAS_F16, qkv, L[i].b_qkv));
    }
    RET(attention_plan(&a_op[i], qkv, B, n_tok[i]));
    // otherwise proj / fc2 write their branch output as fp16 (dbuf); the residual add is fused into the next
    // LayerNorm / prompt kernel, where it is coalesced
    if (L[i].fuse_ln2 || PW[i].fold_on) {
      RET(gemm2_plan_resid_stat(&g_proj[i], att, 768, L[i].w_proj, 768, Ma, 768, 768, x_cur, xn,
                                PW[i].fold_on ? PW[i].b_proj_r8 : L[i].b_proj, lnstat, num_sms / 2));
      if (PW[i].fold_on) RET(gemm2_set_kext(&g_proj[i], F16, PW[i].wx_proj));    // x <- v + U f + b1 + proj(..)
    } else {
      RET(plan_lin(&g_proj[i], att, 768, L[i].w_proj, Ma, 768, 768, EPI_BIAS_F16, dbuf, L[i].b_proj));
    }
    if (L[i].fuse_ln2) {
      RET(gemm2_plan(&g_fc1[i], xn, 768, L[i].w_fc1, 768, Mm, 3072, 768, EPI_LNFOLD_GELU_F16, hbuf, 3072, L[i].b_fc1, num_sms / 2));
      gemm2_set_lnfold(&g_fc1[i], lnstat, 24, L[i].cs_fc1, 1e-6f);
    } else {
      RET(plan_lin(&g_fc1[i], xn, 768, L[i].w_fc1, Mm, 3072, 768, EPI_GELU_F16, hbuf, L[i].b_fc1));
    }
    if (ce_keep[i]) { float* t = x_cur; x_cur = x_oth; x_oth = t; }
    if (i + 1 < kDepth && PW[i + 1].fold_on) {
      RET(gemm2_plan_resid_prompt(&g_fc2[i], hbuf, 3072, L[i].w_fc2, 3072, Mm, 768, 3072, x_cur, xn, L[i].b_fc2, PW[i + 1].colw20, part));
    } else if (i + 1 < kDepth && L[i + 1].fuse_ln1) {
      RET(gemm2_plan_resid_stat(&g_fc2[i], hbuf, 3072, L[i].w_fc2, 3072, Mm, 768, 3072, x_cur, xn, L[i].b_fc2, lnstat, num_sms / 2));
    } else {
      RET(plan_lin(&g_fc2[i], hbuf, 3072, L[i].w_fc2, Mm, 768, 3072, EPI_BIAS_F16, dbuf, L[i].b_fc2));
    }
  }
  const int MH = B * kLensX;
  // center head: implicit 3x3 convolutions (no im2col buffer); BN folded, ReLU fused, towers grouped
  if (B * 3 >= num_sms / 2) RET(gemm2_plan_conv(&g_c1, feat16, 768, 768, w_c1, B, 768, EPI_RELU_F16, h1, 768, b_c1));
  // latency mode: same operand-stream rule as the linear layers (the towers have only 2 row tiles per sequence and
  // K = 9 x C_in, so narrow tiles put them on more SMs and shorten each CTA's stream).  A split of K with a fix-up
  // through L2 (partial tiles, counter, last CTA reduces) was measured and lost: 0.828 -> 1.017 ms at one sequence.
  auto pick_bn = [&](int row_tiles, int N, int max_bn) {
    int bn = max_bn, best = 1 << 30;
    for (int c = max_bn; c >= 32; c >>= 1) {
      const int tiles = row_tiles * (N / c), rounds = (tiles + num_sms - 1) / num_sms, cost = rounds * (128 + c);
      if (cost < best) { best = cost; bn = c; }
    }
    return bn;
  };
  const bool lat = B * 3 < num_sms / 2;
  if (lat) RET(gemm_plan_conv(&g_c1, feat16, 768, 768, 0, w_c1, 1, B, 768, pick_bn(2 * B, 768, 64), EPI_RELU_F16, h1, 768, 0, b_c1));
  RET(gemm_plan_conv(&g_c2, h1, 768, 256, 256, w_c2, 3, B, 128, lat ? pick_bn(6 * B, 128, 128) : 128, EPI_RELU_F16, h2, 384, 128, b_c2));
  RET(gemm_plan_conv(&g_c3, h2, 384, 128, 128, w_c3, 3, B, 64, lat ? pick_bn(6 * B, 64, 64) : 64, EPI_RELU_F16, h3, 192, 64, b_c3));
  RET(gemm_plan_conv(&g_c4, h3, 192, 64, 64, w_c4, 3, B, 32, 32, EPI_RELU_F16, h4, 96, 32, b_c4));
  finalized = true;
  return VIPT_OK;
}

// Runs patch-embed .. head on whatever is in `apatch`.  tracking=true also updates the box state.
// debug_stop > 0 returns after that stage (tests pull intermediate buffers).
int Engine::run_backbone_head(bool tracking, bool want_feat32, int debug_stop, cudaStream_t st) {
  long nl = 0;
  int stage = 0;
  pdl_select(B, num_sms);
#define LAUNCH(cls, expr)   \
  do {                      \
    prof_begin(cls, st);    \
    int r_ = (expr);        \
    if (r_) return r_;      \
    prof_end(st);           \
    ++nl;                   \
  } while (0)
#define STAGE_END()                             \
  do {                                          \
    if (debug_stop > 0 && ++stage == debug_stop) return VIPT_OK; \
  } while (0)
  const size_t RH = (size_t)B * kLensX;
  // No per-step initialisation of the CE bookkeeping (two copy nodes per step until round 2): before the first
  // candidate elimination the slot -> raster map IS the constant identity array, and slot_of_pos is written for all
  // 256 positions by the first ce_topk launch before anything reads it.
  if (pe_pair) {
    LAUNCH(K_GEMM, gemm_launch(g_pe2[0], num_sms, st));
    LAUNCH(K_GEMM, gemm_launch(g_pe2[1], num_sms, st));
  } else {
    LAUNCH(K_GEMM, gemm_launch(g_pe, num_sms, st));
  }
  STAGE_END();   // 1: E
  const float* E0 = E;
  const float* E1 = E + (size_t)B * kLensFull * 768;
  LAUNCH(K_PROMPT, launch_prompt_a(E0, nullptr, kLensFull, nullptr, E1, PW[0].wg_a, PW[0].wg_d, PW[0].sc, XD, B, kLensT, kLensFull, st));
  LAUNCH(K_PROMPT, launch_fovea(XD, PW[0].smooth, FV, B, kLensT, kLensFull, st));
  LAUNCH(K_PROMPT, launch_prompt_b(FV, PW[0].w1, PW[0].b1, nullptr, E0, nullptr, xA, kLensFull, nullptr, pos_embed, L[0].ln1_g,
                      L[0].ln1_b, xn, B, kLensT, kLensFull, st));
  STAGE_END();   // 2: x, xn, P after the input prompt
  float* x = xA;
  float* x_other = xB;
  int* gi = ident;
  int* gi_other = giA;
  int ce_idx = 0;
  for (int i = 0; i < kDepth; ++i) {
    const int n = n_tok[i];
    if (i >= 1) {
      if (deep && PW[i].fold_on) {
        // prompt fold: the 768-wide work is in the epilogues of fc2[i-1], qkv[i], proj[i]; this is the per-sequence part
        LAUNCH(K_PROMPT, launch_prompt_mid(part, n, gi, FV, PW[i].fold, PW[i].fold2, F16, rowstat, B, kLensT, kLensFull, st));
      } else if (deep) {
        // cooperative persistent kernel (prompt_coop.cu)
        LAUNCH(K_PROMPT, launch_prompt_coop(x, dbuf, n, gi, FV, XD, PW[i].wg_a, PW[i].fold, PW[i].w1t, PW[i].b1,
                                            L[i].ln1_g, L[i].ln1_b, xn, B, kLensT, kLensFull, num_sms, st));
      } else if (!L[i].fuse_ln1) {
        LAUNCH(K_NORM, launch_ln_rows(x, n, dbuf, nullptr, 0, x, xn, L[i].ln1_g, L[i].ln1_b, B, n, kLensT, st));
      }
    }
    STAGE_END();   // 3 + 7i: x entering block i, xn = LN1(x)
    LAUNCH(K_GEMM, gemm_launch(g_qkv[i], num_sms, st));
    STAGE_END();   // 4 + 7i: qkv
    LAUNCH(K_ATTN, attention_launch(a_op[i], att, ce_keep[i] ? ce_prob : nullptr, kCeToken, kLensT, kLensX, st));
    STAGE_END();   // 5 + 7i: att, ce_prob
    LAUNCH(K_GEMM, gemm_launch(g_proj[i], num_sms, st));
    STAGE_END();   // 6 + 7i: dbuf = attention branch output
    const int nm = n_mlp[i];
    if (ce_keep[i]) {
      LAUNCH(K_NORM, launch_ce_topk(ce_prob, kHeads, n - kLensT, ce_keep[i], gi, gi_other, perm, removed + (size_t)ce_idx * RH,
                         slot_of_pos, ce_score + (size_t)ce_idx * RH, kLensX, B, st));
      // (prompt fold: attn.proj has already added its branch and the prompt update to x)
      LAUNCH(K_NORM, launch_ln_rows(x, n, PW[i].fold_on ? nullptr : dbuf, perm, kLensX, x_other, xn, L[i].ln2_g, L[i].ln2_b, B, nm, kLensT, st));
      float* t = x; x = x_other; x_other = t;
      int* tg = gi; gi = gi_other; gi_other = tg == ident ? giB : tg;      // the identity array is never a destination
      ++ce_idx;
    } else if (!L[i].fuse_ln2) {
      LAUNCH(K_NORM, launch_ln_rows(x, n, dbuf, nullptr, 0, x, xn, L[i].ln2_g, L[i].ln2_b, B, nm, kLensT, st));
    }
    STAGE_END();   // 7 + 7i: x after CE, xn = LN2(x)
    LAUNCH(K_GEMM, gemm_launch(g_fc1[i], num_sms, st));
    STAGE_END();   // 8 + 7i: hbuf
    LAUNCH(K_GEMM, gemm_launch(g_fc2[i], num_sms, st));
    STAGE_END();   // 9 + 7i: dbuf = MLP branch output
  }
  x_final = x;
  gi_final = gi;
  LAUNCH(K_NORM, launch_final_norm(x, dbuf, n_tok[kDepth], slot_of_pos, norm_g, norm_b, feat16, want_feat32 ? feat32 : nullptr, B, kLensT,
                        kLensFull, st));
  STAGE_END();   // 87: feat16
  LAUNCH(K_GEMM_HEAD, gemm_launch(g_c1, num_sms, st));
  STAGE_END();   // 88: h1
  LAUNCH(K_GEMM_HEAD, gemm_launch(g_c2, num_sms, st));
  LAUNCH(K_GEMM_HEAD, gemm_launch(g_c3, num_sms, st));
  LAUNCH(K_GEMM_HEAD, gemm_launch(g_c4, num_sms, st));
  STAGE_END();   // 89: h4
  LAUNCH(K_HEAD_FINAL, launch_head_final(h4, w5, b5, window, score_map, size_map, offset_map, pred_boxes, tracking ? state : nullptr,
                        geom_x, frame_hw, out_boxes, out_scores, B, st));
  launches_per_step = nl;
#undef STAGE_END
#undef LAUNCH
  return VIPT_OK;
}

int Engine::forward_images(const float* template_dev, const float* search_dev, bool want_feat32, int debug_stop,
                           cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  const int64_t gs = (int64_t)B * kLensFull * 768;
  RET(launch_patchify(template_dev, 128, 0, kLensFull, apatch, gs, B, st));
  RET(launch_patchify(search_dev, 256, kLensT, kLensFull, apatch, gs, B, st));
  return run_backbone_head(false, want_feat32, debug_stop, st);
}

int Engine::init_sequences(const void* const* frames_host_ptrs, const int* hw_host, const double* boxes_host,
                           cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  CK(cudaMemcpyAsync(frame_ptrs, frames_host_ptrs, sizeof(void*) * B, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(frame_hw, hw_host, sizeof(int) * 2 * B, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(state, boxes_host, sizeof(double) * 4 * B, cudaMemcpyHostToDevice, st));
  host_state.assign(boxes_host, boxes_host + 4 * B);
  mirror_valid = true;
  CK(cudaMemsetAsync(err, 0, 4 * (1 + B), st));
  const int64_t gs = (int64_t)B * kLensFull * 768;
  // template: factor 2.0 -> 128 x 128 (vipt.py:43-48); the crop is kept as the resident patch rows 0..63
  RET(launch_crop(frame_ptrs, frame_hw, state, 2.0, 128, 0, kLensFull, apatch, gs, geom_z, err, B, st, err + 1));
  int herr = 0;
  CK(cudaMemcpyAsync(&herr, err, 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  return herr;
}

int Engine::track_enqueue(const void* frame_ptrs_dev, bool use_graph, cudaStream_t st) {
  if (!finalized) return VIPT_ERR_STATE;
  CK(cudaMemcpyAsync(frame_ptrs, frame_ptrs_dev, sizeof(void*) * B, cudaMemcpyDeviceToDevice, st));
  if (use_graph && graph_exec) {
    CK(cudaGraphLaunch(graph_exec, st));
    return VIPT_OK;
  }
  const int64_t gs = (int64_t)B * kLensFull * 768;
  auto body = [&](cudaStream_t s) -> int {
    // search: factor 4.0 -> 256 x 256 around the current state (vipt.py:67-69)
    pdl_select(B, num_sms);
    prof_begin(K_CROP, s);
    RET(launch_crop(frame_ptrs, frame_hw, state, 4.0, 256, kLensT, kLensFull, apatch, gs, geom_x, err, B, s, err + 1));
    prof_end(s);
    return run_backbone_head(true, false, 0, s);
  };
  if (!use_graph) {
    int r = body(st);
    launches_per_step += 1;
    return r;
  }
  // first graph request: capture one step, instantiate, launch
  cudaGraph_t graph = nullptr;
  CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed));
  int r = body(st);
  cudaError_t ce = cudaStreamEndCapture(st, &graph);
  if (r) { if (graph) cudaGraphDestroy(graph); return r; }
  CK(ce);
  launches_per_step += 1;
  CK(cudaGraphInstantiate(&graph_exec, graph, 0));
  cudaGraphDestroy(graph);
  CK(cudaGraphLaunch(graph_exec, st));
  return VIPT_OK;
}

// One un-graphed step with a CUDA-event pair around every launch; sums per kernel class.
int Engine::profile_step(const void* frame_ptrs_dev, float* ms_by_class, int* n_by_class, cudaStream_t st) {
  std::vector<ProfRec> recs;
  prof = &recs;
  int r = track_enqueue(frame_ptrs_dev, false, st);
  prof = nullptr;
  cudaError_t e = cudaStreamSynchronize(st);
  for (int i = 0; i < K_NUM_CLASSES; ++i) { ms_by_class[i] = 0.f; n_by_class[i] = 0; }
  for (auto& p : recs) {
    float ms = 0.f;
    if (e == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
      ms_by_class[p.cls] += ms;
      n_by_class[p.cls] += 1;
    }
    cudaEventDestroy(p.a);
    cudaEventDestroy(p.b);
  }
  if (e != cudaSuccess) return VIPT_ERR_LAUNCH;
  return r;
}

}  // namespace vipt
