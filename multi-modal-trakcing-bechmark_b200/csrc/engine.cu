// Engine: weight packing (state_dict -> fp16 GEMM operands, folded BN / LN), buffers, plans and the
// per-step orchestration of the kernels.  Mirrors the data flow of
// VisionTransformerCE.forward_features (ViPT/lib/models/vipt/vit_ce_prompt.py:184-346) and
// ViPTrack.forward_head (ViPT/lib/models/vipt/ostrack_prompt.py:60-91) without its module tree.
#include "engine.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_fp16.h>

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

// Programmatic dependent launch (sm100_ptx.cuh) is switched per engine before it enqueues or captures a step.
// Latency mode (few sequences: every kernel is short, launch gaps and prologues are a visible part of a step):
// on, batch-1 p50 0.920 -> 0.857 ms.  Throughput mode (64 sequences): the step is power-capped, PDL changes
// nothing on the device (4.151 ms with and without) and the resident-but-waiting CTAs of the next kernels delay
// the concurrent host-frame gather kernel of the pipelined API (e2e 4.25 -> 4.71 ms), so it is off.
// VIPT_PDL=0|1 forces it.
static thread_local bool g_pdl_on = false;
bool pdl_enabled() { return g_pdl_on; }
void pdl_select(int batch, int num_sms) {
  static const char* force = getenv("VIPT_PDL");
  g_pdl_on = force ? force[0] == '1' : batch * 12 < num_sms;
}

Engine::~Engine() {
  if (graph_exec) cudaGraphExecDestroy(graph_exec);
  for (void* p : allocs) cudaFree(p);
  for (void* p : own_frames) cudaFree(p);
  if (res_host) cudaFreeHost(res_host);
  if (side_stream) cudaStreamDestroy(side_stream);
  if (spec_ev) cudaEventDestroy(spec_ev);
  for (cudaEvent_t e : set_free_ev)
    if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : res_ev)
    if (e) cudaEventDestroy(e);
}

void* Engine::dalloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) return nullptr;
  allocs.push_back(p);
  return p;
}

int Engine::set_tensor(const char* name, const float* data, int64_t numel) {
  raw[name].assign(data, data + numel);
  return VIPT_OK;
}

namespace {

struct Packer {
  Engine* e;
  bool ok = true;
  const std::vector<float>* get(const std::string& k, size_t numel) {
    auto it = e->raw.find(k);
    if (it == e->raw.end() || it->second.size() != numel) {
      fprintf(stderr, "vipt: missing or mis-sized tensor %s (want %zu)\n", k.c_str(), numel);
      ok = false;
      static std::vector<float> dummy;
      dummy.assign(numel, 0.f);
      return &dummy;
    }
    return &it->second;
  }
  float* up_f32(const float* h, size_t n) {
    float* d = (float*)e->dalloc(n * 4);
    if (!d || cudaMemcpy(d, h, n * 4, cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
    return d;
  }
  __half* up_f16(const float* h, size_t n) {
    std::vector<__half> t(n);
    for (size_t i = 0; i < n; ++i) t[i] = __float2half_rn(h[i]);
    __half* d = (__half*)e->dalloc(n * 2);
    if (!d || cudaMemcpy(d, t.data(), n * 2, cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
    return d;
  }
};

// conv (cout,cin,3,3)+bias followed by eval BatchNorm -> (cout, 9*cin) with K = tap*cin + c, bias'
void fold_conv_bn(Packer& pk, const std::string& p, int cout, int cin, float* w_out, float* b_out) {
  const auto& w = *pk.get(p + ".0.weight", (size_t)cout * cin * 9);
  const auto& b = *pk.get(p + ".0.bias", cout);
  const auto& g = *pk.get(p + ".1.weight", cout);
  const auto& be = *pk.get(p + ".1.bias", cout);
  const auto& mu = *pk.get(p + ".1.running_mean", cout);
  const auto& var = *pk.get(p + ".1.running_var", cout);
  for (int n = 0; n < cout; ++n) {
    const double sc = (double)g[n] / sqrt((double)var[n] + 1e-5);
    for (int c = 0; c < cin; ++c)
      for (int tap = 0; tap < 9; ++tap)
        w_out[(size_t)n * 9 * cin + tap * cin + c] = (float)(w[((size_t)n * cin + c) * 9 + tap] * sc);
    b_out[n] = (float)(((double)b[n] - mu[n]) * sc + be[n]);
  }
}

}  // namespace

// candidate-elimination schedule: CE_LOC = [3,6,9], keep ratio 0.7 (attn_blocks.py:40); and which LayerNorms travel
// inside the GEMMs (gemm2.cu "LayerNorm fold"): only where both GEMMs of the pair run on the 2-CTA kernel (large
// batches) and no candidate elimination sits between them
void Engine::plan_schedule() {
  int ns = kLensX;
  for (int i = 0; i < kDepth; ++i) {
    n_tok[i] = kLensT + ns;
    ce_keep[i] = 0;
    if (i == 3 || i == 6 || i == 9) {
      ce_keep[i] = (int)ceil(0.7 * (double)ns);
      ns = ce_keep[i];
    }
    n_mlp[i] = kLensT + ns;
  }
  n_tok[kDepth] = kLensT + ns;
  for (int i = 0; i < kDepth; ++i) {
    auto pair = [&](int n) { return ((B * n + 255) / 256) * 3 >= num_sms / 2; };
    L[i].fuse_ln2 = opt_lnfold && !ce_keep[i] && pair(n_tok[i]);
    L[i].fuse_ln1 = opt_lnfold && !deep && i >= 1 && pair(n_tok[i]);
    // prompt fold: the deep-prompt stage of layer i inside fc2[i-1], qkv[i] and proj[i] (proj is then the residual kind,
    // also in the candidate-elimination layers, where the gather + norm2 pass follows it)
    PW[i].fold_on = opt_promptfold && opt_lnfold && deep && i >= 1 && pair(n_tok[i]) && (opt_promptfold_ce || !ce_keep[i]);
  }
}

// W (N, 768) row-major -> gamma * W; returns bias' = b + W beta and the column sums of the fp16-rounded rows
static void fold_ln(const std::vector<float>& w, const std::vector<float>& b, const std::vector<float>& g,
                    const std::vector<float>& be, int N, std::vector<float>& wf, std::vector<float>& bf, std::vector<float>& cs) {
  wf.resize((size_t)N * 768); bf.resize(N); cs.resize(N);
  for (int n = 0; n < N; ++n) {
    double sb = b[n], sc = 0;
    for (int k = 0; k < 768; ++k) {
      const float v = w[(size_t)n * 768 + k] * g[k];
      wf[(size_t)n * 768 + k] = v;
      sc += (double)__half2float(__float2half_rn(v));
      sb += (double)w[(size_t)n * 768 + k] * (double)be[k];
    }
    bf[n] = (float)sb;
    cs[n] = (float)sc;
  }
}

int Engine::finalize(int batch, bool deep_prompt) {
  if (finalized) return VIPT_ERR_STATE;
  B = batch;
  deep = deep_prompt;
  int dev = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  plan_schedule();
  Packer pk{this};
  const std::string bb = "backbone.";
  // ---- patch embeddings: group 0 = RGB (patch_embed), group 1 = X (patch_embed_prompt)
  {
    std::vector<float> w(2 * 768 * 768), b(2 * 768);
    const char* names[2] = {"patch_embed", "patch_embed_prompt"};
    for (int g = 0; g < 2; ++g) {
      const auto& wg = *pk.get(bb + names[g] + ".proj.weight", 768 * 768);
      const auto& bg = *pk.get(bb + names[g] + ".proj.bias", 768);
      memcpy(&w[(size_t)g * 768 * 768], wg.data(), 768 * 768 * 4);
      memcpy(&b[g * 768], bg.data(), 768 * 4);
    }
    w_pe = pk.up_f16(w.data(), w.size());
    b_pe = pk.up_f32(b.data(), b.size());
    std::vector<float> pe(320 * 768);
    memcpy(pe.data(), pk.get(bb + "pos_embed_z", 64 * 768)->data(), 64 * 768 * 4);
    memcpy(pe.data() + 64 * 768, pk.get(bb + "pos_embed_x", 256 * 768)->data(), 256 * 768 * 4);
    pos_embed = pk.up_f32(pe.data(), pe.size());
  }
  // ---- transformer blocks
  for (int i = 0; i < kDepth; ++i) {
    const std::string p = bb + "blocks." + std::to_string(i) + ".";
    // (where a LayerNorm / prompt fold replaces a weight by its gain-scaled form below, the plain copy is not uploaded)
    if (!L[i].fuse_ln1 && !PW[i].fold_on) L[i].w_qkv = pk.up_f16(pk.get(p + "attn.qkv.weight", 2304 * 768)->data(), 2304 * 768);
    L[i].b_qkv = pk.up_f32(pk.get(p + "attn.qkv.bias", 2304)->data(), 2304);
    L[i].w_proj = pk.up_f16(pk.get(p + "attn.proj.weight", 768 * 768)->data(), 768 * 768);
    L[i].b_proj = pk.up_f32(pk.get(p + "attn.proj.bias", 768)->data(), 768);
    if (!L[i].fuse_ln2) L[i].w_fc1 = pk.up_f16(pk.get(p + "mlp.fc1.weight", 3072 * 768)->data(), 3072 * 768);
    L[i].b_fc1 = pk.up_f32(pk.get(p + "mlp.fc1.bias", 3072)->data(), 3072);
    L[i].w_fc2 = pk.up_f16(pk.get(p + "mlp.fc2.weight", 768 * 3072)->data(), 768 * 3072);
    L[i].b_fc2 = pk.up_f32(pk.get(p + "mlp.fc2.bias", 768)->data(), 768);
    L[i].ln1_g = pk.up_f32(pk.get(p + "norm1.weight", 768)->data(), 768);
    L[i].ln1_b = pk.up_f32(pk.get(p + "norm1.bias", 768)->data(), 768);
    L[i].ln2_g = pk.up_f32(pk.get(p + "norm2.weight", 768)->data(), 768);
    L[i].ln2_b = pk.up_f32(pk.get(p + "norm2.bias", 768)->data(), 768);
    std::vector<float> wf, bf, cs;
    if (L[i].fuse_ln2) {
      fold_ln(*pk.get(p + "mlp.fc1.weight", 3072 * 768), *pk.get(p + "mlp.fc1.bias", 3072), *pk.get(p + "norm2.weight", 768),
              *pk.get(p + "norm2.bias", 768), 3072, wf, bf, cs);
      L[i].w_fc1 = pk.up_f16(wf.data(), wf.size());
      L[i].b_fc1 = pk.up_f32(bf.data(), bf.size());
      L[i].cs_fc1 = pk.up_f32(cs.data(), cs.size());
    }
    if (L[i].fuse_ln1) {
      fold_ln(*pk.get(p + "attn.qkv.weight", 2304 * 768), *pk.get(p + "attn.qkv.bias", 2304), *pk.get(p + "norm1.weight", 768),
              *pk.get(p + "norm1.bias", 768), 2304, wf, bf, cs);
      L[i].w_qkv = pk.up_f16(wf.data(), wf.size());
      L[i].b_qkv = pk.up_f32(bf.data(), bf.size());
      L[i].cs_qkv = pk.up_f32(cs.data(), cs.size());
    }
  }
  norm_g = pk.up_f32(pk.get(bb + "norm.weight", 768)->data(), 768);
  norm_b = pk.up_f32(pk.get(bb + "norm.bias", 768)->data(), 768);
  // ---- prompt blocks: LayerNorm folded into the two down-projections (tokens.cu, phase A)
  const int n_prompt = deep ? kDepth : 1;
  for (int i = 0; i < n_prompt; ++i) {
    const std::string p = bb + "prompt_blocks." + std::to_string(i) + ".";
    const int ia = i == 0 ? 0 : i - 1;      // norm of the token stream (vit_ce_prompt.py:206,288)
    const auto& ga = *pk.get(bb + "prompt_norms." + std::to_string(ia) + ".weight", 768);
    const auto& ba = *pk.get(bb + "prompt_norms." + std::to_string(ia) + ".bias", 768);
    const auto& gd = *pk.get(bb + "prompt_norms." + std::to_string(i) + ".weight", 768);
    const auto& bd = *pk.get(bb + "prompt_norms." + std::to_string(i) + ".bias", 768);
    const auto& w00 = *pk.get(p + "conv0_0.weight", 8 * 768);
    const auto& b00 = *pk.get(p + "conv0_0.bias", 8);
    const auto& w01 = *pk.get(p + "conv0_1.weight", 8 * 768);
    const auto& b01 = *pk.get(p + "conv0_1.bias", 8);
    std::vector<float> wga(8 * 768), wgd(8 * 768), sc(32);
    for (int k = 0; k < 8; ++k) {
      double sa = 0, ca = 0, sd = 0, cd = 0;
      for (int c = 0; c < 768; ++c) {
        wga[k * 768 + c] = w00[k * 768 + c] * ga[c];
        wgd[k * 768 + c] = w01[k * 768 + c] * gd[c];
        sa += (double)wga[k * 768 + c]; ca += (double)w00[k * 768 + c] * ba[c];
        sd += (double)wgd[k * 768 + c]; cd += (double)w01[k * 768 + c] * bd[c];
      }
      sc[k] = (float)sa; sc[8 + k] = (float)(ca + b00[k]);
      sc[16 + k] = (float)sd; sc[24 + k] = (float)(cd + b01[k]);
    }
    PW[i].wg_a = pk.up_f32(wga.data(), wga.size());
    PW[i].wg_d = pk.up_f32(wgd.data(), wgd.size());
    PW[i].sc = pk.up_f32(sc.data(), 32);
    PW[i].w1 = pk.up_f32(pk.get(p + "conv1x1.weight", 768 * 8)->data(), 768 * 8);
    PW[i].b1 = pk.up_f32(pk.get(p + "conv1x1.bias", 768)->data(), 768);
    PW[i].smooth = (*pk.get(p + "fovea.smooth", 1))[0];
    PW[i].fold = nullptr; PW[i].w1t = nullptr;
    if (i >= 1) {
      // fused deep-prompt kernel (prompt_coop.cu): conv1x1 transposed to k-major, and the previous
      // prompt P = W1p f + b1p (rank 8) folded through prompt_norms[i] and conv0_1 of this block.
      const auto& w1 = *pk.get(p + "conv1x1.weight", 768 * 8);
      std::vector<float> w1t(8 * 768);
      for (int c = 0; c < 768; ++c)
        for (int k = 0; k < 8; ++k) w1t[k * 768 + c] = w1[c * 8 + k];
      PW[i].w1t = pk.up_f32(w1t.data(), w1t.size());
      const std::string pp = bb + "prompt_blocks." + std::to_string(i - 1) + ".";
      const auto& w1p = *pk.get(pp + "conv1x1.weight", 768 * 8);
      const auto& b1p = *pk.get(pp + "conv1x1.bias", 768);
      double m[8] = {0}, m0 = 0;
      for (int c = 0; c < 768; ++c) {
        for (int k = 0; k < 8; ++k) m[k] += (double)w1p[c * 8 + k];
        m0 += (double)b1p[c];
      }
      for (int k = 0; k < 8; ++k) m[k] /= 768.0;
      m0 /= 768.0;
      double G[64] = {0}, h[8] = {0}, c0 = 0, A[64] = {0}, a0[8] = {0};
      for (int c = 0; c < 768; ++c) {
        double wc[8];
        for (int k = 0; k < 8; ++k) wc[k] = (double)w1p[c * 8 + k] - m[k];
        const double bc = (double)b1p[c] - m0;
        for (int a = 0; a < 8; ++a) {
          for (int q = 0; q < 8; ++q) G[a * 8 + q] += wc[a] * wc[q];
          h[a] += wc[a] * bc;
        }
        c0 += bc * bc;
        for (int k = 0; k < 8; ++k) {
          const double wg = (double)w01[k * 768 + c] * (double)gd[c];
          for (int q = 0; q < 8; ++q) A[k * 8 + q] += wg * wc[q];
          a0[k] += wg * bc;
        }
      }
      std::vector<float> fold(168, 0.f);
      for (int k = 0; k < 8; ++k) {
        fold[0 + k] = sc[8 + k];            // c_a = conv0_0(beta_a) + bias
        fold[72 + k] = (float)a0[k];
        fold[80 + k] = sc[24 + k];          // c_d = conv0_1(beta_d) + bias
        fold[152 + k] = (float)(h[k] / 768.0);
      }
      for (int q = 0; q < 64; ++q) { fold[8 + q] = (float)A[q]; fold[88 + q] = (float)(G[q] / 768.0); }
      fold[160] = (float)(c0 / 768.0);
      fold[161] = PW[i].smooth;
      PW[i].fold = pk.up_f32(fold.data(), fold.size());
      if (PW[i].fold_on) {
        // u = this block's conv1x1 (768, 8) and bias b1: x = v + U f + b1
        std::vector<float> c20((size_t)768 * 20, 0.f), f2(112, 0.f);
        double su[8] = {0}, sb = 0, G9[81] = {0};
        for (int c = 0; c < 768; ++c) {
          double ue[9];
          for (int k = 0; k < 8; ++k) {
            c20[(size_t)c * 20 + k] = wga[k * 768 + c];
            c20[(size_t)c * 20 + 8 + k] = w1[c * 8 + k];
            ue[k] = w1[c * 8 + k];
            su[k] += ue[k];
          }
          const float b1c = (*pk.get(p + "conv1x1.bias", 768))[c];
          c20[(size_t)c * 20 + 16] = b1c;
          ue[8] = b1c;
          sb += b1c;
          for (int a = 0; a < 9; ++a)
            for (int q = 0; q < 9; ++q) G9[a * 9 + q] += ue[a] * ue[q];
        }
        for (int k = 0; k < 8; ++k) { f2[k] = sc[k]; f2[8 + k] = (float)su[k]; }
        f2[16] = (float)sb;
        for (int q = 0; q < 81; ++q) f2[24 + q] = (float)G9[q];
        PW[i].colw20 = pk.up_f32(c20.data(), c20.size());
        PW[i].fold2 = pk.up_f32(f2.data(), f2.size());
        // attn.qkv of this block over x16 = fp16(v): W' = fp16(gamma1 * Wqkv); the rank-8 update enters as W'U, W'b1
        const std::string bp = bb + "blocks." + std::to_string(i) + ".";
        const auto& b1v = *pk.get(p + "conv1x1.bias", 768);
        std::vector<float> wf, bf, cs;
        fold_ln(*pk.get(bp + "attn.qkv.weight", 2304 * 768), *pk.get(bp + "attn.qkv.bias", 2304), *pk.get(bp + "norm1.weight", 768),
                *pk.get(bp + "norm1.bias", 768), 2304, wf, bf, cs);
        L[i].w_qkv = pk.up_f16(wf.data(), wf.size());
        std::vector<float> wxq((size_t)2304 * 64, 0.f), wxp((size_t)768 * 64, 0.f), bq(2304);
        for (int n = 0; n < 2304; ++n) {
          double wu[8] = {0}, wb = 0;
          for (int c = 0; c < 768; ++c) {
            const double wv = (double)__half2float(__float2half_rn(wf[(size_t)n * 768 + c]));
            for (int k = 0; k < 8; ++k) wu[k] += wv * (double)w1[c * 8 + k];
            wb += wv * (double)b1v[c];
          }
          for (int k = 0; k < 8; ++k) wxq[(size_t)n * 64 + k] = (float)wu[k];
          wxq[(size_t)n * 64 + 8] = (float)wb;      // W'b1 is scaled by rstd like the rest of x W'^T: column 8 of Ax is 1
          bq[n] = bf[n];
        }
        for (int c = 0; c < 768; ++c)
          for (int k = 0; k < 8; ++k) wxp[(size_t)c * 64 + k] = w1[c * 8 + k];
        PW[i].wx_qkv = pk.up_f16(wxq.data(), wxq.size());
        PW[i].wx_proj = pk.up_f16(wxp.data(), wxp.size());
        PW[i].b_qkv_r8 = pk.up_f32(bq.data(), bq.size());
        PW[i].cs_qkv_r8 = pk.up_f32(cs.data(), cs.size());
        const auto& bpj = *pk.get(bp + "attn.proj.bias", 768);
        std::vector<float> br(768);
        for (int c = 0; c < 768; ++c) br[c] = bpj[c] + b1v[c];
        PW[i].b_proj_r8 = pk.up_f32(br.data(), 768);
      }
    }
  }
  // ---- center head: towers ctr | offset | size; BN folded
  {
    const char* towers[3] = {"ctr", "offset", "size"};
    const int cin[4] = {768, 256, 128, 64}, cout[4] = {256, 128, 64, 32};
    __half** wdst[4] = {&w_c1, &w_c2, &w_c3, &w_c4};
    float** bdst[4] = {&b_c1, &b_c2, &b_c3, &b_c4};
    for (int li = 0; li < 4; ++li) {
      const size_t per = (size_t)cout[li] * 9 * cin[li];
      std::vector<float> w(3 * per), b(3 * cout[li]);
      for (int t = 0; t < 3; ++t)
        fold_conv_bn(pk, std::string("box_head.conv") + std::to_string(li + 1) + "_" + towers[t], cout[li], cin[li],
                     &w[t * per], &b[t * cout[li]]);
      *wdst[li] = pk.up_f16(w.data(), w.size());
      *bdst[li] = pk.up_f32(b.data(), b.size());
    }
    std::vector<float> w5h(5 * 32), b5h(5);
    const auto& wc = *pk.get("box_head.conv5_ctr.weight", 32);
    const auto& wo = *pk.get("box_head.conv5_offset.weight", 64);
    const auto& ws = *pk.get("box_head.conv5_size.weight", 64);
    memcpy(&w5h[0], wc.data(), 32 * 4); memcpy(&w5h[32], wo.data(), 64 * 4); memcpy(&w5h[96], ws.data(), 64 * 4);
    b5h[0] = (*pk.get("box_head.conv5_ctr.bias", 1))[0];
    b5h[1] = (*pk.get("box_head.conv5_offset.bias", 2))[0]; b5h[2] = (*pk.get("box_head.conv5_offset.bias", 2))[1];
    b5h[3] = (*pk.get("box_head.conv5_size.bias", 2))[0]; b5h[4] = (*pk.get("box_head.conv5_size.bias", 2))[1];
    w5 = pk.up_f32(w5h.data(), 160);
    b5 = pk.up_f32(b5h.data(), 5);
    // hann2d (hann.py:6-16): w[i] = 0.5 (1 - cos(2 pi i / 17)), i = 1..16, fp32 ops like torch
    std::vector<float> win(256), w1(16);
    const float step = (float)(2.0 * M_PI / 17.0);
    for (int i = 0; i < 16; ++i) w1[i] = 0.5f * (1.0f - cosf(step * (float)(i + 1)));
    for (int y = 0; y < 16; ++y)
      for (int x = 0; x < 16; ++x) win[y * 16 + x] = w1[y] * w1[x];
    window = pk.up_f32(win.data(), 256);
  }
  if (!pk.ok) return VIPT_ERR_STATE;
  raw.clear();
  return finalize_buffers();
}

}  // namespace vipt
