// Per-GPU weight replica + per-worker workspaces + the forward the codelet enqueues.
// Replaces load_model/clone_model_to_gpus (/root/reference/src/core/inference_runner.cpp:
// 243-275) and the body of run_inference (/root/reference/src/core/starpu_setup.cpp:594-624).
#include <cmath>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "replica.cuh"

namespace spu {

static thread_local std::string g_last_error;
void set_error(const std::string &msg) { g_last_error = msg; }

namespace {

// Optional per-kernel event trail (measurement hook only).
struct StageTrail {
  std::vector<cudaEvent_t> ev;  // ev[0] = start, ev[i+1] recorded after launch i
  std::vector<int> stage;       // stage id of launch i
  size_t n = 0;
  spu_status mark(int stage_id, cudaStream_t s) {
    if (n + 1 >= ev.size()) {
      set_error("profiling: event trail too short");
      return SPU_ERR_INVALID;
    }
    stage[n] = stage_id;
    SPU_CUDA_TRY(cudaEventRecord(ev[++n], s));
    return SPU_OK;
  }
};

constexpr float kE4m3Max = 448.f;

spu_status quantize_weight(spu_replica *r, uint8_t **dst, float *scale, const __nv_bfloat16 *w_bf16,
                           std::initializer_list<const float *> srcs, size_t n_each) {
  float amax = 0.f;
  for (const float *s : srcs)
    for (size_t i = 0; i < n_each; ++i) amax = std::fmax(amax, std::fabs(s[i]));
  *scale = (amax > 0.f ? amax : 1.f) / kE4m3Max;
  const size_t total = n_each * srcs.size();
  SPU_TRY(dev_alloc(r, dst, total));
  return quantize_e4m3_launch(nullptr, w_bf16, *dst, total, 1.f / *scale, nullptr);
}

spu_status build(spu_replica *r, const float *const *t) {
  const spu_bert_desc &d = r->desc;
  const size_t H = d.hidden, F = d.ffn;
  SPU_TRY(upload_f32(r, &r->word, t[0], static_cast<size_t>(d.vocab) * H));
  SPU_TRY(upload_f32(r, &r->pos, t[1], static_cast<size_t>(d.max_pos) * H));
  SPU_TRY(upload_f32(r, &r->type0, t[2], H));  // token_type_ids are all zero on this path
  SPU_TRY(upload_f32(r, &r->emb_g, t[3], H));
  SPU_TRY(upload_f32(r, &r->emb_b, t[4], H));
  r->layers.resize(d.layers);
  for (int l = 0; l < d.layers; ++l) {
    const float *const *p = t + 5 + 16 * l;
    LayerWeights &w = r->layers[l];
    SPU_TRY(upload_bf16_stacked(r, &w.w_qkv, {p[0], p[2], p[4]}, H * H));
    SPU_TRY(upload_f32_stacked(r, &w.b_qkv, {p[1], p[3], p[5]}, H));
    SPU_TRY(upload_bf16_stacked(r, &w.w_o, {p[6]}, H * H));
    SPU_TRY(upload_f32(r, &w.b_o, p[7], H));
    SPU_TRY(upload_f32(r, &w.ln1_g, p[8], H));
    SPU_TRY(upload_f32(r, &w.ln1_b, p[9], H));
    SPU_TRY(upload_bf16_stacked(r, &w.w_1, {p[10]}, F * H));
    SPU_TRY(upload_f32(r, &w.b_1, p[11], F));
    SPU_TRY(upload_bf16_stacked(r, &w.w_2, {p[12]}, H * F));
    SPU_TRY(upload_f32(r, &w.b_2, p[13], H));
    SPU_TRY(upload_f32(r, &w.ln2_g, p[14], H));
    SPU_TRY(upload_f32(r, &w.ln2_b, p[15], H));
    if (r->flags & SPU_PREC_FP8_E4M3) {
      // per-tensor weight scale = amax / 448 (E4M3 max finite), quantised on the device
      // from the bf16 copy (which the calibration pass also uses)
      SPU_TRY(quantize_weight(r, &w.q_qkv, &w.sw_qkv, w.w_qkv, {p[0], p[2], p[4]}, H * H));
      SPU_TRY(quantize_weight(r, &w.q_o, &w.sw_o, w.w_o, {p[6]}, H * H));
      SPU_TRY(quantize_weight(r, &w.q_1, &w.sw_1, w.w_1, {p[10]}, F * H));
      SPU_TRY(quantize_weight(r, &w.q_2, &w.sw_2, w.w_2, {p[12]}, H * F));
    }
  }
  if (r->flags & SPU_PREC_FP8_E4M3) {
    SPU_TRY(dev_alloc(r, &r->amax, static_cast<size_t>(d.layers) * 4));
    SPU_CUDA_TRY(cudaMemset(r->amax, 0, static_cast<size_t>(d.layers) * 4 * sizeof(float)));
  }
  const size_t rows = r->cap_rows;
  for (Workspace &ws : r->ws) {
    SPU_TRY(dev_alloc(r, &ws.x_f32, rows * H));
    SPU_TRY(dev_alloc(r, &ws.x_bf16, rows * H));
    SPU_TRY(dev_alloc(r, &ws.qkv, rows * 3 * H));
    SPU_TRY(dev_alloc(r, &ws.ctx, rows * H));
    SPU_TRY(dev_alloc(r, &ws.h, rows * F));
    SPU_TRY(dev_alloc(r, &ws.mask_bias, rows));
    SPU_TRY(dev_alloc(r, &ws.cu_seqlens, static_cast<size_t>(kMaxPackedBatch) + 1));
    if (r->flags & SPU_REQUEST_IO) {
      SPU_TRY(dev_alloc(r, &ws.req_ids, rows));
      SPU_TRY(dev_alloc(r, &ws.req_mask, rows));
      SPU_TRY(dev_alloc(r, &ws.req_out, rows * H));
      SPU_TRY(dev_alloc(r, &ws.req_err, 1));
      SPU_CUDA_TRY(cudaMemset(ws.req_err, 0, sizeof(int32_t)));
    }
    if (r->flags & SPU_PREC_FP8_E4M3) {
      SPU_TRY(dev_alloc(r, &ws.x_q, rows * H));
      SPU_TRY(dev_alloc(r, &ws.ctx_q, rows * H));
      SPU_TRY(dev_alloc(r, &ws.h_q, rows * F));
      ws.plans8.resize(d.layers);
      for (int l = 0; l < d.layers; ++l) {
        const LayerWeights &w = r->layers[l];
        const int R = static_cast<int>(rows);
        SPU_TRY(gemm_make_plan(&ws.plans8[l].qkv, ws.x_q, R, w.q_qkv, 3 * d.hidden, d.hidden, kEpiBias,
                               GemmIn::FP8_E4M3));
        SPU_TRY(gemm_make_plan(&ws.plans8[l].attn_out, ws.ctx_q, R, w.q_o, d.hidden, d.hidden, kEpiBiasResid,
                               GemmIn::FP8_E4M3));
        SPU_TRY(gemm_make_plan(&ws.plans8[l].ffn1, ws.x_q, R, w.q_1, d.ffn, d.hidden, kEpiBiasGeluE4m3,
                               GemmIn::FP8_E4M3));
        SPU_TRY(gemm_make_plan(&ws.plans8[l].ffn2, ws.h_q, R, w.q_2, d.hidden, d.ffn, kEpiBiasResid,
                               GemmIn::FP8_E4M3));
        if (r->ln_fused) {
          SPU_TRY(gemm_make_plan(&ws.plans8[l].attn_out_ln, ws.ctx_q, R, w.q_o, d.hidden, d.hidden, kEpiBiasResidLN,
                                 GemmIn::FP8_E4M3));
          SPU_TRY(gemm_make_plan(&ws.plans8[l].ffn2_ln, ws.h_q, R, w.q_2, d.hidden, d.ffn, kEpiBiasResidLN,
                                 GemmIn::FP8_E4M3));
        }
        SPU_TRY(gemm_make_plan(&ws.plans8[l].ffn1_b2q, ws.x_bf16, R, w.w_1, d.ffn, d.hidden, kEpiBiasGeluE4m3,
                               GemmIn::BF16));
        SPU_TRY(gemm_make_plan(&ws.plans8[l].ffn1_q2b, ws.x_q, R, w.q_1, d.ffn, d.hidden, kEpiBiasGelu,
                               GemmIn::FP8_E4M3));
      }
    }
    ws.plans.resize(d.layers);
    for (int l = 0; l < d.layers; ++l) {
      const LayerWeights &w = r->layers[l];
      const int R = static_cast<int>(rows);
      SPU_TRY(gemm_make_plan(&ws.plans[l].qkv, ws.x_bf16, R, w.w_qkv, 3 * d.hidden, d.hidden,
                             SPU_EPI_BIAS, GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].attn_out, ws.ctx, R, w.w_o, d.hidden, d.hidden,
                             SPU_EPI_BIAS_RESID, GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].ffn1, ws.x_bf16, R, w.w_1, d.ffn, d.hidden,
                             SPU_EPI_BIAS_GELU, GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].ffn2, ws.h, R, w.w_2, d.hidden, d.ffn,
                             SPU_EPI_BIAS_RESID, GemmIn::BF16));
      if (r->ln_fused) {
        SPU_TRY(gemm_make_plan(&ws.plans[l].attn_out_ln, ws.ctx, R, w.w_o, d.hidden, d.hidden, kEpiBiasResidLN,
                               GemmIn::BF16));
        SPU_TRY(gemm_make_plan(&ws.plans[l].ffn2_ln, ws.h, R, w.w_2, d.hidden, d.ffn, kEpiBiasResidLN,
                               GemmIn::BF16));
      }
    }
  }
  // every plan consults the replica's table of measured tiles (empty until spu_bert_autotune runs)
  for (Workspace &ws : r->ws) {
    for (std::vector<LayerPlans> *set : {&ws.plans, &ws.plans8})
      for (LayerPlans &p : *set)
        for (GemmPlan *g : {&p.qkv, &p.attn_out, &p.ffn1, &p.ffn2, &p.ffn1_b2q, &p.ffn1_q2b, &p.attn_out_ln, &p.ffn2_ln})
          g->tuned = &r->tuned;
  }
  SPU_CUDA_TRY(cudaDeviceSynchronize());
  return SPU_OK;
}

#define SPU_MARK(id)                                  \
  do {                                                \
    if (trail != nullptr) SPU_TRY(trail->mark((id), stream)); \
  } while (0)
#define SPU_ON(id) ((stage_mask >> (id)) & 1u)

spu_status enqueue_forward_fp8(spu_replica *r, Workspace &ws, const int64_t *ids, const int64_t *mask,
                               int batch, int seq, float *out, cudaStream_t stream, uint32_t stage_mask);

spu_status enqueue_forward(spu_replica *r, Workspace &ws, const int64_t *ids, const int64_t *mask,
                           int batch, int seq, float *out, cudaStream_t stream,
                           StageTrail *trail = nullptr, uint32_t stage_mask = 0xffffffffu) {
  if (r->flags & SPU_PREC_FP8_E4M3) {
    SPU_REQUIRE(r->fp8_calibrated, SPU_ERR_INVALID,
                "forward: FP8 replica has no activation scales yet (call spu_bert_calibrate_fp8 first)");
    SPU_REQUIRE(trail == nullptr, SPU_ERR_UNSUPPORTED, "profiled forward: use spu_bert_stage_chain_ms for FP8");
    return enqueue_forward_fp8(r, ws, ids, mask, batch, seq, out, stream, stage_mask);
  }
  const spu_bert_desc &d = r->desc;
  const int M = batch * seq;
  if (SPU_ON(SPU_STAGE_EMBED)) SPU_TRY(bert_embed_launch(ids, mask, r->word, r->pos, r->type0, r->emb_g, r->emb_b, d.ln_eps,
                            d.vocab, ws.x_f32, ws.x_bf16, ws.mask_bias, batch, seq, d.hidden,
                            stream));
  SPU_MARK(SPU_STAGE_EMBED);
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    const LayerPlans &p = ws.plans[l];
    const bool last = l == d.layers - 1;
    GemmEpilogueArgs e{};
    e.acc_scale = 1.f;
    e.bias = w.b_qkv;
    e.out = ws.qkv;
    if (SPU_ON(SPU_STAGE_GEMM_QKV)) SPU_TRY(gemm_launch(p.qkv, M, e, stream));
    SPU_MARK(SPU_STAGE_GEMM_QKV);
    if (SPU_ON(SPU_STAGE_ATTENTION)) SPU_TRY(attention_launch(ws.qkv, ws.mask_bias, ws.ctx, batch, seq, d.heads, stream));
    SPU_MARK(SPU_STAGE_ATTENTION);
    const bool out16 = last && (r->flags & SPU_OUTPUT_BF16) != 0;  // `out` is a bf16 buffer then
    if (r->ln_fused) {
      // x = LN(ctx Wo^T + bo + x) in ONE kernel: the epilogue normalises over the cluster that spans the row and
      // writes the fp32 residual stream (in place) and the bf16 operand of FFN1
      GemmEpilogueArgs f{};
      f.acc_scale = 1.f;
      f.bias = w.b_o;
      f.resid = ws.x_f32;
      f.out = ws.x_f32;
      f.out2 = ws.x_bf16;
      f.gamma = w.ln1_g;
      f.beta = w.ln1_b;
      f.ln_eps = d.ln_eps;
      if (SPU_ON(SPU_STAGE_GEMM_OUT)) SPU_TRY(gemm_launch(p.attn_out_ln, M, f, stream));
      SPU_MARK(SPU_STAGE_GEMM_OUT);
      e.bias = w.b_1;
      e.resid = nullptr;
      e.out = ws.h;
      if (SPU_ON(SPU_STAGE_GEMM_FFN1)) SPU_TRY(gemm_launch(p.ffn1, M, e, stream));
      SPU_MARK(SPU_STAGE_GEMM_FFN1);
      // the last layer writes last_hidden_state straight into the StarPU output buffer
      f.bias = w.b_2;
      f.gamma = w.ln2_g;
      f.beta = w.ln2_b;
      f.out = out16 ? nullptr : (last ? out : ws.x_f32);
      f.out2 = out16 ? static_cast<void *>(out) : (last ? nullptr : ws.x_bf16);
      if (SPU_ON(SPU_STAGE_GEMM_FFN2)) SPU_TRY(gemm_launch(p.ffn2_ln, M, f, stream));
      SPU_MARK(SPU_STAGE_GEMM_FFN2);
      continue;
    }
    e.bias = w.b_o;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;  // in place: each element is read then written by the same thread
    if (SPU_ON(SPU_STAGE_GEMM_OUT)) SPU_TRY(gemm_launch(p.attn_out, M, e, stream));
    SPU_MARK(SPU_STAGE_GEMM_OUT);
    if (SPU_ON(SPU_STAGE_LN1)) SPU_TRY(layernorm_launch(ws.x_f32, w.ln1_g, w.ln1_b, d.ln_eps, ws.x_f32, ws.x_bf16, M,
                             d.hidden, stream));
    SPU_MARK(SPU_STAGE_LN1);
    e.bias = w.b_1;
    e.resid = nullptr;
    e.out = ws.h;
    if (SPU_ON(SPU_STAGE_GEMM_FFN1)) SPU_TRY(gemm_launch(p.ffn1, M, e, stream));
    SPU_MARK(SPU_STAGE_GEMM_FFN1);
    e.bias = w.b_2;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    if (SPU_ON(SPU_STAGE_GEMM_FFN2)) SPU_TRY(gemm_launch(p.ffn2, M, e, stream));
    SPU_MARK(SPU_STAGE_GEMM_FFN2);
    // the last LayerNorm writes last_hidden_state straight into the StarPU output buffer
    if (SPU_ON(SPU_STAGE_LN2)) SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps,
                             out16 ? nullptr : (last ? out : ws.x_f32),
                             out16 ? static_cast<void *>(out) : (last ? nullptr : ws.x_bf16), M, d.hidden, stream));
    SPU_MARK(SPU_STAGE_LN2);
  }
  return SPU_OK;
}

// Packed forward (bf16): the encoder runs on the batch's TRUE tokens only.  `lens` (host) gives each sequence's
// length; ids is still the padded [batch, seq] array the host gather produced.  Rows of all sequences sit back to
// back in the workspace (T = sum of lengths), so the GEMMs and LayerNorms do T rows of work instead of
// batch * seq, attention reads per-sequence extents from cu_seqlens, and the last LayerNorm writes the padded
// [batch, seq, hidden] output the StarPU handle expects (zeros behind each length).  Every live row equals the
// padded forward's bit for bit: a row's GEMM / LayerNorm result does not depend on its neighbours, and keys past a
// sequence's end get probability exactly 0 either way.
spu_status enqueue_forward_packed(spu_replica *r, Workspace &ws, const int64_t *ids, const int32_t *lens, int batch,
                                  int seq, float *out, cudaStream_t stream, bool compact_out = false) {
  const spu_bert_desc &d = r->desc;
  int T = 0;
  SPU_TRY(write_cu_seqlens_launch(lens, batch, ws.cu_seqlens, &T, stream));
  SPU_REQUIRE(T > 0 && T <= r->cap_rows, SPU_ERR_CAPACITY, "packed forward: token count out of range");
  SPU_TRY(bert_embed_packed_launch(ids, ws.cu_seqlens, batch, seq, T, r->word, r->pos, r->type0, r->emb_g, r->emb_b,
                                   d.ln_eps, d.vocab, ws.x_f32, ws.x_bf16, d.hidden, stream));
  int max_len = 0;
  for (int b = 0; b < batch; ++b) max_len = lens[b] > max_len ? lens[b] : max_len;
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    const LayerPlans &p = ws.plans[l];
    const bool last = l == d.layers - 1;
    GemmEpilogueArgs e{};
    e.acc_scale = 1.f;
    e.bias = w.b_qkv;
    e.out = ws.qkv;
    SPU_TRY(gemm_launch(p.qkv, T, e, stream));
    SPU_TRY(attention_tc_launch(ws.qkv, nullptr, ws.ctx, batch, max_len, d.heads, stream, false, 1.f, ws.cu_seqlens, T));
    e.bias = w.b_o;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.attn_out, T, e, stream));
    SPU_TRY(layernorm_launch(ws.x_f32, w.ln1_g, w.ln1_b, d.ln_eps, ws.x_f32, ws.x_bf16, T, d.hidden, stream));
    e.bias = w.b_1;
    e.resid = nullptr;
    e.out = ws.h;
    SPU_TRY(gemm_launch(p.ffn1, T, e, stream));
    e.bias = w.b_2;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.ffn2, T, e, stream));
    if (!last) {
      SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps, ws.x_f32, ws.x_bf16, T, d.hidden, stream));
    } else {
      const bool out16 = (r->flags & SPU_OUTPUT_BF16) != 0;
      if (compact_out)  // the output keeps the packed layout: [T, hidden], sequence b at row cu_seqlens[b]
        SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps, out16 ? nullptr : out,
                                 out16 ? static_cast<void *>(out) : nullptr, T, d.hidden, stream));
      else
        SPU_TRY(layernorm_unpad_launch(ws.x_f32, ws.cu_seqlens, batch, seq, w.ln2_g, w.ln2_b, d.ln_eps,
                                       out16 ? nullptr : out, out16 ? static_cast<void *>(out) : nullptr, d.hidden, stream));
    }
  }
  return SPU_OK;
}

// FP8 E4M3 variant of enqueue_forward: same kernels; each of a layer's four GEMMs runs either on
// e4m3 operands with per-tensor scales folded into the epilogue (acc * sa * sw + bias) or in bf16,
// as r->fp8_mask[l] says (SPU_FP8_GEMM_* bits).  The producer of every GEMM operand (embedding,
// LayerNorm, attention, the FFN1 GELU epilogue) emits the format its consumer wants; attention,
// LayerNorm statistics and the residual stream stay bf16 / fp32.
spu_status enqueue_forward_fp8(spu_replica *r, Workspace &ws, const int64_t *ids, const int64_t *mask,
                               int batch, int seq, float *out, cudaStream_t stream, uint32_t stage_mask) {
  const spu_bert_desc &d = r->desc;
  const int M = batch * seq;
  auto on = [&](int l, int bit) { return l < d.layers && (r->fp8_mask[l] & bit) != 0; };
  {
    const bool q = on(0, SPU_FP8_GEMM_QKV);
    if (SPU_ON(SPU_STAGE_EMBED))
      SPU_TRY(bert_embed_launch(ids, mask, r->word, r->pos, r->type0, r->emb_g, r->emb_b, d.ln_eps, d.vocab,
                                ws.x_f32, q ? nullptr : ws.x_bf16, ws.mask_bias, batch, seq, d.hidden, stream,
                                q ? ws.x_q : nullptr, 1.f / r->layers[0].sa_x));
  }
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    const LayerPlans &p8 = ws.plans8[l];
    const LayerPlans &p16 = ws.plans[l];
    const bool q_qkv = on(l, SPU_FP8_GEMM_QKV), q_out = on(l, SPU_FP8_GEMM_OUT);
    const bool q_f1 = on(l, SPU_FP8_GEMM_FFN1), q_f2 = on(l, SPU_FP8_GEMM_FFN2);
    const bool last = l == d.layers - 1;
    GemmEpilogueArgs e{};
    e.bias = w.b_qkv;
    e.out = ws.qkv;
    e.acc_scale = q_qkv ? w.sa_x * w.sw_qkv : 1.f;
    if (SPU_ON(SPU_STAGE_GEMM_QKV)) SPU_TRY(gemm_launch(q_qkv ? p8.qkv : p16.qkv, M, e, stream));
    if (SPU_ON(SPU_STAGE_ATTENTION))
      SPU_TRY(attention_launch(ws.qkv, ws.mask_bias, q_out ? static_cast<void *>(ws.ctx_q) : ws.ctx, batch, seq,
                               d.heads, stream, q_out, 1.f / w.sa_ctx));
    e.bias = w.b_o;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    e.acc_scale = q_out ? w.sa_ctx * w.sw_o : 1.f;
    if (r->ln_fused) {
      GemmEpilogueArgs f = e;
      f.gamma = w.ln1_g;
      f.beta = w.ln1_b;
      f.ln_eps = d.ln_eps;
      f.out2 = q_f1 ? static_cast<void *>(ws.x_q) : ws.x_bf16;
      f.out2_e4m3 = q_f1 ? 1 : 0;
      f.out_inv_scale = 1.f / w.sa_x1;
      if (SPU_ON(SPU_STAGE_GEMM_OUT)) SPU_TRY(gemm_launch(q_out ? p8.attn_out_ln : p16.attn_out_ln, M, f, stream));
    } else {
    if (SPU_ON(SPU_STAGE_GEMM_OUT)) SPU_TRY(gemm_launch(q_out ? p8.attn_out : p16.attn_out, M, e, stream));
    if (SPU_ON(SPU_STAGE_LN1))
      SPU_TRY(layernorm_launch(ws.x_f32, w.ln1_g, w.ln1_b, d.ln_eps, ws.x_f32, q_f1 ? nullptr : ws.x_bf16, M,
                               d.hidden, stream, 0, q_f1 ? ws.x_q : nullptr, 1.f / w.sa_x1));
    }
    e.bias = w.b_1;
    e.resid = nullptr;
    e.out = q_f2 ? static_cast<void *>(ws.h_q) : ws.h;
    e.acc_scale = q_f1 ? w.sa_x1 * w.sw_1 : 1.f;
    e.out_inv_scale = 1.f / w.sa_h;
    if (SPU_ON(SPU_STAGE_GEMM_FFN1)) {
      const GemmPlan &pf1 = q_f1 ? (q_f2 ? p8.ffn1 : p8.ffn1_q2b) : (q_f2 ? p8.ffn1_b2q : p16.ffn1);
      SPU_TRY(gemm_launch(pf1, M, e, stream));
    }
    e.bias = w.b_2;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    e.acc_scale = q_f2 ? w.sa_h * w.sw_2 : 1.f;
    if (r->ln_fused) {
      const bool n8 = !last && on(l + 1, SPU_FP8_GEMM_QKV);
      const bool out16 = last && (r->flags & SPU_OUTPUT_BF16) != 0;
      GemmEpilogueArgs f = e;
      f.gamma = w.ln2_g;
      f.beta = w.ln2_b;
      f.ln_eps = d.ln_eps;
      f.out = out16 ? nullptr : (last ? out : ws.x_f32);
      f.out2 = out16 ? static_cast<void *>(out) : (last ? nullptr : (n8 ? static_cast<void *>(ws.x_q) : ws.x_bf16));
      f.out2_e4m3 = (!last && n8) ? 1 : 0;
      f.out_inv_scale = n8 ? 1.f / r->layers[l + 1].sa_x : 1.f;
      if (SPU_ON(SPU_STAGE_GEMM_FFN2)) SPU_TRY(gemm_launch(q_f2 ? p8.ffn2_ln : p16.ffn2_ln, M, f, stream));
      continue;
    }
    if (SPU_ON(SPU_STAGE_GEMM_FFN2)) SPU_TRY(gemm_launch(q_f2 ? p8.ffn2 : p16.ffn2, M, e, stream));
    if (SPU_ON(SPU_STAGE_LN2)) {
      const bool n8 = !last && on(l + 1, SPU_FP8_GEMM_QKV);
      const bool out16 = last && (r->flags & SPU_OUTPUT_BF16) != 0;
      SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps, out16 ? nullptr : (last ? out : ws.x_f32),
                               out16 ? static_cast<void *>(out) : ((last || n8) ? nullptr : ws.x_bf16), M, d.hidden,
                               stream, 0, n8 ? ws.x_q : nullptr,
                               n8 ? 1.f / r->layers[l + 1].sa_x : 1.f));
    }
  }
  return SPU_OK;
}

// bf16 forward with an amax probe after every tensor that feeds an FP8 GEMM.
spu_status enqueue_calibration(spu_replica *r, Workspace &ws, const int64_t *ids, const int64_t *mask, int batch,
                               int seq, cudaStream_t stream) {
  const spu_bert_desc &d = r->desc;
  const int M = batch * seq;
  const uint64_t nH = static_cast<uint64_t>(M) * d.hidden, nF = static_cast<uint64_t>(M) * d.ffn;
  SPU_TRY(bert_embed_launch(ids, mask, r->word, r->pos, r->type0, r->emb_g, r->emb_b, d.ln_eps, d.vocab, ws.x_f32,
                            ws.x_bf16, ws.mask_bias, batch, seq, d.hidden, stream));
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    const LayerPlans &p = ws.plans[l];
    float *am = r->amax + 4 * l;
    SPU_TRY(amax_launch(nullptr, ws.x_bf16, nH, am + 0, stream));
    GemmEpilogueArgs e{};
    e.acc_scale = 1.f;
    e.bias = w.b_qkv;
    e.out = ws.qkv;
    SPU_TRY(gemm_launch(p.qkv, M, e, stream));
    SPU_TRY(attention_launch(ws.qkv, ws.mask_bias, ws.ctx, batch, seq, d.heads, stream));
    SPU_TRY(amax_launch(nullptr, ws.ctx, nH, am + 1, stream));
    e.bias = w.b_o;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.attn_out, M, e, stream));
    SPU_TRY(layernorm_launch(ws.x_f32, w.ln1_g, w.ln1_b, d.ln_eps, ws.x_f32, ws.x_bf16, M, d.hidden, stream));
    SPU_TRY(amax_launch(nullptr, ws.x_bf16, nH, am + 2, stream));
    e.bias = w.b_1;
    e.resid = nullptr;
    e.out = ws.h;
    SPU_TRY(gemm_launch(p.ffn1, M, e, stream));
    SPU_TRY(amax_launch(nullptr, ws.h, nF, am + 3, stream));
    e.bias = w.b_2;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.ffn2, M, e, stream));
    SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps, ws.x_f32, ws.x_bf16, M, d.hidden, stream));
  }
  return SPU_OK;
}

spu_status check_forward_args(spu_replica *r, int ws_idx, const void *ids, const void *mask,
                              int batch, int seq, const void *out) {
  SPU_REQUIRE(r != nullptr, SPU_ERR_INVALID, "forward: null replica");
  SPU_REQUIRE(r->kind == 0, SPU_ERR_INVALID, "forward: not a BERT replica");
  SPU_REQUIRE(ids && mask && out, SPU_ERR_INVALID, "forward: null buffer");
  SPU_REQUIRE(ws_idx >= 0 && ws_idx < static_cast<int>(r->ws.size()), SPU_ERR_INVALID,
              "forward: workspace index out of range");
  SPU_REQUIRE(batch >= 1 && seq >= 1, SPU_ERR_INVALID, "forward: empty batch");
  SPU_REQUIRE(seq <= r->desc.max_pos, SPU_ERR_CAPACITY, "forward: seq exceeds max_position_embeddings");
  SPU_REQUIRE(static_cast<long long>(batch) * seq <= r->cap_rows, SPU_ERR_CAPACITY,
              "forward: batch*seq exceeds the workspace the replica was created with");
  return SPU_OK;
}

}  // namespace

// ---- CUDA-graph helpers shared with vit.cu (declared in replica.cuh) ----

void destroy_graph(CapturedGraph &g) {
  if (g.exec) cudaGraphExecDestroy(g.exec);
  if (g.graph) cudaGraphDestroy(g.graph);
  g.exec = nullptr;
  g.graph = nullptr;
}

// Replace 8-byte parameters of one kernel node in the instantiated graph (and in its template, so the next
// read-back sees what the exec holds).  Launches already enqueued keep their parameters.
spu_status graph_patch_node(CapturedGraph &g, cudaGraphNode_t node, int n_args, std::initializer_list<std::pair<int, const void *>> changes) {
  cudaKernelNodeParams p{};
  SPU_CUDA_TRY(cudaGraphKernelNodeGetParams(node, &p));
  void *args[24];
  const void *vals[24];
  SPU_REQUIRE(n_args <= 24, SPU_ERR_INVALID, "graph patch: too many kernel parameters");
  for (int i = 0; i < n_args; ++i) args[i] = p.kernelParams[i];
  for (const auto &c : changes) {
    vals[c.first] = c.second;
    args[c.first] = &vals[c.first];
  }
  p.kernelParams = args;
  SPU_CUDA_TRY(cudaGraphExecKernelNodeSetParams(g.exec, node, &p));
  SPU_CUDA_TRY(cudaGraphKernelNodeSetParams(node, &p));
  return SPU_OK;
}

// Find the single root and the single leaf of a captured chain; both must be kernel nodes.
bool graph_chain_ends(cudaGraph_t graph, cudaGraphNode_t *root, cudaGraphNode_t *leaf, cudaKernelNodeParams *p0,
                cudaKernelNodeParams *p1) {
  size_t n_nodes = 0, n_roots = 0;
  if (cudaGraphGetNodes(graph, nullptr, &n_nodes) != cudaSuccess || cudaGraphGetRootNodes(graph, nullptr, &n_roots) != cudaSuccess ||
      n_roots != 1 || n_nodes < 2) {
    cudaGetLastError();
    return false;
  }
  std::vector<cudaGraphNode_t> nodes(n_nodes);
  n_roots = 1;
  if (cudaGraphGetNodes(graph, nodes.data(), &n_nodes) != cudaSuccess || cudaGraphGetRootNodes(graph, root, &n_roots) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  int n_leaves = 0;
  for (cudaGraphNode_t n : nodes) {
    size_t deps = 0;
    if (cudaGraphNodeGetDependentNodes(n, nullptr, &deps) != cudaSuccess) {
      cudaGetLastError();
      return false;
    }
    if (deps == 0) {
      *leaf = n;
      ++n_leaves;
    }
  }
  if (n_leaves != 1 || *leaf == *root) return false;
  cudaGraphNodeType t0, t1;
  if (cudaGraphNodeGetType(*root, &t0) != cudaSuccess || cudaGraphNodeGetType(*leaf, &t1) != cudaSuccess ||
      t0 != cudaGraphNodeTypeKernel || t1 != cudaGraphNodeTypeKernel ||
      cudaGraphKernelNodeGetParams(*root, p0) != cudaSuccess || cudaGraphKernelNodeGetParams(*leaf, p1) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return p0->kernelParams != nullptr && p1->kernelParams != nullptr;
}


namespace {

void drop_graphs(Workspace &ws) {
  for (CapturedGraph &g : ws.graphs) destroy_graph(g);
  for (CapturedGraph &g : ws.req_graphs) destroy_graph(g);
  ws.graphs.clear();
  ws.seen.clear();
  ws.req_graphs.clear();
  ws.req_seen.clear();
}

// Capture one forward on `s` into an instantiated graph and find the two nodes that hold the I/O pointers.
spu_status capture_forward(spu_replica *r, Workspace &ws, const int64_t *ids, const int64_t *mask, int batch, int seq,
                           float *out, cudaStream_t s, CapturedGraph *cg) {
  SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  spu_status st = enqueue_forward(r, ws, ids, mask, batch, seq, out, s);
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(s, &graph);
  if (st != SPU_OK) {
    if (graph) cudaGraphDestroy(graph);
    return st;
  }
  SPU_CUDA_TRY(e);
  cg->key = GraphKey{ids, mask, out, batch, seq};
  cg->graph = graph;
  e = cudaGraphInstantiate(&cg->exec, graph, 0);
  if (e != cudaSuccess) {
    cudaGraphDestroy(graph);
    cg->graph = nullptr;
    SPU_CUDA_TRY(e);
  }
  // root = embedding kernel (ids, mask = parameters 0, 1), leaf = final LayerNorm (out_f32 = parameter 4, or
  // out_bf16 = parameter 5 under SPU_OUTPUT_BF16): anything else (LayerNorm-fused epilogues keep `out` inside a
  // tensor map) leaves the graph bound to its pointers
  cg->patchable = false;
  cudaGraphNode_t root = nullptr, leaf = nullptr;
  cudaKernelNodeParams p0{}, p1{};
  if (!graph_chain_ends(graph, &root, &leaf, &p0, &p1)) return SPU_OK;
  if (p0.func != bert_embed_kernel_entry() || p1.func != layernorm_kernel_entry()) return SPU_OK;
  const int out_arg = (r->flags & SPU_OUTPUT_BF16) ? kLayerNormArgOutBf16 : kLayerNormArgOutF32;
  // the captured parameters must be the pointers we are about to treat as patchable
  if (*static_cast<const void *const *>(p0.kernelParams[kEmbedArgIds]) != ids ||
      *static_cast<const void *const *>(p0.kernelParams[kEmbedArgMask]) != mask ||
      *static_cast<const void *const *>(p1.kernelParams[out_arg]) != out)
    return SPU_OK;
  cg->first = root;
  cg->last = leaf;
  cg->out_arg = out_arg;
  cg->patchable = true;
  return SPU_OK;
}

// Point the instantiated graph at other I/O buffers (host-side only; launches already enqueued keep theirs).
spu_status retarget_graph(CapturedGraph &g, const int64_t *ids, const int64_t *mask, float *out) {
  if (g.key.ids != ids || g.key.mask != mask) {
    SPU_TRY(graph_patch_node(g, g.first, kEmbedKernelArgs, {{kEmbedArgIds, ids}, {kEmbedArgMask, mask}}));
    g.key.ids = ids;
    g.key.mask = mask;
  }
  if (g.key.out != out) {
    SPU_TRY(graph_patch_node(g, g.last, kLayerNormKernelArgs, {{g.out_arg, out}}));
    g.key.out = out;
  }
  return SPU_OK;
}

// The graph serving (batch, seq) with these pointers, retargeted if need be; nullptr when none is cached.
spu_status find_graph(Workspace &ws, const int64_t *ids, const int64_t *mask, int batch, int seq, float *out,
                      CapturedGraph **found) {
  *found = nullptr;
  const GraphKey want{ids, mask, out, batch, seq};
  for (CapturedGraph &g : ws.graphs) {
    if (!g.key.same_shape(batch, seq)) continue;
    if (!(g.key == want)) {
      if (!g.patchable) continue;
      SPU_TRY(retarget_graph(g, ids, mask, out));
    }
    g.tick = ++ws.graph_tick;
    *found = &g;
    return SPU_OK;
  }
  return SPU_OK;
}

void insert_graph(Workspace &ws, const CapturedGraph &cg) {
  if (ws.graphs.size() >= kMaxGraphs) {
    size_t lru = 0;
    for (size_t i = 1; i < ws.graphs.size(); ++i)
      if (ws.graphs[i].tick < ws.graphs[lru].tick) lru = i;
    destroy_graph(ws.graphs[lru]);
    ws.graphs.erase(ws.graphs.begin() + static_cast<long>(lru));
  }
  ws.graphs.push_back(cg);
  ws.graphs.back().tick = ++ws.graph_tick;
}

}  // namespace
}  // namespace spu

using namespace spu;

extern "C" spu_status spu_bert_replica_create(const spu_bert_desc *desc, const float *const *tensors,
                                              int32_t n_tensors, int32_t device, int32_t max_batch,
                                              int32_t max_seq, int32_t n_workspaces, uint32_t flags,
                                              spu_replica **out) {
  SPU_REQUIRE(desc && tensors && out, SPU_ERR_INVALID, "replica_create: null argument");
  *out = nullptr;
  SPU_REQUIRE(desc->layers > 0 && desc->hidden > 0 && desc->heads > 0, SPU_ERR_INVALID,
              "replica_create: bad descriptor");
  SPU_REQUIRE(n_tensors == SPU_BERT_NUM_TENSORS(desc->layers), SPU_ERR_INVALID,
              "replica_create: wrong tensor count");
  SPU_REQUIRE(desc->hidden == desc->heads * 64, SPU_ERR_UNSUPPORTED,
              "replica_create: head_dim must be 64");
  SPU_REQUIRE(desc->hidden % 64 == 0 && desc->ffn % 64 == 0 && desc->hidden <= 1024,
              SPU_ERR_UNSUPPORTED, "replica_create: hidden/ffn must be multiples of 64, hidden <= 1024");
  SPU_REQUIRE(max_batch > 0 && max_seq > 0 && max_seq <= desc->max_pos && n_workspaces > 0,
              SPU_ERR_INVALID, "replica_create: bad capacity");
  SPU_REQUIRE((flags & ~(SPU_PREC_FP8_E4M3 | 0xff00u | SPU_FORWARD_AUTO_GRAPH | SPU_REQUEST_IO | SPU_OUTPUT_BF16 | SPU_LN_FUSED)) == 0 &&
                  ((flags & SPU_PREC_FP8_E4M3) || (flags & 0xff00u) == 0),
              SPU_ERR_UNSUPPORTED, "replica_create: unsupported precision flag");
  if (flags & SPU_PREC_FP8_E4M3)
    SPU_REQUIRE(desc->hidden % 128 == 0 && desc->ffn % 128 == 0, SPU_ERR_UNSUPPORTED,
                "replica_create: FP8 needs hidden/ffn multiples of 128 (one 128-byte K tile)");
  for (int i = 0; i < n_tensors; ++i)
    SPU_REQUIRE(tensors[i] != nullptr, SPU_ERR_INVALID, "replica_create: null weight tensor");
  SPU_REQUIRE(spu_device_is_sm100(device) == 1, SPU_ERR_UNSUPPORTED,
              "replica_create: device is not an sm_100 GPU (no CPU fallback exists)");
  DeviceGuard guard(device);
  auto r = std::make_unique<spu_replica>();
  r->desc = *desc;
  r->device = device;
  r->max_batch = max_batch;
  r->max_seq = max_seq;
  r->flags = flags;
  {
    const int n8 = static_cast<int>((flags >> 8) & 0xffu);
    r->fp8_layers = (flags & SPU_PREC_FP8_E4M3) ? ((n8 == 0 || n8 > desc->layers) ? desc->layers : n8) : 0;
    r->fp8_mask.assign(desc->layers, 0);
    if ((flags & SPU_PREC_FP8_E4M3) && n8 == 0) {
      // Default FP8 plan = the largest measured plan inside BASELINE's 6e-2 gate on BERT-base (sweep of the
      // noise every single GEMM adds, profiles/r2_fp8_plan.md): per FLOP moved to FP8 the QKV projection is the
      // cheapest (its error is averaged by the softmax), the attention-output projection the dearest, and
      // the FFN pair fits the budget in the first half of the encoder.  All 48 GEMMs in FP8 sit at the format's
      // floor (7.6e-2) and are only reachable explicitly: SPU_PREC_FP8_LAYERS(layers) or spu_bert_set_fp8_mask.
      for (int l = 0; l < desc->layers; ++l)
        r->fp8_mask[l] = SPU_FP8_GEMM_QKV | (2 * l < desc->layers ? (SPU_FP8_GEMM_FFN1 | SPU_FP8_GEMM_FFN2) : 0u);
      r->fp8_layers = desc->layers;
    } else {
      for (int l = 0; l < r->fp8_layers; ++l) r->fp8_mask[l] = SPU_FP8_GEMM_ALL;
    }
  }
  r->cap_rows = ((max_batch * max_seq + 127) / 128) * 128;
  {
    // SPU_LN_FUSED flag (or SPU_LN_FUSION=1 in the environment): residual + LayerNorm ride in the epilogue of the
    // attention-output / FFN2 GEMMs (cluster of CTA pairs spanning the hidden size, row statistics over DSMEM).
    // Opt-in: a B200 keeps at most 15 such 8-CTA clusters resident, so the 16 row blocks of the benchmarked
    // b32 x s128 batch need two rounds and the separate LayerNorm kernel is faster there
    // (profiles/r2_ln_fusion.md); it pays for batches of <= 15 row blocks.
    const char *env = std::getenv("SPU_LN_FUSION");
    const bool want = (flags & SPU_LN_FUSED) != 0 || (env != nullptr && env[0] == '1');
    r->ln_fused = want && gemm_ln_supported(desc->hidden);
  }
  r->ws.resize(n_workspaces);
  spu_status st = build(r.get(), tensors);
  if (st != SPU_OK) {
    for (void *p : r->allocations) cudaFree(p);
    return st;
  }
  *out = r.release();
  return SPU_OK;
}

extern "C" spu_status spu_replica_destroy(spu_replica *r) {
  if (r == nullptr) return SPU_OK;
  DeviceGuard guard(r->device);
  cudaDeviceSynchronize();
  for (Workspace &ws : r->ws) drop_graphs(ws);
  if (r->vit != nullptr && r->vit_destroy != nullptr) r->vit_destroy(r->vit);
  for (void *p : r->allocations) cudaFree(p);
  delete r;
  return SPU_OK;
}

extern "C" spu_status spu_bert_forward_async(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                             const int64_t *mask, int32_t batch, int32_t seq,
                                             float *out, void *stream) {
  SPU_TRY(check_forward_args(r, ws_idx, ids, mask, batch, seq, out));
  Workspace &ws = r->ws[ws_idx];
  CapturedGraph *g = nullptr;
  SPU_TRY(find_graph(ws, ids, mask, batch, seq, out, &g));
  if (g != nullptr) {
    SPU_CUDA_TRY(cudaGraphLaunch(g->exec, as_stream(stream)));
    return SPU_OK;
  }
  if ((r->flags & SPU_FORWARD_AUTO_GRAPH) != 0) {
    // second sighting of a batch shape (of these very pointers when the graph cannot be retargeted): kernel
    // attributes are set and caches warm from the eager pass
    const GraphKey key{ids, mask, out, batch, seq};
    bool seen = false, shape_unpatchable = false;
    for (const CapturedGraph &c : ws.graphs) shape_unpatchable = shape_unpatchable || (c.key.same_shape(batch, seq) && !c.patchable);
    for (const GraphKey &k : ws.seen) seen = seen || (shape_unpatchable ? k == key : k.same_shape(batch, seq));
    if (seen) {
      cudaStream_t s = as_stream(stream);
      CapturedGraph cg;
      SPU_TRY(capture_forward(r, ws, ids, mask, batch, seq, out, s, &cg));
      insert_graph(ws, cg);
      SPU_CUDA_TRY(cudaGraphLaunch(cg.exec, s));
      return SPU_OK;
    }
    if (ws.seen.size() < 4 * kMaxGraphs) ws.seen.push_back(key);
  }
  return enqueue_forward(r, ws, ids, mask, batch, seq, out, as_stream(stream));
}

// ---- Forward on scratch the library itself holds (no device allocation anywhere on the path) --------------------
// The activations of one forward (residual stream, QKV, attention context, FFN intermediate, mask bias) normally
// live in a workspace that spu_bert_replica_create cudaMalloc-ed.  This variant lays the same six buffers out in
// the array below, which is part of the loaded module on every GPU from the first CUDA call of the process until
// exit.  On the host, `g_static_scratch` is only the handle cudaGetSymbolAddress understands -- never a device
// address -- so static_scratch_base() asks the runtime for the real address once per device and caches it.
static __device__ __align__(256) unsigned char g_static_scratch[SPU_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
unsigned char *g_static_base[kStaticMaxDevices] = {};  // device address of g_static_scratch, per CUDA device

// Caller has made `device` current.
spu_status static_scratch_base(int device, unsigned char **base) {
  SPU_REQUIRE(device >= 0 && device < kStaticMaxDevices, SPU_ERR_INVALID, "forward_static: device index out of range");
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_base[device] == nullptr) {
    void *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error(std::string("forward_static: cudaGetSymbolAddress(g_static_scratch): ") +
                (e != cudaSuccess ? cudaGetErrorString(e) : "null address"));
      return SPU_ERR_CUDA;
    }
    SPU_REQUIRE((reinterpret_cast<uintptr_t>(p) & 255u) == 0, SPU_ERR_CUDA, "forward_static: scratch is not 256-byte aligned");
    g_static_base[device] = static_cast<unsigned char *>(p);
  }
  *base = g_static_base[device];
  return SPU_OK;
}

// rows is a multiple of 128, so every buffer below is a whole number of 256-byte units
size_t static_scratch_need(size_t rows, size_t H, size_t F) { return rows * (14 * H + 2 * F + 4); }

template <typename T>
T *carve(unsigned char *base, size_t *off, size_t count) {
  T *p = reinterpret_cast<T *>(base + *off);
  *off += (count * sizeof(T) + 255u) & ~static_cast<size_t>(255u);
  return p;
}
}  // namespace

extern "C" spu_status spu_bert_forward_static_async(spu_replica *r, const int64_t *ids, const int64_t *mask,
                                                    int32_t batch, int32_t seq, float *out, void *stream) {
  SPU_REQUIRE(r != nullptr, SPU_ERR_INVALID, "forward_static: null replica");
  SPU_REQUIRE(r->kind == 0, SPU_ERR_INVALID, "forward_static: not a BERT replica");
  SPU_REQUIRE(ids && mask && out, SPU_ERR_INVALID, "forward_static: null buffer");
  SPU_REQUIRE((r->flags & SPU_PREC_FP8_E4M3) == 0, SPU_ERR_UNSUPPORTED,
              "forward_static: bf16 replicas only (the FP8 operand buffers are not part of the static scratch)");
  SPU_REQUIRE(batch >= 1 && seq >= 1, SPU_ERR_INVALID, "forward_static: empty batch");
  SPU_REQUIRE(seq <= r->desc.max_pos, SPU_ERR_CAPACITY, "forward_static: seq exceeds max_position_embeddings");
  const spu_bert_desc &d = r->desc;
  const size_t H = d.hidden, F = d.ffn;
  // same row rounding as a replica workspace (cap_rows): whole 128-row tiles
  const size_t rows = (static_cast<size_t>(batch) * static_cast<size_t>(seq) + 127u) / 128u * 128u;
  const size_t need = static_scratch_need(rows, H, F);
  if (need > SPU_STATIC_SCRATCH_BYTES) {
    set_error("forward_static: batch*seq = " + std::to_string(static_cast<long long>(batch) * seq) + " tokens of hidden " +
              std::to_string(d.hidden) + " / ffn " + std::to_string(d.ffn) + " need " + std::to_string(need) +
              " bytes of scratch, the static array holds SPU_STATIC_SCRATCH_BYTES = " +
              std::to_string(static_cast<size_t>(SPU_STATIC_SCRATCH_BYTES)) +
              " (use spu_bert_forward_async on a replica workspace for larger batches)");
    return SPU_ERR_CAPACITY;
  }
  DeviceGuard guard(r->device);
  unsigned char *base = nullptr;
  SPU_TRY(static_scratch_base(r->device, &base));

  // A workspace whose buffers are slices of the static array.  Its GEMM plans (host-side tensor-map encodes of the
  // weight matrices; the kernels take them by value) are rebuilt on every call: this path trades ~100 us of host
  // time for holding no per-replica state, so a destroyed replica can never leave a stale plan behind.
  Workspace ws;
  size_t off = 0;
  ws.x_f32 = carve<float>(base, &off, rows * H);
  ws.x_bf16 = carve<__nv_bfloat16>(base, &off, rows * H);
  ws.qkv = carve<__nv_bfloat16>(base, &off, rows * 3 * H);
  ws.ctx = carve<__nv_bfloat16>(base, &off, rows * H);
  ws.h = carve<__nv_bfloat16>(base, &off, rows * F);
  ws.mask_bias = carve<float>(base, &off, rows);
  SPU_REQUIRE(off == need && off <= SPU_STATIC_SCRATCH_BYTES, SPU_ERR_INVALID, "forward_static: layout does not match its size formula");
  ws.plans.resize(d.layers);
  const int R = static_cast<int>(rows);
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    LayerPlans &p = ws.plans[l];
    SPU_TRY(gemm_make_plan(&p.qkv, ws.x_bf16, R, w.w_qkv, 3 * d.hidden, d.hidden, SPU_EPI_BIAS, GemmIn::BF16));
    SPU_TRY(gemm_make_plan(&p.attn_out, ws.ctx, R, w.w_o, d.hidden, d.hidden, SPU_EPI_BIAS_RESID, GemmIn::BF16));
    SPU_TRY(gemm_make_plan(&p.ffn1, ws.x_bf16, R, w.w_1, d.ffn, d.hidden, SPU_EPI_BIAS_GELU, GemmIn::BF16));
    SPU_TRY(gemm_make_plan(&p.ffn2, ws.h, R, w.w_2, d.hidden, d.ffn, SPU_EPI_BIAS_RESID, GemmIn::BF16));
    if (r->ln_fused) {
      SPU_TRY(gemm_make_plan(&p.attn_out_ln, ws.ctx, R, w.w_o, d.hidden, d.hidden, kEpiBiasResidLN, GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&p.ffn2_ln, ws.h, R, w.w_2, d.hidden, d.ffn, kEpiBiasResidLN, GemmIn::BF16));
    }
    // the same measured tiles a replica workspace would use, so both paths launch identical kernels
    for (GemmPlan *g : {&p.qkv, &p.attn_out, &p.ffn1, &p.ffn2, &p.attn_out_ln, &p.ffn2_ln}) g->tuned = &r->tuned;
  }
  // Nothing in the forward relies on scratch that starts at zero: the embedding kernel writes x / mask_bias before
  // anything reads them and every later buffer is fully produced by the kernel ahead of its consumer, so what the
  // previous call left behind is overwritten, never accumulated into.
  return enqueue_forward(r, ws, ids, mask, batch, seq, out, as_stream(stream));
}

extern "C" int32_t spu_bert_vocab_size(const spu_replica *r) {
  return (r != nullptr && r->kind == 0) ? r->desc.vocab : 0;
}

extern "C" int32_t spu_bert_packed_supported(const spu_replica *r, int32_t batch) {
  return (r != nullptr && r->kind == 0 && (r->flags & SPU_PREC_FP8_E4M3) == 0 && !r->ln_fused && batch >= 1 &&
          batch <= kMaxPackedBatch)
             ? 1
             : 0;
}

static spu_status forward_packed(spu_replica *r, int32_t ws_idx, const int64_t *ids, const int32_t *lens, int32_t batch,
                                 int32_t seq, float *out, void *stream, bool compact_out);

extern "C" spu_status spu_bert_forward_packed_async(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                                    const int32_t *lens, int32_t batch, int32_t seq, float *out,
                                                    void *stream) {
  return forward_packed(r, ws_idx, ids, lens, batch, seq, out, stream, false);
}

extern "C" spu_status spu_bert_forward_packed_compact_async(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                                            const int32_t *lens, int32_t batch, int32_t seq,
                                                            float *out, void *stream) {
  return forward_packed(r, ws_idx, ids, lens, batch, seq, out, stream, true);
}

static spu_status forward_packed(spu_replica *r, int32_t ws_idx, const int64_t *ids, const int32_t *lens, int32_t batch,
                                 int32_t seq, float *out, void *stream, bool compact_out) {
  SPU_TRY(check_forward_args(r, ws_idx, ids, lens, batch, seq, out));
  SPU_REQUIRE((r->flags & SPU_PREC_FP8_E4M3) == 0 && !r->ln_fused, SPU_ERR_UNSUPPORTED,
              "forward_packed: bf16 replicas with the separate LayerNorm kernel only");
  SPU_REQUIRE(batch <= kMaxPackedBatch, SPU_ERR_CAPACITY, "forward_packed: at most 256 sequences per batch");
  for (int b = 0; b < batch; ++b)
    SPU_REQUIRE(lens[b] >= 1 && lens[b] <= seq, SPU_ERR_INVALID, "forward_packed: every length must be in [1, seq]");
  return enqueue_forward_packed(r, r->ws[ws_idx], ids, lens, batch, seq, out, as_stream(stream), compact_out);
}

extern "C" spu_status spu_bert_forward_requests_async(spu_replica *r, int32_t ws_idx,
                                                      const int64_t *const *src_ids,
                                                      const int64_t *const *src_mask, const int32_t *len,
                                                      int32_t n_req, int32_t seq, float *const *dsts,
                                                      int32_t full_rows, void *stream) {
  return spu_bert_forward_requests_checked_async(r, ws_idx, src_ids, src_mask, len, n_req, seq, dsts, full_rows, nullptr,
                                                 stream);
}

extern "C" spu_status spu_bert_forward_requests_checked_async(spu_replica *r, int32_t ws_idx,
                                                              const int64_t *const *src_ids,
                                                              const int64_t *const *src_mask, const int32_t *len,
                                                              int32_t n_req, int32_t seq, float *const *dsts,
                                                              int32_t full_rows, int32_t *status, void *stream) {
  SPU_REQUIRE(r != nullptr && r->kind == 0, SPU_ERR_INVALID, "forward_requests: not a BERT replica");
  SPU_REQUIRE((r->flags & SPU_REQUEST_IO) != 0, SPU_ERR_UNSUPPORTED,
              "forward_requests: replica was created without SPU_REQUEST_IO");
  SPU_REQUIRE((r->flags & SPU_OUTPUT_BF16) == 0, SPU_ERR_UNSUPPORTED,
              "forward_requests: fp32 responses only (SPU_OUTPUT_BF16 is set)");
  SPU_REQUIRE(ws_idx >= 0 && ws_idx < static_cast<int>(r->ws.size()), SPU_ERR_INVALID,
              "forward_requests: workspace index out of range");
  SPU_REQUIRE(src_ids && len && dsts, SPU_ERR_INVALID, "forward_requests: null table");
  Workspace &ws = r->ws[ws_idx];
  SPU_TRY(check_forward_args(r, ws_idx, ws.req_ids, ws.req_mask, n_req, seq, ws.req_out));
  cudaStream_t s = as_stream(stream);
  auto enqueue = [&]() -> spu_status {
    // without a status word nobody would read (or re-arm) the error word: the gather then skips the id check
    SPU_TRY(gather_pad_mask_launch(src_ids, src_mask, len, n_req, seq, 0, ws.req_ids, ws.req_mask, r->desc.vocab,
                                   status != nullptr ? ws.req_err : nullptr, s));
    SPU_TRY(enqueue_forward(r, ws, ws.req_ids, ws.req_mask, n_req, seq, ws.req_out, s));
    return scatter_requests_launch(ws.req_out, len, n_req, seq, r->desc.hidden, full_rows != 0, dsts, s,
                                   status != nullptr ? ws.req_err : nullptr, status);
  };
  // one graph per (n_req, seq, full_rows): the tables' addresses (they live in whatever slot the task uses) are
  // parameters of the first kernel (gather) and the last one (scatter), patched per launch
  GraphKey want{src_ids, len, dsts, n_req, seq};
  want.aux = src_mask;
  want.full_rows = full_rows != 0 ? 1 : 0;
  want.status = status;
  // a graph captured with the id check serves only checked calls (its gather node holds the error word) and vice versa
  auto same = [&](const GraphKey &k) {
    return k.same_shape(n_req, seq) && k.full_rows == want.full_rows && (k.status != nullptr) == (status != nullptr);
  };
  for (CapturedGraph &g : ws.req_graphs) {
    if (!same(g.key) || !g.patchable) continue;
    if (g.key.ids != want.ids || g.key.aux != want.aux || g.key.mask != want.mask) {
      // gather_pad_mask_kernel(src_ids, src_mask, len, ...)
      SPU_TRY(graph_patch_node(g, g.first, kGatherKernelArgs, {{0, want.ids}, {1, want.aux}, {2, want.mask}}));
    }
    if (g.key.out != want.out || g.key.mask != want.mask || g.key.status != want.status) {
      // scatter_requests_kernel(src, len, seq, hidden, full_rows, dsts, err, status)
      SPU_TRY(graph_patch_node(g, g.last, kScatterKernelArgs,
                               {{kScatterArgLen, want.mask}, {kScatterArgDsts, want.out}, {kScatterArgStatus, want.status}}));
    }
    g.key = want;
    g.tick = ++ws.graph_tick;
    SPU_CUDA_TRY(cudaGraphLaunch(g.exec, s));
    return SPU_OK;
  }
  if ((r->flags & SPU_FORWARD_AUTO_GRAPH) != 0 && s != nullptr) {
    bool seen = false;
    for (const GraphKey &k : ws.req_seen) seen = seen || same(k);
    if (seen) {
      SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      spu_status st = enqueue();
      cudaGraph_t graph = nullptr;
      cudaError_t e = cudaStreamEndCapture(s, &graph);
      if (st != SPU_OK) {
        if (graph) cudaGraphDestroy(graph);
        return st;
      }
      SPU_CUDA_TRY(e);
      CapturedGraph cg;
      cg.key = want;
      cg.graph = graph;
      e = cudaGraphInstantiate(&cg.exec, graph, 0);
      if (e != cudaSuccess) {
        cudaGraphDestroy(graph);
        SPU_CUDA_TRY(e);
      }
      cudaKernelNodeParams p0{}, p1{};
      cg.patchable = graph_chain_ends(graph, &cg.first, &cg.last, &p0, &p1) &&
                     p0.func == gather_pad_mask_kernel_entry() && p1.func == scatter_requests_kernel_entry();
      // the captured parameters must be the very pointers that will be treated as patchable
      auto held = [](const cudaKernelNodeParams &p, int i) { return *static_cast<const void *const *>(p.kernelParams[i]); };
      cg.patchable = cg.patchable && held(p0, 0) == want.ids && held(p0, 1) == want.aux && held(p0, 2) == want.mask &&
                     held(p1, kScatterArgLen) == want.mask && held(p1, kScatterArgDsts) == want.out &&
                     held(p1, kScatterArgStatus) == want.status;
      if (!cg.patchable) {  // cannot be retargeted: useless for a slot-cycling server
        SPU_CUDA_TRY(cudaGraphLaunch(cg.exec, s));
        cudaStreamSynchronize(s);
        destroy_graph(cg);
        return SPU_OK;
      }
      if (ws.req_graphs.size() >= kMaxGraphs) {
        size_t lru = 0;
        for (size_t i = 1; i < ws.req_graphs.size(); ++i)
          if (ws.req_graphs[i].tick < ws.req_graphs[lru].tick) lru = i;
        destroy_graph(ws.req_graphs[lru]);
        ws.req_graphs.erase(ws.req_graphs.begin() + static_cast<long>(lru));
      }
      cg.tick = ++ws.graph_tick;
      ws.req_graphs.push_back(cg);
      SPU_CUDA_TRY(cudaGraphLaunch(cg.exec, s));
      return SPU_OK;
    }
    if (ws.req_seen.size() < 4 * kMaxGraphs) ws.req_seen.push_back(want);
  }
  return enqueue();
}

extern "C" spu_status spu_bert_forward_capture(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                               const int64_t *mask, int32_t batch, int32_t seq,
                                               float *out, void *stream) {
  SPU_TRY(check_forward_args(r, ws_idx, ids, mask, batch, seq, out));
  Workspace &ws = r->ws[ws_idx];
  CapturedGraph *g = nullptr;
  SPU_TRY(find_graph(ws, ids, mask, batch, seq, out, &g));
  if (g != nullptr) return SPU_OK;
  cudaStream_t s = as_stream(stream);
  // one eager pass first: sets kernel attributes and warms caches outside the capture
  SPU_TRY(enqueue_forward(r, ws, ids, mask, batch, seq, out, s));
  SPU_CUDA_TRY(cudaStreamSynchronize(s));
  CapturedGraph cg;
  SPU_TRY(capture_forward(r, ws, ids, mask, batch, seq, out, s, &cg));
  insert_graph(ws, cg);
  return SPU_OK;
}

extern "C" spu_status spu_bert_forward_profiled(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                                const int64_t *mask, int32_t batch, int32_t seq,
                                                float *out, void *stream, float *stage_ms) {
  SPU_TRY(check_forward_args(r, ws_idx, ids, mask, batch, seq, out));
  SPU_REQUIRE(stage_ms != nullptr, SPU_ERR_INVALID, "profiled: null stage_ms");
  DeviceGuard guard(r->device);
  StageTrail trail;
  const size_t n_launch = static_cast<size_t>(spu_bert_forward_launch_count(r));
  trail.ev.resize(n_launch + 1, nullptr);
  trail.stage.resize(n_launch, 0);
  spu_status st = SPU_OK;
  for (cudaEvent_t &e : trail.ev)
    if (cudaEventCreate(&e) != cudaSuccess) st = SPU_ERR_CUDA;
  cudaStream_t s = as_stream(stream);
  if (st == SPU_OK && cudaEventRecord(trail.ev[0], s) != cudaSuccess) st = SPU_ERR_CUDA;
  if (st == SPU_OK) st = enqueue_forward(r, r->ws[ws_idx], ids, mask, batch, seq, out, s, &trail);
  if (st == SPU_OK && cudaStreamSynchronize(s) != cudaSuccess) {
    set_error(std::string("profiled forward: ") + cudaGetErrorString(cudaGetLastError()));
    st = SPU_ERR_CUDA;
  }
  if (st == SPU_OK)
    for (size_t i = 0; i < trail.n; ++i) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, trail.ev[i], trail.ev[i + 1]) == cudaSuccess)
        stage_ms[trail.stage[i]] += ms;
    }
  for (cudaEvent_t e : trail.ev)
    if (e) cudaEventDestroy(e);
  return st;
}

extern "C" spu_status spu_bert_calibrate_fp8(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                             const int64_t *mask, int32_t batch, int32_t seq, void *stream) {
  float dummy = 0.f;
  SPU_TRY(check_forward_args(r, ws_idx, ids, mask, batch, seq, &dummy));
  SPU_REQUIRE(r->flags & SPU_PREC_FP8_E4M3, SPU_ERR_INVALID, "calibrate: not an FP8 replica");
  DeviceGuard guard(r->device);
  cudaStream_t s = as_stream(stream);
  SPU_TRY(enqueue_calibration(r, r->ws[ws_idx], ids, mask, batch, seq, s));
  std::vector<float> am(static_cast<size_t>(r->desc.layers) * 4);
  SPU_CUDA_TRY(cudaMemcpyAsync(am.data(), r->amax, am.size() * sizeof(float), cudaMemcpyDeviceToHost, s));
  SPU_CUDA_TRY(cudaStreamSynchronize(s));
  for (int l = 0; l < r->desc.layers; ++l) {
    LayerWeights &w = r->layers[l];
    auto scale = [](float a) { return (a > 1e-12f ? a : 1.f) / kE4m3Max; };
    w.sa_x = scale(am[4 * l + 0]);
    w.sa_ctx = scale(am[4 * l + 1]);
    w.sa_x1 = scale(am[4 * l + 2]);
    w.sa_h = scale(am[4 * l + 3]);
  }
  r->fp8_calibrated = true;
  // captured graphs bake the scales in: drop them
  for (Workspace &ws : r->ws) drop_graphs(ws);
  return SPU_OK;
}

extern "C" spu_status spu_bert_set_fp8_mask(spu_replica *r, const uint8_t *layer_masks, int32_t n_layers) {
  SPU_REQUIRE(r != nullptr && r->kind == 0 && (r->flags & SPU_PREC_FP8_E4M3), SPU_ERR_INVALID,
              "set_fp8_mask: not an FP8 BERT replica");
  SPU_REQUIRE(layer_masks != nullptr && n_layers == r->desc.layers, SPU_ERR_INVALID,
              "set_fp8_mask: one mask byte per encoder layer is required");
  for (int l = 0; l < n_layers; ++l)
    SPU_REQUIRE((layer_masks[l] & ~SPU_FP8_GEMM_ALL) == 0, SPU_ERR_INVALID, "set_fp8_mask: unknown GEMM bit");
  DeviceGuard guard(r->device);
  SPU_CUDA_TRY(cudaDeviceSynchronize());  // start-up only: no forward of the old plan may be in flight
  r->fp8_mask.assign(layer_masks, layer_masks + n_layers);
  for (Workspace &ws : r->ws) drop_graphs(ws);  // captured graphs bake the plan in
  return SPU_OK;
}

extern "C" int32_t spu_bert_get_fp8_mask(const spu_replica *r, uint8_t *out, int32_t cap) {
  if (r == nullptr || r->kind != 0 || !(r->flags & SPU_PREC_FP8_E4M3)) return 0;
  for (int l = 0; l < r->desc.layers && out != nullptr && l < cap; ++l) out[l] = r->fp8_mask[l];
  return r->desc.layers;
}

extern "C" int32_t spu_bert_fp8_scales(const spu_replica *r, float *out, int32_t cap) {
  if (r == nullptr || r->kind != 0 || !(r->flags & SPU_PREC_FP8_E4M3)) return 0;
  const int n = r->desc.layers * 8;
  for (int l = 0; l < r->desc.layers && out != nullptr; ++l) {
    const LayerWeights &w = r->layers[l];
    const float v[8] = {w.sa_x, w.sa_ctx, w.sa_x1, w.sa_h, w.sw_qkv, w.sw_o, w.sw_1, w.sw_2};
    for (int i = 0; i < 8; ++i)
      if (l * 8 + i < cap) out[l * 8 + i] = v[i];
  }
  return n;
}

extern "C" spu_status spu_gemm_fp8_async(const void *A, const void *W, const float *bias, const float *resid,
                                         void *out, int32_t M, int32_t N, int32_t K, int32_t epilogue,
                                         float acc_scale, float out_inv_scale, void *stream) {
  SPU_REQUIRE(epilogue == SPU_EPI_BIAS || epilogue == SPU_EPI_BIAS_RESID || epilogue == SPU_EPI_BIAS_GELU_E4M3,
              SPU_ERR_INVALID, "fp8 gemm: unknown epilogue");
  SPU_REQUIRE(M > 0, SPU_ERR_INVALID, "gemm: empty shape");
  GemmPlan plan;
  SPU_TRY(gemm_make_plan(&plan, A, M, W, N, K, epilogue, GemmIn::FP8_E4M3));
  GemmEpilogueArgs e{};
  e.bias = bias;
  e.resid = resid;
  e.out = out;
  e.acc_scale = acc_scale;
  e.out_inv_scale = out_inv_scale;
  return gemm_launch(plan, M, e, as_stream(stream));
}

extern "C" spu_status spu_bert_stage_chain_ms(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                              const int64_t *mask, int32_t batch, int32_t seq, float *out,
                                              uint32_t stage_mask, int32_t reps, void *stream, float *ms_per_forward) {
  SPU_TRY(check_forward_args(r, ws_idx, ids, mask, batch, seq, out));
  SPU_REQUIRE(ms_per_forward != nullptr && reps > 0, SPU_ERR_INVALID, "stage_chain: bad argument");
  DeviceGuard guard(r->device);
  cudaStream_t s = as_stream(stream);
  Workspace &ws = r->ws[ws_idx];
  // capture `reps` forwards' worth of the selected kernel classes into one graph so the
  // chain runs back to back (programmatic dependent launch in effect, no host gaps)
  SPU_TRY(enqueue_forward(r, ws, ids, mask, batch, seq, out, s, nullptr, stage_mask));  // warm-up, attributes
  SPU_CUDA_TRY(cudaStreamSynchronize(s));
  SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  spu_status st = SPU_OK;
  for (int i = 0; i < reps && st == SPU_OK; ++i)
    st = enqueue_forward(r, ws, ids, mask, batch, seq, out, s, nullptr, stage_mask);
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(s, &graph);
  if (st != SPU_OK) {
    if (graph) cudaGraphDestroy(graph);
    return st;
  }
  SPU_CUDA_TRY(e);
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  SPU_CUDA_TRY(e);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaGraphLaunch(exec, s);  // warm
  cudaEventRecord(e0, s);
  cudaGraphLaunch(exec, s);
  cudaEventRecord(e1, s);
  e = cudaStreamSynchronize(s);
  float ms = 0.f;
  if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaGraphExecDestroy(exec);
  SPU_CUDA_TRY(e);
  *ms_per_forward = ms / static_cast<float>(reps);
  return SPU_OK;
}

extern "C" uint64_t spu_bert_output_bytes(const spu_replica *r, int32_t batch, int32_t seq) {
  if (r == nullptr || r->kind != 0 || batch <= 0 || seq <= 0) return 0;
  return static_cast<uint64_t>(batch) * static_cast<uint64_t>(seq) * static_cast<uint64_t>(r->desc.hidden) *
         ((r->flags & SPU_OUTPUT_BF16) ? 2u : 4u);
}

namespace {
// Microseconds per launch of `plan` forced to tile `cfg`: `reps` launches back to back in one CUDA graph
// (programmatic dependent launch in effect, as in the captured forward), best of three replays.
spu_status time_tile(const GemmPlan &plan, GemmConfig cfg, int M, const GemmEpilogueArgs &e, cudaStream_t s,
                     int reps, float *us) {
  GemmPlan p = plan;
  p.force_cg = cfg.half ? 3 : cfg.cta_group;
  p.force_bn = cfg.block_n;
  p.tuned = nullptr;
  SPU_TRY(gemm_launch(p, M, e, s));  // kernel attributes, caches
  SPU_CUDA_TRY(cudaStreamSynchronize(s));
  SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  spu_status st = SPU_OK;
  for (int i = 0; i < reps && st == SPU_OK; ++i) st = gemm_launch(p, M, e, s);
  cudaGraph_t graph = nullptr;
  cudaError_t err = cudaStreamEndCapture(s, &graph);
  if (st != SPU_OK) {
    if (graph) cudaGraphDestroy(graph);
    return st;
  }
  SPU_CUDA_TRY(err);
  cudaGraphExec_t exec = nullptr;
  err = cudaGraphInstantiate(&exec, graph, 0);
  cudaGraphDestroy(graph);
  SPU_CUDA_TRY(err);
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  cudaGraphLaunch(exec, s);
  for (int k = 0; k < 3 && err == cudaSuccess; ++k) {
    cudaEventRecord(e0, s);
    cudaGraphLaunch(exec, s);
    cudaEventRecord(e1, s);
    err = cudaStreamSynchronize(s);
    float ms = 0.f;
    if (err == cudaSuccess && cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) best = std::fmin(best, ms);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaGraphExecDestroy(exec);
  SPU_CUDA_TRY(err);
  *us = best * 1e3f / static_cast<float>(reps);
  return SPU_OK;
}
}  // namespace

// Warm-up autotune (the reference's hook is the fixed-worker warm-up pass,
// src/core/warmup.cpp:226-299): for the GEMM problems a forward of (batch, seq) launches, time every
// (cta_group, BLOCK_N) tile the kernel is built for on the replica's own weights and workspace, and let the winner
// replace the static cost model's choice for that (N, K, epilogue, format, M).  Start-up only: no forward may be in
// flight; captured graphs (which bake tiles in) are dropped.
extern "C" spu_status spu_bert_autotune(spu_replica *r, int32_t ws_idx, int32_t batch, int32_t seq, void *stream,
                                        int32_t *n_changed) {
  float dummy = 0.f;
  SPU_TRY(check_forward_args(r, ws_idx, &dummy, &dummy, batch, seq, &dummy));
  DeviceGuard guard(r->device);
  // the measurements are captured into CUDA graphs, which the legacy default stream cannot do: without a
  // caller stream, use a private one
  struct OwnStream {
    cudaStream_t s = nullptr;
    ~OwnStream() {
      if (s) cudaStreamDestroy(s);
    }
  } own;
  cudaStream_t s = as_stream(stream);
  if (s == nullptr) {
    SPU_CUDA_TRY(cudaStreamCreateWithFlags(&own.s, cudaStreamNonBlocking));
    s = own.s;
  }
  SPU_CUDA_TRY(cudaDeviceSynchronize());
  Workspace &ws = r->ws[ws_idx];
  const spu_bert_desc &d = r->desc;
  const int M = batch * seq;
  struct Problem {
    const GemmPlan *plan;
    GemmEpilogueArgs e;
  };
  std::vector<Problem> problems;
  auto add = [&](const GemmPlan &p, const float *bias, const float *resid, void *out) {
    for (const Problem &q : problems)
      if (q.plan->N == p.N && q.plan->K == p.K && q.plan->epilogue == p.epilogue && q.plan->in == p.in) return;
    GemmEpilogueArgs e{};
    e.bias = bias;
    e.resid = resid;
    e.out = out;
    e.acc_scale = 1.f;
    e.out_inv_scale = 1.f;
    problems.push_back(Problem{&p, e});
  };
  const bool fp8 = (r->flags & SPU_PREC_FP8_E4M3) != 0;
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = r->layers[l];
    const LayerPlans &p16 = ws.plans[l];
    const uint8_t m = fp8 ? r->fp8_mask[l] : 0;
    const LayerPlans *p8 = fp8 ? &ws.plans8[l] : nullptr;
    const bool q_qkv = m & SPU_FP8_GEMM_QKV, q_out = m & SPU_FP8_GEMM_OUT, q_f1 = m & SPU_FP8_GEMM_FFN1,
               q_f2 = m & SPU_FP8_GEMM_FFN2;
    add(q_qkv ? p8->qkv : p16.qkv, w.b_qkv, nullptr, ws.qkv);
    if (!r->ln_fused) add(q_out ? p8->attn_out : p16.attn_out, w.b_o, ws.x_f32, ws.x_f32);
    add(q_f1 ? (q_f2 ? p8->ffn1 : p8->ffn1_q2b) : (q_f2 ? p8->ffn1_b2q : p16.ffn1), w.b_1, nullptr,
        q_f2 ? static_cast<void *>(ws.h_q) : ws.h);
    if (!r->ln_fused) add(q_f2 ? p8->ffn2 : p16.ffn2, w.b_2, ws.x_f32, ws.x_f32);
  }
  int changed = 0;
  for (const Problem &pb : problems) {
    const GemmPlan &p = *pb.plan;
    GemmConfig cands[16];
    const int n = gemm_candidates(p.N, p.epilogue, cands, 16);
    GemmConfig best{0, 0, 0};
    float best_us = 1e30f;
    for (int i = 0; i < n; ++i) {
      float us = 0.f;
      if (time_tile(p, cands[i], M, pb.e, s, 20, &us) != SPU_OK) {
        cudaGetLastError();  // a tile this problem cannot use (no weight map, footprint): skip it
        continue;
      }
      if (us < best_us) {
        best_us = us;
        best = cands[i];
      }
    }
    SPU_REQUIRE(best.block_n != 0, SPU_ERR_CUDA, "autotune: no candidate tile could be timed");
    const GemmConfig model = choose_config(M, p.N, p.K, p.in == GemmIn::FP8_E4M3 ? 1 : 2);
    changed += (model.cta_group != best.cta_group || model.block_n != best.block_n) ? 1 : 0;
    TunedTable::Entry entry{p.N, p.K, p.epilogue, static_cast<int>(p.in), M, best, best_us};
    bool replaced = false;
    for (TunedTable::Entry &e : r->tuned.entries)
      if (e.M == M && e.N == p.N && e.K == p.K && e.epilogue == p.epilogue && e.in == entry.in) {
        e = entry;
        replaced = true;
      }
    if (!replaced) r->tuned.entries.push_back(entry);
  }
  for (Workspace &w2 : r->ws) drop_graphs(w2);
  if (n_changed != nullptr) *n_changed = changed;
  return SPU_OK;
}

// Rows of {N, K, epilogue, operand format (0 bf16, 1 e4m3), M, cta_group, BLOCK_N, nanoseconds per launch}.
extern "C" int32_t spu_bert_tuned_configs(const spu_replica *r, int32_t *out, int32_t cap_rows) {
  if (r == nullptr) return 0;
  const int n = static_cast<int>(r->tuned.entries.size());
  for (int i = 0; i < n && i < cap_rows && out != nullptr; ++i) {
    const TunedTable::Entry &e = r->tuned.entries[static_cast<size_t>(i)];
    const int row[8] = {e.N, e.K, e.epilogue, e.in, e.M, e.cfg.cta_group, e.cfg.block_n, static_cast<int>(e.us * 1e3f)};
    for (int k = 0; k < 8; ++k) out[i * 8 + k] = row[k];
  }
  return n;
}

extern "C" int32_t spu_bert_graph_count(const spu_replica *r, int32_t ws_idx) {
  if (r == nullptr || r->kind != 0 || ws_idx < 0 || ws_idx >= static_cast<int>(r->ws.size())) return -1;
  return static_cast<int32_t>(r->ws[ws_idx].graphs.size());
}

extern "C" int32_t spu_bert_forward_launch_count(const spu_replica *r) {
  return r == nullptr ? 0 : 1 + (r->ln_fused ? 5 : 7) * r->desc.layers;
}

extern "C" spu_status spu_gemm_bf16_cfg_async(const void *A, const void *W, const float *bias,
                                              const float *resid, void *out, int32_t M, int32_t N,
                                              int32_t K, int32_t epilogue, int32_t cta_group,
                                              int32_t block_n, void *stream) {
  SPU_REQUIRE(epilogue >= SPU_EPI_BIAS && epilogue <= SPU_EPI_BIAS_RESID, SPU_ERR_INVALID,
              "gemm: unknown epilogue");
  SPU_REQUIRE(M > 0, SPU_ERR_INVALID, "gemm: empty shape");
  GemmPlan plan;
  SPU_TRY(gemm_make_plan(&plan, A, M, W, N, K, epilogue, GemmIn::BF16, cta_group, block_n));
  GemmEpilogueArgs e{};
  e.bias = bias;
  e.resid = resid;
  e.out = out;
  e.acc_scale = 1.f;
  return gemm_launch(plan, M, e, as_stream(stream));
}

extern "C" spu_status spu_gemm_ln_bf16_async(const void *A, const void *W, const float *bias, const float *resid,
                                             const float *gamma, const float *beta, float eps, float *out_f32,
                                             void *out_bf16, int32_t M, int32_t N, int32_t K, int32_t cta_group,
                                             int32_t block_n, void *stream) {
  SPU_REQUIRE(M > 0, SPU_ERR_INVALID, "gemm: empty shape");
  GemmPlan plan;
  SPU_TRY(gemm_make_plan(&plan, A, M, W, N, K, kEpiBiasResidLN, GemmIn::BF16, cta_group, block_n));
  GemmEpilogueArgs e{};
  e.bias = bias;
  e.resid = resid;
  e.out = out_f32;
  e.out2 = out_bf16;
  e.gamma = gamma;
  e.beta = beta;
  e.ln_eps = eps;
  e.acc_scale = 1.f;
  return gemm_launch(plan, M, e, as_stream(stream));
}

extern "C" spu_status spu_gemm_bf16_async(const void *A, const void *W, const float *bias,
                                          const float *resid, void *out, int32_t M, int32_t N,
                                          int32_t K, int32_t epilogue, void *stream) {
  return spu_gemm_bf16_cfg_async(A, W, bias, resid, out, M, N, K, epilogue, 0, 0, stream);
}

extern "C" spu_status spu_attention_async(const void *qkv, const float *mask_bias, void *ctx,
                                          int32_t batch, int32_t seq, int32_t heads, void *stream) {
  return attention_launch(qkv, mask_bias, ctx, batch, seq, heads, as_stream(stream));
}

extern "C" spu_status spu_attention_cfg_async(const void *qkv, const float *mask_bias, void *ctx,
                                              int32_t batch, int32_t seq, int32_t heads, int32_t impl,
                                              int32_t out_e4m3, float q_inv_scale, void *stream) {
  if (impl == 1)
    return attention_tc_launch(qkv, mask_bias, ctx, batch, seq, heads, as_stream(stream), out_e4m3 != 0, q_inv_scale);
  if (impl == 2)
    return attention_mma_launch(qkv, mask_bias, ctx, batch, seq, heads, as_stream(stream), out_e4m3 != 0, q_inv_scale);
  return attention_launch(qkv, mask_bias, ctx, batch, seq, heads, as_stream(stream), out_e4m3 != 0, q_inv_scale);
}

#ifdef SPU_DEBUG_HOOKS  // include/spu_b200_debug.h: measurement builds only
extern "C" void spu_debug_gemm_trace(void *device_buf, int32_t cta) {
  gemm_set_trace(static_cast<unsigned long long *>(device_buf), cta);
}

extern "C" void spu_debug_gemm_mode(int32_t mode) { gemm_set_debug_mode(mode); }
#endif

extern "C" const char *spu_last_error(void) { return g_last_error.c_str(); }
extern "C" int32_t spu_abi_version(void) { return SPU_ABI_VERSION; }

extern "C" int32_t spu_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int32_t spu_device_is_sm100(int32_t device) {
  int major = 0;
  if (device < 0 || device >= spu_device_count()) return 0;
  if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return major == 10 ? 1 : 0;
}
