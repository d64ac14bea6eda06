// ViT encoder forward (torchvision VisionTransformer graph, pre-LN blocks) on the same
// kernels as BERT: patch embedding as a tcgen05 GEMM over im2col'ed patches, fused
// attention without a mask, bias+GELU / bias+residual GEMM epilogues, warp-per-row
// LayerNorm.  Replaces model->forward of the scripted vit_l_16
// (/root/reference/models/import_vit.py:10-13; models/vit_l_16.yml:13-16) inside
// run_inference (/root/reference/src/core/starpu_setup.cpp:594-624).
// Oracle: oracle/vit_oracle.py:vit_forward.
#include "replica.cuh"

namespace spu {

struct VitLayerPlans {
  GemmPlan qkv, attn_out, mlp1, mlp2;
};

struct VitWorkspace {
  __nv_bfloat16 *patches = nullptr;  // [B*np, C*P*P]   im2col of the pixels
  float *patch_out = nullptr;        // [B*np, H]       conv_proj output + bias + pos[1+p]
  float *x_f32 = nullptr;            // [B*S, H]        residual stream
  __nv_bfloat16 *y_bf16 = nullptr;   // [B*S, H]        LayerNorm output = GEMM operand
  __nv_bfloat16 *qkv = nullptr, *ctx = nullptr, *h = nullptr;
  __nv_bfloat16 *cls_bf16 = nullptr;  // [B, H]          final LN of the class tokens
  float *logits_pad = nullptr;        // [B, classes_pad]
  GemmPlan patch_plan, head_plan;
  std::vector<VitLayerPlans> plans;
  // SPU_FORWARD_AUTO_GRAPH: one captured forward per batch size; pixels / logits are parameters of the first
  // (im2col) and last (logit slice) kernel and are patched per launch (key.ids = pixels, key.out = logits)
  std::vector<CapturedGraph> graphs;
  std::vector<int> seen_batches;
  unsigned long long graph_tick = 0;
};

struct VitState {
  spu_vit_desc d{};
  int n_patches = 0, seq = 0, patch_dim = 0, classes_pad = 0;
  __nv_bfloat16 *w_patch = nullptr, *w_head = nullptr;
  float *b_patch = nullptr, *b_head = nullptr, *cls_pos0 = nullptr, *pos_tiled = nullptr;
  float *lnf_g = nullptr, *lnf_b = nullptr, *zeros = nullptr;
  std::vector<LayerWeights> layers;
  std::vector<VitWorkspace> ws;
};

namespace {

// fp32 NCHW pixels -> bf16 patch rows.  One thread per float4 of the image: reads are fully
// coalesced (consecutive threads walk an image row), every 4 threads fill one 32-byte sector
// of the patch matrix.  Row = b*np + py*nw + px, column = c*P*P + i*P + j.
__global__ void __launch_bounds__(256)
vit_im2col_kernel(const float *__restrict__ pixels, __nv_bfloat16 *__restrict__ patches, int batch,
                  int channels, int image, int patch) {
  const int w4 = image >> 2;
  const long long total = static_cast<long long>(batch) * channels * image * w4;
  const int nw = image / patch;
  const int pp = patch * patch;
  const int patch_dim = channels * pp;
  for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int x4 = static_cast<int>(idx % w4);
    long long t = idx / w4;
    const int y = static_cast<int>(t % image);
    t /= image;
    const int c = static_cast<int>(t % channels);
    const int b = static_cast<int>(t / channels);
    const float4 v = __ldg(reinterpret_cast<const float4 *>(pixels) + idx);
    const int x = x4 << 2;
    const int py = y / patch, i = y - py * patch;
    const int px = x / patch, j = x - px * patch;
    const size_t row = static_cast<size_t>(b) * nw * nw + static_cast<size_t>(py) * nw + px;
    const size_t col = static_cast<size_t>(c) * pp + static_cast<size_t>(i) * patch + j;
    *reinterpret_cast<uint2 *>(patches + row * patch_dim + col) =
        make_uint2(pack_bf16(v.x, v.y), pack_bf16(v.z, v.w));
  }
}

constexpr int kRowThreads = 256;
constexpr int kMaxChunksPerLane = 8;

// Token assembly + first LayerNorm: x[b,0] = class_token + pos[0]; x[b,1+p] = patch_out[b,p]
// (bias and pos already added by the GEMM epilogue); y = LN1_layer0(x) as bf16.
__global__ void __launch_bounds__(kRowThreads)
vit_assemble_ln_kernel(const float *__restrict__ patch_out, const float *__restrict__ cls_pos0,
                       const float *__restrict__ gamma, const float *__restrict__ beta, float eps,
                       float *__restrict__ x_f32, __nv_bfloat16 *__restrict__ y_bf16, int rows, int seq,
                       int hidden) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int row = blockIdx.x * (kRowThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunk = hidden >> 2;
  const int nchunk_lane = (nchunk - lane + 31) >> 5;
  const int b = row / seq, t = row - b * seq;
  const float4 *src = t == 0 ? reinterpret_cast<const float4 *>(cls_pos0)
                             : reinterpret_cast<const float4 *>(
                                   patch_out + (static_cast<size_t>(b) * (seq - 1) + (t - 1)) * hidden);
  float4 v[kMaxChunksPerLane];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) {
      v[i] = src[lane + i * 32];
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  const float mean = warp_sum(s) / static_cast<float>(hidden);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) {
      const float a = v[i].x - mean, bb = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + bb * bb) + (c * c + d * d);
    }
  const float rstd = rsqrtf(warp_sum(q) / static_cast<float>(hidden) + eps);
  const size_t off = static_cast<size_t>(row) * hidden;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i) {
    if (i >= nchunk_lane) break;
    const int c = lane + i * 32;
    reinterpret_cast<float4 *>(x_f32 + off)[c] = v[i];
    const float4 g = __ldg(reinterpret_cast<const float4 *>(gamma) + c);
    const float4 be = __ldg(reinterpret_cast<const float4 *>(beta) + c);
    reinterpret_cast<uint2 *>(y_bf16 + off)[c] =
        make_uint2(pack_bf16((v[i].x - mean) * rstd * g.x + be.x, (v[i].y - mean) * rstd * g.y + be.y),
                   pack_bf16((v[i].z - mean) * rstd * g.z + be.z, (v[i].w - mean) * rstd * g.w + be.w));
  }
}

// logits[b, 0:classes) = logits_pad[b, 0:classes)
__global__ void __launch_bounds__(256)
vit_slice_logits_kernel(const float *__restrict__ padded, float *__restrict__ logits, int batch, int classes,
                        int classes_pad) {
  const int total = batch * classes;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / classes, c = i - b * classes;
    logits[i] = padded[static_cast<size_t>(b) * classes_pad + c];
  }
}

// the two kernels whose parameters spu_vit_forward_async patches in an instantiated graph
constexpr int kIm2colKernelArgs = 6, kIm2colArgPixels = 0, kSliceLogitsKernelArgs = 5, kSliceLogitsArgLogits = 1;
SPU_PIN_KERNEL_ARGS(vit_im2col_kernel, kIm2colKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(vit_im2col_kernel, kIm2colArgPixels);
SPU_PIN_KERNEL_ARGS(vit_slice_logits_kernel, kSliceLogitsKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(vit_slice_logits_kernel, kSliceLogitsArgLogits);

void vit_destroy(VitState *v) {
  if (v != nullptr)
    for (VitWorkspace &ws : v->ws)
      for (CapturedGraph &g : ws.graphs) destroy_graph(g);
  delete v;
}

spu_status vit_build(spu_replica *r, VitState *v, const float *const *t, int max_batch, int n_workspaces) {
  const spu_vit_desc &d = v->d;
  const size_t H = d.hidden, M = d.mlp, K = v->patch_dim, S = v->seq, NP = v->n_patches;
  SPU_TRY(upload_bf16_stacked(r, &v->w_patch, {t[0]}, H * K));
  SPU_TRY(upload_f32(r, &v->b_patch, t[1], H));
  {  // class_token + pos[0]; pos[1..] tiled over the batch as the patch GEMM's residual operand
    std::vector<float> c0(H);
    for (size_t i = 0; i < H; ++i) c0[i] = t[2][i] + t[3][i];
    SPU_TRY(upload_f32(r, &v->cls_pos0, c0.data(), H));
    std::vector<float> tiled(static_cast<size_t>(max_batch) * NP * H);
    for (int b = 0; b < max_batch; ++b)
      std::memcpy(tiled.data() + static_cast<size_t>(b) * NP * H, t[3] + H, NP * H * sizeof(float));
    SPU_TRY(upload_f32(r, &v->pos_tiled, tiled.data(), tiled.size()));
  }
  v->layers.resize(d.layers);
  for (int l = 0; l < d.layers; ++l) {
    const float *const *p = t + 4 + 12 * l;
    LayerWeights &w = v->layers[l];
    SPU_TRY(upload_f32(r, &w.ln1_g, p[0], H));
    SPU_TRY(upload_f32(r, &w.ln1_b, p[1], H));
    SPU_TRY(upload_bf16_stacked(r, &w.w_qkv, {p[2]}, 3 * H * H));
    SPU_TRY(upload_f32(r, &w.b_qkv, p[3], 3 * H));
    SPU_TRY(upload_bf16_stacked(r, &w.w_o, {p[4]}, H * H));
    SPU_TRY(upload_f32(r, &w.b_o, p[5], H));
    SPU_TRY(upload_f32(r, &w.ln2_g, p[6], H));
    SPU_TRY(upload_f32(r, &w.ln2_b, p[7], H));
    SPU_TRY(upload_bf16_stacked(r, &w.w_1, {p[8]}, M * H));
    SPU_TRY(upload_f32(r, &w.b_1, p[9], M));
    SPU_TRY(upload_bf16_stacked(r, &w.w_2, {p[10]}, H * M));
    SPU_TRY(upload_f32(r, &w.b_2, p[11], H));
  }
  const float *const *tail = t + 4 + 12 * d.layers;
  SPU_TRY(upload_f32(r, &v->lnf_g, tail[0], H));
  SPU_TRY(upload_f32(r, &v->lnf_b, tail[1], H));
  {  // head padded to a multiple of 64 output columns with zero rows
    const size_t CP = v->classes_pad;
    std::vector<float> wh(CP * H, 0.f), bh(CP, 0.f);
    std::memcpy(wh.data(), tail[2], static_cast<size_t>(d.classes) * H * sizeof(float));
    std::memcpy(bh.data(), tail[3], static_cast<size_t>(d.classes) * sizeof(float));
    SPU_TRY(upload_bf16_stacked(r, &v->w_head, {wh.data()}, CP * H));
    SPU_TRY(upload_f32(r, &v->b_head, bh.data(), CP));
    SPU_TRY(dev_alloc(r, &v->zeros, static_cast<size_t>(max_batch) * CP));
    SPU_CUDA_TRY(cudaMemset(v->zeros, 0, static_cast<size_t>(max_batch) * CP * sizeof(float)));
  }
  const size_t rows = r->cap_rows;                                  // >= max_batch * S, multiple of 128
  const size_t prow = ((static_cast<size_t>(max_batch) * NP + 127) / 128) * 128;
  const size_t brow = ((static_cast<size_t>(max_batch) + 127) / 128) * 128;
  v->ws.resize(n_workspaces);
  for (VitWorkspace &ws : v->ws) {
    SPU_TRY(dev_alloc(r, &ws.patches, prow * K));
    SPU_TRY(dev_alloc(r, &ws.patch_out, prow * H));
    SPU_TRY(dev_alloc(r, &ws.x_f32, rows * H));
    SPU_TRY(dev_alloc(r, &ws.y_bf16, rows * H));
    SPU_TRY(dev_alloc(r, &ws.qkv, rows * 3 * H));
    SPU_TRY(dev_alloc(r, &ws.ctx, rows * H));
    SPU_TRY(dev_alloc(r, &ws.h, rows * M));
    SPU_TRY(dev_alloc(r, &ws.cls_bf16, brow * H));
    SPU_TRY(dev_alloc(r, &ws.logits_pad, brow * v->classes_pad));
    SPU_TRY(gemm_make_plan(&ws.patch_plan, ws.patches, static_cast<int>(prow), v->w_patch, d.hidden, v->patch_dim,
                           SPU_EPI_BIAS_RESID, GemmIn::BF16));
    SPU_TRY(gemm_make_plan(&ws.head_plan, ws.cls_bf16, static_cast<int>(brow), v->w_head, v->classes_pad, d.hidden,
                           SPU_EPI_BIAS_RESID, GemmIn::BF16));
    ws.plans.resize(d.layers);
    const int R = static_cast<int>(rows);
    for (int l = 0; l < d.layers; ++l) {
      const LayerWeights &w = v->layers[l];
      SPU_TRY(gemm_make_plan(&ws.plans[l].qkv, ws.y_bf16, R, w.w_qkv, 3 * d.hidden, d.hidden, SPU_EPI_BIAS,
                             GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].attn_out, ws.ctx, R, w.w_o, d.hidden, d.hidden, SPU_EPI_BIAS_RESID,
                             GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].mlp1, ws.y_bf16, R, w.w_1, d.mlp, d.hidden, SPU_EPI_BIAS_GELU,
                             GemmIn::BF16));
      SPU_TRY(gemm_make_plan(&ws.plans[l].mlp2, ws.h, R, w.w_2, d.hidden, d.mlp, SPU_EPI_BIAS_RESID,
                             GemmIn::BF16));
    }
  }
  SPU_CUDA_TRY(cudaDeviceSynchronize());
  return SPU_OK;
}

// gemm_only (measurement hook, spu_vit_gemm_chain_ms): enqueue the dense GEMMs only -- patch embedding, QKV,
// attention-output, MLP1, MLP2, head -- on whatever the workspace holds
spu_status vit_enqueue(spu_replica *r, VitState *v, VitWorkspace &ws, const float *pixels, int batch,
                       float *logits, cudaStream_t stream, bool gemm_only = false) {
  const spu_vit_desc &d = v->d;
  const int S = v->seq, M = batch * S, MP = batch * v->n_patches;
  if (!gemm_only) {
    const long long total = static_cast<long long>(batch) * d.channels * d.image * (d.image / 4);
    const long long blocks = (total + 255) / 256;
    const int grid = static_cast<int>(blocks > 148LL * 16 ? 148LL * 16 : blocks);
    vit_im2col_kernel<<<grid, 256, 0, stream>>>(pixels, ws.patches, batch, d.channels, d.image, d.patch);
    SPU_CUDA_TRY(cudaGetLastError());
  }
  GemmEpilogueArgs e{};
  e.acc_scale = 1.f;
  e.bias = v->b_patch;
  e.resid = v->pos_tiled;
  e.out = ws.patch_out;
  SPU_TRY(gemm_launch(ws.patch_plan, MP, e, stream));
  if (!gemm_only) {
    const int grid = (M + kRowThreads / 32 - 1) / (kRowThreads / 32);
    SPU_CUDA_TRY(launch_pdl(vit_assemble_ln_kernel, dim3(grid), dim3(kRowThreads), 0, stream, ws.patch_out,
                            v->cls_pos0, v->layers[0].ln1_g, v->layers[0].ln1_b, d.ln_eps, ws.x_f32, ws.y_bf16, M, S,
                            d.hidden));
  }
  for (int l = 0; l < d.layers; ++l) {
    const LayerWeights &w = v->layers[l];
    const VitLayerPlans &p = ws.plans[l];
    e.bias = w.b_qkv;
    e.resid = nullptr;
    e.out = ws.qkv;
    SPU_TRY(gemm_launch(p.qkv, M, e, stream));
    if (!gemm_only) SPU_TRY(attention_launch(ws.qkv, nullptr, ws.ctx, batch, S, d.heads, stream));
    e.bias = w.b_o;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.attn_out, M, e, stream));
    if (!gemm_only) SPU_TRY(layernorm_launch(ws.x_f32, w.ln2_g, w.ln2_b, d.ln_eps, nullptr, ws.y_bf16, M, d.hidden, stream));
    e.bias = w.b_1;
    e.resid = nullptr;
    e.out = ws.h;
    SPU_TRY(gemm_launch(p.mlp1, M, e, stream));
    e.bias = w.b_2;
    e.resid = ws.x_f32;
    e.out = ws.x_f32;
    SPU_TRY(gemm_launch(p.mlp2, M, e, stream));
    if (l + 1 < d.layers && !gemm_only) {
      const LayerWeights &n = v->layers[l + 1];
      SPU_TRY(layernorm_launch(ws.x_f32, n.ln1_g, n.ln1_b, d.ln_eps, nullptr, ws.y_bf16, M, d.hidden, stream));
    }
  }
  // encoder.ln on the class tokens only (rows b*S), then the classification head
  if (!gemm_only)
    SPU_TRY(layernorm_launch(ws.x_f32, v->lnf_g, v->lnf_b, d.ln_eps, nullptr, ws.cls_bf16, batch, d.hidden, stream,
                             static_cast<long long>(S) * d.hidden));
  e.bias = v->b_head;
  e.resid = v->zeros;
  e.out = ws.logits_pad;
  SPU_TRY(gemm_launch(ws.head_plan, batch, e, stream));
  if (!gemm_only) {
    const int total = batch * d.classes;
    vit_slice_logits_kernel<<<(total + 255) / 256, 256, 0, stream>>>(ws.logits_pad, logits, batch, d.classes,
                                                                    v->classes_pad);
    SPU_CUDA_TRY(cudaGetLastError());
  }
  (void)r;
  return SPU_OK;
}

}  // namespace
}  // namespace spu

using namespace spu;

extern "C" spu_status spu_vit_replica_create(const spu_vit_desc *desc, const float *const *tensors, int32_t n_tensors,
                                             int32_t device, int32_t max_batch, int32_t n_workspaces, uint32_t flags,
                                             spu_replica **out) {
  SPU_REQUIRE(desc && tensors && out, SPU_ERR_INVALID, "vit_replica_create: null argument");
  *out = nullptr;
  SPU_REQUIRE(desc->layers > 0 && desc->hidden > 0 && desc->heads > 0 && desc->mlp > 0 && desc->patch > 0 &&
                  desc->image > 0 && desc->channels > 0 && desc->classes > 0,
              SPU_ERR_INVALID, "vit_replica_create: bad descriptor");
  SPU_REQUIRE(n_tensors == SPU_VIT_NUM_TENSORS(desc->layers), SPU_ERR_INVALID,
              "vit_replica_create: wrong tensor count");
  SPU_REQUIRE(desc->hidden == desc->heads * 64, SPU_ERR_UNSUPPORTED, "vit_replica_create: head_dim must be 64");
  SPU_REQUIRE(desc->hidden % 64 == 0 && desc->mlp % 64 == 0 && desc->hidden <= 1024, SPU_ERR_UNSUPPORTED,
              "vit_replica_create: hidden/mlp must be multiples of 64, hidden <= 1024");
  SPU_REQUIRE(desc->image % desc->patch == 0 && desc->image % 4 == 0 && desc->patch % 4 == 0, SPU_ERR_UNSUPPORTED,
              "vit_replica_create: image must be a multiple of the patch, both multiples of 4");
  SPU_REQUIRE((desc->channels * desc->patch * desc->patch) % 64 == 0, SPU_ERR_UNSUPPORTED,
              "vit_replica_create: C*P*P must be a multiple of 64");
  SPU_REQUIRE(max_batch > 0 && n_workspaces > 0, SPU_ERR_INVALID, "vit_replica_create: bad capacity");
  SPU_REQUIRE((flags & ~SPU_FORWARD_AUTO_GRAPH) == SPU_PREC_BF16, SPU_ERR_UNSUPPORTED,
              "vit_replica_create: unsupported precision flag");
  for (int i = 0; i < n_tensors; ++i)
    SPU_REQUIRE(tensors[i] != nullptr, SPU_ERR_INVALID, "vit_replica_create: null weight tensor");
  SPU_REQUIRE(spu_device_is_sm100(device) == 1, SPU_ERR_UNSUPPORTED,
              "vit_replica_create: device is not an sm_100 GPU (no CPU fallback exists)");
  DeviceGuard guard(device);
  auto r = std::make_unique<spu_replica>();
  auto v = std::make_unique<VitState>();
  v->d = *desc;
  const int nw = desc->image / desc->patch;
  v->n_patches = nw * nw;
  v->seq = v->n_patches + 1;
  v->patch_dim = desc->channels * desc->patch * desc->patch;
  v->classes_pad = ((desc->classes + 63) / 64) * 64;
  r->kind = 1;
  r->device = device;
  r->max_batch = max_batch;
  r->max_seq = v->seq;
  r->flags = flags;
  r->cap_rows = ((max_batch * v->seq + 127) / 128) * 128;
  spu_status st = vit_build(r.get(), v.get(), tensors, max_batch, n_workspaces);
  if (st != SPU_OK) {
    for (void *p : r->allocations) cudaFree(p);
    return st;
  }
  r->vit = v.release();
  r->vit_destroy = &vit_destroy;
  *out = r.release();
  return SPU_OK;
}

extern "C" spu_status spu_vit_forward_async(spu_replica *r, int32_t ws_idx, const float *pixels, int32_t batch,
                                            float *logits, void *stream) {
  SPU_REQUIRE(r != nullptr && r->kind == 1 && r->vit != nullptr, SPU_ERR_INVALID, "vit_forward: not a ViT replica");
  SPU_REQUIRE(pixels && logits, SPU_ERR_INVALID, "vit_forward: null buffer");
  SPU_REQUIRE(ws_idx >= 0 && ws_idx < static_cast<int>(r->vit->ws.size()), SPU_ERR_INVALID,
              "vit_forward: workspace index out of range");
  SPU_REQUIRE(batch >= 1, SPU_ERR_INVALID, "vit_forward: empty batch");
  SPU_REQUIRE(batch <= r->max_batch, SPU_ERR_CAPACITY,
              "vit_forward: batch exceeds the workspace the replica was created with");
  SPU_REQUIRE((reinterpret_cast<uintptr_t>(pixels) & 15u) == 0, SPU_ERR_INVALID, "vit_forward: pixels must be 16-byte aligned");
  VitWorkspace &ws = r->vit->ws[ws_idx];
  cudaStream_t s = as_stream(stream);
  if ((r->flags & SPU_FORWARD_AUTO_GRAPH) == 0 || s == nullptr) return vit_enqueue(r, r->vit, ws, pixels, batch, logits, s);
  // one graph per batch size, retargeted to the task's buffers (as the BERT forward, replica.cu)
  for (CapturedGraph &g : ws.graphs) {
    if (g.key.batch != batch) continue;
    if (g.key.ids != pixels) {
      SPU_TRY(graph_patch_node(g, g.first, kIm2colKernelArgs, {{kIm2colArgPixels, pixels}}));
      g.key.ids = pixels;
    }
    if (g.key.out != logits) {
      SPU_TRY(graph_patch_node(g, g.last, kSliceLogitsKernelArgs, {{kSliceLogitsArgLogits, logits}}));
      g.key.out = logits;
    }
    g.tick = ++ws.graph_tick;
    SPU_CUDA_TRY(cudaGraphLaunch(g.exec, s));
    return SPU_OK;
  }
  bool seen = false;
  for (int b : ws.seen_batches) seen = seen || b == batch;
  if (!seen) {  // first sighting: eager (sets kernel attributes, warms caches)
    if (ws.seen_batches.size() < 4 * kMaxGraphs) ws.seen_batches.push_back(batch);
    return vit_enqueue(r, r->vit, ws, pixels, batch, logits, s);
  }
  SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  spu_status st = vit_enqueue(r, r->vit, ws, pixels, batch, logits, s);
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(s, &graph);
  if (st != SPU_OK) {
    if (graph) cudaGraphDestroy(graph);
    return st;
  }
  SPU_CUDA_TRY(e);
  CapturedGraph cg;
  cg.key.ids = pixels;
  cg.key.out = logits;
  cg.key.batch = batch;
  cg.graph = graph;
  e = cudaGraphInstantiate(&cg.exec, graph, 0);
  if (e != cudaSuccess) {
    cudaGraphDestroy(graph);
    SPU_CUDA_TRY(e);
  }
  cudaKernelNodeParams p0{}, p1{};
  cg.patchable = graph_chain_ends(graph, &cg.first, &cg.last, &p0, &p1) &&
                 p0.func == reinterpret_cast<const void *>(&vit_im2col_kernel) &&
                 p1.func == reinterpret_cast<const void *>(&vit_slice_logits_kernel);
  // ... and the parameters captured there must be the very pointers that will be patched later (as
  // replica.cu's capture_forward checks for the BERT chain)
  cg.patchable = cg.patchable && *static_cast<const void *const *>(p0.kernelParams[kIm2colArgPixels]) == pixels &&
                 *static_cast<const void *const *>(p1.kernelParams[kSliceLogitsArgLogits]) == logits;
  SPU_CUDA_TRY(cudaGraphLaunch(cg.exec, s));
  if (!cg.patchable) {  // cannot follow the task's buffers: not worth keeping
    cudaStreamSynchronize(s);
    destroy_graph(cg);
    return SPU_OK;
  }
  if (ws.graphs.size() >= kMaxGraphs) {
    size_t lru = 0;
    for (size_t i = 1; i < ws.graphs.size(); ++i)
      if (ws.graphs[i].tick < ws.graphs[lru].tick) lru = i;
    destroy_graph(ws.graphs[lru]);
    ws.graphs.erase(ws.graphs.begin() + static_cast<long>(lru));
  }
  cg.tick = ++ws.graph_tick;
  ws.graphs.push_back(cg);
  return SPU_OK;
}

// Measurement hook (bench.py's roofline of config E): device ms per forward of the dense GEMMs alone -- `reps`
// forwards' worth of them back to back in one CUDA graph, CUDA events on `stream`.  Clobbers the workspace.
extern "C" spu_status spu_vit_gemm_chain_ms(spu_replica *r, int32_t ws_idx, const float *pixels, int32_t batch,
                                            float *logits, int32_t reps, void *stream, float *ms_per_forward) {
  SPU_REQUIRE(r != nullptr && r->kind == 1 && r->vit != nullptr, SPU_ERR_INVALID, "vit_gemm_chain: not a ViT replica");
  SPU_REQUIRE(pixels && logits && ms_per_forward && reps > 0 && stream != nullptr, SPU_ERR_INVALID,
              "vit_gemm_chain: bad argument (a non-default stream is needed for the capture)");
  SPU_REQUIRE(ws_idx >= 0 && ws_idx < static_cast<int>(r->vit->ws.size()) && batch >= 1 && batch <= r->max_batch,
              SPU_ERR_INVALID, "vit_gemm_chain: bad workspace or batch");
  cudaStream_t s = as_stream(stream);
  VitWorkspace &ws = r->vit->ws[ws_idx];
  SPU_TRY(vit_enqueue(r, r->vit, ws, pixels, batch, logits, s, true));  // warm-up, kernel attributes
  SPU_CUDA_TRY(cudaStreamSynchronize(s));
  SPU_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
  spu_status st = SPU_OK;
  for (int i = 0; i < reps && st == SPU_OK; ++i) st = vit_enqueue(r, r->vit, ws, pixels, batch, logits, s, true);
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
  cudaGraphLaunch(exec, s);
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

extern "C" uint64_t spu_vit_input_bytes(const spu_replica *r, int32_t batch) {
  if (r == nullptr || r->kind != 1 || r->vit == nullptr || batch <= 0) return 0;
  const spu_vit_desc &d = r->vit->d;
  return static_cast<uint64_t>(batch) * d.channels * d.image * d.image * 4u;
}

extern "C" uint64_t spu_vit_output_bytes(const spu_replica *r, int32_t batch) {
  if (r == nullptr || r->kind != 1 || r->vit == nullptr || batch <= 0) return 0;
  return static_cast<uint64_t>(batch) * r->vit->d.classes * 4u;
}
