// Shared definitions of the weight replica (BERT fields here; the ViT state hangs off
// `vit`, owned by vit.cu).  Internal header of libspu_b200.so.
#pragma once
#include <cstring>
#include <initializer_list>
#include <memory>
#include <string>
#include <tuple>
#include <type_traits>
#include <vector>

#include "common.cuh"
#include "gemm_sm100.cuh"

namespace spu {

// ctx is bf16 [rows, heads*64], or e4m3 of ctx * q_inv_scale when out_e4m3 is set
spu_status attention_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq,
                            int heads, cudaStream_t stream, bool out_e4m3 = false,
                            float q_inv_scale = 1.f);
// the two implementations behind attention_launch (attention_sm100.cu: tcgen05; attention.cu: mma.sync)
spu_status attention_tc_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq,
                               int heads, cudaStream_t stream, bool out_e4m3, float q_inv_scale,
                               const int32_t *cu_seqlens = nullptr, int total_tokens = 0);
spu_status attention_mma_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq,
                                int heads, cudaStream_t stream, bool out_e4m3, float q_inv_scale);
spu_status bert_embed_launch(const int64_t *ids, const int64_t *mask, const float *word,
                             const float *pos, const float *type0, const float *gamma,
                             const float *beta, float eps, int vocab, float *x_f32, void *x_bf16,
                             float *mask_bias, int batch, int seq, int hidden, cudaStream_t stream,
                             void *x_q = nullptr, float q_inv_scale = 1.f);
// in_row_stride: elements between consecutive input rows (0 = hidden, i.e. contiguous)
spu_status layernorm_launch(const float *x, const float *gamma, const float *beta, float eps,
                            float *out_f32, void *out_bf16, int rows, int hidden,
                            cudaStream_t stream, long long in_row_stride = 0, void *out_q = nullptr,
                            float q_inv_scale = 1.f);
// ---- packed (unpadded) forward: sequences back to back, cu_seqlens[b] = first row of sequence b ----
constexpr int kMaxPackedBatch = 256;  // cu_seqlens rides in a kernel parameter
// host lengths -> device cu_seqlens [batch + 1] (as a kernel parameter: no staging buffer to race on); *total = sum
spu_status write_cu_seqlens_launch(const int32_t *lens_host, int batch, int32_t *cu_dev, int *total,
                                   cudaStream_t stream);
spu_status bert_embed_packed_launch(const int64_t *ids, const int32_t *cu, int batch, int seq, int total,
                                    const float *word, const float *pos, const float *type0, const float *gamma,
                                    const float *beta, float eps, int vocab, float *x_f32, void *x_bf16, int hidden,
                                    cudaStream_t stream);
// final LayerNorm reading packed rows and writing the padded [batch, seq, hidden] output (zeros behind each length)
spu_status layernorm_unpad_launch(const float *x, const int32_t *cu, int batch, int seq, const float *gamma,
                                  const float *beta, float eps, float *out_f32, void *out_bf16, int hidden,
                                  cudaStream_t stream);
// kernel entry points of the first / last launch of a forward (graph parameter patching, replica.cu)
const void *gather_pad_mask_kernel_entry();
const void *scatter_requests_kernel_entry();
const void *bert_embed_kernel_entry();
const void *layernorm_kernel_entry();
// request i <- rows [i*seq, i*seq + (full_rows ? seq : len[i])) of src [n_req*seq, hidden]
// err (may be null): the word the gather kernel reported into; status (may be null): where this task's finding
// goes (0 = every token id inside the embedding table, else 1 + a row that carried one outside); err is re-armed
spu_status scatter_requests_launch(const float *src, const int32_t *len, int n_req, int seq, int hidden,
                                   bool full_rows, float *const *dsts, cudaStream_t stream, int32_t *err = nullptr,
                                   int32_t *status = nullptr);
// spu_gather_pad_mask_async plus the id check: err != null -> atomicMax(err, row + 1) for a live id outside [0, vocab)
spu_status gather_pad_mask_launch(const int64_t *const *src_ids, const int64_t *const *src_mask, const int32_t *len,
                                  int n_req, int seq, int64_t pad_id, int64_t *ids_out, int64_t *mask_out,
                                  int64_t vocab, int32_t *err, cudaStream_t stream);
// parameter counts / positions of the two kernels whose arguments replica.cu patches in instantiated graphs
// (rowops.cu: gather_pad_mask_kernel(src_ids, src_mask, len, seq, pad_id, ids_out, mask_out, vocab, err);
//  scatter_requests_kernel(src, len, seq, hidden, full_rows, dsts, err, status))
constexpr int kGatherKernelArgs = 9, kScatterKernelArgs = 8;
constexpr int kScatterArgLen = 1, kScatterArgDsts = 5, kScatterArgStatus = 7;
// bert_embed_ln_kernel(ids, mask, ...17 in all); layernorm_kernel(x, gamma, beta, eps, out_f32 [4], out_bf16 [5], ...11)
constexpr int kEmbedKernelArgs = 17, kEmbedArgIds = 0, kEmbedArgMask = 1;
constexpr int kLayerNormKernelArgs = 11, kLayerNormArgOutF32 = 4, kLayerNormArgOutBf16 = 5;

// The file that defines a patched kernel pins those numbers to its real signature (static_assert beside the kernel),
// so a changed parameter list fails the build instead of patching the wrong argument of an instantiated graph.
template <typename F>
struct KernelSig;
template <typename... A>
struct KernelSig<void(A...)> {
  static constexpr int count = static_cast<int>(sizeof...(A));
  template <int I>
  static constexpr bool pointer_at = std::is_pointer<typename std::tuple_element<I, std::tuple<A...>>::type>::value;
};
#define SPU_PIN_KERNEL_ARGS(kernel, n) static_assert(::spu::KernelSig<decltype(kernel)>::count == (n), #kernel ": parameter count changed; update the graph-patching constants")
#define SPU_PIN_KERNEL_PTR_ARG(kernel, i) static_assert(::spu::KernelSig<decltype(kernel)>::pointer_at<(i)>, #kernel ": patched parameter is not a pointer")
// e4m3 = sat(in * inv_scale); exactly one of in_f32 / in_bf16 is non-null; n % 4 == 0
spu_status quantize_e4m3_launch(const float *in_f32, const void *in_bf16, void *out, uint64_t n,
                                float inv_scale, cudaStream_t stream);
// amax[0] = max(amax[0], max|in|)
spu_status amax_launch(const float *in_f32, const void *in_bf16, uint64_t n, float *amax,
                       cudaStream_t stream);

inline uint16_t f32_to_bf16_rne(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40u);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return static_cast<uint16_t>(u >> 16);
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

struct LayerWeights {
  __nv_bfloat16 *w_qkv = nullptr, *w_o = nullptr, *w_1 = nullptr, *w_2 = nullptr;
  float *b_qkv = nullptr, *b_o = nullptr, *b_1 = nullptr, *b_2 = nullptr;
  float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
  // FP8 E4M3 variant: per-tensor scaled weights (value = q * sw) and the calibrated
  // per-tensor activation scales of the four GEMM inputs of this layer
  uint8_t *q_qkv = nullptr, *q_o = nullptr, *q_1 = nullptr, *q_2 = nullptr;
  float sw_qkv = 1.f, sw_o = 1.f, sw_1 = 1.f, sw_2 = 1.f;
  float sa_x = 1.f, sa_ctx = 1.f, sa_x1 = 1.f, sa_h = 1.f;
};

struct LayerPlans {
  GemmPlan qkv, attn_out, ffn1, ffn2;
  // FP8 plan set only: FFN1 with mixed operand / output formats (per-GEMM FP8 mask)
  GemmPlan ffn1_b2q;  // bf16 operands -> e4m3 GELU output (FFN2 runs in FP8, FFN1 does not)
  GemmPlan ffn1_q2b;  // e4m3 operands -> bf16 GELU output (FFN1 runs in FP8, FFN2 does not)
  // residual + LayerNorm fused into the epilogue (replica.ln_fused): replace attn_out / ffn2 + layernorm_kernel
  GemmPlan attn_out_ln, ffn2_ln;
};

// One captured forward per (batch, seq) and workspace.  The I/O pointers (ids, mask, out) are NOT part of the key:
// they only appear in the first kernel (embedding) and the last one (the final LayerNorm), whose parameters are
// patched in the instantiated graph before a launch with other pointers (cudaGraphExecKernelNodeSetParams).  A
// worker therefore needs one graph per batch shape, not one per (slot, shape) tuple, and a dynamic-batch server
// cannot fill the cache with cold entries; beyond kMaxGraphs the least recently used one is evicted.
struct GraphKey {
  const void *ids = nullptr, *mask = nullptr, *out = nullptr;  // the pointers the graph currently holds
  int batch = 0, seq = 0;
  const void *aux = nullptr;  // requests path: the source-mask table
  const void *status = nullptr;  // requests path: the task's status word
  int full_rows = 0;          // requests path
  bool same_shape(int b, int s) const { return batch == b && seq == s; }
  bool operator==(const GraphKey &o) const {
    return ids == o.ids && mask == o.mask && out == o.out && batch == o.batch && seq == o.seq;
  }
};

struct CapturedGraph {
  GraphKey key;
  cudaGraphExec_t exec = nullptr;
  cudaGraph_t graph = nullptr;  // kept alive: the patched nodes' unchanged parameters point into it
  // patchable = the graph's single root is the embedding kernel and its single leaf the LayerNorm kernel
  bool patchable = false;
  cudaGraphNode_t first = nullptr, last = nullptr;
  int out_arg = 4;              // index of the output pointer among the last kernel's parameters
  unsigned long long tick = 0;  // LRU stamp
};

// CUDA-graph helpers (replica.cu): find the single root / leaf kernel nodes of a captured chain; replace 8-byte
// parameters of one of them in the instantiated graph (and its template); free a captured graph.
constexpr size_t kMaxGraphs = 64;  // per workspace; least recently used goes first
bool graph_chain_ends(cudaGraph_t graph, cudaGraphNode_t *root, cudaGraphNode_t *leaf, cudaKernelNodeParams *p0,
                      cudaKernelNodeParams *p1);
spu_status graph_patch_node(CapturedGraph &g, cudaGraphNode_t node, int n_args,
                            std::initializer_list<std::pair<int, const void *>> changes);
void destroy_graph(CapturedGraph &g);

struct Workspace {
  float *x_f32 = nullptr;
  __nv_bfloat16 *x_bf16 = nullptr, *qkv = nullptr, *ctx = nullptr, *h = nullptr;
  float *mask_bias = nullptr;
  uint8_t *x_q = nullptr, *ctx_q = nullptr, *h_q = nullptr;  // FP8 variant operands
  // SPU_REQUEST_IO: the assembled [batch, seq] ids / mask and the batched output of
  // spu_bert_forward_requests_async (stage 1 writes, stage 4 reads)
  int64_t *req_ids = nullptr, *req_mask = nullptr;
  float *req_out = nullptr;
  int32_t *req_err = nullptr;  // requests path: 1 + a row whose token id lies outside the table (0 = none), per task
  int32_t *cu_seqlens = nullptr;  // packed forward: [kMaxPackedBatch + 1]
  std::vector<LayerPlans> plans;
  std::vector<LayerPlans> plans8;
  std::vector<CapturedGraph> graphs;
  std::vector<GraphKey> seen;  // SPU_FORWARD_AUTO_GRAPH: shapes run eagerly once so far
  // spu_bert_forward_requests_async (gather -> forward -> scatter): one graph per (n_req, seq, full_rows); the
  // key's ids / mask / out fields hold the source-id table, the length array and the destination table
  std::vector<CapturedGraph> req_graphs;
  std::vector<GraphKey> req_seen;
  unsigned long long graph_tick = 0;
};

struct VitState;  // vit.cu

}  // namespace spu

struct spu_replica {
  int kind = 0;  // 0 = BERT, 1 = ViT
  spu_bert_desc desc{};
  int device = 0;
  int max_batch = 0, max_seq = 0, cap_rows = 0;
  uint32_t flags = 0;
  float *word = nullptr, *pos = nullptr, *type0 = nullptr, *emb_g = nullptr, *emb_b = nullptr;
  std::vector<spu::LayerWeights> layers;
  std::vector<spu::Workspace> ws;
  std::vector<void *> allocations;
  float *amax = nullptr;      // [layers * 4] running calibration maxima (FP8 variant)
  bool fp8_calibrated = false;
  bool ln_fused = false;      // N = hidden GEMMs carry the residual + LayerNorm in their epilogue
  int fp8_layers = 0;         // leading layers whose GEMMs run in FP8 (rest bf16)
  // per layer: which of its four GEMMs run in FP8 (SPU_FP8_GEMM_* bits); others run in bf16
  std::vector<uint8_t> fp8_mask;
  spu::TunedTable tuned;      // spu_bert_autotune: measured tile winners, read by every GEMM plan of this replica
  spu::VitState *vit = nullptr;
  void (*vit_destroy)(spu::VitState *) = nullptr;
};

namespace spu {

template <typename T>
spu_status dev_alloc(spu_replica *r, T **out, size_t count) {
  void *p = nullptr;
  cudaError_t e = cudaMalloc(&p, (count ? count : 1) * sizeof(T));
  if (e != cudaSuccess) {
    set_error(std::string("cudaMalloc: ") + cudaGetErrorString(e));
    return SPU_ERR_NOMEM;
  }
  r->allocations.push_back(p);
  *out = static_cast<T *>(p);
  return SPU_OK;
}

inline spu_status upload_f32(spu_replica *r, float **dst, const float *src, size_t n) {
  spu_status st = dev_alloc(r, dst, n);
  if (st != SPU_OK) return st;
  SPU_CUDA_TRY(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice));
  return SPU_OK;
}

// Upload fp32 host matrices, stacked, as one bf16 device matrix.
inline spu_status upload_bf16_stacked(spu_replica *r, __nv_bfloat16 **dst,
                                      std::initializer_list<const float *> srcs, size_t n_each) {
  const size_t total = n_each * srcs.size();
  std::vector<uint16_t> tmp(total);
  size_t o = 0;
  for (const float *s : srcs)
    for (size_t i = 0; i < n_each; ++i) tmp[o++] = f32_to_bf16_rne(s[i]);
  spu_status st = dev_alloc(r, dst, total);
  if (st != SPU_OK) return st;
  SPU_CUDA_TRY(cudaMemcpy(*dst, tmp.data(), total * 2, cudaMemcpyHostToDevice));
  return SPU_OK;
}

inline spu_status upload_f32_stacked(spu_replica *r, float **dst, std::initializer_list<const float *> srcs,
                                     size_t n_each) {
  std::vector<float> tmp;
  tmp.reserve(n_each * srcs.size());
  for (const float *s : srcs) tmp.insert(tmp.end(), s, s + n_each);
  return upload_f32(r, dst, tmp.data(), tmp.size());
}

#define SPU_TRY(expr)                 \
  do {                                \
    spu_status _s = (expr);           \
    if (_s != SPU_OK) return _s;      \
  } while (0)

}  // namespace spu
