/*
 * spu_b200.h -- C-ABI of the B200-native inference hot path (libspu_b200.so).
 *
 * Drop-in boundary for ONE path of daxmawal/StarPU-Inference-Server: the body of
 * the StarPU inference codelet's cuda_func.  In the reference that body is
 *   InferenceCodelet::cuda_inference_func      src/core/starpu_setup.cpp:807-846
 *     -> run_codelet_inference                 src/core/starpu_setup.cpp:638-723
 *       -> run_inference                       src/core/starpu_setup.cpp:594-624
 *            TensorBuilder::prepare_input_ivalues  src/core/tensor_builder.cpp:148-155
 *            model->forward(ivalues)               src/core/starpu_setup.cpp:610
 *            wrapper.copy_(out, non_blocking)      src/core/starpu_setup.cpp:834-844
 * Everything below replaces "wrap StarPU device buffers as torch tensors, call
 * LibTorch forward, copy the result into the STARPU_W buffer" with hand-written
 * sm_100a kernels.  Plain pointers and sizes only; no torch / C++ types; never
 * throws; every call returns an spu_status and records a message retrievable
 * with spu_last_error().  All *_async calls only enqueue work on `stream`
 * (a cudaStream_t passed as void*) -- no allocation, no device/stream
 * synchronisation, no default-stream use -- which is the STARPU_CUDA_ASYNC
 * contract (codelet cuda_flags[0] = 1U, src/core/starpu_setup.cpp:559-568).
 *
 * There is NO CPU fallback: if no sm_100 device is usable the create calls fail.
 */
#ifndef SPU_B200_H
#define SPU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPU_ABI_VERSION 2 /* 2: packed / compact forward, autotune, graph count, vocab size */

typedef enum spu_status {
  SPU_OK = 0,
  SPU_ERR_INVALID = 1,     /* bad argument (null pointer, shape out of range)   */
  SPU_ERR_CUDA = 2,        /* a CUDA runtime / driver call failed               */
  SPU_ERR_CAPACITY = 3,    /* batch/seq exceeds what the replica was sized for  */
  SPU_ERR_UNSUPPORTED = 4, /* device is not sm_100, or unsupported model shape  */
  SPU_ERR_NOMEM = 5
} spu_status;

/* Opaque per-GPU weight replica + per-worker workspaces.  Replaces one entry of
 * InferenceParams::models.models_gpu (src/core/inference_params.hpp:46-51), i.e.
 * one `model_cpu.clone().to(cuda:id)` of clone_model_to_gpus
 * (src/core/inference_runner.cpp:251-275). */
typedef struct spu_replica spu_replica;

/* ---- BERT encoder (models/bert.yml: input_ids,attention_mask int64 [B,S] ->
 *      output0 fp32 [B,S,hidden] = last_hidden_state) ------------------------ */
typedef struct spu_bert_desc {
  int32_t layers, hidden, heads, ffn, vocab, max_pos, type_vocab;
  float ln_eps;
} spu_bert_desc;

/* precision flags for spu_*_replica_create */
#define SPU_PREC_BF16 0u
/* Per-tensor scaled FP8 E4M3 GEMM operands.  Alone, this selects the default FP8 plan: the QKV projection of
 * every layer and both FFN GEMMs of the first half of the encoder run in FP8, the rest in bf16 -- the largest
 * measured plan inside the 6e-2 gate on BERT-base (58 % of the GEMM FLOPs; profiles/r2_fp8_plan.md). */
#define SPU_PREC_FP8_E4M3 1u
/* All four GEMMs of the first n encoder layers in FP8, bf16 GEMMs in the rest (n >= layers: every GEMM,
 * which sits at the E4M3 floor of 7.6e-2 on BERT-base, outside the gate).  spu_bert_set_fp8_mask sets any plan. */
#define SPU_PREC_FP8_LAYERS(n) (SPU_PREC_FP8_E4M3 | ((uint32_t)(n) << 8))
/* Warm-up capture in place (src/core/warmup.cpp:225-299 runs fixed-worker jobs so that every
 * worker has executed once before traffic arrives): with this flag spu_bert_forward_async
 * captures a CUDA graph the second time it sees the same (workspace, ids, mask, out, batch,
 * seq) tuple -- StarPU hands a worker the same device buffers for the same slot handles, so
 * tuples recur -- and replays it from then on.  Capture is thread-local stream capture on the
 * caller's stream: no synchronisation, no allocation visible to other workers. */
#define SPU_FORWARD_AUTO_GRAPH (1u << 16)
/* Allocate the per-workspace staging (ids, mask, batched output) that
 * spu_bert_forward_requests_async needs.  Without it that call returns SPU_ERR_UNSUPPORTED. */
#define SPU_REQUEST_IO (1u << 17)
/* SURVEY 8(f) row 3 (response path): output0 of spu_bert_forward_async / _capture is bf16
 * [batch, seq, hidden] (the fp32 last_hidden_state rounded to nearest even by the last LayerNorm)
 * instead of fp32: half the D2H and response bytes.  spu_bert_output_bytes follows.  Declare the
 * model's output as TYPE_BF16 in models/<name>.yml.  Not combinable with SPU_REQUEST_IO. */
#define SPU_OUTPUT_BF16 (1u << 18)

/* BERT replicas: fuse the residual add + LayerNorm that follow the attention-output and FFN2 projections into
 * those GEMMs' epilogues (see spu_gemm_ln_bf16_async): 5 launches per layer instead of 7.  Ignored when no
 * cluster tiling spans the hidden size.  Results differ from the unfused path only by fp32 summation order. */
#define SPU_LN_FUSED (1u << 19)

/* Number of weight tensors spu_bert_replica_create expects: 5 + 16*layers, in the
 * order of the HF state dict restated in oracle/bert_oracle.py:weight_names()
 * (word, position, token_type embeddings, embeddings.LayerNorm.{weight,bias},
 * then per layer: q.w q.b k.w k.b v.w v.b attn_out.w attn_out.b attn_LN.g
 * attn_LN.b ffn1.w ffn1.b ffn2.w ffn2.b out_LN.g out_LN.b).  All fp32, host
 * memory, nn.Linear layout [out_features, in_features]. */
#define SPU_BERT_NUM_TENSORS(layers) (5 + 16 * (layers))

/* Create one replica on CUDA device `device`: converts the fp32 host tensors to
 * bf16 (packing q/k/v into one [3H,H] matrix), uploads them, and pre-allocates
 * `n_workspaces` activation arenas sized for (max_batch, max_seq) -- one per
 * StarPU CUDA worker that may run on this device (STARPU_NWORKER_PER_CUDA,
 * models/bert.yml:6-8), so forward calls never allocate.
 * Replaces load_model + clone_model_to_gpus (src/core/inference_runner.cpp:243-275). */
spu_status spu_bert_replica_create(const spu_bert_desc *desc,
                                   const float *const *tensors, int32_t n_tensors,
                                   int32_t device, int32_t max_batch,
                                   int32_t max_seq, int32_t n_workspaces,
                                   uint32_t flags, spu_replica **out);

/* Destroy a replica (synchronises its device). */
spu_status spu_replica_destroy(spu_replica *r);

/* The forward the codelet enqueues.  ids/mask: DEVICE pointers to contiguous
 * int64 [batch, seq] (exactly what StarPU hands the reference codelet as
 * buffers[0..1]->ptr after the host gather, slot_manager_component.cpp:649-727);
 * mask != 0 means "keep".  out: DEVICE fp32 [batch, seq, hidden] (buffers[2]->ptr).
 * ws_idx selects the workspace (worker-private; two in-flight calls must not
 * share one). */
spu_status spu_bert_forward_async(spu_replica *r, int32_t ws_idx,
                                  const int64_t *ids, const int64_t *mask,
                                  int32_t batch, int32_t seq, float *out,
                                  void *stream);

/* The same forward for RIGHT-PADDED batches whose true lengths the host knows (a variable-length
 * batch padded by the host gather): lens[b] in [1, seq] (HOST array, read before the call returns) is the number of
 * real tokens of sequence b; ids is the padded [batch, seq] DEVICE array; no mask is needed.  The encoder runs on
 * the sum(lens) real tokens only -- GEMMs, LayerNorms and attention skip the padding instead of multiplying it --
 * and `out` receives the padded [batch, seq, hidden] layout with rows t >= lens[b] set to ZERO (the reference
 * leaves the model's values for [PAD] tokens there; no consumer reads them).  Rows t < lens[b] are bit-identical
 * to spu_bert_forward_async with the equivalent mask.  bf16 replicas, batch <= 256; launches eagerly (the token
 * count changes from batch to batch, so there is no graph to replay). */
spu_status spu_bert_forward_packed_async(spu_replica *r, int32_t ws_idx, const int64_t *ids, const int32_t *lens,
                                         int32_t batch, int32_t seq, float *out, void *stream);
/* The same forward as spu_bert_forward_async, for a caller that has no workspace to name and no way to get one.
 * (No counterpart in the reference: LibTorch allocates its activations from the caching allocator inside
 * model->forward, and the reference never needed a path that cannot.)
 *
 * Who this is for.  spu_bert_replica_create sizes its workspaces once, at start-up, for the CUDA workers it was
 * told about; after that, code running on the serving path must not obtain device memory any more: inside a
 * StarPU codelet body a cudaMalloc / cudaFree stalls every stream of the device (all other workers wait) and takes
 * memory StarPU's own accounting (STARPU_LIMIT_CUDA_MEM, its allocation cache) does not know about, and while any
 * thread is capturing a stream into a CUDA graph an allocation call fails and breaks that capture.  A worker the
 * replica was not sized for, a health probe, or a late warm-up of one more shape is in exactly that position: it
 * holds a replica, input and output buffers, and nothing else.  This entry point gives it the forward anyway.
 *
 * How.  The activations live in ONE array of SPU_STATIC_SCRATCH_BYTES that is compiled into the library as a
 * device global: every GPU that loads libspu_b200.so sets that much memory aside until the process exits, whether
 * or not this function is ever called.  The call itself performs no cudaMalloc, cudaFree or any other allocating
 * CUDA / library routine; it only enqueues the same kernels, with the same tiles, as spu_bert_forward_async, so
 * `out` is bit-identical to that call's.  ids / mask / out are the caller's DEVICE buffers, as there.
 *
 * Limits.  bf16 replicas only (SPU_ERR_UNSUPPORTED for an FP8 replica).  With rows = batch*seq rounded up to a
 * multiple of 128, the forward needs rows * (14*hidden + 2*ffn + 4) bytes; more than SPU_STATIC_SCRATCH_BYTES
 * returns SPU_ERR_CAPACITY with the numbers in spu_last_error() and enqueues nothing.  For BERT-base that is 512
 * tokens (e.g. 4 x 128 or 1 x 512).
 *
 * Sharing.  There is one array per GPU and every call on that GPU uses it, from any thread, stream or replica:
 * two calls whose kernels can overlap on the device (different streams, or no ordering between them) race on the
 * activations and both results are garbage -- order them on one stream or with events.  The array keeps whatever
 * the previous call left in it; the forward does not depend on that (every buffer is written by its producer
 * kernel before any consumer reads it; nothing is accumulated across calls), so no clearing pass is needed or run.
 *
 * First call.  The first call on each device looks the array's device address up (cudaGetSymbolAddress, with
 * that device current) and remembers it; make that first call outside any stream capture, e.g. during warm-up. */
#define SPU_STATIC_SCRATCH_BYTES (9u << 20) /* 9 MiB per GPU */
spu_status spu_bert_forward_static_async(spu_replica *r, const int64_t *ids, const int64_t *mask, int32_t batch,
                                         int32_t seq, float *out, void *stream);

/* Rows of the replica's word-embedding table (0 for a non-BERT replica).  The embedding kernel clamps ids into
 * [0, vocab) for memory safety where LibTorch's embedding would raise; a host layer that wants the reference's
 * failure instead validates ids against this before submitting (core/task_runner.cpp does). */
int32_t spu_bert_vocab_size(const spu_replica *r);
/* The same with a COMPACT output: `out` receives the packed rows only, [sum(lens), hidden] with sequence b at row
 * lens[0] + ... + lens[b-1], nothing for the padding.  A host layer that knows the lengths slices its responses
 * from that (core/task_runner.cpp does on variable_length servers): the device-to-host copy of the output shrinks
 * by the padding fraction, which is what bounds 8 GPUs feeding one host. */
spu_status spu_bert_forward_packed_compact_async(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                                 const int32_t *lens, int32_t batch, int32_t seq, float *out,
                                                 void *stream);
/* 1 when spu_bert_forward_packed_async serves this replica for batches of `batch` sequences. */
int32_t spu_bert_packed_supported(const spu_replica *r, int32_t batch);

/* Number of kernel launches one forward of (batch, seq) enqueues. */
int32_t spu_bert_forward_launch_count(const spu_replica *r);

/* Warm-up autotune (the reference's hook is its fixed-worker warm-up pass,
 * src/core/warmup.cpp:226-299).  For every distinct GEMM problem a forward of (batch, seq) launches -- QKV,
 * attention-output, FFN1, FFN2 in the operand formats the replica's plan uses -- times each (cta_group, BLOCK_N)
 * tile the kernel is built for, on the replica's own weights and workspace ws_idx (20 launches back to back in a
 * CUDA graph, best of 3), and lets the winner replace the static cost model's choice for exactly that
 * (N, K, epilogue, format, M = batch*seq).  Other M keep the cost model.  START-UP ONLY: synchronises the
 * device, clobbers the workspace activations and drops captured graphs.  *n_changed (may be NULL) = problems
 * whose winner differs from the cost model's tile. */
spu_status spu_bert_autotune(spu_replica *r, int32_t ws_idx, int32_t batch, int32_t seq, void *stream,
                             int32_t *n_changed);
/* The measured winners: rows of 8 int32 {N, K, epilogue, operand format (0 bf16 / 1 e4m3), M, cta_group, BLOCK_N,
 * nanoseconds per launch}; returns the number of rows the table holds. */
int32_t spu_bert_tuned_configs(const spu_replica *r, int32_t *out, int32_t cap_rows);

/* CUDA graphs cached for workspace ws_idx (spu_bert_forward_capture / SPU_FORWARD_AUTO_GRAPH).  A graph is keyed
 * by (batch, seq) only: launching it with other ids / mask / out pointers patches the two kernels that hold
 * them, so a worker keeps ONE graph per batch shape however many slots it serves; at most 64 per workspace,
 * least recently used evicted.  -1 on a bad argument. */
int32_t spu_bert_graph_count(const spu_replica *r, int32_t ws_idx);

/* Stages 1-4 in one call, on PER-REQUEST buffers (SURVEY 8(f) rows 1+2 at the codelet level;
 * replaces the host staging copy of SlotManager::copy_job_inputs_to_slot,
 * slot_manager_component.cpp:649-727, and the slot slicing of batching_helpers.hpp:74-127).
 * Request i supplies len[i] <= seq token ids at src_ids[i] (and optionally len[i] mask words at
 * src_mask[i]; src_mask or src_mask[i] NULL = all ones).  Stage 1 pads every row to `seq` with
 * id 0 / mask 0 on the device; stages 2-3 run the encoder on the assembled [n_req, seq] batch
 * (identical to spu_bert_forward_async on that batch); stage 4 writes request i's hidden states
 * straight into dsts[i]: `seq` rows when full_rows != 0 (the reference's fixed [1, seq, H]
 * output), else only its len[i] live rows.  src tables, len and dsts are DEVICE arrays of
 * device-visible pointers (device memory or pinned, mapped host memory: the kernels then read
 * the requests and write the responses over PCIe directly, with no staging copy at all).
 * The replica must have been created with SPU_REQUEST_IO. */
spu_status spu_bert_forward_requests_async(spu_replica *r, int32_t ws_idx,
                                           const int64_t *const *src_ids,
                                           const int64_t *const *src_mask,
                                           const int32_t *len, int32_t n_req, int32_t seq,
                                           float *const *dsts, int32_t full_rows,
                                           void *stream);

/* The same call with the token ids CHECKED against the embedding table, which is what LibTorch's embedding does on
 * the reference path (an id outside [0, vocab) raises inside model->forward, src/core/starpu_setup.cpp:610, and the
 * whole job fails).  The host cannot do that check here without reading every request (the point of this path is
 * that it does not), so the gather kernel does: `status` is a DEVICE int32 the caller reads once the stream has
 * reached this call's end -- 0 when every live id was inside the table, else 1 + the index of a request row that
 * carried one outside (the largest such row).  The forward still runs (the embedding clamps, nothing reads out
 * of bounds) and the responses of such a batch are written but meaningless: fail the batch.  status == NULL is
 * spu_bert_forward_requests_async. */
spu_status spu_bert_forward_requests_checked_async(spu_replica *r, int32_t ws_idx,
                                                   const int64_t *const *src_ids,
                                                   const int64_t *const *src_mask,
                                                   const int32_t *len, int32_t n_req, int32_t seq,
                                                   float *const *dsts, int32_t full_rows, int32_t *status,
                                                   void *stream);

/* Capture the forward of (batch, seq) on ws_idx into a CUDA graph (warm-up hook,
 * src/core/warmup.cpp:225-299) so later spu_bert_forward_async calls with the
 * same (ws_idx, batch, seq, ids, mask, out) pointers replay it. */
spu_status spu_bert_forward_capture(spu_replica *r, int32_t ws_idx,
                                    const int64_t *ids, const int64_t *mask,
                                    int32_t batch, int32_t seq, float *out,
                                    void *stream);

/* Measurement hook (bench.py, not the codelet path): runs one forward with a CUDA
 * event recorded after every kernel, SYNCHRONISES `stream`, and adds the device time
 * of each kernel class (summed over layers, milliseconds) into stage_ms[SPU_STAGE_COUNT].
 * Never replays a captured graph. */
#define SPU_STAGE_EMBED 0      /* embedding gather + LayerNorm + mask-bias builder   */
#define SPU_STAGE_GEMM_QKV 1   /* [M,H] x [3H,H]^T + bias                            */
#define SPU_STAGE_ATTENTION 2  /* fused masked-softmax attention                     */
#define SPU_STAGE_GEMM_OUT 3   /* [M,H] x [H,H]^T + bias + residual                  */
#define SPU_STAGE_LN1 4
#define SPU_STAGE_GEMM_FFN1 5  /* [M,H] x [F,H]^T + bias, erf GELU                   */
#define SPU_STAGE_GEMM_FFN2 6  /* [M,F] x [H,F]^T + bias + residual                  */
#define SPU_STAGE_LN2 7        /* last layer: writes fp32 output0                    */
#define SPU_STAGE_COUNT 8
spu_status spu_bert_forward_profiled(spu_replica *r, int32_t ws_idx,
                                     const int64_t *ids, const int64_t *mask,
                                     int32_t batch, int32_t seq, float *out,
                                     void *stream, float *stage_ms);

/* FP8 E4M3 variant (replica created with SPU_PREC_FP8_E4M3): per-tensor weight scales are
 * fixed at creation (amax/448); the per-tensor activation scales of the four GEMM inputs of
 * every layer come from this calibration pass: it runs the bf16 forward of (ids, mask) with
 * an amax probe after each such tensor, SYNCHRONISES the stream and installs scale =
 * running amax / 448.  May be called several times (maxima accumulate); must be called at
 * least once before spu_bert_forward_async.  Start-up only (warm-up, warmup.cpp:225-299). */
spu_status spu_bert_calibrate_fp8(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                  const int64_t *mask, int32_t batch, int32_t seq, void *stream);
/* Which GEMMs of which encoder layer run on FP8 operands (the rest run in bf16): one byte per
 * layer, OR of SPU_FP8_GEMM_*.  Replaces the mask the creation flags implied.  Start-up only:
 * synchronises the device and drops captured graphs.  The producer of each GEMM operand
 * (embedding, LayerNorm, attention, the FFN1 GELU epilogue) emits the format its consumer wants. */
#define SPU_FP8_GEMM_QKV 1u
#define SPU_FP8_GEMM_OUT 2u
#define SPU_FP8_GEMM_FFN1 4u
#define SPU_FP8_GEMM_FFN2 8u
#define SPU_FP8_GEMM_ALL 15u
spu_status spu_bert_set_fp8_mask(spu_replica *r, const uint8_t *layer_masks, int32_t n_layers);
/* Current mask into out[0..cap); returns the layer count (0 for a bf16 replica). */
int32_t spu_bert_get_fp8_mask(const spu_replica *r, uint8_t *out, int32_t cap);
/* Scales of an FP8 replica, 8 per layer: activation x, ctx, x1, h then weight qkv, out,
 * ffn1, ffn2.  Returns the count (0 for a bf16 replica). */
int32_t spu_bert_fp8_scales(const spu_replica *r, float *out, int32_t cap);

/* Measurement hook: device time (ms per forward) of ONLY the kernel classes whose bit
 * (1u << SPU_STAGE_*) is set in stage_mask, run as `reps` forwards' worth of launches back
 * to back inside one CUDA graph (no host gaps, programmatic dependent launch in effect) and
 * bracketed by two CUDA events on `stream`.  Synchronises the stream.  stage_mask = ~0u
 * times the whole forward. */
spu_status spu_bert_stage_chain_ms(spu_replica *r, int32_t ws_idx, const int64_t *ids,
                                   const int64_t *mask, int32_t batch, int32_t seq, float *out,
                                   uint32_t stage_mask, int32_t reps, void *stream,
                                   float *ms_per_forward);

/* Bytes of output0 for (batch, seq): batch*seq*hidden*4 (0 for a null / non-BERT replica).
 * The codelet checks the STARPU_W buffer against it, like the reference's
 * TORCH_CHECK(buffer.size() == out.nbytes()) (src/core/starpu_setup.cpp:840-842). */
uint64_t spu_bert_output_bytes(const spu_replica *r, int32_t batch, int32_t seq);

/* ---- ViT encoder (models/vit_l_16.yml: input fp32 [B,3,224,224] -> output fp32
 *      [B,num_classes]; torchvision vit_{b,l}_16 graph, models/import_vit.py:10-13) ---- */
typedef struct spu_vit_desc {
  int32_t layers, hidden, heads, mlp, patch, image, channels, classes;
  float ln_eps;
} spu_vit_desc;

/* Weight tensors, fp32 host, torchvision state-dict order restated in
 * oracle/vit_oracle.py:weight_names(): conv_proj.weight [H,C,P,P], conv_proj.bias,
 * class_token [1,1,H], encoder.pos_embedding [1,S,H], then per layer: ln_1.{weight,bias},
 * self_attention.in_proj_{weight [3H,H],bias}, self_attention.out_proj.{weight,bias},
 * ln_2.{weight,bias}, mlp.0.{weight,bias}, mlp.3.{weight,bias}; then encoder.ln.{weight,
 * bias}, heads.head.{weight [classes,H],bias}. */
#define SPU_VIT_NUM_TENSORS(layers) (4 + 12 * (layers) + 4)

spu_status spu_vit_replica_create(const spu_vit_desc *desc, const float *const *tensors,
                                  int32_t n_tensors, int32_t device, int32_t max_batch,
                                  int32_t n_workspaces, uint32_t flags, spu_replica **out);

/* pixels: DEVICE fp32 [batch, C, image, image] (NCHW, what StarPU hands the codelet as
 * buffers[0]->ptr); logits: DEVICE fp32 [batch, classes]. */
spu_status spu_vit_forward_async(spu_replica *r, int32_t ws_idx, const float *pixels,
                                 int32_t batch, float *logits, void *stream);
/* Measurement hook: device ms per forward of the ViT's dense GEMMs alone (patch embedding, QKV, attention-output,
 * MLP1, MLP2 of every layer, head): `reps` forwards' worth back to back in one CUDA graph, CUDA events on `stream`
 * (non-default).  Clobbers the workspace. */
spu_status spu_vit_gemm_chain_ms(spu_replica *r, int32_t ws_idx, const float *pixels, int32_t batch, float *logits,
                                 int32_t reps, void *stream, float *ms_per_forward);
uint64_t spu_vit_input_bytes(const spu_replica *r, int32_t batch);
uint64_t spu_vit_output_bytes(const spu_replica *r, int32_t batch);

/* ---- Stage 1: gather + pad + mask (device) --------------------------------
 * Ragged batch assembly: request i supplies len[i] int64 token ids at
 * src_ids[i] (and optionally len[i] mask words at src_mask[i]; src_mask or
 * src_mask[i] may be NULL = all ones).  Writes ids_out/mask_out int64 [n_req,seq]:
 * tokens then pad_id / 0.  src tables and len are DEVICE arrays of device-visible
 * pointers (device or pinned-mapped host memory).  With every len == seq this is
 * byte-identical to SlotManager::copy_job_inputs_to_slot
 * (src/starpu_task_worker/slot_manager_component.cpp:649-727). */
spu_status spu_gather_pad_mask_async(const int64_t *const *src_ids,
                                     const int64_t *const *src_mask,
                                     const int32_t *len, int32_t n_req,
                                     int32_t seq, int64_t pad_id,
                                     int64_t *ids_out, int64_t *mask_out,
                                     void *stream);

/* Dense byte gather: segment i (nbytes[i] bytes at srcs[i]) is copied to
 * dst + dst_off[i].  Used for fixed-shape inputs (ViT pixels, full-length ids).
 * All table arguments are DEVICE arrays. */
spu_status spu_gather_segments_async(const void *const *srcs,
                                     const uint64_t *dst_off,
                                     const uint64_t *nbytes, int32_t n_seg,
                                     void *dst, void *stream);

/* ---- Stage 4: scatter to per-request buffers ------------------------------
 * Segment i: nbytes[i] bytes from src + src_off[i] to dsts[i].  Replaces the
 * dim-0 slicing of ResultDispatcher::propagate_completion_to_sub_jobs
 * (result_dispatcher_component.cpp:678-739; batching_helpers.hpp:74-127) with a
 * device-side write into each request's own output buffer. */
spu_status spu_scatter_segments_async(const void *src, const uint64_t *src_off,
                                      const uint64_t *nbytes,
                                      void *const *dsts, int32_t n_seg,
                                      void *stream);

/* ---- Stage-level entry points (parity tests, benchmarks) ------------------ */
#define SPU_EPI_BIAS 0        /* out_bf16 = A W^T + bias                       */
#define SPU_EPI_BIAS_GELU 1   /* out_bf16 = gelu_erf(A W^T + bias)             */
#define SPU_EPI_BIAS_RESID 2  /* out_f32  = A W^T + bias + resid_f32           */

/* C = A[M,K](bf16,row-major) * W[N,K]^T(bf16,row-major) with fused epilogue on
 * tcgen05/TMEM, operands staged by TMA, output staged in smem and TMA-stored.
 * K % 64 == 0, N % 64 == 0, all pointers 16-byte aligned.  `out` is bf16 [M,N] for
 * BIAS/BIAS_GELU, fp32 [M,N] for BIAS_RESID (resid may alias out).
 * Creates its tensor maps per call (test path; the replica caches the weight maps). */
spu_status spu_gemm_bf16_async(const void *A, const void *W, const float *bias,
                               const float *resid, void *out, int32_t M,
                               int32_t N, int32_t K, int32_t epilogue,
                               void *stream);

/* FP8 E4M3 GEMM: A [M,K] and W [N,K] hold e4m3 bytes; C = (A W^T) * acc_scale + bias with
 * epilogue SPU_EPI_BIAS (bf16 out), SPU_EPI_BIAS_RESID (fp32 out = ... + resid) or
 * SPU_EPI_BIAS_GELU_E4M3 (e4m3 out = gelu(...) * out_inv_scale).  K % 128 == 0. */
#define SPU_EPI_BIAS_GELU_E4M3 3
spu_status spu_gemm_fp8_async(const void *A, const void *W, const float *bias,
                              const float *resid, void *out, int32_t M, int32_t N,
                              int32_t K, int32_t epilogue, float acc_scale,
                              float out_inv_scale, void *stream);
/* out_e4m3[i] = sat(in[i] * inv_scale), n % 4 == 0 (round to nearest even, saturating at 448) */
spu_status spu_quantize_e4m3_async(const float *in_f32, void *out, uint64_t n, float inv_scale,
                                   void *stream);

/* Same GEMM with the tile shape forced (tests / tile sweeps): cta_group 1 = one CTA per
 * 128 x block_n tile, 2 = a CTA pair per 256 x block_n tile; block_n in {64,128,192,256}
 * dividing N.  cta_group == 0 lets the library choose as spu_gemm_bf16_async does. */
spu_status spu_gemm_bf16_cfg_async(const void *A, const void *W, const float *bias,
                                   const float *resid, void *out, int32_t M,
                                   int32_t N, int32_t K, int32_t epilogue,
                                   int32_t cta_group, int32_t block_n, void *stream);

/* GEMM with the residual add AND the row LayerNorm fused into its epilogue (BERT's attention-output and FFN2
 * projections): y = LayerNorm(A W^T + bias + resid) * gamma + beta over the N columns of every row.  The CTAs
 * (cta_group 1) or CTA pairs (cta_group 2) that hold the N / block_n column blocks of one row block form ONE
 * thread-block cluster and exchange per-row (mean, M2) over distributed shared memory, so the pre-LayerNorm sum
 * never reaches HBM.  Writes y as fp32 (out_f32, may alias resid, may be NULL) and as bf16 (out_bf16, may be
 * NULL).  N / block_n <= 4 and block_n in {64,128,192} (cta_group 0 / block_n 0: the library chooses);
 * SPU_ERR_UNSUPPORTED when no such tiling spans N. */
spu_status spu_gemm_ln_bf16_async(const void *A, const void *W, const float *bias, const float *resid,
                                  const float *gamma, const float *beta, float eps, float *out_f32,
                                  void *out_bf16, int32_t M, int32_t N, int32_t K, int32_t cta_group,
                                  int32_t block_n, void *stream);

/* Fused masked-softmax attention.  qkv: bf16 [batch*seq, 3*heads*64] (q|k|v);
 * mask_bias: fp32 [batch, seq] additive (0 keep, large negative drop);
 * ctx: bf16 [batch*seq, heads*64].  head_dim is 64. */
spu_status spu_attention_async(const void *qkv, const float *mask_bias, void *ctx,
                               int32_t batch, int32_t seq, int32_t heads,
                               void *stream);
/* Same, with the implementation and output format selectable (tests / tools):
 * impl 0 = default, 1 = tcgen05/TMEM kernel, 2 = mma.sync kernel; out_e4m3 != 0 writes
 * ctx as e4m3 bytes of ctx * q_inv_scale (the FP8 variant's attention-output operand). */
spu_status spu_attention_cfg_async(const void *qkv, const float *mask_bias, void *ctx,
                                   int32_t batch, int32_t seq, int32_t heads, int32_t impl,
                                   int32_t out_e4m3, float q_inv_scale, void *stream);

/* Row LayerNorm: y = LN(x) * gamma + beta over the last dim (fp32 statistics).
 * x fp32 [rows, hidden]; writes fp32 out_f32 (may be NULL, may alias x) and bf16
 * out_bf16 (may be NULL). */
spu_status spu_layernorm_async(const float *x, const float *gamma,
                               const float *beta, float eps, float *out_f32,
                               void *out_bf16, int32_t rows, int32_t hidden,
                               void *stream);

/* Embedding lookup + LayerNorm + mask-bias builder (BERT stage 1b):
 * x = LN(word[ids] + pos[t] + type[0]); mask_bias = mask != 0 ? 0 : -1e30.
 * Tables are fp32 device arrays. */
spu_status spu_bert_embed_async(const int64_t *ids, const int64_t *mask,
                                const float *word, const float *pos,
                                const float *type0, const float *gamma,
                                const float *beta, float eps, int32_t vocab,
                                float *x_f32, void *x_bf16, float *mask_bias,
                                int32_t batch, int32_t seq, int32_t hidden,
                                void *stream);

/* Toy "model" used by the codelet known-answer tests the reference runs with
 * TorchScript x+1 / x+1.5 modules (tests/integration/starpu/
 * integration_starpu_setup.cpp:42-60, tests/unit/core/unit_starpu_setup.cpp:
 * 2332-2433): out[i] = in[i] * scale + shift, fp32, on the device. */
spu_status spu_affine_f32_async(const float *in, float *out, uint64_t n,
                                float scale, float shift, void *stream);

/* (The GEMM pipeline-trace / debug-mode hooks the measurement tools use are NOT part of this ABI: they exist
 * only in libraries built with -DSPU_DEBUG_HOOKS and are declared in include/spu_b200_debug.h.) */

/* ---- Misc ------------------------------------------------------------------ */
const char *spu_last_error(void);       /* thread-local message of the last failure */
int32_t spu_abi_version(void);
int32_t spu_device_count(void);         /* CUDA devices visible (0 without a GPU)   */
int32_t spu_device_is_sm100(int32_t device);

#ifdef __cplusplus
}
#endif
#endif /* SPU_B200_H */
