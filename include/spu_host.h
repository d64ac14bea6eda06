/*
 * spu_host.h -- C-ABI of libspu_host.so: the C++ host side of the inference path
 * (StarPU-shaped codelet, slot pools, batch assembly, result slicing) driven through plain
 * C calls so that tests and bench.py can exercise exactly what a StarPU worker would run.
 *
 * What sits behind it mirrors the reference one to one (names kept):
 *   InferenceCodelet          src/core/starpu_setup.hpp:21-37, starpu_setup.cpp:559-846
 *   InputSlotPool/OutputSlotPool  src/core/input_slot_pool.cpp:235-328, output_slot_pool.cpp:226-330
 *   copy_job_inputs_to_slot   src/starpu_task_worker/slot_manager_component.cpp:649-727
 *   resize_starpu_vector_handle   src/starpu_task_worker/starpu_vector_resize_utils.hpp:19-89
 *   can_merge_jobs            src/starpu_task_worker/batch_composition_policy.cpp:102-151
 *   slice_outputs_for_sub_job src/starpu_task_worker/batching_helpers.hpp:74-127
 *   BatchCollector            src/starpu_task_worker/batch_collector_component.cpp:405-473
 * The runtime underneath is real StarPU when built with -DSPU_USE_REAL_STARPU, else the
 * in-repo stand-in (core/mini_starpu.cpp).  Nothing here computes on the CPU: every
 * inference goes through the codelet's cuda_func -> include/spu_b200.h.
 */
#ifndef SPU_HOST_H
#define SPU_HOST_H

#include <stddef.h>
#include <stdint.h>

#include "spu_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

#define SPU_HOST_MAX_IO 16
#define SPU_HOST_MAX_DIMS 8
#define SPU_HOST_MAX_DEVICES 32

/* element types of models/<name>.yml (TYPE_FP32, TYPE_INT64, ...) */
#define SPU_DT_F32 0
#define SPU_DT_F64 1
#define SPU_DT_I32 2
#define SPU_DT_I64 3
#define SPU_DT_I8 4
#define SPU_DT_U8 5
#define SPU_DT_F16 6
#define SPU_DT_BF16 7
#define SPU_DT_BOOL 8

#define SPU_MODEL_BERT 0
#define SPU_MODEL_VIT 1
#define SPU_MODEL_AFFINE 2 /* out = in*scale + shift: the reference tests' x+1 / x+1.5 / x*2 toys */

typedef struct spu_tensor_cfg {
  char name[64];
  int32_t dtype;
  int32_t ndims;
  int64_t dims[SPU_HOST_MAX_DIMS]; /* dims[0] = per-request batch (1 in the shipped configs) */
} spu_tensor_cfg;

typedef struct spu_server_cfg {
  /* use_cuda: [{device_ids: [...]}]  (config_loader_device.hpp:104-149) */
  int32_t n_devices;
  int32_t device_ids[SPU_HOST_MAX_DEVICES];
  int32_t nworker_per_cuda; /* STARPU_NWORKER_PER_CUDA (models/bert.yml:6-8); 0 = env / 1 */
  int32_t cuda_pipeline;    /* STARPU_CUDA_PIPELINE; 0 = env / 2 */
  /* model + batching (models/bert.yml: max_batch_size, pool_size, batch_coalesce_timeout_ms) */
  int32_t pool_size;
  int32_t max_batch_size;
  int32_t batch_coalesce_timeout_ms;
  int32_t n_inputs, n_outputs;
  spu_tensor_cfg inputs[SPU_HOST_MAX_IO], outputs[SPU_HOST_MAX_IO];
  int32_t model_kind;
  spu_replica *replicas[SPU_HOST_MAX_DEVICES]; /* replicas[i] serves device_ids[i] (BERT / ViT) */
  int32_t n_workspaces;                        /* arenas each replica was created with */
  float affine_scale, affine_shift;
  /* SURVEY 8(f) row 1 (variable-length admission): != 0 lets a request carry any sequence length
   * 1 <= seq_len <= inputs[*].dims[1]; requests of different lengths merge into one batch that the host
   * gather pads with zeros (token id 0 = [PAD], mask 0) to its longest request, slots stay sized by the
   * configured dims[1], and a single-sample request gets only its own seq_len rows back. */
  int32_t variable_length;
  /* variable_length servers, dynamic batching (spu_server_enqueue): > 0 buckets requests by
   * ceil(seq_len / length_bucket): a batch only takes requests of its oldest request's bucket (the collector
   * looks past other buckets' requests instead of closing the batch at the first one), so a batch pads to at
   * most length_bucket - 1 extra tokens per row instead of to the longest request in arrival order.  The oldest
   * queued request always opens the next batch, so no bucket starves.  0 = arrival order (the reference). */
  int32_t length_bucket;
  /* gpu_model_replication (config_loader_device.hpp): 0 = per_device: replicas[i] serves device_ids[i];
   * 1 = per_worker: one replica per CUDA worker, replicas[] ordered device by device then by StarPU worker id
   * (n_devices x nworker_per_cuda entries), as build_gpu_replica_assignments orders them
   * (src/core/inference_runner.cpp:128-183). */
  int32_t gpu_model_replication;
  /* SURVEY 8(f) row 4 (tile-aware batching): > 0 = row tile of the codelet's GEMMs (256 for the CTA-pair tiles of
   * libspu_b200).  When the coalesce window of a dynamic batch closes on rows x seq token rows that end inside
   * a tile, the collector waits a bounded grace (a quarter of batch_coalesce_timeout_ms) for the requests that
   * complete the tile and takes no more than complete it.  0 = the reference's timer-only rule. */
  int32_t tile_rows;
  /* SURVEY 8(f) row 2 (device-resident gather / scatter): != 0 declares that every request's input buffers AND its
   * output buffer are DEVICE-VISIBLE (cudaHostAlloc / cudaHostRegister'ed host memory, or device memory).  The
   * host then copies nothing: the task's first input slot carries a table of row addresses, and the codelet runs
   * gather + pad + mask -> encoder -> per-request scatter on the GPU (spu_bert_forward_requests_async), writing
   * each response straight into the request's own buffer ([batch, own seq, hidden] fp32).  BERT models whose
   * replicas were created with SPU_REQUEST_IO; every request must bring an output buffer.  Replaces
   * copy_job_inputs_to_slot (slot_manager_component.cpp:649-727) and slice_outputs_for_sub_job
   * (batching_helpers.hpp:74-127) on that path. */
  int32_t device_visible_requests;
  /* Warm-up autotune (src/core/warmup.cpp:226-299 is the reference's warm-up hook): != 0 runs
   * spu_bert_autotune on every BERT replica for the full batch shape (max_batch_size x inputs[0].dims[1]) before
   * the first task, so the GEMM tiles of the steady-state batch are measured winners, not cost-model guesses. */
  int32_t warmup_autotune;
} spu_server_cfg;

typedef struct spu_request {
  int32_t request_id;
  int64_t batch;                         /* rows this request carries (dims[0]) */
  const void *inputs[SPU_HOST_MAX_IO];   /* host, contiguous [batch, ...] per model input */
  void *outputs[SPU_HOST_MAX_IO];        /* host destination per model output (may be NULL); sized by THIS
                                            request's shape: [batch, seq_len, ...] on a variable_length server
                                            (the response is compacted to the request's own seq_len even when
                                            it rode in a longer padded batch), [batch, dims[1..]] otherwise */
  int64_t seq_len;                       /* variable_length servers: dims[1] of this request's inputs
                                            (0 = the configured length); ignored otherwise */
} spu_request;

typedef struct spu_result {
  int32_t request_id;
  int32_t ok;           /* 1 = outputs delivered */
  int64_t rows;         /* processed_length of this request's slice */
  int64_t batch_rows;   /* effective batch of the task it rode in */
  int32_t worker_id, device_id, executed_on; /* executed_on: 0 CPU, 1 CUDA, 2 unknown */
  double codelet_ms;    /* codelet_end - codelet_start (host time of the async enqueue) */
  double task_ms;       /* before_starpu_submitted -> callback_start (task latency) */
  double latency_ms;    /* enqueued -> callback_end */
  float first_value;    /* outputs[0] row 0 element 0 read from the landed slot (fp32 outputs) */
  char error[192];
  int64_t batch_seq;    /* variable_length servers: padded sequence length of the task it rode in (else 0) */
} spu_result;

typedef struct spu_server spu_server;

/* starpu_init + codelet + slot pools + replica table.  One server per process (StarPU is a
 * process-wide singleton). */
int32_t spu_server_create(const spu_server_cfg *cfg, spu_server **out);
void spu_server_destroy(spu_server *s);

/* One StarPU task from reqs[0] (master) + reqs[1..n) (aggregated sub-jobs, row order =
 * request order).  fixed_worker >= 0 pins the task (warm-up).  Returns a ticket >= 0, or
 * -1 with spu_host_last_error() set. */
int64_t spu_server_submit_batch(spu_server *s, const spu_request *reqs, int32_t n, int32_t fixed_worker);
/* Dynamic batching: the collector thread merges compatible requests up to max_batch_size
 * rows or batch_coalesce_timeout_ms.  Ticket covers this one request. */
int64_t spu_server_enqueue(spu_server *s, const spu_request *req);
/* Blocks until every request of the ticket has completed; fills up to cap results (in
 * request order) and returns how many the ticket held.  A ticket can be waited once. */
int32_t spu_server_wait(spu_server *s, int64_t ticket, spu_result *results, int32_t cap);
int32_t spu_server_wait_all(spu_server *s);
/* Measurement (bench.py's offered-load leg): an OPEN-LOOP load generator in front of the batch collector.  n_req
 * requests arrive as a Poisson process of `rate_per_s` (seeded), request i is rows[i % n_rows] (per-request host
 * buffers the caller keeps alive) passed to spu_server_enqueue; afterwards every ticket is waited.  Per request
 * (arrays of n_req, may be NULL): codelet_ms = codelet_end - codelet_start of the task it rode in
 * (starpu_setup.cpp:645-649,718-722; the host time of the asynchronous enqueue), task_ms = task submitted ->
 * output callback (device work + D2H included), latency_ms = enqueued -> callback end (response copied),
 * batch_rows = the effective batch of that task.  Returns the number of failed requests, -1 on a usage / submission error. */
int32_t spu_server_offered_load(spu_server *s, const spu_request *rows, int32_t n_rows, int64_t n_req,
                                double rate_per_s, uint64_t seed, double *codelet_ms, double *task_ms,
                                double *latency_ms, int32_t *batch_rows, double *elapsed_s);
/* counters: [0] tasks submitted, [1] requests completed, [2] free input slots, [3] free
 * output slots, [4] StarPU workers, [5] codelet failures */
void spu_server_stats(spu_server *s, int64_t out[6]);

const char *spu_host_last_error(void);

/* ---- unit-test seams (no runtime needed; mirror the reference's unit tests) ---------- */
/* InferenceCodelet field values (unit_starpu_setup.cpp:4065-4074):
 * out = {nbuffers, type, max_parallelism, cpu_funcs[0]!=0, cuda_funcs[0]!=0, cuda_flags[0], where} */
void spu_host_codelet_fields(int64_t out[7]);
/* buffer_byte_size (starpu_setup.cpp:515-542); returns 0 ok, 1 exception (message in last_error) */
int32_t spu_host_buffer_byte_size(int32_t interface_id, uint32_t nx, uint64_t elemsize, int32_t null_buffer,
                                  uint64_t *out);
/* select_gpu_module (starpu_setup.cpp:725-778) over a table of n_models dummy replicas:
 * returns the chosen index, or -1 on StarPUCodeletException (message in last_error). */
int32_t spu_host_select_replica(const int32_t *device_ids, int32_t n_device_ids, const int32_t *worker_ids,
                                int32_t n_worker_ids, int32_t n_models, int32_t worker_id, int32_t device_id);
/* NvtxRange around the codelet body (starpu_setup.cpp:817; src/utils/nvtx.hpp SetHooks): enable != 0 routes the
 * ranges to a recorder instead of NVTX, 0 restores NVTX.  Reports the last label pushed and {pushes, pops}. */
int32_t spu_host_nvtx_capture(int32_t enable, char *last_label, uint64_t cap, int64_t counts[2]);
/* build_gpu_replica_assignments (inference_runner.cpp:128-183) + ReplicaRegistry::fill_model_pointers
 * (inference_task.cpp:469-516) + select_gpu_module, without a runtime: the CUDA workers are described by
 * worker_dev[w] = device of StarPU worker w.  Writes the assignments (device, worker or -1) and returns their
 * count; `pick_*` (when >= 0) additionally selects the replica for that worker / device and stores its
 * assignment index in *picked.  -1 on an exception (message in last_error). */
int32_t spu_host_replica_assignments(const int32_t *device_ids, int32_t n_device_ids, int32_t per_worker,
                                     const int32_t *worker_dev, int32_t n_workers, int32_t *out_device,
                                     int32_t *out_worker, int32_t cap, int32_t pick_worker, int32_t pick_device,
                                     int32_t *picked);
/* copy_job_inputs_to_slot over one input: n_jobs host arrays back to back into a slot of
 * `capacity` bytes registered with elemsize; returns 0 ok / 1 exception.  nx_out = the
 * handle's nx afterwards on node 0. */
int32_t spu_host_gather_to_slot(const void *const *srcs, const uint64_t *nbytes, int32_t n_jobs, uint64_t elemsize,
                                void *slot, uint64_t capacity, uint64_t *nx_out, uint64_t *bytes_out);
/* slice_outputs_for_sub_job over one aggregated [rows, row_bytes] output: returns the byte
 * offset and length of the view through *off_out / *len_out and the processed length. */
int64_t spu_host_slice_rows(uint64_t rows, uint64_t row_bytes, int64_t offset, int64_t batch_size, uint64_t *off_out,
                            uint64_t *len_out);
/* the collector's tile-aware rule: 1 when a batch of rows x seq token rows ends inside a GEMM row tile and may
 * still grow (rows < max_rows) */
int32_t spu_host_tile_aware_wants_more(int64_t rows, int64_t seq, int64_t max_rows, int32_t tile_rows);
/* write_request_table (spu_server_cfg.device_visible_requests) over n_jobs requests of batch[j] samples x len[j]
 * tokens whose ids / mask / response buffers live at the given addresses (dst 0 = no response buffer): writes the
 * table into `slot`; out = {rows, off_mask, off_dst, off_len, table_bytes, full_rows, nx of the table handle, nx of
 * the second input handle}.  0 ok / 1 exception (message in last_error). */
int32_t spu_host_request_table(const uint64_t *ids_ptrs, const uint64_t *mask_ptrs, const uint64_t *dst_ptrs,
                               const int64_t *batch, const int64_t *len, int32_t n_jobs, int64_t seq,
                               uint64_t token_bytes, void *slot, uint64_t capacity_bytes, int64_t out[8]);
/* masks_all_ones over n_jobs int64 mask arrays: 1 when the packed forward may be used for the batch */
int32_t spu_host_masks_all_ones(const int64_t *const *masks, const int64_t *numel, int32_t n_jobs);
/* can_merge_jobs over two single-input requests described by dims/dtype */
int32_t spu_host_can_merge(const int64_t *dims_a, int32_t nd_a, int32_t dt_a, const int64_t *dims_b, int32_t nd_b,
                           int32_t dt_b);
/* the same with the variable-length relaxation (dims[1] may differ) */
int32_t spu_host_can_merge_varlen(const int64_t *dims_a, int32_t nd_a, int32_t dt_a, const int64_t *dims_b,
                                  int32_t nd_b, int32_t dt_b);
/* copy_job_inputs_to_slot_padded over one int64 input: request j is one [1, len[j]] row, every row is padded
 * to `seq`; returns 0 ok / 1 exception; nx_out = the handle's nx afterwards. */
int32_t spu_host_gather_padded(const int64_t *const *srcs, const int64_t *len, int32_t n_jobs, int64_t seq,
                               int64_t *slot, uint64_t capacity_bytes, uint64_t *nx_out);
/* Calls codelet.cuda_funcs[0] directly, like unit_starpu_setup.cpp:2332-2433: DEVICE fp32
 * in/out of n elements, an affine "model" on `device`, mocked worker id / device id / stream.
 * rec = {device_id, worker_id, executed_on, start!=0, end!=0, inflight gauge after}.
 * Returns 0, or 1 when the codelet threw (message in last_error). */
int32_t spu_host_call_cuda_func_affine(const float *in_dev, float *out_dev, uint64_t n, float scale, float shift,
                                       int32_t device, int32_t worker_id, void *stream, int32_t n_models,
                                       int32_t use_vector_iface, int64_t rec[6]);

#ifdef __cplusplus
}
#endif
#endif /* SPU_HOST_H */
