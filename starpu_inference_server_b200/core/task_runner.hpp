// Batch assembly (host gather), task submission and result slicing (host scatter): the
// callers either side of the codelet.  Restates, without torch::Tensor:
//   SlotManager::validate_batch_and_copy_inputs / copy_job_inputs_to_slot
//       /root/reference/src/starpu_task_worker/slot_manager_component.cpp:392-515, 649-727
//   resize_starpu_vector_handle   starpu_vector_resize_utils.hpp:19-89
//   can_merge_jobs                batch_composition_policy.cpp:102-151
//   aggregate_batch_metadata / resize_outputs_for_batch   batching_helpers.hpp:129-191
//   InferenceTask::create_task / fill_task_buffers        src/core/inference_task.cpp:690-822
//   starpu_output_callback -> acquire -> finalize          inference_task.cpp:850-1037
//   propagate_completion_to_sub_jobs / slice_outputs_for_sub_job
//       result_dispatcher_component.cpp:678-739, batching_helpers.hpp:74-127
#pragma once

#include <algorithm>
#include <atomic>
#include <functional>
#include <memory>
#include <optional>
#include <string>
#include <vector>

#include "inference_codelet.hpp"
#include "replica_registry.hpp"
#include "slot_pools.hpp"

namespace starpu_server {

// A borrowed, contiguous, row-major host array (what at::Tensor is to the reference here).
struct HostTensor {
  std::byte *data = nullptr;
  std::vector<int64_t> dims;
  ScalarType type = ScalarType::Float32;
  [[nodiscard]] auto numel() const -> size_t {
    size_t n = 1;
    for (int64_t d : dims) n *= static_cast<size_t>(d);
    return dims.empty() ? 0 : n;
  }
  [[nodiscard]] auto nbytes() const -> size_t { return numel() * element_size(type); }
  [[nodiscard]] auto defined() const -> bool { return data != nullptr; }
};

struct JobTiming {  // detail::TimingInfo subset (inference_job.hpp)
  MonotonicClock::time_point enqueued_time{}, before_starpu_submitted_time{}, codelet_start_time{},
      inference_start_time{}, codelet_end_time{}, callback_start_time{}, callback_end_time{};
};

struct JobResult {
  int request_id = 0;
  bool ok = false;
  std::string error;                // failure_info message
  std::vector<HostTensor> outputs;  // views of rows [off, off+b) of the output slot; valid inside on_complete
  int64_t rows = 0;                 // processed_length
  int worker_id = -1, device_id = -1;
  DeviceType executed_on = DeviceType::Unknown;
  JobTiming timing;
  int64_t batch_rows = 0;  // effective batch size of the task this request rode in
  int64_t batch_seq = 0;   // variable_length: padded sequence length of that task (0 = fixed shapes)
};

struct InferenceJob {
  int request_id = 0;
  std::vector<HostTensor> inputs;  // inputs[i].dims[0] = this request's batch
  std::optional<int> fixed_worker_id;
  // ModelConfig::device_visible_requests: where this request's response goes, one DEVICE-VISIBLE buffer per
  // model output laid out [batch, own seq, ...]; the codelet's scatter kernel writes it directly
  std::vector<void *> output_dsts;
  std::function<void(const JobResult &)> on_complete;
  JobTiming timing;
  [[nodiscard]] auto batch_size() const -> int64_t {
    return inputs.empty() || inputs[0].dims.empty() ? 1 : std::max<int64_t>(1, inputs[0].dims[0]);
  }
};

// batch_composition_policy.cpp:102-151.  variable_length: dims[1], the sequence
// length, may differ between the two requests; every other rule of the reference stands.
auto can_merge_jobs(const InferenceJob *lhs, const InferenceJob *rhs, bool variable_length = false) -> bool;

struct VectorResizeSpec {
  std::size_t element_count;
  std::size_t byte_count;
};
void resize_starpu_vector_interface(starpu_vector_interface *vi, VectorResizeSpec spec, bool is_input_handle);
void resize_starpu_vector_handle(starpu_data_handle_t handle, VectorResizeSpec spec, bool is_input_handle);

// Host gather of a master job + pending sub-jobs into a slot; returns bytes copied and
// resizes every handle to nx = sum(numel).
auto copy_job_inputs_to_slot(const InferenceJob *job, const std::vector<const InferenceJob *> &pending_jobs,
                             const std::vector<starpu_data_handle_t> &handles,
                             const std::vector<std::byte *> &base_ptrs,
                             const std::vector<HostBufferInfo> &buffer_infos) -> std::size_t;

// Variable-length host gather: like copy_job_inputs_to_slot, but every sample row of
// every request (dims [b, L, ...], L <= seq) is padded with zero bytes to `seq` (token id 0 = [PAD], mask 0),
// so the slot holds one contiguous [sum b, seq, ...] array per input; nx := sum(b) * seq * rest.
auto copy_job_inputs_to_slot_padded(const InferenceJob *job, const std::vector<const InferenceJob *> &pending_jobs,
                                    const std::vector<starpu_data_handle_t> &handles,
                                    const std::vector<std::byte *> &base_ptrs,
                                    const std::vector<HostBufferInfo> &buffer_infos, int64_t seq) -> std::size_t;

// SURVEY 8(f) row 2 (ModelConfig::device_visible_requests): describe the requests instead of gathering them.
// Writes into the first input slot one entry per sample row, in request order: address of its token ids, of its
// mask words, of its destination block ([own seq, hidden...] of that request's output buffer) and its length;
// resizes the two input handles to the table / to one element.  `io` receives the layout the codelet needs.
void write_request_table(const std::vector<std::shared_ptr<InferenceJob>> &jobs,
                         const std::vector<starpu_data_handle_t> &handles, const std::vector<std::byte *> &base_ptrs,
                         const std::vector<HostBufferInfo> &buffer_infos, int64_t rows, int64_t seq,
                         size_t token_bytes, detail::RequestIo *io);
// Is every request's attention_mask (input 1, int64) all non-zero?  Then the batch's only padding is the host
// gather's own and the rows' true lengths are the requests' dims[1] (packed forward, compact output).
auto masks_all_ones(const std::vector<std::shared_ptr<InferenceJob>> &jobs) -> bool;

struct SubJobSliceResult {
  std::vector<HostTensor> outputs;
  int64_t processed_length = 0;
};
// batching_helpers.hpp:74-127 (zero-copy views into the aggregated outputs)
auto slice_outputs_for_sub_job(const std::vector<HostTensor> &aggregated_outputs, int64_t offset,
                               int64_t batch_size) -> SubJobSliceResult;

struct ModelConfig {  // the slice of models/*.yml the path needs (config_loader.cpp)
  std::string name;
  std::vector<TensorConfig> inputs, outputs;
  int max_batch_size = 1;
  int pool_size = 12;  // models/bert.yml: pool_size
  // requests may carry any sequence length L <= inputs[*].dims[1]; a batch is padded to
  // the longest request it holds (mask synthesised by the zero padding), slots stay sized by the configured
  // dims[1] (= max_seq), and a single-sample request gets only its own L rows back.
  bool variable_length = false;
  // request inputs AND response buffers are device-visible (pinned host or device memory).
  // The host then gathers nothing: it writes a table of row addresses into the input slot and the codelet runs
  // gather/pad/mask -> encoder -> per-request scatter on the device (BERT models, fp32 responses).
  bool device_visible_requests = false;
  // > 0: input 0 holds int64 token ids that must lie in [0, id_vocab).  LibTorch's embedding raises on an id outside
  // its table and the reference turns that into a failed job; the device kernel clamps for memory safety, so the
  // host gather (which touches every id anyway) refuses the batch instead.
  int64_t id_vocab = 0;
};

// StarPUSetup + StarPUTaskRunner rolled into what the path needs: owns the codelet and the
// slot pools, knows the per-device replicas, submits one StarPU task per batch.
class TaskRunner {
 public:
  // `replicas` says which replica serves which device / worker (core/replica_registry.hpp); it must outlive
  // the runner.  Every task's InferenceParams::models is filled from it (fill_model_pointers).
  TaskRunner(ModelConfig model, const ReplicaRegistry *replicas);
  ~TaskRunner();

  // One task from jobs[0] (master) + jobs[1..] (aggregated sub-jobs, in order).  Blocks while
  // no slot is free.  Throws the reference's exception types on contract violations.
  void submit_batch(const std::vector<std::shared_ptr<InferenceJob>> &jobs);
  void wait_for_all();

  [[nodiscard]] auto codelet() -> starpu_codelet * { return codelet_.get_codelet(); }
  [[nodiscard]] auto input_pool() -> InputSlotPool & { return *input_pool_; }
  [[nodiscard]] auto output_pool() -> OutputSlotPool & { return *output_pool_; }
  [[nodiscard]] auto model() const -> const ModelConfig & { return model_; }
  [[nodiscard]] auto tasks_submitted() const -> long { return submitted_.load(); }
  // Completion fan-out: the per-request on_complete calls of one finished task (each makes that request's response
  // copy) are independent; an executor may run them on several threads.  It must return only when all have run --
  // the output slot is released right after.  Default: the callback thread runs them one after the other.
  using CompletionExecutor = std::function<void(std::vector<std::function<void()>> &)>;
  void set_completion_executor(CompletionExecutor ex) { completion_executor_ = std::move(ex); }

 private:
  struct CallbackContext;
  static void starpu_output_callback(void *arg);
  static void finalize(CallbackContext *ctx);
  [[nodiscard]] auto request_token_bytes() const -> size_t;
  [[nodiscard]] auto packed_forward_everywhere(int64_t batch) const -> bool;

  ModelConfig model_;
  const ReplicaRegistry *replicas_;
  InferenceCodelet codelet_;
  std::unique_ptr<InputSlotPool> input_pool_;
  std::unique_ptr<OutputSlotPool> output_pool_;
  std::atomic<long> submitted_{0};
  CompletionExecutor completion_executor_;
};

}  // namespace starpu_server
