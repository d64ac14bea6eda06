#include "task_runner.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>

namespace starpu_server {

auto can_merge_jobs(const InferenceJob *lhs, const InferenceJob *rhs, bool variable_length) -> bool {
  if (!lhs || !rhs) return false;
  if (lhs->inputs.size() != rhs->inputs.size()) return false;
  for (size_t i = 0; i < lhs->inputs.size(); ++i)
    if (lhs->inputs[i].type != rhs->inputs[i].type) return false;
  for (size_t i = 0; i < lhs->inputs.size(); ++i) {
    const HostTensor &a = lhs->inputs[i], &b = rhs->inputs[i];
    if (!a.defined() || !b.defined()) return false;
    if (a.dims.size() != b.dims.size()) return false;
    if (a.dims.empty()) return false;
    for (size_t d = 1; d < a.dims.size(); ++d)
      if (a.dims[d] != b.dims[d] && !(variable_length && d == 1)) return false;
  }
  return true;
}

void resize_starpu_vector_interface(starpu_vector_interface *vi, VectorResizeSpec spec, bool is_input_handle) {
  if (vi == nullptr) return;
  const auto elem_size = vi->elemsize;
  if (elem_size == 0) throw StarPUDataAcquireException("StarPU vector interface reported zero element size");
  if (spec.byte_count % elem_size != 0) {
    const std::string msg = std::string(is_input_handle ? "Input" : "Output") + " tensor byte size (" +
                            std::to_string(spec.byte_count) + ") is not divisible by element size (" +
                            std::to_string(elem_size) + ")";
    if (is_input_handle) throw InvalidInputTensorException(msg);
    throw InvalidInferenceJobException(msg);
  }
  spec.element_count = spec.byte_count / elem_size;
  const auto alloc_size = vi->allocsize;
  if (alloc_size != std::numeric_limits<size_t>::max() && alloc_size != 0 && spec.byte_count > alloc_size) {
    const std::string msg = std::string(is_input_handle ? "Input" : "Output") + " tensor requires " +
                            std::to_string(spec.byte_count) + " bytes but slot capacity is " +
                            std::to_string(alloc_size) + " bytes";
    if (is_input_handle) throw InputPoolCapacityException(msg);
    throw InvalidInferenceJobException(msg);
  }
  vi->nx = static_cast<uint32_t>(spec.element_count);
}

void resize_starpu_vector_handle(starpu_data_handle_t handle, VectorResizeSpec spec, bool is_input_handle) {
  if (handle == nullptr) throw StarPUDataAcquireException("StarPU vector handle is null");
  if (starpu_data_get_interface_id(handle) != STARPU_VECTOR_INTERFACE_ID)
    throw StarPUDataAcquireException("Expected StarPU vector interface for handle");
  const unsigned memory_nodes = starpu_memory_nodes_get_count();
  for (unsigned node = 0; node < memory_nodes; ++node) {
    auto *raw = starpu_data_get_interface_on_node(handle, node);
    if (raw == nullptr) continue;
    resize_starpu_vector_interface(static_cast<starpu_vector_interface *>(raw), spec, is_input_handle);
  }
}

auto copy_job_inputs_to_slot(const InferenceJob *job, const std::vector<const InferenceJob *> &pending_jobs,
                             const std::vector<starpu_data_handle_t> &handles,
                             const std::vector<std::byte *> &base_ptrs,
                             const std::vector<HostBufferInfo> &buffer_infos) -> std::size_t {
  if (!job) return 0;
  std::size_t total_bytes_copied = 0;
  for (std::size_t input_idx = 0; input_idx < job->inputs.size(); ++input_idx) {
    std::byte *dst = base_ptrs[input_idx];
    const std::size_t capacity = buffer_infos[input_idx].bytes;
    std::size_t offset = 0, total_numel = 0, total_bytes = 0;
    auto copy_tensor = [&](const HostTensor &t) {
      if (!t.defined())
        throw InvalidInputTensorException("input[" + std::to_string(input_idx) + "] is not a defined host tensor");
      const size_t bytes = t.nbytes();
      if (capacity < offset + bytes)
        throw InputPoolCapacityException("Input tensor exceeds allocated slot buffer capacity");
      std::memcpy(dst + offset, t.data, bytes);
      offset += bytes;
      total_bytes += bytes;
      total_numel += t.numel();
      total_bytes_copied += bytes;
    };
    copy_tensor(job->inputs[input_idx]);
    for (const InferenceJob *pending : pending_jobs) {
      if (!pending) continue;
      if (input_idx >= pending->inputs.size())
        throw InconsistentInputTensorCountException("Inconsistent input tensor count");
      copy_tensor(pending->inputs[input_idx]);
    }
    resize_starpu_vector_handle(handles[input_idx], VectorResizeSpec{total_numel, total_bytes}, true);
  }
  return total_bytes_copied;
}

auto copy_job_inputs_to_slot_padded(const InferenceJob *job, const std::vector<const InferenceJob *> &pending_jobs,
                                    const std::vector<starpu_data_handle_t> &handles,
                                    const std::vector<std::byte *> &base_ptrs,
                                    const std::vector<HostBufferInfo> &buffer_infos, int64_t seq) -> std::size_t {
  if (!job) return 0;
  std::size_t total_bytes_copied = 0;
  for (std::size_t input_idx = 0; input_idx < job->inputs.size(); ++input_idx) {
    std::byte *dst = base_ptrs[input_idx];
    const std::size_t capacity = buffer_infos[input_idx].bytes;
    std::size_t offset = 0, total_numel = 0;
    auto copy_tensor = [&](const HostTensor &t) {
      if (!t.defined())
        throw InvalidInputTensorException("input[" + std::to_string(input_idx) + "] is not a defined host tensor");
      if (t.dims.size() < 2 || t.dims[0] < 1 || t.dims[1] < 1 || t.dims[1] > seq)
        throw InvalidInputTensorException("variable-length input needs dims [batch, seq <= " + std::to_string(seq) +
                                          ", ...]");
      const size_t samples = static_cast<size_t>(t.dims[0]);
      const size_t row_bytes = t.nbytes() / samples;                                  // L * rest * elem
      const size_t padded_bytes = row_bytes / static_cast<size_t>(t.dims[1]) * static_cast<size_t>(seq);
      if (capacity < offset + samples * padded_bytes)
        throw InputPoolCapacityException("Input tensor exceeds allocated slot buffer capacity");
      for (size_t r = 0; r < samples; ++r) {
        std::memcpy(dst + offset, t.data + r * row_bytes, row_bytes);
        std::memset(dst + offset + row_bytes, 0, padded_bytes - row_bytes);
        offset += padded_bytes;
      }
      total_numel += padded_bytes / element_size(t.type) * samples;
      total_bytes_copied += samples * row_bytes;
    };
    copy_tensor(job->inputs[input_idx]);
    for (const InferenceJob *pending : pending_jobs) {
      if (!pending) continue;
      if (input_idx >= pending->inputs.size())
        throw InconsistentInputTensorCountException("Inconsistent input tensor count");
      copy_tensor(pending->inputs[input_idx]);
    }
    resize_starpu_vector_handle(handles[input_idx], VectorResizeSpec{total_numel, offset}, true);
  }
  return total_bytes_copied;
}

auto slice_outputs_for_sub_job(const std::vector<HostTensor> &aggregated_outputs, int64_t offset,
                               int64_t batch_size) -> SubJobSliceResult {
  SubJobSliceResult result;
  const int64_t slice_size = std::max<int64_t>(1, batch_size);
  result.processed_length = slice_size;
  if (aggregated_outputs.empty()) return result;
  bool determined_length = false;
  for (const HostTensor &t : aggregated_outputs) {
    if (!t.defined() || t.dims.empty()) {
      result.outputs.push_back(t);
      continue;
    }
    const int64_t available = t.dims[0];
    const int64_t slice_end = std::min<int64_t>(available, offset + slice_size);
    const int64_t length = std::max<int64_t>(0, slice_end - offset);
    if (length <= 0 || offset < 0) {
      result.outputs.emplace_back();
      continue;
    }
    if (!determined_length) {
      result.processed_length = length;
      determined_length = true;
    }
    const size_t sample_bytes = t.nbytes() / static_cast<size_t>(available);
    HostTensor view = t;
    view.data = t.data + static_cast<size_t>(offset) * sample_bytes;
    view.dims[0] = length;
    result.outputs.push_back(std::move(view));
  }
  return result;
}

// Is every request's second input (attention_mask, int64 [batch, L]) all non-zero?  Then the only padding of the
// batch is the host gather's own and the rows' true lengths are the requests' L.
auto masks_all_ones(const std::vector<std::shared_ptr<InferenceJob>> &jobs) -> bool {
  for (const auto &j : jobs) {
    if (!j) continue;
    if (j->inputs.size() != 2 || j->inputs[1].type != ScalarType::Int64 || !j->inputs[1].defined()) return false;
    const auto *m = reinterpret_cast<const int64_t *>(j->inputs[1].data);
    const size_t n = j->inputs[1].numel();
    for (size_t i = 0; i < n; ++i)
      if (m[i] == 0) return false;
  }
  return true;
}

// instead of gathering the requests' bytes, describe them.  One entry per sample row of the
// batch, in request order: where its token ids and mask words live, how many there are, and where its hidden
// states go ([own seq, hidden] fp32 on a variable_length server, [seq, hidden] otherwise).  The table rides in the
// first input slot (what StarPU copies to the device is these few bytes); the second slot shrinks to one element.
void write_request_table(const std::vector<std::shared_ptr<InferenceJob>> &jobs,
                         const std::vector<starpu_data_handle_t> &handles, const std::vector<std::byte *> &base_ptrs,
                         const std::vector<HostBufferInfo> &buffer_infos, int64_t rows, int64_t seq,
                         size_t token_bytes, detail::RequestIo *io) {
  if (handles.size() != 2 || base_ptrs.size() != 2)
    throw InputPoolMismatchException("device_visible_requests needs the two BERT inputs (input_ids, attention_mask)");
  const size_t n = static_cast<size_t>(rows);
  io->enabled = true;
  io->rows = rows;
  io->seq = seq;
  io->off_mask = n * sizeof(void *);
  io->off_dst = 2 * n * sizeof(void *);
  io->off_len = 3 * n * sizeof(void *);
  io->table_bytes = (io->off_len + n * sizeof(int32_t) + 7) / 8 * 8;
  if (buffer_infos[0].bytes < io->table_bytes)
    throw InputPoolCapacityException("Request table exceeds allocated slot buffer capacity");
  std::byte *base = base_ptrs[0];
  auto **ids_tab = reinterpret_cast<const void **>(base);
  auto **mask_tab = reinterpret_cast<const void **>(base + io->off_mask);
  auto **dst_tab = reinterpret_cast<void **>(base + io->off_dst);
  auto *len_tab = reinterpret_cast<int32_t *>(base + io->off_len);
  bool ragged = false;
  size_t r = 0;
  for (const auto &job : jobs) {
    if (!job) continue;
    if (job->inputs.size() != 2 || job->inputs[0].dims.size() != 2 || job->inputs[0].type != ScalarType::Int64 ||
        job->inputs[1].type != ScalarType::Int64 || job->inputs[1].dims != job->inputs[0].dims)
      throw InvalidInputTensorException("device_visible_requests: inputs must be two int64 [batch, seq] arrays");
    if (job->output_dsts.empty() || job->output_dsts[0] == nullptr)
      throw InvalidInferenceJobException("device_visible_requests: every request needs a device-visible output buffer");
    const int64_t b = job->inputs[0].dims[0], L = job->inputs[0].dims[1];
    if (L < 1 || L > seq) throw InvalidInputTensorException("device_visible_requests: request longer than the batch");
    ragged = ragged || L != seq;
    for (int64_t s = 0; s < b; ++s, ++r) {
      if (r >= n) throw InputPoolCapacityException("Batch size exceeds input pool capacity");
      ids_tab[r] = job->inputs[0].data + static_cast<size_t>(s * L) * sizeof(int64_t);
      mask_tab[r] = job->inputs[1].data + static_cast<size_t>(s * L) * sizeof(int64_t);
      len_tab[r] = static_cast<int32_t>(L);
      dst_tab[r] = static_cast<std::byte *>(job->output_dsts[0]) + static_cast<size_t>(s * L) * token_bytes;
    }
  }
  if (r != n) throw InvalidInferenceJobException("device_visible_requests: row count mismatch");
  // sample s of a request lands at its own [s] slice of that request's [batch, own seq, hidden] buffer
  io->full_rows = !ragged;
  resize_starpu_vector_handle(handles[0], VectorResizeSpec{io->table_bytes / 8, io->table_bytes}, true);
  resize_starpu_vector_handle(handles[1], VectorResizeSpec{1, sizeof(int64_t)}, true);
}

// ---------------------------------------------------------------------------------
struct TaskRunner::CallbackContext {
  TaskRunner *runner = nullptr;
  std::vector<std::shared_ptr<InferenceJob>> jobs;  // [0] = master
  std::unique_ptr<InferenceParams> params;
  int input_slot = -1, output_slot = -1;
  std::vector<starpu_data_handle_t> outputs_handles;
  int64_t effective_batch = 0;
  int64_t seq = 0;  // variable_length: the padded sequence length of this batch (0 = fixed shapes)
  bool request_io = false;  // responses were written by the codelet into the requests' own buffers
  bool compact = false;     // output 0 holds the batch's true tokens back to back (packed forward)
  int worker_id = -1, device_id = -1;
  DeviceType executed_on = DeviceType::Unknown;
  JobTiming timing;
  std::vector<bool> acquired;
};

TaskRunner::TaskRunner(ModelConfig model, const ReplicaRegistry *replicas)
    : model_(std::move(model)), replicas_(replicas) {
  if (replicas_ == nullptr) throw std::invalid_argument("TaskRunner needs a replica registry");
  if (replicas_->size() > InferLimits::MaxModelsGPU) throw std::invalid_argument("too many GPU replicas");
  if (model_.inputs.empty() || model_.inputs.size() > InferLimits::MaxInputs)
    throw std::invalid_argument("model must declare 1..16 inputs");
  if (model_.outputs.empty()) throw std::invalid_argument("model must declare outputs");
  if (model_.device_visible_requests &&
      (model_.inputs.size() != 2 || model_.outputs.size() != 1 || model_.outputs[0].dims.size() < 3 ||
       model_.outputs[0].type != ScalarType::Float32 || model_.inputs[0].dims.size() != 2))
    throw std::invalid_argument("device_visible_requests serves BERT-shaped models: int64 [1, S] x 2 -> fp32 [1, S, H]");
  const bool pinned = replicas_->size() > 0;
  input_pool_ = std::make_unique<InputSlotPool>(model_.inputs, model_.max_batch_size, model_.pool_size, pinned);
  output_pool_ = std::make_unique<OutputSlotPool>(model_.outputs, model_.max_batch_size, model_.pool_size, pinned);
  // Output slots are only ever consumed on the host (starpu_output_callback acquires them, inference_task.cpp:
  // 907-935): mark them write-through to main RAM, so the runtime starts the D2H as soon as the codelet's
  // kernels finish instead of when the callback asks for it.  SPU_NO_WRITE_THROUGH=1 restores fetch-on-acquire.
  if (pinned && std::getenv("SPU_NO_WRITE_THROUGH") == nullptr)
    for (int s = 0; s < output_pool_->num_slots(); ++s)
      for (starpu_data_handle_t h : output_pool_->handles(s)) starpu_data_set_wt_mask(h, 1u << STARPU_MAIN_RAM);
}

TaskRunner::~TaskRunner() { wait_for_all(); }

// Does every replica a task may land on implement the packed forward for this batch size?  (The host decides the
// output layout before it knows which worker takes the task.)
auto TaskRunner::packed_forward_everywhere(int64_t batch) const -> bool {
  if (replicas_->size() == 0 || batch > INT32_MAX) return false;
  for (size_t i = 0; i < replicas_->size(); ++i) {
    const DeviceModel *m = replicas_->model(i);
    if (m == nullptr || m->kind != DeviceModel::Kind::Bert ||
        spu_bert_packed_supported(m->replica, static_cast<int32_t>(batch)) != 1)
      return false;
  }
  return true;
}

// bytes of one token row of output 0 ([batch, seq, hidden...]): hidden... x element size
auto TaskRunner::request_token_bytes() const -> size_t {
  size_t n = element_size(model_.outputs[0].type);
  for (size_t d = 2; d < model_.outputs[0].dims.size(); ++d) n *= static_cast<size_t>(model_.outputs[0].dims[d]);
  return n;
}

void TaskRunner::wait_for_all() { starpu_task_wait_for_all(); }

void TaskRunner::submit_batch(const std::vector<std::shared_ptr<InferenceJob>> &jobs) {
  if (jobs.empty() || !jobs.front()) throw InvalidInferenceJobException("empty batch");
  const InferenceJob *master = jobs.front().get();
  if (master->inputs.size() != model_.inputs.size())
    throw InputPoolMismatchException("Input count mismatch between job and slot");
  int64_t effective_batch = 0;
  std::vector<const InferenceJob *> pending;
  for (size_t k = 0; k < jobs.size(); ++k) {
    if (!jobs[k]) continue;
    if (k > 0) {
      if (!can_merge_jobs(master, jobs[k].get(), model_.variable_length))
        throw InvalidInferenceJobException("sub-job cannot be merged with the master job (shape/type mismatch)");
      pending.push_back(jobs[k].get());
    }
    effective_batch += jobs[k]->batch_size();  // aggregate_batch_metadata: sum of sub batch sizes
  }
  if (effective_batch < 1 || effective_batch > input_pool_->max_batch_size())
    throw InputPoolCapacityException("Batch size exceeds input pool capacity");
  // variable_length: the batch is padded to its longest request; the configured dims[1] is the cap
  int64_t seq = 0, seq_cap = 0;
  if (model_.variable_length) {
    if (model_.inputs[0].dims.size() < 2) throw std::invalid_argument("variable_length needs inputs of rank >= 2");
    seq_cap = model_.inputs[0].dims[1];
    for (const auto &j : jobs) {
      if (!j) continue;
      for (const HostTensor &t : j->inputs) {
        if (t.dims.size() < 2 || t.dims[1] < 1 || t.dims[1] > seq_cap || t.dims[1] != j->inputs[0].dims[1])
          throw InvalidInputTensorException("variable-length request: every input needs the same seq in [1, " +
                                            std::to_string(seq_cap) + "]");
      }
      seq = std::max(seq, j->inputs[0].dims[1]);
    }
  }

  auto ctx = std::make_unique<CallbackContext>();
  ctx->runner = this;
  ctx->jobs = jobs;
  ctx->effective_batch = effective_batch;
  ctx->seq = seq;
  ctx->input_slot = input_pool_->acquire();
  ctx->output_slot = output_pool_->acquire();
  struct SlotGuard {
    TaskRunner *r;
    CallbackContext *c;
    bool armed = true;
    ~SlotGuard() {
      if (!armed) return;
      r->input_pool_->release(c->input_slot);
      r->output_pool_->release(c->output_slot);
    }
  } guard{this, ctx.get()};

  const auto &in_handles = input_pool_->handles(ctx->input_slot);
  const auto &out_handles = output_pool_->handles(ctx->output_slot);
  detail::RequestIo request_io{};
  // SlotHandleLease(handles, STARPU_W) around the host gather (slot_manager_component.cpp:425)
  size_t leased = 0;
  try {
    for (; leased < in_handles.size(); ++leased)
      if (starpu_data_acquire(in_handles[leased], STARPU_W) != 0)
        throw StarPUDataAcquireException("starpu_data_acquire failed");
    if (model_.device_visible_requests)
      write_request_table(jobs, in_handles, input_pool_->base_ptrs(ctx->input_slot),
                          input_pool_->host_buffer_infos(ctx->input_slot), effective_batch,
                          seq > 0 ? seq : model_.inputs[0].dims.at(1), request_token_bytes(), &request_io);
    else if (model_.variable_length)
      copy_job_inputs_to_slot_padded(master, pending, in_handles, input_pool_->base_ptrs(ctx->input_slot),
                                     input_pool_->host_buffer_infos(ctx->input_slot), seq);
    else
      copy_job_inputs_to_slot(master, pending, in_handles, input_pool_->base_ptrs(ctx->input_slot),
                              input_pool_->host_buffer_infos(ctx->input_slot));
    if (model_.id_vocab > 0 && !model_.device_visible_requests && model_.inputs[0].type == ScalarType::Int64)
      for (const auto &j : jobs) {
        if (!j) continue;
        const auto *ids = reinterpret_cast<const int64_t *>(j->inputs[0].data);
        const size_t n = j->inputs[0].numel();
        for (size_t i = 0; i < n; ++i)
          if (ids[i] < 0 || ids[i] >= model_.id_vocab)
            throw InvalidInputTensorException("token id " + std::to_string(ids[i]) + " of request " +
                                              std::to_string(j->request_id) + " is outside the embedding table [0, " +
                                              std::to_string(model_.id_vocab) + ")");
      }
  } catch (...) {
    for (size_t i = 0; i < leased; ++i) starpu_data_release(in_handles[i]);
    throw;
  }
  for (size_t i = 0; i < leased; ++i) starpu_data_release(in_handles[i]);

  // variable_length + every mask all ones + replicas that have the packed forward: the codelet gets the true
  // lengths and returns output 0 in the compact layout
  std::vector<int32_t> seq_lens;
  if (model_.variable_length && !request_io.enabled && model_.inputs.size() == 2 && model_.outputs.size() == 1 &&
      model_.outputs[0].dims.size() >= 2 && model_.outputs[0].dims[1] == seq_cap && packed_forward_everywhere(effective_batch) &&
      masks_all_ones(jobs))
    for (const auto &j : jobs)
      if (j)
        for (int64_t s = 0; s < j->batch_size(); ++s) seq_lens.push_back(static_cast<int32_t>(j->inputs[0].dims[1]));
  size_t total_tokens = 0;
  for (int32_t l : seq_lens) total_tokens += static_cast<size_t>(l);

  // outputs: prototype resized to [effective_batch, ...] (resize_outputs_for_batch)
  for (size_t o = 0; o < out_handles.size(); ++o) {
    size_t numel = output_pool_->per_sample_numel(o) * static_cast<size_t>(effective_batch);
    if (!seq_lens.empty())  // compact: one row of hidden... per true token
      numel = output_pool_->per_sample_numel(o) / static_cast<size_t>(seq_cap) * total_tokens;
    else
    // variable_length: an output whose dims[1] is the configured sequence length shrinks with the batch's seq
    if (seq > 0 && model_.outputs[o].dims.size() >= 2 && model_.outputs[o].dims[1] == seq_cap)
      numel = numel / static_cast<size_t>(seq_cap) * static_cast<size_t>(seq);
    if (request_io.enabled) numel = 1;  // the response bypasses the slot: nothing worth a D2H lands in it
    resize_starpu_vector_handle(out_handles[o],
                                VectorResizeSpec{numel, numel * element_size(output_pool_->tensor_type(o))}, false);
  }

  auto params = std::make_unique<InferenceParams>();
  replicas_->fill_model_pointers(params->models);  // InferenceTask::fill_model_pointers
  params->num_inputs = model_.inputs.size();
  params->num_outputs = model_.outputs.size();
  params->batch_size = effective_batch;
  params->request_id = master->request_id;
  params->request_io = request_io;
  params->layout.compact_output = !seq_lens.empty();
  ctx->compact = !seq_lens.empty();
  params->layout.seq_lens = std::move(seq_lens);
  ctx->request_io = request_io.enabled;
  params->limits.max_inputs = InferLimits::MaxInputs;
  params->limits.max_dims = InferLimits::MaxDims;
  for (size_t i = 0; i < master->inputs.size(); ++i) {
    std::vector<int64_t> dims = master->inputs[i].dims;
    if (!dims.empty()) dims[0] = effective_batch;  // inference_task.cpp:606-613
    if (seq > 0 && dims.size() >= 2) dims[1] = seq;
    params->layout.num_dims.push_back(static_cast<int64_t>(dims.size()));
    params->layout.dims.push_back(std::move(dims));
    params->layout.input_types.push_back(master->inputs[i].type);
  }
  CallbackContext *raw = ctx.get();
  params->device.device_id = &raw->device_id;
  params->device.worker_id = &raw->worker_id;
  params->device.executed_on = &raw->executed_on;
  params->timing.codelet_start_time = &raw->timing.codelet_start_time;
  params->timing.codelet_end_time = &raw->timing.codelet_end_time;
  params->timing.inference_start_time = &raw->timing.inference_start_time;
  ctx->params = std::move(params);
  ctx->outputs_handles = out_handles;

  starpu_task *task = starpu_task_create();
  if (task == nullptr) throw StarPUTaskCreationException("Failed to create StarPU task.");
  const size_t num_inputs = in_handles.size(), num_buffers = num_inputs + out_handles.size();
  task->nbuffers = static_cast<int>(num_buffers);
  task->cl = codelet_.get_codelet();
  task->synchronous = 0;
  task->cl_arg = ctx->params.get();
  task->cl_arg_size = sizeof(InferenceParams);
  task->priority = std::max(starpu_sched_get_min_priority(), starpu_sched_get_max_priority() - master->request_id);
  task->destroy = 1;
  task->dyn_handles = static_cast<starpu_data_handle_t *>(std::malloc(num_buffers * sizeof(starpu_data_handle_t)));
  task->dyn_modes =
      static_cast<starpu_data_access_mode *>(std::malloc(num_buffers * sizeof(starpu_data_access_mode)));
  for (size_t i = 0; i < num_inputs; ++i) {
    task->dyn_handles[i] = in_handles[i];
    task->dyn_modes[i] = STARPU_R;
  }
  for (size_t o = 0; o < out_handles.size(); ++o) {
    task->dyn_handles[num_inputs + o] = out_handles[o];
    task->dyn_modes[num_inputs + o] = STARPU_W;
  }
  if (master->fixed_worker_id.has_value()) {
    const int wid = *master->fixed_worker_id;
    if (wid < 0) {
      starpu_task_destroy(task);
      throw std::invalid_argument("Fixed worker ID must be non-negative");
    }
    if (wid >= static_cast<int>(starpu_worker_get_count())) {
      starpu_task_destroy(task);
      throw std::out_of_range("Fixed worker ID exceeds available workers");
    }
    task->workerid = static_cast<unsigned>(wid);
    task->execute_on_a_specific_worker = 1;
  }
  task->callback_func = &TaskRunner::starpu_output_callback;
  task->callback_arg = raw;
  raw->timing.before_starpu_submitted_time = MonotonicClock::now();
  const int ret = starpu_task_submit(task);
  if (ret != 0) {
    starpu_task_destroy(task);
    throw StarPUTaskSubmissionException("[ERROR] StarPU task submission failed (code " + std::to_string(ret) + ")");
  }
  guard.armed = false;
  (void)ctx.release();  // owned by the callback from here on
  ++submitted_;
}

void TaskRunner::starpu_output_callback(void *arg) {
  auto *ctx = static_cast<CallbackContext *>(arg);
  if (ctx == nullptr) return;
  ctx->timing.callback_start_time = MonotonicClock::now();
  ctx->acquired.assign(ctx->outputs_handles.size(), false);
  // acquire every STARPU_W handle on the host: this is what lands the D2H copy into the
  // pinned output slot (inference_task.cpp:907-935)
  bool ok = true;
  for (size_t i = 0; i < ctx->outputs_handles.size(); ++i) {
    if (starpu_data_acquire(ctx->outputs_handles[i], STARPU_R) != 0) {
      ok = false;
      break;
    }
    ctx->acquired[i] = true;
  }
  (void)ok;
  finalize(ctx);
}

void TaskRunner::finalize(CallbackContext *ctx) {
  TaskRunner *r = ctx->runner;
  const bool all_acquired = std::all_of(ctx->acquired.begin(), ctx->acquired.end(), [](bool b) { return b; });
  // the codelet stamps codelet_end only on success (starpu_setup.cpp:709-722)
  const bool codelet_ok = ctx->timing.codelet_end_time != MonotonicClock::time_point{};
  // bind_output_views_from_pool (inference_task.cpp:949-989): slot buffers viewed as [B, ...]
  std::vector<HostTensor> aggregated;
  const auto &bases = r->output_pool_->base_ptrs(ctx->output_slot);
  for (size_t o = 0; o < bases.size(); ++o) {
    HostTensor t;
    t.data = bases[o];
    t.type = r->output_pool_->tensor_type(o);
    t.dims = r->model_.outputs[o].dims;
    if (t.dims.empty()) t.dims.push_back(ctx->effective_batch);
    else t.dims[0] = ctx->effective_batch;
    if (ctx->seq > 0 && t.dims.size() >= 2 && t.dims[1] == r->model_.inputs[0].dims[1]) t.dims[1] = ctx->seq;
    aggregated.push_back(std::move(t));
  }
  ctx->timing.callback_end_time = MonotonicClock::now();
  // device_visible_requests: nobody on the host read the token ids, the gather kernel checked them against the
  // embedding table and the codelet routed its finding into the first word of output 0 (inference_codelet.cpp).
  // A non-zero word fails the whole batch, which is what the reference does when LibTorch's embedding raises.
  int32_t bad_id_row = 0;
  if (codelet_ok && all_acquired && ctx->request_io && !bases.empty() && bases[0] != nullptr &&
      r->output_pool_->per_sample_numel(0) * element_size(r->output_pool_->tensor_type(0)) >= sizeof(int32_t))
    std::memcpy(&bad_id_row, bases[0], sizeof(int32_t));
  // propagate_completion_to_sub_jobs (result_dispatcher_component.cpp:678-739)
  int64_t offset = 0;
  std::vector<std::function<void()>> completions;
  completions.reserve(ctx->jobs.size());
  for (const auto &job : ctx->jobs) {
    if (!job) continue;
    JobResult res;
    res.request_id = job->request_id;
    res.worker_id = ctx->worker_id;
    res.device_id = ctx->device_id;
    res.executed_on = ctx->executed_on;
    res.batch_rows = ctx->effective_batch;
    res.batch_seq = ctx->seq;
    res.timing = ctx->timing;
    res.timing.enqueued_time = job->timing.enqueued_time;
    if (codelet_ok && all_acquired && ctx->compact) {
      // compact output: this request's [batch, own seq, hidden...] block starts at its token offset
      const int64_t L = job->inputs[0].dims[1], b = job->batch_size();
      HostTensor v = aggregated[0];
      v.dims[0] = b;
      v.dims[1] = L;
      const size_t token_bytes = r->request_token_bytes();
      v.data = aggregated[0].data + static_cast<size_t>(offset) * token_bytes;
      res.outputs.push_back(std::move(v));
      res.rows = b;
      res.ok = true;
      offset += b * L;  // `offset` counts TOKENS on this path
    } else if (codelet_ok && all_acquired && ctx->request_io && bad_id_row != 0) {
      res.ok = false;
      res.error = "token id of batch row " + std::to_string(bad_id_row - 1) + " is outside the embedding table" +
                  (r->model_.id_vocab > 0 ? " [0, " + std::to_string(r->model_.id_vocab) + ")" : std::string());
      offset += job->batch_size();
    } else if (codelet_ok && all_acquired && ctx->request_io) {
      // the codelet's scatter kernel already wrote the response into the request's own buffers
      for (size_t o = 0; o < aggregated.size(); ++o) {
        HostTensor v = aggregated[o];
        v.data = o < job->output_dsts.size() ? static_cast<std::byte *>(job->output_dsts[o]) : nullptr;
        v.dims[0] = job->batch_size();
        if (ctx->seq > 0 && v.dims.size() >= 2 && !job->inputs.empty() && job->inputs[0].dims.size() >= 2)
          v.dims[1] = job->inputs[0].dims[1];  // variable_length: compact [batch, own seq, ...]
        res.outputs.push_back(std::move(v));
      }
      res.rows = job->batch_size();
      res.ok = true;
      offset += job->batch_size();
    } else if (codelet_ok && all_acquired) {
      auto slice = slice_outputs_for_sub_job(aggregated, offset, job->batch_size());
      res.outputs = std::move(slice.outputs);
      // variable_length: a single-sample request gets its own L rows (a contiguous prefix of its padded sample)
      if (ctx->seq > 0 && job->batch_size() == 1 && !job->inputs.empty() && job->inputs[0].dims.size() >= 2)
        for (HostTensor &v : res.outputs)
          if (v.defined() && v.dims.size() >= 2 && v.dims[1] == ctx->seq) v.dims[1] = job->inputs[0].dims[1];
      res.rows = slice.processed_length;
      res.ok = true;
      offset += slice.processed_length;
    } else {
      res.ok = false;
      res.error = codelet_ok ? "Failed to acquire output data handle from StarPU." : "Inference failed";
      offset += job->batch_size();
    }
    if (job->on_complete)
      completions.emplace_back([cb = &job->on_complete, res = std::move(res)] { (*cb)(res); });
  }
  if (r->completion_executor_ && completions.size() > 1) r->completion_executor_(completions);
  else
    for (auto &c : completions) c();
  for (size_t i = 0; i < ctx->outputs_handles.size(); ++i)
    if (ctx->acquired[i]) starpu_data_release(ctx->outputs_handles[i]);
  r->input_pool_->release(ctx->input_slot);
  r->output_pool_->release(ctx->output_slot);
  delete ctx;
}

}  // namespace starpu_server
