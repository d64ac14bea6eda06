// Codelet body: /root/reference/src/core/starpu_setup.cpp:559-568 (constructor),
// :594-624 (run_inference), :638-723 (run_codelet_inference), :725-778 (select_gpu_module),
// :807-846 (cuda_inference_func) -- restated without LibTorch on top of include/spu_b200.h.
#include "inference_codelet.hpp"

#include "nvtx_range.hpp"

#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <climits>
#include <limits>
#include <map>
#include <mutex>
#include <string>
#include <thread>

namespace starpu_server {
namespace {

WorkerContextHooks g_hooks{};
std::array<std::atomic<int>, STARPU_NMAXWORKERS> g_inflight{};

int ctx_worker_id() { return g_hooks.worker_get_id ? g_hooks.worker_get_id() : starpu_worker_get_id(); }
int ctx_worker_devid(int w) { return g_hooks.worker_get_devid ? g_hooks.worker_get_devid(w) : starpu_worker_get_devid(w); }
void *ctx_stream() {
  return g_hooks.cuda_get_local_stream ? g_hooks.cuda_get_local_stream()
                                       : static_cast<void *>(starpu_cuda_get_local_stream());
}

void set_gauge(int worker_id, int v) {
  if (worker_id >= 0 && worker_id < STARPU_NMAXWORKERS) g_inflight[static_cast<size_t>(worker_id)].store(v);
}

// One workspace arena per worker thread per replica: a StarPU worker is one thread with
// one stream, so its (up to STARPU_CUDA_PIPELINE) in-flight tasks are ordered on that
// stream and may share an arena; two workers of one device never share one.
std::mutex g_ws_mu;
std::map<std::pair<const spu_replica *, std::thread::id>, int> g_ws_of_thread;
std::map<const spu_replica *, int> g_ws_next;

int workspace_for_this_thread(const DeviceModel *m) {
  std::lock_guard<std::mutex> lk(g_ws_mu);
  const auto key = std::make_pair(static_cast<const spu_replica *>(m->replica), std::this_thread::get_id());
  auto it = g_ws_of_thread.find(key);
  if (it != g_ws_of_thread.end()) return it->second;
  int &next = g_ws_next[m->replica];
  if (next >= m->n_workspaces)
    throw InferenceExecutionException(
        "[ERROR] more worker threads run on this replica than it has workspaces (raise n_workspaces to "
        "STARPU_NWORKER_PER_CUDA)");
  const int idx = next++;
  g_ws_of_thread.emplace(key, idx);
  return idx;
}

int64_t dim_or(const InferenceParams *p, size_t input, size_t d, int64_t dflt) {
  if (input >= p->layout.dims.size() || d >= p->layout.dims[input].size()) return dflt;
  return p->layout.dims[input][d];
}

void check_spu(spu_status st) {
  if (st != SPU_OK) throw InferenceExecutionException(std::string("[ERROR] ") + spu_last_error());
}

void require_bytes(const StarpuBufferInterface *b, size_t want, const char *what) {
  const size_t have = buffer_byte_size(b);
  if (have != want)
    throw InferenceExecutionException(std::string(what) + " buffer size mismatch in bytes (have " +
                                      std::to_string(have) + ", expected " + std::to_string(want) + ")");
}

}  // namespace

void set_worker_context_hooks(const WorkerContextHooks &hooks) { g_hooks = hooks; }

void forget_replica_workspaces(const spu_replica *replica) {
  std::lock_guard<std::mutex> lk(g_ws_mu);
  for (auto it = g_ws_of_thread.begin(); it != g_ws_of_thread.end();)
    it = it->first.first == replica ? g_ws_of_thread.erase(it) : std::next(it);
  g_ws_next.erase(replica);
}

int worker_inflight_gauge(int worker_id) {
  if (worker_id < 0 || worker_id >= STARPU_NMAXWORKERS) return 0;
  return g_inflight[static_cast<size_t>(worker_id)].load();
}

auto buffer_byte_size(const StarpuBufferInterface *buffer_iface) -> size_t {
  if (buffer_iface == nullptr) throw InferenceExecutionException("[ERROR] StarPU buffer is null");
  switch (buffer_iface->id) {
    case STARPU_VARIABLE_INTERFACE_ID:
      return buffer_iface->elemsize;
    case STARPU_VECTOR_INTERFACE_ID: {
      const auto *vec = reinterpret_cast<const starpu_vector_interface *>(buffer_iface);
      if (vec->elemsize != 0U && vec->nx > std::numeric_limits<size_t>::max() / vec->elemsize)
        throw InferenceExecutionException("[ERROR] StarPU buffer size exceeds size_t capacity");
      return static_cast<size_t>(vec->nx) * vec->elemsize;
    }
    default:
      throw InferenceExecutionException("[ERROR] Unsupported StarPU buffer interface id " +
                                        std::to_string(static_cast<int>(buffer_iface->id)));
  }
}

InferenceCodelet::InferenceCodelet() : codelet_{} {
  starpu_codelet_init(&codelet_);
  codelet_.nbuffers = STARPU_VARIABLE_NBUFFERS;
  codelet_.type = STARPU_FORKJOIN;
  codelet_.max_parallelism = INT_MAX;
  // The reference also installs cpu_funcs[0] (LibTorch on the host).  This codelet is the
  // CUDA implementation only: there is deliberately no CPU path to fall back to.
  codelet_.cuda_funcs[0] = &InferenceCodelet::cuda_inference_func;
  codelet_.cuda_flags[0] = 1U;  // STARPU_CUDA_ASYNC
  codelet_.where = STARPU_CUDA;
  codelet_.name = "inference_b200";
}

auto InferenceCodelet::get_codelet() -> struct starpu_codelet * { return &codelet_; }

auto select_gpu_module(const InferenceParams &params, const int worker_id, const int device_id) -> DeviceModel * {
  const auto invalid_index = std::numeric_limits<size_t>::max();
  const auto fetch_model = [&](size_t index) -> DeviceModel * {
    return index >= params.models.models_gpu.size() ? nullptr : params.models.models_gpu[index];
  };
  if (!params.models.worker_ids.empty()) {
    if (worker_id >= 0) {
      const auto found = std::find(params.models.worker_ids.begin(), params.models.worker_ids.end(), worker_id);
      if (found != params.models.worker_ids.end())
        if (auto *m = fetch_model(static_cast<size_t>(found - params.models.worker_ids.begin()))) return m;
    }
    throw StarPUCodeletException("[ERROR] No GPU model replica available for worker " + std::to_string(worker_id) +
                                 " on device " + std::to_string(device_id));
  }
  size_t module_index = invalid_index;
  if (device_id >= 0) {
    if (!params.models.device_ids.empty()) {
      const auto found = std::find(params.models.device_ids.begin(), params.models.device_ids.end(), device_id);
      if (found != params.models.device_ids.end())
        module_index = static_cast<size_t>(found - params.models.device_ids.begin());
    } else {
      module_index = static_cast<size_t>(device_id);
    }
  }
  if (module_index != invalid_index)
    if (auto *m = fetch_model(module_index)) return m;
  throw StarPUCodeletException("[ERROR] No GPU model replica available for device " + std::to_string(device_id));
}

void run_inference(InferenceParams *params, StarpuBufferSpan buffers, int device_id, DeviceModel *model,
                   void *stream) {
  (void)device_id;
  // TensorBuilder::prepare_input_ivalues' checks (tensor_builder.cpp:93-123,148-155)
  if (params->limits.max_inputs != 0 || params->num_inputs != 0) {
    if (params->num_inputs > (params->limits.max_inputs ? params->limits.max_inputs : InferLimits::MaxInputs))
      throw InferenceExecutionException("[ERROR] Too many input tensors");
  }
  if (buffers.size() < params->num_inputs + params->num_outputs)
    throw InferenceExecutionException("[ERROR] Mismatch between model buffers and StarPU buffers");
  for (size_t i = 0; i < params->num_inputs + params->num_outputs; ++i)
    if (buffers[i] == nullptr || buffers[i]->ptr == 0)
      throw InferenceExecutionException("[ERROR] StarPU buffer is null");

  if (params->timing.set_inference_start_time) params->timing.set_inference_start_time(MonotonicClock::now());
  else if (params->timing.inference_start_time != nullptr) *params->timing.inference_start_time = MonotonicClock::now();

  switch (model->kind) {
    case DeviceModel::Kind::Bert: {
      // models/bert.yml:11-15: input_ids, attention_mask int64 [B,S] -> output0 fp32 [B,S,H]
      if (params->num_inputs != 2 || params->num_outputs != 1)
        throw InferenceExecutionException("Mismatch between model outputs and StarPU buffers");
      const int64_t B = dim_or(params, 0, 0, params->batch_size);
      const int64_t S = dim_or(params, 0, 1, 0);
      if (B <= 0 || S <= 0 || B > INT32_MAX || S > INT32_MAX)
        throw InferenceExecutionException("[ERROR] Invalid BERT input dims");
      if (params->request_io.enabled) {
        // stages 1-4 on per-request buffers: buffers[0] carries the address table the host wrote
        const detail::RequestIo &io = params->request_io;
        if (io.rows != B || io.seq != S || io.table_bytes == 0)
          throw InferenceExecutionException("[ERROR] Request table does not match the task layout");
        require_bytes(buffers[0], io.table_bytes, "Input");
        const auto base = static_cast<uintptr_t>(buffers[0]->ptr);
        const int ws = workspace_for_this_thread(model);
        // The responses bypass the STARPU_W buffer on this path; its first word carries the task's verdict on the
        // token ids instead (0 = all inside the embedding table).  StarPU brings it to the host with the handle
        // like any output, and the output callback fails the batch on a non-zero word, as the reference's job
        // fails when LibTorch's embedding raises (task_runner.cpp: finalize).
        int32_t *status = buffer_byte_size(buffers[2]) >= sizeof(int32_t) ? reinterpret_cast<int32_t *>(buffers[2]->ptr) : nullptr;
        check_spu(spu_bert_forward_requests_checked_async(
            model->replica, ws, reinterpret_cast<const int64_t *const *>(base),
            reinterpret_cast<const int64_t *const *>(base + io.off_mask), reinterpret_cast<const int32_t *>(base + io.off_len),
            static_cast<int32_t>(B), static_cast<int32_t>(S), reinterpret_cast<float *const *>(base + io.off_dst),
            io.full_rows ? 1 : 0, status, stream));
        break;
      }
      const size_t in_bytes = static_cast<size_t>(B) * static_cast<size_t>(S) * sizeof(int64_t);
      require_bytes(buffers[0], in_bytes, "Input");
      require_bytes(buffers[1], in_bytes, "Input");
      const int ws = workspace_for_this_thread(model);
      if (params->layout.seq_lens.size() == static_cast<size_t>(B)) {
        // right-padded batch with host-known lengths: skip the padding instead of multiplying it
        if (spu_bert_packed_supported(model->replica, static_cast<int32_t>(B)) != 1)
          throw InferenceExecutionException("[ERROR] Replica has no packed forward for a task laid out with seq_lens");
        int64_t tokens = 0;
        for (int32_t l : params->layout.seq_lens) tokens += l;
        const size_t token_bytes = spu_bert_output_bytes(model->replica, 1, 1);
        auto *ids = reinterpret_cast<const int64_t *>(buffers[0]->ptr);
        auto *out = reinterpret_cast<float *>(buffers[2]->ptr);
        if (params->layout.compact_output) {
          require_bytes(buffers[2], static_cast<size_t>(tokens) * token_bytes, "Output");
          check_spu(spu_bert_forward_packed_compact_async(model->replica, ws, ids, params->layout.seq_lens.data(),
                                                          static_cast<int32_t>(B), static_cast<int32_t>(S), out, stream));
        } else {
          require_bytes(buffers[2], static_cast<size_t>(B) * static_cast<size_t>(S) * token_bytes, "Output");
          check_spu(spu_bert_forward_packed_async(model->replica, ws, ids, params->layout.seq_lens.data(),
                                                  static_cast<int32_t>(B), static_cast<int32_t>(S), out, stream));
        }
        break;
      }
      require_bytes(buffers[2], spu_bert_output_bytes(model->replica, static_cast<int32_t>(B), static_cast<int32_t>(S)),
                    "Output");
      check_spu(spu_bert_forward_async(model->replica, ws, reinterpret_cast<const int64_t *>(buffers[0]->ptr),
                                       reinterpret_cast<const int64_t *>(buffers[1]->ptr), static_cast<int32_t>(B),
                                       static_cast<int32_t>(S), reinterpret_cast<float *>(buffers[2]->ptr), stream));
      break;
    }
    case DeviceModel::Kind::Vit: {
      // models/vit_l_16.yml:13-16: input fp32 [B,3,224,224] -> logits fp32 [B,classes]
      if (params->num_inputs != 1 || params->num_outputs != 1)
        throw InferenceExecutionException("Mismatch between model outputs and StarPU buffers");
      const int64_t B = dim_or(params, 0, 0, params->batch_size);
      if (B <= 0 || B > INT32_MAX) throw InferenceExecutionException("[ERROR] Invalid ViT input dims");
      require_bytes(buffers[0], spu_vit_input_bytes(model->replica, static_cast<int32_t>(B)), "Input");
      require_bytes(buffers[1], spu_vit_output_bytes(model->replica, static_cast<int32_t>(B)), "Output");
      const int ws = workspace_for_this_thread(model);
      check_spu(spu_vit_forward_async(model->replica, ws, reinterpret_cast<const float *>(buffers[0]->ptr),
                                      static_cast<int32_t>(B), reinterpret_cast<float *>(buffers[1]->ptr), stream));
      break;
    }
    case DeviceModel::Kind::Affine: {
      if (params->num_inputs != 1 || params->num_outputs != 1)
        throw InferenceExecutionException("Mismatch between model outputs and StarPU buffers");
      const size_t n_in = buffer_byte_size(buffers[0]), n_out = buffer_byte_size(buffers[1]);
      if (n_in != n_out) throw InferenceExecutionException("Output buffer size mismatch in bytes");
      check_spu(spu_affine_f32_async(reinterpret_cast<const float *>(buffers[0]->ptr),
                                     reinterpret_cast<float *>(buffers[1]->ptr), n_in / sizeof(float), model->scale,
                                     model->shift, stream));
      break;
    }
  }
}

namespace {

void run_codelet_inference(InferenceParams *params, StarpuBufferSpan buffers, int worker_id, int device_id,
                           DeviceModel *model, void *stream, DeviceType executed_on_type) {
  if (params->timing.set_codelet_start_time) params->timing.set_codelet_start_time(MonotonicClock::now());
  else if (params->timing.codelet_start_time) *params->timing.codelet_start_time = MonotonicClock::now();

  struct Guard {
    int worker;
    ~Guard() { set_gauge(worker, 0); }
  } guard{worker_id};
  set_gauge(worker_id, 1);

  if (params->verbosity >= VerbosityLevel::Trace)  // log_trace (starpu_setup.cpp:685-692)
    std::fprintf(stderr, "[TRACE] %s device id %d, worker id %d, job id %d\n",
                 executed_on_type == DeviceType::CPU ? "CPU" : "GPU", device_id, worker_id, params->request_id);

  if (params->device.set_executed_on) params->device.set_executed_on(executed_on_type);
  else if (params->device.executed_on) *params->device.executed_on = executed_on_type;
  if (params->device.set_worker_id) params->device.set_worker_id(worker_id);
  else if (params->device.worker_id) *params->device.worker_id = worker_id;
  if (params->device.set_device_id) params->device.set_device_id(device_id);
  else if (params->device.device_id) *params->device.device_id = device_id;

  try {
    run_inference(params, buffers, device_id, model, stream);
  } catch (const std::exception &e) {
    throw StarPUCodeletException(std::string("[ERROR] Codelet failure: ") + e.what());
  }

  if (params->timing.set_codelet_end_time) params->timing.set_codelet_end_time(MonotonicClock::now());
  else if (params->timing.codelet_end_time) *params->timing.codelet_end_time = MonotonicClock::now();
}

}  // namespace

void InferenceCodelet::cuda_inference_func(void **buffers, void *cl_arg) {
  auto *params = static_cast<InferenceParams *>(cl_arg);
  const auto total_buffers = params->num_inputs + params->num_outputs;
  const auto *typed = reinterpret_cast<StarpuBufferPtr const *>(buffers);
  const StarpuBufferSpan span(typed, total_buffers);
  const int worker_id = ctx_worker_id();
  const int device_id = ctx_worker_devid(worker_id);
  char label[64];
  std::snprintf(label, sizeof(label), "codelet.cuda job %d dev %d", params->request_id, device_id);
  const NvtxRange nvtx_scope(label);  // starpu_setup.cpp:817
  void *stream = ctx_stream();
  DeviceModel *model = select_gpu_module(*params, worker_id, device_id);
  // every kernel goes on StarPU's local stream and the function returns without
  // synchronising: STARPU_CUDA_ASYNC (starpu_setup.cpp:820-823)
  int prev = -1;
  cudaGetDevice(&prev);
  if (prev != device_id && device_id >= 0) cudaSetDevice(device_id);
  run_codelet_inference(params, span, worker_id, device_id, model, stream, DeviceType::CUDA);
}

}  // namespace starpu_server
