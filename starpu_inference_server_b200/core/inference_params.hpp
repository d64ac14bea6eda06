// The cl_arg contract of the inference codelet, LibTorch-free.
// Mirrors /root/reference/src/core/inference_params.hpp:14-92 field for field; the only
// type that changes is what a "model" is: a torch::jit::script::Module* becomes a
// DeviceModel* (a weight replica created through the C-ABI of include/spu_b200.h).
#pragma once

#include <chrono>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/spu_b200.h"

namespace starpu_server {

using MonotonicClock = std::chrono::steady_clock;  // utils/monotonic_clock.hpp

enum class DeviceType : uint8_t { CPU, CUDA, Unknown };  // src/core/device_type.hpp
inline const char *to_string(DeviceType t) {
  return t == DeviceType::CPU ? "CPU" : (t == DeviceType::CUDA ? "CUDA" : "Unknown");
}

enum class VerbosityLevel : uint8_t { Silent = 0, Info, Stats, Debug, Trace };

// Element types models/*.yml may declare (config_loader maps "TYPE_INT64" etc.).
enum class ScalarType : uint8_t { Float32, Float64, Int32, Int64, Int8, UInt8, Float16, BFloat16, Bool };
inline size_t element_size(ScalarType t) {
  switch (t) {
    case ScalarType::Float64: case ScalarType::Int64: return 8;
    case ScalarType::Float32: case ScalarType::Int32: return 4;
    case ScalarType::Float16: case ScalarType::BFloat16: return 2;
    default: return 1;
  }
}

// src/core/inference_limits.hpp:8
struct InferLimits {
  static constexpr size_t MaxModelsGPU = 32;
  static constexpr size_t MaxInputs = 16;
  static constexpr size_t MaxDims = 8;
};

// What runs for `model->forward(...)` on a CUDA worker.  The entries live in the replica registry
// (core/replica_registry.hpp: ReplicaRegistry); the codelet only reads them.
struct DeviceModel {
  enum class Kind : uint8_t { Bert, Vit, Affine } kind = Kind::Affine;
  spu_replica *replica = nullptr;  // Bert / Vit: weights + per-worker workspaces on one GPU
  int n_workspaces = 0;
  float scale = 1.f, shift = 0.f;  // Affine: out = in*scale + shift (the reference tests' x+1 / x+1.5 / x*2 toys)
  int device_id = -1;
};

namespace detail {
struct Timing {
  MonotonicClock::time_point *codelet_start_time = nullptr;
  MonotonicClock::time_point *codelet_end_time = nullptr;
  MonotonicClock::time_point *inference_start_time = nullptr;
  std::function<void(MonotonicClock::time_point)> set_codelet_start_time;
  std::function<void(MonotonicClock::time_point)> set_codelet_end_time;
  std::function<void(MonotonicClock::time_point)> set_inference_start_time;
};

struct DeviceInfo {
  int *device_id = nullptr;
  int *worker_id = nullptr;
  DeviceType *executed_on = nullptr;
  std::function<void(int)> set_device_id;
  std::function<void(int)> set_worker_id;
  std::function<void(DeviceType)> set_executed_on;
};

struct ModelPointers {
  // no model_cpu: the new codelet has no CPU implementation (north_star: no CPU fallback)
  std::vector<int> device_ids;
  std::vector<int> worker_ids;
  std::vector<DeviceModel *> models_gpu;
};

struct TensorLayout {
  std::vector<std::vector<int64_t>> dims;  // dims[i][0] = effective batch (inference_task.cpp:606-613)
  std::vector<int64_t> num_dims;
  std::vector<ScalarType> input_types;
  // variable_length batches whose every request mask is all ones (the usual case: the padding is the host
  // gather's own): true token count of each sample row.  Lets the codelet run the encoder on the real tokens
  // only (spu_bert_forward_packed_async).  Empty = unknown, the padded forward runs.
  std::vector<int32_t> seq_lens;
  // with seq_lens: output 0 is laid out COMPACT, [sum(seq_lens), hidden...] (sample row r at token offset
  // seq_lens[0] + ... + seq_lens[r-1]) instead of padded [batch, seq, hidden...]: no padding crosses PCIe
  bool compact_output = false;
};

struct Limits {
  size_t max_inputs = 0;
  size_t max_dims = 0;
};

// SURVEY 8(f) row 2 (device-resident gather / scatter over per-request buffers).  When `enabled`, buffers[0]
// of the task is not the gathered [B, S] ids array but a table of DEVICE-VISIBLE addresses the host wrote into
// the input slot -- one entry per sample row -- and the codelet runs stages 1-4 in one call
// (spu_bert_forward_requests_async): it replaces copy_job_inputs_to_slot
// (slot_manager_component.cpp:649-727) and the slot slicing (batching_helpers.hpp:74-127).  Layout of the table
// (byte offsets from buffers[0]->ptr): ids pointers at 0, mask pointers at off_mask, destination pointers at
// off_dst (8 bytes each), int32 true lengths at off_len.
struct RequestIo {
  bool enabled = false;
  int64_t rows = 0;  // sample rows of the batch
  int64_t seq = 0;   // padded sequence length of the batch
  bool full_rows = true;  // destinations take `seq` rows each (false: only their own length)
  size_t off_mask = 0, off_dst = 0, off_len = 0, table_bytes = 0;
};
}  // namespace detail

struct InferenceParams {
  detail::ModelPointers models;
  detail::TensorLayout layout;
  detail::DeviceInfo device;
  detail::Timing timing;
  detail::Limits limits{};
  detail::RequestIo request_io{};
  size_t num_inputs = 0;
  size_t num_outputs = 0;
  int64_t batch_size = 1;
  int request_id = 0;
  VerbosityLevel verbosity = VerbosityLevel::Silent;
};

// ---- exceptions (src/utils/exceptions.hpp names) ---------------------------------
struct InferenceEngineException : std::runtime_error { using std::runtime_error::runtime_error; };
struct InferenceExecutionException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct StarPUCodeletException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct StarPURegistrationException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct StarPUDataAcquireException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct StarPUTaskCreationException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct StarPUTaskSubmissionException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct InputPoolCapacityException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct InputPoolMismatchException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct InvalidInputTensorException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct InvalidInferenceJobException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };
struct InconsistentInputTensorCountException : InferenceEngineException { using InferenceEngineException::InferenceEngineException; };

}  // namespace starpu_server
