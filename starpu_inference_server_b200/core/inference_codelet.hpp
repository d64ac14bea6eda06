// The StarPU inference codelet, B200-native body.
// Same class, same codelet field values, same cl_arg / buffer contract as
// /root/reference/src/core/starpu_setup.hpp:21-37 + starpu_setup.cpp:515-846; what changes
// is the body of cuda_inference_func: instead of wrapping the StarPU device buffers as
// torch tensors and calling LibTorch's Module::forward, it hands the raw device pointers,
// (batch, seq) and the worker's CUDA stream to the C-ABI of include/spu_b200.h.
#pragma once

#include <cstddef>
#include <span>

#include "inference_params.hpp"
#include "starpu_compat.h"

namespace starpu_server {

// tensor_builder.hpp:12-14 -- the codelet reads every buffer through this alias; the
// handles are really vector interfaces (same leading fields), sized by buffer_byte_size.
using StarpuBufferInterface = starpu_variable_interface;
using StarpuBufferPtr = StarpuBufferInterface *;
using StarpuBufferSpan = std::span<const StarpuBufferPtr>;

class InferenceCodelet {
 public:
  InferenceCodelet();
  ~InferenceCodelet() = default;
  InferenceCodelet(const InferenceCodelet &) = delete;
  auto operator=(const InferenceCodelet &) -> InferenceCodelet & = delete;

  auto get_codelet() -> struct starpu_codelet *;

 private:
  static void cuda_inference_func(void **buffers, void *cl_arg);
  struct starpu_codelet codelet_;
};

// starpu_setup.cpp:515-542
auto buffer_byte_size(const StarpuBufferInterface *buffer_iface) -> size_t;

// starpu_setup.cpp:725-778 (same lookup order and messages; returns the replica record)
auto select_gpu_module(const InferenceParams &params, int worker_id, int device_id) -> DeviceModel *;

// starpu_setup.cpp:594-624 without LibTorch: validates the buffers against
// params->layout, stamps inference_start, enqueues the forward on `stream`.
void run_inference(InferenceParams *params, StarpuBufferSpan buffers, int device_id, DeviceModel *model,
                   void *stream);

// Test seam equivalent to the reference's ScopedWorkerContextOverride
// (tests/unit/core/unit_starpu_setup.cpp:2349-2354): lets a unit test call cuda_func
// directly with a chosen worker id / device id / stream, without a running runtime.
struct WorkerContextHooks {
  int (*worker_get_id)() = nullptr;
  int (*worker_get_devid)(int) = nullptr;
  void *(*cuda_get_local_stream)() = nullptr;
};
void set_worker_context_hooks(const WorkerContextHooks &hooks);  // {} restores StarPU's
void forget_replica_workspaces(const spu_replica *replica);       // call before destroying a replica

// Gauge the reference exports per worker (set_worker_inflight_gauge, starpu_setup.cpp:651-683).
int worker_inflight_gauge(int worker_id);

}  // namespace starpu_server
