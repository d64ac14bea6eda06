// Per-GPU replica registration: which weight replica serves which CUDA device (or worker).
// Restates, for spu_replica handles instead of torch::jit::Module clones:
//   build_gpu_replica_assignments   /root/reference/src/core/inference_runner.cpp:128-183
//   validate_device_ids             /root/reference/src/core/inference_runner.cpp:104-125
//   clone_model_to_gpus             /root/reference/src/core/inference_runner.cpp:251-275 (one replica per assignment)
//   InferenceTask::fill_model_pointers   /root/reference/src/core/inference_task.cpp:469-516
// The registry OWNS the DeviceModel entries the codelet reads (InferenceParams::models) and, for replicas it was
// asked to own, destroys them -- after starpu_shutdown, like `std::vector<Module> models_gpu` living in main()
// (/root/reference/src/server_main.cpp:248).
#pragma once

#include <functional>
#include <map>
#include <memory>
#include <vector>

#include "inference_params.hpp"

namespace starpu_server {

enum class GpuModelReplicationPolicy : uint8_t { PerDevice, PerWorker };

struct GpuReplicaAssignment {
  int device_id = -1;
  int worker_id = -1;  // -1 under PerDevice
};

struct InvalidGpuDeviceException : InferenceEngineException {
  using InferenceEngineException::InferenceEngineException;
};
struct StarPUWorkerQueryException : InferenceEngineException {
  using InferenceEngineException::InferenceEngineException;
};

// device id -> ids of the StarPU CUDA workers bound to it (StarPUSetup::get_cuda_workers_by_device)
auto get_cuda_workers_by_device(const std::vector<int> &device_ids) -> std::map<int, std::vector<int>>;

void validate_device_ids(const std::vector<int> &device_ids, int available_device_count);

// PerDevice: one assignment per non-negative device id, in order.  PerWorker: one per CUDA worker of every
// listed device, devices in order, workers in StarPU id order; throws if a device has no worker.
auto build_gpu_replica_assignments(const std::vector<int> &device_ids, GpuModelReplicationPolicy policy,
                                   const std::function<std::map<int, std::vector<int>>(const std::vector<int> &)>
                                       &workers_by_device = get_cuda_workers_by_device)
    -> std::vector<GpuReplicaAssignment>;

class ReplicaRegistry {
 public:
  ReplicaRegistry() = default;
  ~ReplicaRegistry();
  ReplicaRegistry(const ReplicaRegistry &) = delete;
  auto operator=(const ReplicaRegistry &) -> ReplicaRegistry & = delete;

  // Register the replica serving `where`.  `owned`: the registry destroys model.replica (spu_replica_destroy).
  auto add(GpuReplicaAssignment where, DeviceModel model, bool owned) -> DeviceModel *;

  // clone_model_to_gpus: one replica per assignment from `make` (e.g. the TorchScript loader's create call);
  // a failing create destroys what was created so far and rethrows.
  void create(const std::vector<GpuReplicaAssignment> &assignments, DeviceModel prototype,
              const std::function<spu_replica *(const GpuReplicaAssignment &)> &make);

  // fill_model_pointers: device_ids (+ worker_ids under PerWorker) and models_gpu, index-aligned
  void fill_model_pointers(detail::ModelPointers &models) const;

  [[nodiscard]] auto size() const -> size_t { return entries_.size(); }
  [[nodiscard]] auto per_worker() const -> bool { return per_worker_; }
  [[nodiscard]] auto find_by_device(int device_id) const -> DeviceModel *;
  [[nodiscard]] auto find_by_worker(int worker_id) const -> DeviceModel *;
  [[nodiscard]] auto assignment(size_t i) const -> const GpuReplicaAssignment & { return entries_.at(i).where; }
  [[nodiscard]] auto model(size_t i) const -> DeviceModel * { return entries_.at(i).model.get(); }

  // Destroy owned replicas now (idempotent).  Call after starpu_shutdown: workers may still hold the pointers.
  void destroy_owned();

 private:
  struct Entry {
    GpuReplicaAssignment where;
    std::unique_ptr<DeviceModel> model;  // stable address: InferenceParams keeps raw pointers
    bool owned = false;
  };
  std::vector<Entry> entries_;
  bool per_worker_ = false;
};

}  // namespace starpu_server
