#include "replica_registry.hpp"

#include "starpu_compat.h"

#include <algorithm>
#include <string>

namespace starpu_server {

auto get_cuda_workers_by_device(const std::vector<int> &device_ids) -> std::map<int, std::vector<int>> {
  std::map<int, std::vector<int>> out;
  int ids[STARPU_NMAXWORKERS];
  const unsigned n = starpu_worker_get_ids_by_type(STARPU_CUDA_WORKER, ids, STARPU_NMAXWORKERS);
  for (unsigned i = 0; i < n; ++i) {
    const int dev = starpu_worker_get_devid(ids[i]);
    if (std::find(device_ids.begin(), device_ids.end(), dev) != device_ids.end()) out[dev].push_back(ids[i]);
  }
  for (auto &kv : out) std::sort(kv.second.begin(), kv.second.end());
  return out;
}

void validate_device_ids(const std::vector<int> &device_ids, int available_device_count) {
  for (const int device_id : device_ids)
    if (device_id < 0 || device_id >= available_device_count)
      throw InvalidGpuDeviceException("Invalid GPU device ID " + std::to_string(device_id) + " (" +
                                      std::to_string(available_device_count) + " device(s) available).");
}

auto build_gpu_replica_assignments(const std::vector<int> &device_ids, GpuModelReplicationPolicy policy,
                                   const std::function<std::map<int, std::vector<int>>(const std::vector<int> &)>
                                       &workers_by_device) -> std::vector<GpuReplicaAssignment> {
  std::vector<GpuReplicaAssignment> assignments;
  if (device_ids.empty()) return assignments;
  if (policy == GpuModelReplicationPolicy::PerDevice) {
    for (const int device_id : device_ids)
      if (device_id >= 0) assignments.push_back({device_id, -1});
    return assignments;
  }
  const auto by_device = workers_by_device(device_ids);
  for (const int device_id : device_ids) {
    if (device_id < 0) continue;
    const auto it = by_device.find(device_id);
    if (it == by_device.end() || it->second.empty())
      throw StarPUWorkerQueryException("No CUDA workers available for device " + std::to_string(device_id) +
                                       " while gpu_model_replication='per_worker'");
  }
  for (const int device_id : device_ids) {
    if (device_id < 0) continue;
    for (const int worker_id : by_device.at(device_id)) assignments.push_back({device_id, worker_id});
  }
  return assignments;
}

ReplicaRegistry::~ReplicaRegistry() { destroy_owned(); }

auto ReplicaRegistry::add(GpuReplicaAssignment where, DeviceModel model, bool owned) -> DeviceModel * {
  if (entries_.size() >= InferLimits::MaxModelsGPU) throw std::invalid_argument("too many GPU replicas");
  if (where.device_id < 0) throw InvalidGpuDeviceException("replica registered for a negative device id");
  if (!entries_.empty() && per_worker_ != (where.worker_id >= 0))
    throw std::invalid_argument("per-device and per-worker replicas cannot be mixed in one registry");
  per_worker_ = where.worker_id >= 0;
  model.device_id = where.device_id;
  Entry e;
  e.where = where;
  e.model = std::make_unique<DeviceModel>(model);
  e.owned = owned;
  entries_.push_back(std::move(e));
  return entries_.back().model.get();
}

void ReplicaRegistry::create(const std::vector<GpuReplicaAssignment> &assignments, DeviceModel prototype,
                             const std::function<spu_replica *(const GpuReplicaAssignment &)> &make) {
  const size_t first = entries_.size();
  try {
    for (const GpuReplicaAssignment &a : assignments) {
      DeviceModel m = prototype;
      m.replica = make(a);
      if (m.kind != DeviceModel::Kind::Affine && m.replica == nullptr)
        throw InferenceEngineException("replica creation failed for device " + std::to_string(a.device_id));
      add(a, m, /*owned=*/true);
    }
  } catch (...) {
    for (size_t i = first; i < entries_.size(); ++i)
      if (entries_[i].owned && entries_[i].model->replica) spu_replica_destroy(entries_[i].model->replica);
    entries_.resize(first);
    throw;
  }
}

void ReplicaRegistry::fill_model_pointers(detail::ModelPointers &models) const {
  models.models_gpu.clear();
  models.device_ids.clear();
  models.worker_ids.clear();
  for (const Entry &e : entries_) {
    models.device_ids.push_back(e.where.device_id);
    if (per_worker_) models.worker_ids.push_back(e.where.worker_id);
    models.models_gpu.push_back(e.model.get());
  }
}

auto ReplicaRegistry::find_by_device(int device_id) const -> DeviceModel * {
  for (const Entry &e : entries_)
    if (e.where.device_id == device_id) return e.model.get();
  return nullptr;
}

auto ReplicaRegistry::find_by_worker(int worker_id) const -> DeviceModel * {
  for (const Entry &e : entries_)
    if (e.where.worker_id == worker_id && worker_id >= 0) return e.model.get();
  return nullptr;
}

void ReplicaRegistry::destroy_owned() {
  for (Entry &e : entries_)
    if (e.owned && e.model && e.model->replica != nullptr) {
      spu_replica_destroy(e.model->replica);
      e.model->replica = nullptr;
      e.owned = false;
    }
}

}  // namespace starpu_server
