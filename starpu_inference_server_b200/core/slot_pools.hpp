// Pinned host slot pools registered as StarPU vector handles.
// Mirrors /root/reference/src/core/slot_pool_base.hpp:32-92, input_slot_pool.cpp:235-328 and
// output_slot_pool.cpp:226-330: `slots` x one buffer per model tensor, each sized
// max_batch x per-sample bytes, registered on STARPU_MAIN_RAM with
// starpu_vector_data_register(ptr, max_batch*numel, elemsize); acquire() blocks until a slot
// is free, release() returns it.
#pragma once

#include <condition_variable>
#include <cstddef>
#include <mutex>
#include <optional>
#include <string>
#include <vector>

#include "inference_params.hpp"
#include "starpu_compat.h"

namespace starpu_server {

// One model tensor as models/*.yml declares it (config_loader: name, dims incl. batch dim, type).
struct TensorConfig {
  std::string name;
  std::vector<int64_t> dims;  // dims[0] is the per-request batch (1 in the shipped configs)
  ScalarType type = ScalarType::Float32;
};

struct HostBufferInfo {
  bool cuda_pinned = false;
  bool starpu_pinned = false;
  size_t bytes = 0;  // capacity
};

struct SlotPoolSlot {
  int id = 0;
  std::vector<std::byte *> base_ptrs;
  std::vector<starpu_data_handle_t> handles;
};

class SlotPool {
 public:
  // per-sample shape = dims without dims[0]; capacity rows = max(max_batch, dims[0])
  SlotPool(const std::vector<TensorConfig> &tensors, int max_batch, int slots, bool want_pinned);
  ~SlotPool();
  SlotPool(const SlotPool &) = delete;
  auto operator=(const SlotPool &) -> SlotPool & = delete;

  auto acquire() -> int;
  [[nodiscard]] auto try_acquire() -> std::optional<int>;
  void release(int slot_id);

  [[nodiscard]] auto handles(int slot_id) const -> const std::vector<starpu_data_handle_t> & {
    return slots_.at(static_cast<size_t>(slot_id)).handles;
  }
  [[nodiscard]] auto base_ptrs(int slot_id) const -> const std::vector<std::byte *> & {
    return slots_.at(static_cast<size_t>(slot_id)).base_ptrs;
  }
  [[nodiscard]] auto host_buffer_infos(int slot_id) const -> const std::vector<HostBufferInfo> & {
    return infos_.at(static_cast<size_t>(slot_id));
  }
  [[nodiscard]] auto max_batch_size() const -> int { return bmax_; }
  [[nodiscard]] auto num_slots() const -> int { return static_cast<int>(slots_.size()); }
  [[nodiscard]] auto num_tensors() const -> size_t { return per_sample_numel_.size(); }
  [[nodiscard]] auto per_sample_numel(size_t i) const -> size_t { return per_sample_numel_[i]; }
  [[nodiscard]] auto per_sample_bytes(size_t i) const -> size_t { return per_sample_bytes_[i]; }
  [[nodiscard]] auto tensor_type(size_t i) const -> ScalarType { return types_[i]; }
  [[nodiscard]] auto free_slots() const -> int;

 private:
  int bmax_;
  std::vector<ScalarType> types_;
  std::vector<size_t> per_sample_numel_, per_sample_bytes_;
  std::vector<SlotPoolSlot> slots_;
  std::vector<std::vector<HostBufferInfo>> infos_;
  std::vector<int> free_ids_;
  mutable std::mutex mtx_;
  std::condition_variable cv_;
};

using InputSlotPool = SlotPool;
using OutputSlotPool = SlotPool;

}  // namespace starpu_server
