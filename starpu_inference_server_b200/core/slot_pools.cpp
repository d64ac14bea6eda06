#include "slot_pools.hpp"

#include <cuda_runtime.h>

#include <cstdlib>
#include <limits>
#include <stdexcept>

namespace starpu_server {
namespace {

size_t product_tail(const std::vector<int64_t> &dims) {
  // per-sample element count: every dim but dims[0] (input_slot_pool.cpp:250-268 computes
  // product_dims over all dims with dims[0] == 1 in the shipped configs)
  size_t n = 1;
  for (size_t i = 1; i < dims.size(); ++i) {
    if (dims[i] <= 0) throw std::invalid_argument("tensor dims must be positive");
    n *= static_cast<size_t>(dims[i]);
  }
  return n;
}

std::byte *alloc_host(size_t bytes, bool want_pinned, HostBufferInfo *info) {
  info->bytes = bytes;
  void *p = nullptr;
  if (want_pinned && cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) == cudaSuccess) {
    info->cuda_pinned = true;
    return static_cast<std::byte *>(p);
  }
  cudaGetLastError();
  // no CUDA device (CPU-only unit tests): pageable, 64-byte aligned like the reference's
  // aligned_alloc fallback (input_slot_pool.cpp allocate_and_pin_buffer)
  const size_t rounded = ((bytes ? bytes : 1) + 63) / 64 * 64;
  p = std::aligned_alloc(64, rounded);
  if (p == nullptr) throw std::bad_alloc();
  return static_cast<std::byte *>(p);
}

void free_host(std::byte *p, const HostBufferInfo &info) {
  if (p == nullptr) return;
  if (info.cuda_pinned) cudaFreeHost(p);
  else std::free(p);
}

}  // namespace

SlotPool::SlotPool(const std::vector<TensorConfig> &tensors, int max_batch, int slots, bool want_pinned)
    : bmax_(max_batch > 0 ? max_batch : 1) {
  if (slots <= 0) throw std::invalid_argument("slot pool needs at least one slot");
  for (size_t i = 0; i < tensors.size(); ++i) {
    const TensorConfig &t = tensors[i];
    if (t.dims.size() >= 2) {
      if (t.dims[0] <= 0) throw std::invalid_argument("dims[0] (batch) must be positive");
      if (t.dims[0] > std::numeric_limits<int>::max()) throw std::invalid_argument("dims[0] (batch) exceeds int max");
      bmax_ = std::max(bmax_, static_cast<int>(t.dims[0]));
    }
    const size_t numel = product_tail(t.dims);
    const size_t es = element_size(t.type);
    if (es != 0 && numel > std::numeric_limits<size_t>::max() / es)
      throw std::overflow_error("SlotPool: per-sample bytes overflow for tensor " + std::to_string(i));
    types_.push_back(t.type);
    per_sample_numel_.push_back(numel);
    per_sample_bytes_.push_back(numel * es);
  }
  slots_.resize(static_cast<size_t>(slots));
  infos_.resize(static_cast<size_t>(slots));
  for (int s = 0; s < slots; ++s) {
    SlotPoolSlot &slot = slots_[static_cast<size_t>(s)];
    slot.id = s;
    slot.base_ptrs.resize(tensors.size());
    slot.handles.resize(tensors.size());
    infos_[static_cast<size_t>(s)].resize(tensors.size());
    for (size_t i = 0; i < tensors.size(); ++i) {
      const size_t total_numel = per_sample_numel_[i] * static_cast<size_t>(bmax_);
      if (total_numel > std::numeric_limits<uint32_t>::max())
        throw std::overflow_error("SlotPool: element count exceeds StarPU's uint32 nx");
      const size_t total_bytes = per_sample_bytes_[i] * static_cast<size_t>(bmax_);
      slot.base_ptrs[i] = alloc_host(total_bytes, want_pinned, &infos_[static_cast<size_t>(s)][i]);
      starpu_data_handle_t h = nullptr;
      starpu_vector_data_register(&h, STARPU_MAIN_RAM, reinterpret_cast<uintptr_t>(slot.base_ptrs[i]),
                                  static_cast<uint32_t>(total_numel), element_size(types_[i]));
      if (h == nullptr) throw StarPURegistrationException("Failed to register StarPU vector handle");
      slot.handles[i] = h;
    }
    free_ids_.push_back(s);
  }
}

SlotPool::~SlotPool() {
  for (size_t s = 0; s < slots_.size(); ++s)
    for (size_t i = 0; i < slots_[s].handles.size(); ++i) {
      if (slots_[s].handles[i]) starpu_data_unregister(slots_[s].handles[i]);
      free_host(slots_[s].base_ptrs[i], infos_[s][i]);
    }
}

auto SlotPool::acquire() -> int {
  std::unique_lock<std::mutex> lock(mtx_);
  cv_.wait(lock, [this] { return !free_ids_.empty(); });
  const int id = free_ids_.back();
  free_ids_.pop_back();
  return id;
}

auto SlotPool::try_acquire() -> std::optional<int> {
  std::scoped_lock lock(mtx_);
  if (free_ids_.empty()) return std::nullopt;
  const int id = free_ids_.back();
  free_ids_.pop_back();
  return id;
}

void SlotPool::release(int slot_id) {
  {
    const std::scoped_lock lock(mtx_);
    free_ids_.push_back(slot_id);
  }
  cv_.notify_one();
}

auto SlotPool::free_slots() const -> int {
  const std::scoped_lock lock(mtx_);
  return static_cast<int>(free_ids_.size());
}

}  // namespace starpu_server
