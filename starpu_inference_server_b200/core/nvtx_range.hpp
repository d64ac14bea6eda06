// RAII NVTX range around the codelet body, as the reference keeps one around its own
// (/root/reference/src/core/starpu_setup.cpp:817, class NvtxRange in src/utils/nvtx.hpp:73-95: push in the
// constructor, pop in the destructor, hooks replaceable by tests).  NVTX v3 is header-only: without a profiler
// attached the calls return immediately, with Nsight Systems attached every task shows as one named range
// ("codelet.cuda job <request id> dev <device>") above its kernels.
#pragma once

#include <atomic>
#include <cstdio>

#if defined(__has_include)
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
#define SPU_HAVE_NVTX 1
#endif
#endif

namespace starpu_server {

class NvtxRange {
 public:
  using PushHook = void (*)(const char *label);
  using PopHook = void (*)();

  explicit NvtxRange(const char *label) {
    if (PushHook h = push_hook().load(std::memory_order_acquire)) h(label);
    else default_push(label);
  }
  ~NvtxRange() {
    if (PopHook h = pop_hook().load(std::memory_order_acquire)) h();
    else default_pop();
  }
  NvtxRange(const NvtxRange &) = delete;
  auto operator=(const NvtxRange &) -> NvtxRange & = delete;

  // Tests observe the ranges through hooks (nullptr restores the NVTX calls).
  static void SetHooks(PushHook push, PopHook pop) {
    push_hook().store(push, std::memory_order_release);
    pop_hook().store(pop, std::memory_order_release);
  }
  static void ResetHooks() { SetHooks(nullptr, nullptr); }

 private:
  static auto push_hook() -> std::atomic<PushHook> & {
    static std::atomic<PushHook> h{nullptr};
    return h;
  }
  static auto pop_hook() -> std::atomic<PopHook> & {
    static std::atomic<PopHook> h{nullptr};
    return h;
  }
  static void default_push(const char *label) {
#ifdef SPU_HAVE_NVTX
    nvtxRangePushA(label);
#else
    (void)label;
#endif
  }
  static void default_pop() {
#ifdef SPU_HAVE_NVTX
    nvtxRangePop();
#endif
  }
};

}  // namespace starpu_server
