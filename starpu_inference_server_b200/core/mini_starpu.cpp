// mini_starpu.cpp -- in-process stand-in for the StarPU 1.4 runtime calls declared in
// starpu_compat.h (used only when the real library is absent).
//
// Shape of the real thing it keeps (what the inference path relies on):
//  * one worker THREAD + one CUDA stream per (device x STARPU_NWORKER_PER_CUDA), worker ids
//    dense from 0, CUDA workers first; starpu_worker_get_id() is thread-local
//  * a central "eager" queue ordered by task priority, FIFO inside a priority
//    (models/bert.yml:5 `scheduler: eager`); a task pinned with
//    execute_on_a_specific_worker only runs on that worker (warm-up, warmup.cpp:225-299)
//  * vector handles registered on STARPU_MAIN_RAM with one lazily allocated copy per CUDA
//    memory node and MSI-style validity: STARPU_R buffers are copied H2D on the worker's
//    stream before the codelet when the device copy is stale, STARPU_W buffers become the
//    only valid copy, starpu_data_acquire(R) on the host fetches the valid copy back (D2H)
//  * cuda_flags & STARPU_CUDA_ASYNC: the worker does not synchronise after cuda_func; it
//    records an event and keeps up to STARPU_CUDA_PIPELINE tasks in flight; the task's
//    callback_func runs (on a per-worker completion thread) once the event has fired
//  * per-interface `nx` can be rewritten on every node by the submitter
//    (starpu_vector_resize_utils.hpp:66-89); transfers move nx*elemsize bytes.
// It is NOT a scheduler study: no performance models, no data prefetch, no MPI.
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <exception>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "starpu_compat.h"

#if !defined(SPU_USE_REAL_STARPU)

struct NodeCopy {
  starpu_vector_interface iface{};
  bool allocated = false;  // device memory owned by the runtime
  bool valid = false;
};

struct _starpu_data_state {
  starpu_data_interface_id id = STARPU_VECTOR_INTERFACE_ID;
  std::mutex mu;
  std::condition_variable cv;
  std::vector<NodeCopy> nodes;  // [0] = main RAM, [1 + i] = CUDA device i of the runtime
  int pending = 0;              // submitted, not yet completed tasks using this handle
  bool acquired = false;
  bool host_pinned_by_us = false;
  uint32_t wt_mask = 0;         // starpu_data_set_wt_mask: nodes every write is propagated to
};

namespace {

struct Runtime;
Runtime *g_rt = nullptr;
std::mutex g_rt_mu;
thread_local int tls_worker = -1;

struct Worker {
  int id = 0;
  starpu_worker_archtype type = STARPU_CUDA_WORKER;
  int dev_index = 0;  // index into Runtime::gpuids (CUDA) / cpu index
  int devid = 0;      // CUDA ordinal
  cudaStream_t stream = nullptr;
  cudaStream_t xfer = nullptr;  // write-through D2H of STARPU_W buffers, behind the compute event
  std::thread thread, completer;
  std::mutex cmu;
  std::condition_variable ccv;
  struct InFlight {
    starpu_task *task = nullptr;
    cudaEvent_t compute = nullptr;  // recorded on `stream` after the codelet returned
    cudaEvent_t landed = nullptr;   // recorded on `xfer` after the write-through copies (null: none issued)
    std::vector<_starpu_data_state *> written_through;
  };
  std::deque<InFlight> inflight;
  std::vector<cudaEvent_t> free_events;  // created once, recycled (guarded by cmu)
  bool cstop = false;

  cudaEvent_t take_event();
  void give_event(cudaEvent_t e) {
    if (e == nullptr) return;
    std::lock_guard<std::mutex> lk(cmu);
    free_events.push_back(e);
  }
};

// (stream, event) pairs for transfers issued by non-worker threads (starpu_data_acquire from a
// callback or the submitter): created on first use, recycled, one free list per device.
struct XferSlot {
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;
};

struct Runtime {
  std::vector<unsigned> gpuids;
  std::vector<std::unique_ptr<Worker>> workers;
  int pipeline = 1;
  std::mutex qmu;
  std::condition_variable qcv, idle_cv;
  std::multimap<int, starpu_task *, std::greater<int>> queue;  // priority desc, FIFO within
  long outstanding = 0;  // submitted tasks whose callback has not returned yet
  bool stop = false;
  std::atomic<long> codelet_failures{0};
  std::string last_failure;
  std::mutex xfer_mu;
  std::vector<std::vector<XferSlot>> host_xfer;  // per device: free transfer slots
};

unsigned node_of_dev(int dev_index) { return 1u + static_cast<unsigned>(dev_index); }

starpu_data_handle_t *task_handles(starpu_task *t) { return t->dyn_handles ? t->dyn_handles : t->handles; }
starpu_data_access_mode task_mode(starpu_task *t, int i) {
  if (t->dyn_modes) return t->dyn_modes[i];
  if (t->cl && t->cl->nbuffers != STARPU_VARIABLE_NBUFFERS) {
    if (t->cl->dyn_modes) return t->cl->dyn_modes[i];
    return t->cl->modes[i];
  }
  return t->modes[i];
}
int task_nbuffers(starpu_task *t) {
  if (t->cl && t->cl->nbuffers != STARPU_VARIABLE_NBUFFERS) return t->cl->nbuffers;
  return t->nbuffers;
}

size_t live_bytes(const starpu_vector_interface &v) { return static_cast<size_t>(v.nx) * v.elemsize; }

// Make the copy on `node` usable for `mode` by a task about to run on `stream`.
bool make_valid_on_device(_starpu_data_state *h, unsigned node, starpu_data_access_mode mode,
                          cudaStream_t stream, std::string *err) {
  std::lock_guard<std::mutex> lk(h->mu);
  NodeCopy &dst = h->nodes[node];
  NodeCopy &host = h->nodes[0];
  if (!dst.allocated) {
    void *p = nullptr;
    const size_t cap = host.iface.allocsize ? host.iface.allocsize : live_bytes(host.iface);
    if (cudaMalloc(&p, cap ? cap : 1) != cudaSuccess) {
      *err = "mini_starpu: cudaMalloc of a device copy failed";
      cudaGetLastError();
      return false;
    }
    dst.iface = host.iface;
    dst.iface.ptr = reinterpret_cast<uintptr_t>(p);
    dst.iface.dev_handle = dst.iface.ptr;
    dst.allocated = true;
    dst.valid = false;
  }
  dst.iface.nx = host.iface.nx;  // the submitter resizes node 0 at least
  if ((mode & STARPU_R) && !dst.valid) {
    if (!host.valid) {
      // fetch through the host from whichever device holds it (not needed by this server:
      // inputs are always host-produced), keep it simple and explicit
      *err = "mini_starpu: no valid host copy to read from";
      return false;
    }
    if (cudaMemcpyAsync(reinterpret_cast<void *>(dst.iface.ptr), reinterpret_cast<void *>(host.iface.ptr),
                        live_bytes(host.iface), cudaMemcpyHostToDevice, stream) != cudaSuccess) {
      *err = "mini_starpu: H2D copy failed";
      cudaGetLastError();
      return false;
    }
    dst.valid = true;
  }
  if (mode & STARPU_W) {
    for (NodeCopy &n : h->nodes) n.valid = false;
    dst.valid = true;
  }
  return true;
}

void finish_task(Runtime *rt, starpu_task *task) {
  const int nb = task_nbuffers(task);
  starpu_data_handle_t *hs = task_handles(task);
  for (int i = 0; i < nb; ++i) {
    _starpu_data_state *h = hs[i];
    {
      std::lock_guard<std::mutex> lk(h->mu);
      --h->pending;
    }
    h->cv.notify_all();
  }
  if (task->callback_func) task->callback_func(task->callback_arg);
  const bool destroy = task->destroy;
  const bool dyn = task->dyn_handles != nullptr;
  (void)dyn;
  if (destroy) starpu_task_destroy(task);
  {
    std::lock_guard<std::mutex> lk(rt->qmu);
    --rt->outstanding;
  }
  rt->idle_cv.notify_all();
}

// Completion waits (one completer thread per worker, one per output acquire) sleep on the event
// instead of spinning: with one process per GPU and 4 workers each, 8 ranks would otherwise keep
// 32+ host cores busy polling, which is what capped the end-to-end rate at 8 GPUs.  SPU_SPIN_SYNC=1
// restores the spinning waits (lowest latency when cores are plentiful).
unsigned sync_event_flags() {
  static const bool spin = std::getenv("SPU_SPIN_SYNC") != nullptr;
  return spin ? cudaEventDisableTiming : (cudaEventDisableTiming | cudaEventBlockingSync);
}

cudaEvent_t Worker::take_event() {
  {
    std::lock_guard<std::mutex> lk(cmu);
    if (!free_events.empty()) {
      cudaEvent_t e = free_events.back();
      free_events.pop_back();
      return e;
    }
  }
  cudaEvent_t e = nullptr;  // pool exhausted (never in steady state: 2 per in-flight task are pre-created)
  if (cudaEventCreateWithFlags(&e, sync_event_flags()) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return e;
}

void completer_loop(Runtime *rt, Worker *w) {
  tls_worker = -1;  // callbacks do not run "on" a worker (starpu_worker_get_id() == -1)
  if (w->type == STARPU_CUDA_WORKER) cudaSetDevice(w->devid);
  for (;;) {
    Worker::InFlight f{};
    {
      std::unique_lock<std::mutex> lk(w->cmu);
      w->ccv.wait(lk, [&] { return w->cstop || !w->inflight.empty(); });
      if (w->inflight.empty()) return;
      f = w->inflight.front();
    }
    cudaError_t e = cudaSuccess;
    if (f.landed) e = cudaEventSynchronize(f.landed);  // implies the compute event
    else if (f.compute) e = cudaEventSynchronize(f.compute);
    if (e != cudaSuccess) {
      // an asynchronous fault of the task's kernels or copies: count it as a codelet failure and leave
      // the host copies invalid, so the output acquire of the callback reports the task as failed
      cudaGetLastError();
      ++rt->codelet_failures;
      std::lock_guard<std::mutex> lk(rt->qmu);
      rt->last_failure = std::string("asynchronous CUDA failure after the codelet: ") + cudaGetErrorString(e);
    } else {
      for (_starpu_data_state *h : f.written_through) {
        std::lock_guard<std::mutex> lk(h->mu);
        h->nodes[0].valid = true;  // shared: the device copy stays valid too
      }
    }
    w->give_event(f.compute);
    w->give_event(f.landed);
    {
      std::lock_guard<std::mutex> lk(w->cmu);
      w->inflight.pop_front();
    }
    w->ccv.notify_all();
    finish_task(rt, f.task);
  }
}

void worker_loop(Runtime *rt, Worker *w) {
  tls_worker = w->id;
  if (w->type == STARPU_CUDA_WORKER) cudaSetDevice(w->devid);
  for (;;) {
    starpu_task *task = nullptr;
    // pipeline back-pressure BEFORE taking a task: a task must not wait behind this worker's full
    // pipeline while another worker is idle (only this thread adds to `inflight`, so room stays room)
    {
      std::unique_lock<std::mutex> lk(w->cmu);
      w->ccv.wait(lk, [&] { return static_cast<int>(w->inflight.size()) < rt->pipeline; });
    }
    {
      std::unique_lock<std::mutex> lk(rt->qmu);
      for (;;) {
        if (rt->stop) return;
        for (auto it = rt->queue.begin(); it != rt->queue.end(); ++it) {
          starpu_task *t = it->second;
          if (t->execute_on_a_specific_worker && static_cast<int>(t->workerid) != w->id) continue;
          const bool can = w->type == STARPU_CUDA_WORKER ? (t->cl && t->cl->cuda_funcs[0] != nullptr)
                                                         : (t->cl && t->cl->cpu_funcs[0] != nullptr);
          if (!can) continue;
          task = t;
          rt->queue.erase(it);
          break;
        }
        if (task) break;
        rt->qcv.wait(lk);
      }
    }
    const int nb = task_nbuffers(task);
    starpu_data_handle_t *hs = task_handles(task);
    std::vector<void *> ifaces(static_cast<size_t>(nb));
    std::string err;
    bool ok = true;
    for (int i = 0; i < nb && ok; ++i) {
      if (w->type == STARPU_CUDA_WORKER) {
        const unsigned node = node_of_dev(w->dev_index);
        ok = make_valid_on_device(hs[i], node, task_mode(task, i), w->stream, &err);
        ifaces[i] = &hs[i]->nodes[node].iface;
      } else {
        ifaces[i] = &hs[i]->nodes[0].iface;
      }
    }
    if (ok) {
      try {
        if (w->type == STARPU_CUDA_WORKER) task->cl->cuda_funcs[0](ifaces.data(), task->cl_arg);
        else task->cl->cpu_funcs[0](ifaces.data(), task->cl_arg);
      } catch (const std::exception &e) {
        // real StarPU has C frames here; the stand-in records the failure and still
        // completes the task so that the submitter's callback can report it
        ok = false;
        err = e.what();
      } catch (...) {
        ok = false;
        err = "unknown exception out of the codelet";
      }
    }
    if (!ok) {
      ++rt->codelet_failures;
      std::lock_guard<std::mutex> lk(rt->qmu);
      rt->last_failure = err;
    }
    Worker::InFlight f{};
    f.task = task;
    if (w->type == STARPU_CUDA_WORKER) {
      const bool async = task->cl->cuda_flags[0] & STARPU_CUDA_ASYNC;
      if (!async) cudaStreamSynchronize(w->stream);
      f.compute = w->take_event();
      if (f.compute) cudaEventRecord(f.compute, w->stream);
      // write-through (starpu_data_set_wt_mask with the main-RAM bit): the D2H of every STARPU_W buffer is
      // enqueued on the worker's transfer stream behind the compute event right here, so it overlaps the
      // next task's kernels and the callback's acquire finds the host copy valid instead of issuing it
      if (ok && f.compute) {
        const unsigned node = node_of_dev(w->dev_index);
        bool waited = false;
        for (int i = 0; i < nb; ++i) {
          _starpu_data_state *h = hs[i];
          if (!(task_mode(task, i) & STARPU_W)) continue;
          std::lock_guard<std::mutex> lk(h->mu);
          if (!(h->wt_mask & (1u << STARPU_MAIN_RAM))) continue;
          NodeCopy &src = h->nodes[node];
          NodeCopy &host = h->nodes[0];
          if (!src.valid || host.iface.ptr == 0) continue;
          if (!waited) {
            cudaStreamWaitEvent(w->xfer, f.compute, 0);
            waited = true;
          }
          if (cudaMemcpyAsync(reinterpret_cast<void *>(host.iface.ptr), reinterpret_cast<void *>(src.iface.ptr),
                              live_bytes(src.iface), cudaMemcpyDeviceToHost, w->xfer) == cudaSuccess) {
            host.iface.nx = src.iface.nx;
            f.written_through.push_back(h);
          } else {
            cudaGetLastError();
          }
        }
        if (!f.written_through.empty()) {
          f.landed = w->take_event();
          if (f.landed) cudaEventRecord(f.landed, w->xfer);
          else f.written_through.clear();  // no event to wait on: fall back to fetch-on-acquire
        }
      }
    }
    {
      std::lock_guard<std::mutex> lk(w->cmu);
      w->inflight.push_back(std::move(f));
    }
    w->ccv.notify_all();
  }
}

int env_int(const char *name, int dflt) {
  const char *s = std::getenv(name);
  if (!s || !*s) return dflt;
  const int v = std::atoi(s);
  return v > 0 ? v : dflt;
}

// Bring the valid copy of `h` to main RAM (caller holds h->mu).
int fetch_to_host_locked(Runtime *rt, _starpu_data_state *h) {
  NodeCopy &host = h->nodes[0];
  if (host.valid) return 0;
  for (size_t n = 1; n < h->nodes.size(); ++n) {
    NodeCopy &src = h->nodes[n];
    if (!src.valid) continue;
    const int dev_index = static_cast<int>(n) - 1;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(static_cast<int>(rt->gpuids[dev_index]));
    // a recycled (stream, event) pair per concurrent acquire: concurrent acquires do not serialise on one
    // stream, and nothing is created or destroyed per call in steady state
    XferSlot slot;
    {
      std::lock_guard<std::mutex> lk(rt->xfer_mu);
      auto &pool = rt->host_xfer[dev_index];
      if (!pool.empty()) {
        slot = pool.back();
        pool.pop_back();
      }
    }
    cudaError_t e = cudaSuccess;
    if (slot.stream == nullptr) {
      e = cudaStreamCreateWithFlags(&slot.stream, cudaStreamNonBlocking);
      if (e == cudaSuccess) e = cudaEventCreateWithFlags(&slot.done, sync_event_flags());
    }
    if (e == cudaSuccess)
      e = cudaMemcpyAsync(reinterpret_cast<void *>(host.iface.ptr), reinterpret_cast<void *>(src.iface.ptr),
                          live_bytes(src.iface), cudaMemcpyDeviceToHost, slot.stream);
    // wait on an event so the wait can block (see sync_event_flags) instead of spinning for the whole D2H
    if (e == cudaSuccess) e = cudaEventRecord(slot.done, slot.stream);
    if (e == cudaSuccess) e = cudaEventSynchronize(slot.done);
    if (slot.stream != nullptr && slot.done != nullptr) {
      std::lock_guard<std::mutex> lk(rt->xfer_mu);
      rt->host_xfer[dev_index].push_back(slot);
    } else {
      if (slot.done) cudaEventDestroy(slot.done);
      if (slot.stream) cudaStreamDestroy(slot.stream);
    }
    if (prev >= 0) cudaSetDevice(prev);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return -1;
    }
    host.iface.nx = src.iface.nx;
    host.valid = true;
    return 0;
  }
  host.valid = true;  // never written anywhere: the registered host buffer is the data
  return 0;
}

}  // namespace

extern "C" {

int starpu_conf_init(struct starpu_conf *conf) {
  if (!conf) return -1;
  std::memset(conf, 0, sizeof(*conf));
  conf->ncpus = -1;
  conf->ncuda = -1;
  conf->sched_policy_name = "eager";
  return 0;
}

int starpu_init(struct starpu_conf *conf) {
  std::lock_guard<std::mutex> lk(g_rt_mu);
  if (g_rt) return -17;  // already initialised (one runtime per process, like StarPU)
  struct starpu_conf dflt;
  if (!conf) {
    starpu_conf_init(&dflt);
    conf = &dflt;
  }
  auto rt = std::make_unique<Runtime>();
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess) {
    cudaGetLastError();
    ndev = 0;
  }
  int ncuda = conf->ncuda < 0 ? ndev : conf->ncuda;
  if (ncuda > ndev && !conf->use_explicit_workers_cuda_gpuid) ncuda = ndev;
  for (int i = 0; i < ncuda; ++i) {
    const unsigned id = conf->use_explicit_workers_cuda_gpuid ? conf->workers_cuda_gpuid[i] : static_cast<unsigned>(i);
    if (static_cast<int>(id) >= ndev) return -19;  // ENODEV
    rt->gpuids.push_back(id);
  }
  const int per = conf->nworker_per_cuda > 0 ? conf->nworker_per_cuda : env_int("STARPU_NWORKER_PER_CUDA", 1);
  rt->pipeline = conf->cuda_pipeline > 0 ? conf->cuda_pipeline : env_int("STARPU_CUDA_PIPELINE", 2);
  int next_id = 0;
  for (size_t d = 0; d < rt->gpuids.size(); ++d) {
    cudaSetDevice(static_cast<int>(rt->gpuids[d]));
    rt->host_xfer.emplace_back();
    for (int k = 0; k < per; ++k) {
      auto w = std::make_unique<Worker>();
      w->id = next_id++;
      w->type = STARPU_CUDA_WORKER;
      w->dev_index = static_cast<int>(d);
      w->devid = static_cast<int>(rt->gpuids[d]);
      cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking);
      cudaStreamCreateWithFlags(&w->xfer, cudaStreamNonBlocking);
      for (int e = 0; e < 2 * (rt->pipeline + 1); ++e) {
        cudaEvent_t ev = nullptr;
        if (cudaEventCreateWithFlags(&ev, sync_event_flags()) == cudaSuccess) w->free_events.push_back(ev);
      }
      rt->workers.push_back(std::move(w));
    }
  }
  const int ncpus = conf->ncpus < 0 ? 0 : conf->ncpus;  // the new codelet has no cpu_func
  for (int c = 0; c < ncpus; ++c) {
    auto w = std::make_unique<Worker>();
    w->id = next_id++;
    w->type = STARPU_CPU_WORKER;
    w->dev_index = c;
    w->devid = c;
    rt->workers.push_back(std::move(w));
  }
  g_rt = rt.release();
  for (auto &w : g_rt->workers) {
    Worker *wp = w.get();
    wp->completer = std::thread(completer_loop, g_rt, wp);
    wp->thread = std::thread(worker_loop, g_rt, wp);
  }
  return 0;
}

int starpu_is_initialized(void) {
  std::lock_guard<std::mutex> lk(g_rt_mu);
  return g_rt != nullptr ? 1 : 0;
}

int starpu_task_wait_for_all(void) {
  Runtime *rt = g_rt;
  if (!rt) return 0;
  std::unique_lock<std::mutex> lk(rt->qmu);
  rt->idle_cv.wait(lk, [&] { return rt->outstanding == 0; });
  return 0;
}

void starpu_shutdown(void) {
  Runtime *rt;
  {
    std::lock_guard<std::mutex> lk(g_rt_mu);
    rt = g_rt;
  }
  if (!rt) return;
  starpu_task_wait_for_all();
  {
    std::lock_guard<std::mutex> lk(rt->qmu);
    rt->stop = true;
  }
  rt->qcv.notify_all();
  for (auto &w : rt->workers) {
    if (w->thread.joinable()) w->thread.join();
    {
      std::lock_guard<std::mutex> lk(w->cmu);
      w->cstop = true;
    }
    w->ccv.notify_all();
    if (w->completer.joinable()) w->completer.join();
    if (w->type == STARPU_CUDA_WORKER) {
      cudaSetDevice(w->devid);
      if (w->stream) cudaStreamDestroy(w->stream);
      if (w->xfer) cudaStreamDestroy(w->xfer);
      for (cudaEvent_t e : w->free_events) cudaEventDestroy(e);
    }
  }
  for (size_t d = 0; d < rt->host_xfer.size(); ++d) {
    cudaSetDevice(static_cast<int>(rt->gpuids[d]));
    for (XferSlot &x : rt->host_xfer[d]) {
      cudaEventDestroy(x.done);
      cudaStreamDestroy(x.stream);
    }
  }
  {
    std::lock_guard<std::mutex> lk(g_rt_mu);
    g_rt = nullptr;
  }
  delete rt;
}

void starpu_codelet_init(struct starpu_codelet *cl) {
  if (cl) std::memset(cl, 0, sizeof(*cl));
}

struct starpu_task *starpu_task_create(void) {
  auto *t = static_cast<starpu_task *>(std::calloc(1, sizeof(starpu_task)));
  if (t) {
    t->destroy = 1;
    t->priority = STARPU_DEFAULT_PRIO;
  }
  return t;
}

void starpu_task_destroy(struct starpu_task *task) {
  if (!task) return;
  // dyn_handles / dyn_modes are malloc'ed by the submitter and freed with the task, as
  // StarPU does for tasks with destroy = 1 (inference_task.cpp:758-775)
  std::free(task->dyn_handles);
  std::free(task->dyn_modes);
  std::free(task);
}

int starpu_task_submit(struct starpu_task *task) {
  Runtime *rt = g_rt;
  if (!rt) return -19;
  if (!task || !task->cl) return -22;
  const int nb = task_nbuffers(task);
  if (nb < 0) return -22;
  const bool has_cuda = task->cl->cuda_funcs[0] != nullptr;
  const bool has_cpu = task->cl->cpu_funcs[0] != nullptr;
  bool runnable = false;
  for (auto &w : rt->workers) {
    if (task->execute_on_a_specific_worker && static_cast<int>(task->workerid) != w->id) continue;
    if ((w->type == STARPU_CUDA_WORKER && has_cuda) || (w->type == STARPU_CPU_WORKER && has_cpu)) runnable = true;
  }
  if (!runnable) return -19;  // -ENODEV: no worker can execute this task (StarPU semantics)
  starpu_data_handle_t *hs = task_handles(task);
  for (int i = 0; i < nb; ++i) {
    if (!hs[i]) return -22;
    std::lock_guard<std::mutex> lk(hs[i]->mu);
    ++hs[i]->pending;
  }
  const bool sync = task->synchronous;
  {
    std::lock_guard<std::mutex> lk(rt->qmu);
    ++rt->outstanding;
    rt->queue.emplace(task->priority, task);
  }
  rt->qcv.notify_all();
  if (sync) starpu_task_wait_for_all();
  return 0;
}

static void register_common(starpu_data_handle_t *handle, starpu_data_interface_id id, uintptr_t ptr,
                            uint32_t nx, size_t elemsize) {
  if (!handle) return;
  auto *h = new _starpu_data_state();
  h->id = id;
  const size_t nnodes = 1 + (g_rt ? g_rt->gpuids.size() : 0);
  h->nodes.resize(nnodes);
  starpu_vector_interface &v = h->nodes[0].iface;
  v.id = id;
  v.ptr = ptr;
  v.dev_handle = ptr;
  v.offset = 0;
  v.nx = nx;
  v.elemsize = elemsize;
  v.allocsize = static_cast<size_t>(nx) * elemsize;
  v.slice_base = 0;
  h->nodes[0].valid = true;
  for (size_t n = 1; n < nnodes; ++n) {
    h->nodes[n].iface = v;
    h->nodes[n].iface.ptr = 0;
    h->nodes[n].iface.dev_handle = 0;
  }
  *handle = h;
}

void starpu_vector_data_register(starpu_data_handle_t *handle, int home_node, uintptr_t ptr, uint32_t nx,
                                 size_t elemsize) {
  (void)home_node;
  register_common(handle, STARPU_VECTOR_INTERFACE_ID, ptr, nx, elemsize);
}

void starpu_variable_data_register(starpu_data_handle_t *handle, int home_node, uintptr_t ptr, size_t size) {
  (void)home_node;
  register_common(handle, STARPU_VARIABLE_INTERFACE_ID, ptr, 1, size);
}

void starpu_data_unregister(starpu_data_handle_t h) {
  if (!h) return;
  {
    std::unique_lock<std::mutex> lk(h->mu);
    h->cv.wait(lk, [&] { return h->pending == 0 && !h->acquired; });
    if (g_rt) fetch_to_host_locked(g_rt, h);  // unregister brings the data home, like StarPU
    for (size_t n = 1; n < h->nodes.size(); ++n)
      if (h->nodes[n].allocated) {
        int prev = -1;
        cudaGetDevice(&prev);
        if (g_rt) cudaSetDevice(static_cast<int>(g_rt->gpuids[n - 1]));
        cudaFree(reinterpret_cast<void *>(h->nodes[n].iface.ptr));
        if (prev >= 0) cudaSetDevice(prev);
      }
  }
  delete h;
}

void starpu_data_unregister_submit(starpu_data_handle_t h) { starpu_data_unregister(h); }

int starpu_data_acquire(starpu_data_handle_t h, enum starpu_data_access_mode mode) {
  if (!h) return -22;
  std::unique_lock<std::mutex> lk(h->mu);
  h->cv.wait(lk, [&] { return h->pending == 0 && !h->acquired; });
  if ((mode & STARPU_R) && g_rt) {
    if (fetch_to_host_locked(g_rt, h) != 0) return -5;
  }
  if (mode & STARPU_W) {
    for (NodeCopy &n : h->nodes) n.valid = false;
    h->nodes[0].valid = true;
  }
  h->acquired = true;
  return 0;
}

int starpu_data_acquire_cb(starpu_data_handle_t h, enum starpu_data_access_mode mode, void (*callback)(void *),
                           void *arg) {
  const int rc = starpu_data_acquire(h, mode);
  if (rc != 0) return rc;
  if (callback) callback(arg);
  return 0;
}

void starpu_data_set_wt_mask(starpu_data_handle_t h, uint32_t wt_mask) {
  if (!h) return;
  std::lock_guard<std::mutex> lk(h->mu);
  h->wt_mask = wt_mask;
}

void starpu_data_release(starpu_data_handle_t h) {
  if (!h) return;
  {
    std::lock_guard<std::mutex> lk(h->mu);
    h->acquired = false;
  }
  h->cv.notify_all();
}

enum starpu_data_interface_id starpu_data_get_interface_id(starpu_data_handle_t h) {
  return h ? h->id : STARPU_UNKNOWN_INTERFACE_ID;
}

void *starpu_data_get_interface_on_node(starpu_data_handle_t h, unsigned node) {
  if (!h || node >= h->nodes.size()) return nullptr;
  return &h->nodes[node].iface;
}

unsigned starpu_memory_nodes_get_count(void) { return 1u + (g_rt ? static_cast<unsigned>(g_rt->gpuids.size()) : 0u); }

int starpu_memory_pin(void *addr, size_t size) {
  if (cudaHostRegister(addr, size, cudaHostRegisterPortable) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return 0;
}
int starpu_memory_unpin(void *addr, size_t size) {
  (void)size;
  if (cudaHostUnregister(addr) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return 0;
}

int starpu_worker_get_id(void) { return tls_worker; }
int starpu_worker_get_devid(int id) {
  if (!g_rt || id < 0 || id >= static_cast<int>(g_rt->workers.size())) return -1;
  return g_rt->workers[id]->devid;
}
unsigned starpu_worker_get_count(void) { return g_rt ? static_cast<unsigned>(g_rt->workers.size()) : 0u; }
enum starpu_worker_archtype starpu_worker_get_type(int id) {
  if (!g_rt || id < 0 || id >= static_cast<int>(g_rt->workers.size())) return STARPU_ANY_WORKER;
  return g_rt->workers[id]->type;
}
unsigned starpu_worker_get_ids_by_type(enum starpu_worker_archtype type, int *workerids, unsigned maxsize) {
  unsigned n = 0;
  if (!g_rt) return 0;
  for (auto &w : g_rt->workers)
    if (w->type == type || type == STARPU_ANY_WORKER) {
      if (workerids && n < maxsize) workerids[n] = w->id;
      ++n;
    }
  return n < maxsize || !workerids ? n : maxsize;
}
unsigned starpu_worker_get_memory_node(unsigned workerid) {
  if (!g_rt || workerid >= g_rt->workers.size()) return 0;
  Worker *w = g_rt->workers[workerid].get();
  return w->type == STARPU_CUDA_WORKER ? node_of_dev(w->dev_index) : 0u;
}
int starpu_sched_get_min_priority(void) { return -1000000; }
int starpu_sched_get_max_priority(void) { return 1000000; }
void *starpu_cuda_get_local_stream_ptr(void) {
  if (!g_rt || tls_worker < 0) return nullptr;
  return g_rt->workers[tls_worker]->stream;
}

/* mini-runtime diagnostics (not StarPU API) */
long spu_mini_starpu_codelet_failures(void) { return g_rt ? g_rt->codelet_failures.load() : 0; }
int spu_mini_starpu_last_failure(char *buf, size_t cap) {
  if (!g_rt || !buf || cap == 0) return 0;
  std::lock_guard<std::mutex> lk(g_rt->qmu);
  std::snprintf(buf, cap, "%s", g_rt->last_failure.c_str());
  return static_cast<int>(g_rt->last_failure.size());
}

}  // extern "C"
#endif  // !SPU_USE_REAL_STARPU
