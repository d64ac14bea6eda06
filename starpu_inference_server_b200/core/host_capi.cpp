// extern "C" surface of libspu_host.so (include/spu_host.h): server object = StarPU runtime
// + TaskRunner + BatchCollector; tickets instead of foreign-thread callbacks into Python.
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <random>
#include <thread>

#include "../../include/spu_host.h"
#include "task_runner.hpp"

#if !defined(SPU_USE_REAL_STARPU)
extern "C" long spu_mini_starpu_codelet_failures(void);
#endif

using namespace starpu_server;

namespace {
thread_local std::string g_host_error;
void set_host_error(const std::string &m) { g_host_error = m; }

ScalarType to_scalar(int32_t dt) {
  switch (dt) {
    case SPU_DT_F32: return ScalarType::Float32;
    case SPU_DT_F64: return ScalarType::Float64;
    case SPU_DT_I32: return ScalarType::Int32;
    case SPU_DT_I64: return ScalarType::Int64;
    case SPU_DT_I8: return ScalarType::Int8;
    case SPU_DT_U8: return ScalarType::UInt8;
    case SPU_DT_F16: return ScalarType::Float16;
    case SPU_DT_BF16: return ScalarType::BFloat16;
    default: return ScalarType::Bool;
  }
}

TensorConfig to_cfg(const spu_tensor_cfg &c) {
  TensorConfig t;
  t.name = c.name;
  t.type = to_scalar(c.dtype);
  for (int i = 0; i < c.ndims && i < SPU_HOST_MAX_DIMS; ++i) t.dims.push_back(c.dims[i]);
  return t;
}

double ms_between(MonotonicClock::time_point a, MonotonicClock::time_point b) {
  if (a == MonotonicClock::time_point{} || b == MonotonicClock::time_point{}) return 0.0;
  return std::chrono::duration<double, std::milli>(b - a).count();
}

struct Ticket {
  int expected = 0, done = 0;
  std::vector<spu_result> results;
};

// Does a batch of `rows` requests x `seq` tokens end inside a GEMM row tile, with room left to grow?
bool tile_aware_wants_more(int64_t rows, int64_t seq, int64_t max_rows, int tile_rows) {
  if (tile_rows <= 0 || rows >= max_rows) return false;
  return (rows * std::max<int64_t>(seq, 1)) % tile_rows != 0;
}
}  // namespace

struct spu_server {
  ReplicaRegistry registry;  // device / worker -> replica; outlives the runner and the runtime
  std::unique_ptr<TaskRunner> runner;
  ModelConfig model;
  bool starpu_up = false;
  // tickets
  std::mutex tmu;
  std::condition_variable tcv;
  std::map<int64_t, Ticket> tickets;
  int64_t next_ticket = 0;
  std::atomic<long> completed{0};
  // BatchCollector (batch_collector_component.cpp): one thread forming batches
  std::mutex qmu;
  std::condition_variable qcv;
  std::deque<std::shared_ptr<InferenceJob>> queue;
  std::thread collector;
  bool stop = false;
  int coalesce_ms = 0;
  int length_bucket = 0;  // variable_length: requests per ceil(L / length_bucket) bucket batch together
  int tile_rows = 0;      // > 0: GEMM row tile of the codelet's kernels; see tile_aware_close()
  int forming = 0;  // batches popped from the queue but not yet submitted (guarded by qmu)
  // Response-copy helpers (the per-request copy into caller buffers, fill_output_tensor in the reference,
  // must not serialise on the one completer thread of a worker): a finished task's on_complete calls are spread
  // over these threads plus the calling one.
  std::vector<std::thread> helpers;
  std::mutex hmu;
  std::condition_variable hcv;
  std::deque<std::function<void()>> hq;
  bool hstop = false;

  void helper_loop() {
    for (;;) {
      std::function<void()> fn;
      {
        std::unique_lock<std::mutex> lk(hmu);
        hcv.wait(lk, [&] { return hstop || !hq.empty(); });
        if (hq.empty()) return;
        fn = std::move(hq.front());
        hq.pop_front();
      }
      fn();
    }
  }

  void run_completions(std::vector<std::function<void()>> &fns) {
    struct Sync {
      std::mutex m;
      std::condition_variable cv;
      size_t left;
    } sync;
    const size_t n = fns.size(), parts = std::min(n, helpers.size() + 1);
    sync.left = parts - 1;
    auto run_part = [&fns, n, parts](size_t p) {
      for (size_t i = p * n / parts; i < (p + 1) * n / parts; ++i) fns[i]();
    };
    {
      std::lock_guard<std::mutex> lk(hmu);
      for (size_t p = 1; p < parts; ++p)
        hq.emplace_back([&sync, run_part, p] {
          run_part(p);
          std::lock_guard<std::mutex> l2(sync.m);
          if (--sync.left == 0) sync.cv.notify_one();
        });
    }
    hcv.notify_all();
    run_part(0);
    std::unique_lock<std::mutex> lk(sync.m);
    sync.cv.wait(lk, [&] { return sync.left == 0; });
  }

  std::shared_ptr<InferenceJob> make_job(const spu_request &rq, int64_t ticket, int index) {
    auto job = std::make_shared<InferenceJob>();
    job->request_id = rq.request_id;
    job->timing.enqueued_time = MonotonicClock::now();
    for (size_t i = 0; i < model.inputs.size(); ++i) {
      HostTensor t;
      t.data = static_cast<std::byte *>(const_cast<void *>(rq.inputs[i]));
      t.type = model.inputs[i].type;
      t.dims = model.inputs[i].dims;
      if (t.dims.empty()) t.dims.push_back(rq.batch);
      else t.dims[0] = rq.batch;
      // variable-length admission: the request states its own sequence length (checked in submit_batch)
      if (model.variable_length && rq.seq_len > 0 && t.dims.size() >= 2) t.dims[1] = rq.seq_len;
      job->inputs.push_back(std::move(t));
    }
    std::vector<void *> dsts(rq.outputs, rq.outputs + model.outputs.size());
    if (model.device_visible_requests) job->output_dsts = dsts;  // the codelet scatters straight into them
    // variable_length: the sequence length this request's response is laid out with
    const int64_t own_seq = model.variable_length && !job->inputs.empty() && job->inputs[0].dims.size() >= 2
                                ? job->inputs[0].dims[1]
                                : 0;
    job->on_complete = [this, ticket, index, dsts, own_seq](const JobResult &res) {
      spu_result out{};
      out.request_id = res.request_id;
      out.ok = res.ok ? 1 : 0;
      out.rows = res.rows;
      out.batch_rows = res.batch_rows;
      out.batch_seq = res.batch_seq;
      out.worker_id = res.worker_id;
      out.device_id = res.device_id;
      out.executed_on = static_cast<int32_t>(res.executed_on);
      out.codelet_ms = ms_between(res.timing.codelet_start_time, res.timing.codelet_end_time);
      out.task_ms = ms_between(res.timing.before_starpu_submitted_time, res.timing.callback_start_time);
      std::snprintf(out.error, sizeof(out.error), "%s", res.error.c_str());
      if (res.ok && !res.outputs.empty() && res.outputs[0].defined() && res.outputs[0].type == ScalarType::Float32)
        std::memcpy(&out.first_value, res.outputs[0].data, sizeof(float));
      else if (res.ok && !res.outputs.empty() && res.outputs[0].defined() &&
               res.outputs[0].type == ScalarType::BFloat16) {
        uint16_t h = 0;
        std::memcpy(&h, res.outputs[0].data, sizeof(h));
        const uint32_t u = static_cast<uint32_t>(h) << 16;
        std::memcpy(&out.first_value, &u, sizeof(float));
      }
      if (res.ok)
        // the response copy the reference makes per request (fill_output_tensor,
        // src/grpc/server/inference_service_io.cpp:377-419): slot view -> caller's buffer.  The caller sizes
        // its buffer by its OWN shape: on a variable_length server a multi-sample request that rode in a longer
        // batch is compacted sample by sample to [batch, own_seq, ...], never the padded [batch, batch_seq, ...].
        for (size_t o = 0; o < res.outputs.size() && o < dsts.size(); ++o) {
          const HostTensor &v = res.outputs[o];
          if (!dsts[o] || !v.defined() || v.data == dsts[o]) continue;  // == : written in place by the codelet
          if (own_seq > 0 && res.batch_seq > own_seq && v.dims.size() >= 2 && v.dims[1] == res.batch_seq) {
            const size_t samples = static_cast<size_t>(v.dims[0]);
            const size_t padded = v.nbytes() / samples;
            const size_t live = padded / static_cast<size_t>(res.batch_seq) * static_cast<size_t>(own_seq);
            for (size_t r = 0; r < samples; ++r)
              std::memcpy(static_cast<std::byte *>(dsts[o]) + r * live, v.data + r * padded, live);
          } else {
            std::memcpy(dsts[o], v.data, v.nbytes());
          }
        }
      out.latency_ms = ms_between(res.timing.enqueued_time, MonotonicClock::now());
      {
        std::lock_guard<std::mutex> lk(tmu);
        Ticket &t = tickets[ticket];
        if (static_cast<int>(t.results.size()) <= index) t.results.resize(static_cast<size_t>(index) + 1);
        t.results[static_cast<size_t>(index)] = out;
        ++t.done;
      }
      ++completed;
      tcv.notify_all();
    };
    return job;
  }

  int64_t open_ticket(int n) {
    std::lock_guard<std::mutex> lk(tmu);
    const int64_t id = next_ticket++;
    tickets[id].expected = n;
    tickets[id].results.resize(static_cast<size_t>(n));
    return id;
  }

  void fail_ticket_entry(int64_t ticket, int index, int request_id, const std::string &why) {
    {
      std::lock_guard<std::mutex> lk(tmu);
      Ticket &t = tickets[ticket];
      spu_result r{};
      r.request_id = request_id;
      r.ok = 0;
      r.worker_id = r.device_id = -1;
      r.executed_on = 2;
      std::snprintf(r.error, sizeof(r.error), "%s", why.c_str());
      t.results[static_cast<size_t>(index)] = r;
      ++t.done;
    }
    tcv.notify_all();
  }

  void collector_loop() {
    for (;;) {
      std::vector<std::shared_ptr<InferenceJob>> batch;
      {
        std::unique_lock<std::mutex> lk(qmu);
        qcv.wait(lk, [&] { return stop || !queue.empty(); });
        if (queue.empty()) return;
        batch.push_back(queue.front());
        queue.pop_front();
        ++forming;
        int64_t rows = batch[0]->batch_size();
        const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(coalesce_ms);
        // variable-length bucketing: the batch belongs to the bucket of its oldest request; the collector
        // looks past other buckets' requests (they keep their queue order and open later batches)
        const bool bucketing = model.variable_length && length_bucket > 0 && !batch[0]->inputs.empty() &&
                               batch[0]->inputs[0].dims.size() >= 2;
        auto bucket_of = [&](const InferenceJob &j) {
          return j.inputs.empty() || j.inputs[0].dims.size() < 2 ? int64_t{-1}
                                                                 : (j.inputs[0].dims[1] - 1) / length_bucket;
        };
        while (bucketing && rows < model.max_batch_size) {
          const int64_t mine = bucket_of(*batch[0]);
          for (auto it = queue.begin(); it != queue.end() && rows < model.max_batch_size;) {
            if (bucket_of(**it) == mine && can_merge_jobs(batch[0].get(), it->get(), true) &&
                rows + (*it)->batch_size() <= model.max_batch_size) {
              rows += (*it)->batch_size();
              batch.push_back(*it);
              it = queue.erase(it);
            } else {
              ++it;
            }
          }
          if (rows >= model.max_batch_size || coalesce_ms <= 0 || stop) break;
          if (qcv.wait_until(lk, deadline) == std::cv_status::timeout) break;  // else: new arrivals, rescan
        }
        while (!bucketing && rows < model.max_batch_size) {
          if (queue.empty()) {
            if (coalesce_ms <= 0 || stop) break;
            if (qcv.wait_until(lk, deadline) == std::cv_status::timeout && queue.empty()) break;
            if (queue.empty()) continue;
          }
          auto &cand = queue.front();
          if (!can_merge_jobs(batch[0].get(), cand.get(), model.variable_length)) break;  // incompatible: starts the next batch
          if (rows + cand->batch_size() > model.max_batch_size) break;
          rows += cand->batch_size();
          batch.push_back(cand);
          queue.pop_front();
        }
        // Tile-aware preference (the reference's hook for batch-size policy is
        // AdaptiveBatchingStrategy::decide, batching_strategy.cpp:195-236): the encoder GEMMs run over
        // M = rows x seq token rows in tiles of `tile_rows`; a batch that ends inside a tile pays for the whole
        // tile.  When the coalesce window closed on such a batch, spend a bounded grace (a quarter of the window)
        // on arrivals that complete the tile -- and take no more than complete it.
        if (tile_rows > 0 && coalesce_ms > 0 && !bucketing && !stop) {
          const auto grace_end = deadline + std::chrono::microseconds(250 * coalesce_ms);
          auto seq_of = [&](const InferenceJob &j) {
            return j.inputs.empty() || j.inputs[0].dims.size() < 2 ? int64_t{1} : j.inputs[0].dims[1];
          };
          int64_t seq = seq_of(*batch[0]);
          for (const auto &j : batch) seq = std::max(seq, seq_of(*j));
          while (tile_aware_wants_more(rows, seq, model.max_batch_size, tile_rows)) {
            if (queue.empty()) {
              if (qcv.wait_until(lk, grace_end) == std::cv_status::timeout && queue.empty()) break;
              if (queue.empty()) continue;
            }
            auto &cand = queue.front();
            if (!can_merge_jobs(batch[0].get(), cand.get(), model.variable_length)) break;
            if (rows + cand->batch_size() > model.max_batch_size) break;
            seq = std::max(seq, seq_of(*cand));
            rows += cand->batch_size();
            batch.push_back(cand);
            queue.pop_front();
          }
        }
      }
      try {
        runner->submit_batch(batch);
      } catch (const std::exception &e) {
        for (auto &j : batch) {
          JobResult r;
          r.request_id = j->request_id;
          r.ok = false;
          r.error = e.what();
          r.timing.enqueued_time = j->timing.enqueued_time;
          if (j->on_complete) j->on_complete(r);
        }
      }
      {
        std::lock_guard<std::mutex> lk(qmu);
        --forming;
      }
    }
  }
};

extern "C" {

const char *spu_host_last_error(void) { return g_host_error.c_str(); }
void spu_host_set_error_for_hooks(const char *m) { set_host_error(m ? m : ""); }

int32_t spu_server_create(const spu_server_cfg *cfg, spu_server **out) {
  if (!cfg || !out) {
    set_host_error("server_create: null argument");
    return 1;
  }
  *out = nullptr;
  try {
    if (cfg->n_devices < 0 || cfg->n_devices > SPU_HOST_MAX_DEVICES) throw std::invalid_argument("bad device count");
    if (cfg->n_inputs <= 0 || cfg->n_inputs > SPU_HOST_MAX_IO || cfg->n_outputs <= 0 ||
        cfg->n_outputs > SPU_HOST_MAX_IO)
      throw std::invalid_argument("bad tensor count");
    auto s = std::make_unique<spu_server>();
    starpu_conf conf;
    starpu_conf_init(&conf);
    conf.sched_policy_name = "eager";
    conf.ncpus = 0;
    conf.ncuda = cfg->n_devices;
    conf.use_explicit_workers_cuda_gpuid = 1;
    for (int i = 0; i < cfg->n_devices; ++i) conf.workers_cuda_gpuid[i] = static_cast<unsigned>(cfg->device_ids[i]);
#if !defined(SPU_USE_REAL_STARPU)
    conf.nworker_per_cuda = cfg->nworker_per_cuda;
    conf.cuda_pipeline = cfg->cuda_pipeline;
#endif
    const int rc = starpu_init(&conf);
    if (rc != 0) throw std::runtime_error("starpu_init failed with code " + std::to_string(rc));
    s->starpu_up = true;
    try {
      s->model.name = "model";
      for (int i = 0; i < cfg->n_inputs; ++i) s->model.inputs.push_back(to_cfg(cfg->inputs[i]));
      for (int i = 0; i < cfg->n_outputs; ++i) s->model.outputs.push_back(to_cfg(cfg->outputs[i]));
      s->model.max_batch_size = cfg->max_batch_size > 0 ? cfg->max_batch_size : 1;
      s->model.pool_size = cfg->pool_size > 0 ? cfg->pool_size : 12;
      s->model.variable_length = cfg->variable_length != 0;
      s->model.device_visible_requests = cfg->device_visible_requests != 0;
      if (s->model.device_visible_requests && cfg->model_kind != SPU_MODEL_BERT)
        throw std::invalid_argument("device_visible_requests is a BERT-path feature");
      s->length_bucket = cfg->length_bucket > 0 ? cfg->length_bucket : 0;
      s->tile_rows = cfg->tile_rows > 0 ? cfg->tile_rows : 0;
      s->coalesce_ms = cfg->batch_coalesce_timeout_ms;
      // build_gpu_replica_assignments (inference_runner.cpp:128-183): one replica per device, or per CUDA worker
      // (gpu_model_replication: per_worker); cfg->replicas[i] serves assignment i.  The caller created them
      // (spu_bert_replica_create / libspu_loader.so) and keeps ownership.
      std::vector<int> dev_ids(cfg->device_ids, cfg->device_ids + cfg->n_devices);
      const auto policy = cfg->gpu_model_replication != 0 ? GpuModelReplicationPolicy::PerWorker
                                                          : GpuModelReplicationPolicy::PerDevice;
      const auto assignments = build_gpu_replica_assignments(dev_ids, policy);
      if (assignments.size() > SPU_HOST_MAX_DEVICES) throw std::invalid_argument("too many replica assignments");
      for (size_t i = 0; i < assignments.size(); ++i) {
        DeviceModel m;
        m.kind = cfg->model_kind == SPU_MODEL_BERT  ? DeviceModel::Kind::Bert
                 : cfg->model_kind == SPU_MODEL_VIT ? DeviceModel::Kind::Vit
                                                    : DeviceModel::Kind::Affine;
        m.replica = cfg->replicas[i];
        m.n_workspaces = cfg->n_workspaces;
        m.scale = cfg->affine_scale;
        m.shift = cfg->affine_shift;
        if (m.kind != DeviceModel::Kind::Affine && m.replica == nullptr)
          throw std::invalid_argument("missing replica for device " + std::to_string(assignments[i].device_id) +
                                      (assignments[i].worker_id >= 0
                                           ? " worker " + std::to_string(assignments[i].worker_id)
                                           : std::string()));
        s->registry.add(assignments[i], m, /*owned=*/false);
      }
      if (cfg->warmup_autotune != 0 && cfg->model_kind == SPU_MODEL_BERT && s->model.inputs[0].dims.size() >= 2)
        for (size_t i = 0; i < s->registry.size(); ++i) {
          spu_replica *rep = s->registry.model(i)->replica;
          if (spu_bert_autotune(rep, 0, s->model.max_batch_size, static_cast<int32_t>(s->model.inputs[0].dims[1]),
                                nullptr, nullptr) != SPU_OK)
            throw std::runtime_error(std::string("warm-up autotune failed: ") + spu_last_error());
        }
      if (cfg->model_kind == SPU_MODEL_BERT && s->registry.size() > 0)
        s->model.id_vocab = spu_bert_vocab_size(s->registry.model(0)->replica);
      s->runner = std::make_unique<TaskRunner>(s->model, &s->registry);
      spu_server *raw = s.get();
      {
        const char *env = std::getenv("SPU_RESPONSE_THREADS");
        const int n_helpers = env ? std::atoi(env) : 3;
        for (int i = 0; i < n_helpers; ++i) s->helpers.emplace_back([raw] { raw->helper_loop(); });
        if (n_helpers > 0)
          s->runner->set_completion_executor([raw](std::vector<std::function<void()>> &fns) { raw->run_completions(fns); });
      }
      s->collector = std::thread([raw] { raw->collector_loop(); });
    } catch (...) {
      s->runner.reset();
      starpu_shutdown();
      throw;
    }
    *out = s.release();
    return 0;
  } catch (const std::exception &e) {
    set_host_error(e.what());
    return 1;
  }
}

void spu_server_destroy(spu_server *s) {
  if (!s) return;
  {
    std::lock_guard<std::mutex> lk(s->qmu);
    s->stop = true;
  }
  s->qcv.notify_all();
  if (s->collector.joinable()) s->collector.join();
  if (s->runner) s->runner->wait_for_all();
  {
    std::lock_guard<std::mutex> lk(s->hmu);
    s->hstop = true;
  }
  s->hcv.notify_all();
  for (auto &t : s->helpers)
    if (t.joinable()) t.join();
  for (size_t i = 0; i < s->registry.size(); ++i)
    if (s->registry.model(i)->replica) forget_replica_workspaces(s->registry.model(i)->replica);
  s->runner.reset();  // unregisters the slot handles while the runtime is still up
  if (s->starpu_up) starpu_shutdown();
  delete s;
}

int64_t spu_server_submit_batch(spu_server *s, const spu_request *reqs, int32_t n, int32_t fixed_worker) {
  if (!s || !reqs || n <= 0) {
    set_host_error("submit_batch: bad argument");
    return -1;
  }
  const int64_t ticket = s->open_ticket(n);
  std::vector<std::shared_ptr<InferenceJob>> jobs;
  for (int i = 0; i < n; ++i) jobs.push_back(s->make_job(reqs[i], ticket, i));
  if (fixed_worker >= 0) jobs[0]->fixed_worker_id = fixed_worker;
  try {
    s->runner->submit_batch(jobs);
  } catch (const std::exception &e) {
    set_host_error(e.what());
    std::lock_guard<std::mutex> lk(s->tmu);
    s->tickets.erase(ticket);
    return -1;
  }
  return ticket;
}

int64_t spu_server_enqueue(spu_server *s, const spu_request *req) {
  if (!s || !req) {
    set_host_error("enqueue: bad argument");
    return -1;
  }
  if (req->batch < 1 || req->batch > s->model.max_batch_size) {
    set_host_error("enqueue: request batch exceeds max_batch_size");
    return -1;
  }
  const int64_t ticket = s->open_ticket(1);
  auto job = s->make_job(*req, ticket, 0);
  {
    std::lock_guard<std::mutex> lk(s->qmu);
    s->queue.push_back(std::move(job));
  }
  s->qcv.notify_all();
  return ticket;
}

int32_t spu_server_wait(spu_server *s, int64_t ticket, spu_result *results, int32_t cap) {
  if (!s) return -1;
  std::unique_lock<std::mutex> lk(s->tmu);
  auto it = s->tickets.find(ticket);
  if (it == s->tickets.end()) {
    set_host_error("wait: unknown ticket");
    return -1;
  }
  s->tcv.wait(lk, [&] {
    auto f = s->tickets.find(ticket);
    return f != s->tickets.end() && f->second.done >= f->second.expected;
  });
  it = s->tickets.find(ticket);
  const int n = it->second.expected;
  for (int i = 0; i < n && i < cap && results; ++i) results[i] = it->second.results[static_cast<size_t>(i)];
  s->tickets.erase(it);
  return n;
}

int32_t spu_server_wait_all(spu_server *s) {
  if (!s) return -1;
  for (;;) {
    {
      std::lock_guard<std::mutex> lk(s->qmu);
      if (s->queue.empty() && s->forming == 0) break;
    }
    std::this_thread::sleep_for(std::chrono::microseconds(200));
  }
  s->runner->wait_for_all();
  return 0;
}

int32_t spu_host_tile_aware_wants_more(int64_t rows, int64_t seq, int64_t max_rows, int32_t tile_rows) {
  return tile_aware_wants_more(rows, seq, max_rows, tile_rows) ? 1 : 0;
}

int32_t spu_server_offered_load(spu_server *s, const spu_request *rows, int32_t n_rows, int64_t n_req,
                                double rate_per_s, uint64_t seed, double *codelet_ms, double *task_ms,
                                double *latency_ms, int32_t *batch_rows, double *elapsed_s) {
  if (!s || !rows || n_rows <= 0 || n_req <= 0 || !(rate_per_s > 0.0)) {
    set_host_error("offered_load: bad argument");
    return -1;
  }
  // open loop: arrival i happens at the i-th point of a Poisson process whatever the server does
  std::mt19937_64 rng(seed);
  std::exponential_distribution<double> gap(rate_per_s);
  std::vector<int64_t> tickets(static_cast<size_t>(n_req), -1);
  const auto t0 = std::chrono::steady_clock::now();
  double t_arrive = 0.0;
  for (int64_t i = 0; i < n_req; ++i) {
    t_arrive += gap(rng);
    const auto due = t0 + std::chrono::duration_cast<std::chrono::steady_clock::duration>(
                              std::chrono::duration<double>(t_arrive));
    for (;;) {
      const auto now = std::chrono::steady_clock::now();
      if (now >= due) break;
      if (due - now > std::chrono::microseconds(200)) std::this_thread::sleep_for(due - now - std::chrono::microseconds(100));
      else std::this_thread::yield();
    }
    spu_request rq = rows[i % n_rows];
    rq.request_id = static_cast<int32_t>(i);
    tickets[static_cast<size_t>(i)] = spu_server_enqueue(s, &rq);
    if (tickets[static_cast<size_t>(i)] < 0) return -1;
  }
  int32_t failed = 0;
  for (int64_t i = 0; i < n_req; ++i) {
    spu_result r{};
    if (spu_server_wait(s, tickets[static_cast<size_t>(i)], &r, 1) != 1) return -1;
    failed += r.ok == 1 ? 0 : 1;
    if (codelet_ms) codelet_ms[i] = r.codelet_ms;
    if (task_ms) task_ms[i] = r.task_ms;
    if (latency_ms) latency_ms[i] = r.latency_ms;
    if (batch_rows) batch_rows[i] = static_cast<int32_t>(r.batch_rows);
  }
  if (elapsed_s) *elapsed_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return failed;
}

void spu_server_stats(spu_server *s, int64_t out[6]) {
  if (!s || !out) return;
  out[0] = s->runner->tasks_submitted();
  out[1] = s->completed.load();
  out[2] = s->runner->input_pool().free_slots();
  out[3] = s->runner->output_pool().free_slots();
  out[4] = starpu_worker_get_count();
#if !defined(SPU_USE_REAL_STARPU)
  out[5] = spu_mini_starpu_codelet_failures();
#else
  out[5] = 0;
#endif
}

}  // extern "C"
