// Unit-test seams of libspu_host.so (second half of include/spu_host.h): each one drives a
// single host-side function the way the reference's own unit tests do, so that
// tests/test_host_*.py read like tests/unit/core/unit_starpu_setup.cpp and
// tests/unit/starpu/unit_starpu_task_runner_*.cpp.
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

#include "../../include/spu_host.h"
#include "nvtx_range.hpp"
#include "task_runner.hpp"

using namespace starpu_server;

// shares the thread-local error slot of host_capi.cpp
extern "C" void spu_host_set_error_for_hooks(const char *m);

namespace {
void fail(const std::exception &e) { spu_host_set_error_for_hooks(e.what()); }

int g_mock_worker = 0, g_mock_device = 0;
void *g_mock_stream = nullptr;
int mock_worker_id() { return g_mock_worker; }
int mock_worker_devid(int) { return g_mock_device; }
void *mock_stream() { return g_mock_stream; }
}  // namespace

extern "C" {

void spu_host_codelet_fields(int64_t out[7]) {
  InferenceCodelet cl;
  starpu_codelet *c = cl.get_codelet();
  out[0] = c->nbuffers;
  out[1] = static_cast<int64_t>(c->type);
  out[2] = c->max_parallelism;
  out[3] = c->cpu_funcs[0] != nullptr;
  out[4] = c->cuda_funcs[0] != nullptr;
  out[5] = static_cast<int64_t>(c->cuda_flags[0]);
  out[6] = static_cast<int64_t>(c->where);
}

int32_t spu_host_buffer_byte_size(int32_t interface_id, uint32_t nx, uint64_t elemsize, int32_t null_buffer,
                                  uint64_t *out) {
  try {
    if (null_buffer) {
      *out = buffer_byte_size(nullptr);
      return 0;
    }
    if (interface_id == STARPU_VECTOR_INTERFACE_ID) {
      starpu_vector_interface v{};
      v.id = STARPU_VECTOR_INTERFACE_ID;
      v.nx = nx;
      v.elemsize = static_cast<size_t>(elemsize);
      *out = buffer_byte_size(reinterpret_cast<const StarpuBufferInterface *>(&v));
      return 0;
    }
    starpu_variable_interface v{};
    v.id = static_cast<starpu_data_interface_id>(interface_id);
    v.elemsize = static_cast<size_t>(elemsize);
    *out = buffer_byte_size(&v);
    return 0;
  } catch (const std::exception &e) {
    fail(e);
    return 1;
  }
}

int32_t spu_host_select_replica(const int32_t *device_ids, int32_t n_device_ids, const int32_t *worker_ids,
                                int32_t n_worker_ids, int32_t n_models, int32_t worker_id, int32_t device_id) {
  InferenceParams p;
  std::vector<DeviceModel> models(static_cast<size_t>(n_models > 0 ? n_models : 0));
  for (auto &m : models) p.models.models_gpu.push_back(&m);
  for (int i = 0; i < n_device_ids; ++i) p.models.device_ids.push_back(device_ids[i]);
  for (int i = 0; i < n_worker_ids; ++i) p.models.worker_ids.push_back(worker_ids[i]);
  try {
    DeviceModel *m = select_gpu_module(p, worker_id, device_id);
    return static_cast<int32_t>(m - models.data());
  } catch (const StarPUCodeletException &e) {
    fail(e);
    return -1;
  }
}

namespace {
std::mutex g_nvtx_mu;
std::string g_nvtx_last;
int64_t g_nvtx_pushes = 0, g_nvtx_pops = 0;
void nvtx_capture_push(const char *label) {
  std::lock_guard<std::mutex> lk(g_nvtx_mu);
  g_nvtx_last = label ? label : "";
  ++g_nvtx_pushes;
}
void nvtx_capture_pop() {
  std::lock_guard<std::mutex> lk(g_nvtx_mu);
  ++g_nvtx_pops;
}
}  // namespace

int32_t spu_host_nvtx_capture(int32_t enable, char *last_label, uint64_t cap, int64_t counts[2]) {
  if (enable) NvtxRange::SetHooks(&nvtx_capture_push, &nvtx_capture_pop);
  else NvtxRange::ResetHooks();
  std::lock_guard<std::mutex> lk(g_nvtx_mu);
  if (last_label && cap) std::snprintf(last_label, cap, "%s", g_nvtx_last.c_str());
  if (counts) {
    counts[0] = g_nvtx_pushes;
    counts[1] = g_nvtx_pops;
  }
  return 0;
}

int32_t spu_host_replica_assignments(const int32_t *device_ids, int32_t n_device_ids, int32_t per_worker,
                                     const int32_t *worker_dev, int32_t n_workers, int32_t *out_device,
                                     int32_t *out_worker, int32_t cap, int32_t pick_worker, int32_t pick_device,
                                     int32_t *picked) {
  try {
    std::vector<int> devs(device_ids, device_ids + n_device_ids);
    auto workers = [&](const std::vector<int> &ids) {
      std::map<int, std::vector<int>> by_dev;
      for (int w = 0; w < n_workers; ++w)
        if (std::find(ids.begin(), ids.end(), worker_dev[w]) != ids.end()) by_dev[worker_dev[w]].push_back(w);
      return by_dev;
    };
    const auto assignments = build_gpu_replica_assignments(
        devs, per_worker ? GpuModelReplicationPolicy::PerWorker : GpuModelReplicationPolicy::PerDevice, workers);
    ReplicaRegistry reg;
    for (const auto &a : assignments) reg.add(a, DeviceModel{}, false);
    for (size_t i = 0; i < assignments.size() && static_cast<int32_t>(i) < cap; ++i) {
      if (out_device) out_device[i] = assignments[i].device_id;
      if (out_worker) out_worker[i] = assignments[i].worker_id;
    }
    if (picked != nullptr && (pick_worker >= 0 || pick_device >= 0)) {
      InferenceParams p;
      reg.fill_model_pointers(p.models);
      DeviceModel *m = select_gpu_module(p, pick_worker, pick_device);
      *picked = -1;
      for (size_t i = 0; i < reg.size(); ++i)
        if (reg.model(i) == m) *picked = static_cast<int32_t>(i);
    }
    return static_cast<int32_t>(assignments.size());
  } catch (const std::exception &e) {
    fail(e);
    return -1;
  }
}

int32_t spu_host_gather_to_slot(const void *const *srcs, const uint64_t *nbytes, int32_t n_jobs, uint64_t elemsize,
                                void *slot, uint64_t capacity, uint64_t *nx_out, uint64_t *bytes_out) {
  // a slot of one input registered exactly like InputSlotPool does (input_slot_pool.cpp:195-214)
  starpu_data_handle_t h = nullptr;
  starpu_vector_data_register(&h, STARPU_MAIN_RAM, reinterpret_cast<uintptr_t>(slot),
                              static_cast<uint32_t>(capacity / (elemsize ? elemsize : 1)), static_cast<size_t>(elemsize));
  std::vector<InferenceJob> jobs(static_cast<size_t>(n_jobs));
  for (int j = 0; j < n_jobs; ++j) {
    HostTensor t;
    t.data = static_cast<std::byte *>(const_cast<void *>(srcs[j]));
    t.type = elemsize == 8 ? ScalarType::Int64 : (elemsize == 4 ? ScalarType::Float32 : ScalarType::UInt8);
    t.dims = {static_cast<int64_t>(nbytes[j] / (elemsize ? elemsize : 1))};
    jobs[static_cast<size_t>(j)].inputs.push_back(std::move(t));
  }
  std::vector<const InferenceJob *> pending;
  for (int j = 1; j < n_jobs; ++j) pending.push_back(&jobs[static_cast<size_t>(j)]);
  int32_t rc = 0;
  try {
    HostBufferInfo info;
    info.bytes = static_cast<size_t>(capacity);
    const size_t copied = copy_job_inputs_to_slot(n_jobs > 0 ? &jobs[0] : nullptr, pending, {h},
                                                  {static_cast<std::byte *>(slot)}, {info});
    if (bytes_out) *bytes_out = copied;
    if (nx_out) *nx_out = static_cast<starpu_vector_interface *>(starpu_data_get_interface_on_node(h, 0))->nx;
  } catch (const std::exception &e) {
    fail(e);
    rc = 1;
  }
  starpu_data_unregister(h);
  return rc;
}

int64_t spu_host_slice_rows(uint64_t rows, uint64_t row_bytes, int64_t offset, int64_t batch_size, uint64_t *off_out,
                            uint64_t *len_out) {
  static std::byte base_byte;
  HostTensor t;
  t.data = &base_byte;  // only address arithmetic is performed
  t.type = ScalarType::UInt8;
  t.dims = {static_cast<int64_t>(rows), static_cast<int64_t>(row_bytes)};
  std::vector<HostTensor> agg;
  if (rows > 0) agg.push_back(t);
  const SubJobSliceResult r = slice_outputs_for_sub_job(agg, offset, batch_size);
  *off_out = 0;
  *len_out = 0;
  if (!r.outputs.empty() && r.outputs[0].defined()) {
    *off_out = static_cast<uint64_t>(r.outputs[0].data - &base_byte);
    *len_out = r.outputs[0].nbytes();
  }
  return r.processed_length;
}

int32_t spu_host_can_merge(const int64_t *dims_a, int32_t nd_a, int32_t dt_a, const int64_t *dims_b, int32_t nd_b,
                           int32_t dt_b) {
  static std::byte dummy;
  auto mk = [](const int64_t *d, int32_t nd, int32_t dt) {
    InferenceJob j;
    HostTensor t;
    t.data = &dummy;
    t.type = static_cast<ScalarType>(dt);
    t.dims.assign(d, d + nd);
    j.inputs.push_back(std::move(t));
    return j;
  };
  const InferenceJob a = mk(dims_a, nd_a, dt_a), b = mk(dims_b, nd_b, dt_b);
  return can_merge_jobs(&a, &b) ? 1 : 0;
}

int32_t spu_host_can_merge_varlen(const int64_t *dims_a, int32_t nd_a, int32_t dt_a, const int64_t *dims_b,
                                  int32_t nd_b, int32_t dt_b) {
  static std::byte dummy;
  auto mk = [](const int64_t *d, int32_t nd, int32_t dt) {
    InferenceJob j;
    HostTensor t;
    t.data = &dummy;
    t.type = static_cast<ScalarType>(dt);
    t.dims.assign(d, d + nd);
    j.inputs.push_back(std::move(t));
    return j;
  };
  const InferenceJob a = mk(dims_a, nd_a, dt_a), b = mk(dims_b, nd_b, dt_b);
  return can_merge_jobs(&a, &b, true) ? 1 : 0;
}

int32_t spu_host_gather_padded(const int64_t *const *srcs, const int64_t *len, int32_t n_jobs, int64_t seq,
                               int64_t *slot, uint64_t capacity_bytes, uint64_t *nx_out) {
  starpu_data_handle_t h = nullptr;
  starpu_vector_data_register(&h, STARPU_MAIN_RAM, reinterpret_cast<uintptr_t>(slot),
                              static_cast<uint32_t>(capacity_bytes / 8), 8);
  std::vector<InferenceJob> jobs(static_cast<size_t>(n_jobs));
  for (int j = 0; j < n_jobs; ++j) {
    HostTensor t;
    t.data = reinterpret_cast<std::byte *>(const_cast<int64_t *>(srcs[j]));
    t.type = ScalarType::Int64;
    t.dims = {1, len[j]};
    jobs[static_cast<size_t>(j)].inputs.push_back(std::move(t));
  }
  std::vector<const InferenceJob *> pending;
  for (int j = 1; j < n_jobs; ++j) pending.push_back(&jobs[static_cast<size_t>(j)]);
  int32_t rc = 0;
  try {
    HostBufferInfo info;
    info.bytes = static_cast<size_t>(capacity_bytes);
    copy_job_inputs_to_slot_padded(n_jobs > 0 ? &jobs[0] : nullptr, pending, {h},
                                   {reinterpret_cast<std::byte *>(slot)}, {info}, seq);
    if (nx_out) *nx_out = static_cast<starpu_vector_interface *>(starpu_data_get_interface_on_node(h, 0))->nx;
  } catch (const std::exception &e) {
    fail(e);
    rc = 1;
  }
  starpu_data_unregister(h);
  return rc;
}

int32_t spu_host_request_table(const uint64_t *ids_ptrs, const uint64_t *mask_ptrs, const uint64_t *dst_ptrs,
                               const int64_t *batch, const int64_t *len, int32_t n_jobs, int64_t seq,
                               uint64_t token_bytes, void *slot, uint64_t capacity_bytes, int64_t out[8]) {
  starpu_data_handle_t h0 = nullptr, h1 = nullptr;
  std::vector<int64_t> second(8, 0);
  starpu_vector_data_register(&h0, STARPU_MAIN_RAM, reinterpret_cast<uintptr_t>(slot),
                              static_cast<uint32_t>(capacity_bytes / 8), 8);
  starpu_vector_data_register(&h1, STARPU_MAIN_RAM, reinterpret_cast<uintptr_t>(second.data()), 8, 8);
  std::vector<std::shared_ptr<InferenceJob>> jobs;
  int64_t rows = 0;
  for (int j = 0; j < n_jobs; ++j) {
    auto job = std::make_shared<InferenceJob>();
    job->request_id = j;
    for (const uint64_t *tab : {ids_ptrs, mask_ptrs}) {
      HostTensor t;
      t.data = reinterpret_cast<std::byte *>(static_cast<uintptr_t>(tab[j]));
      t.type = ScalarType::Int64;
      t.dims = {batch[j], len[j]};
      job->inputs.push_back(std::move(t));
    }
    if (dst_ptrs[j] != 0) job->output_dsts.push_back(reinterpret_cast<void *>(static_cast<uintptr_t>(dst_ptrs[j])));
    rows += batch[j];
    jobs.push_back(std::move(job));
  }
  int32_t rc = 0;
  try {
    HostBufferInfo i0, i1;
    i0.bytes = static_cast<size_t>(capacity_bytes);
    i1.bytes = 64;
    detail::RequestIo io;
    write_request_table(jobs, {h0, h1}, {reinterpret_cast<std::byte *>(slot), reinterpret_cast<std::byte *>(second.data())},
                        {i0, i1}, rows, seq, static_cast<size_t>(token_bytes), &io);
    out[0] = io.rows;
    out[1] = static_cast<int64_t>(io.off_mask);
    out[2] = static_cast<int64_t>(io.off_dst);
    out[3] = static_cast<int64_t>(io.off_len);
    out[4] = static_cast<int64_t>(io.table_bytes);
    out[5] = io.full_rows ? 1 : 0;
    out[6] = static_cast<starpu_vector_interface *>(starpu_data_get_interface_on_node(h0, 0))->nx;
    out[7] = static_cast<starpu_vector_interface *>(starpu_data_get_interface_on_node(h1, 0))->nx;
  } catch (const std::exception &e) {
    fail(e);
    rc = 1;
  }
  starpu_data_unregister(h0);
  starpu_data_unregister(h1);
  return rc;
}

int32_t spu_host_masks_all_ones(const int64_t *const *masks, const int64_t *numel, int32_t n_jobs) {
  std::vector<std::shared_ptr<InferenceJob>> jobs;
  for (int j = 0; j < n_jobs; ++j) {
    auto job = std::make_shared<InferenceJob>();
    for (int k = 0; k < 2; ++k) {
      HostTensor t;
      t.data = reinterpret_cast<std::byte *>(const_cast<int64_t *>(masks[j]));
      t.type = ScalarType::Int64;
      t.dims = {1, numel[j]};
      job->inputs.push_back(std::move(t));
    }
    jobs.push_back(std::move(job));
  }
  return masks_all_ones(jobs) ? 1 : 0;
}

int32_t spu_host_call_cuda_func_affine(const float *in_dev, float *out_dev, uint64_t n, float scale, float shift,
                                       int32_t device, int32_t worker_id, void *stream, int32_t n_models,
                                       int32_t use_vector_iface, int64_t rec[6]) {
  g_mock_worker = worker_id;
  g_mock_device = device;
  g_mock_stream = stream;
  WorkerContextHooks hooks;
  hooks.worker_get_id = &mock_worker_id;
  hooks.worker_get_devid = &mock_worker_devid;
  hooks.cuda_get_local_stream = &mock_stream;
  set_worker_context_hooks(hooks);

  DeviceModel model;
  model.kind = DeviceModel::Kind::Affine;
  model.scale = scale;
  model.shift = shift;
  model.device_id = device;
  InferenceParams params;  // make_basic_params(n) of the reference tests
  params.num_inputs = 1;
  params.num_outputs = 1;
  params.layout.dims = {{static_cast<int64_t>(n)}};
  params.layout.num_dims = {1};
  params.layout.input_types = {ScalarType::Float32};
  params.limits.max_inputs = InferLimits::MaxInputs;
  params.limits.max_dims = InferLimits::MaxDims;
  params.models.models_gpu.resize(static_cast<size_t>(n_models > 0 ? n_models : 0), nullptr);
  if (device >= 0 && device < n_models) params.models.models_gpu[static_cast<size_t>(device)] = &model;

  int rec_dev = -1, rec_worker = -1;
  DeviceType rec_on = DeviceType::Unknown;
  MonotonicClock::time_point t0{}, t1{};
  params.device.device_id = &rec_dev;
  params.device.worker_id = &rec_worker;
  params.device.executed_on = &rec_on;
  params.timing.codelet_start_time = &t0;
  params.timing.codelet_end_time = &t1;

  starpu_variable_interface vin{}, vout{};
  starpu_vector_interface win{}, wout{};
  void *buffers[2];
  if (use_vector_iface) {
    win.id = wout.id = STARPU_VECTOR_INTERFACE_ID;
    win.ptr = reinterpret_cast<uintptr_t>(in_dev);
    wout.ptr = reinterpret_cast<uintptr_t>(out_dev);
    win.nx = wout.nx = static_cast<uint32_t>(n);
    win.elemsize = wout.elemsize = sizeof(float);
    buffers[0] = &win;
    buffers[1] = &wout;
  } else {
    // make_variable_interface (tests/common/test_utils.hpp:88-135): elemsize = n * sizeof(T)
    vin.id = vout.id = STARPU_VARIABLE_INTERFACE_ID;
    vin.ptr = reinterpret_cast<uintptr_t>(in_dev);
    vout.ptr = reinterpret_cast<uintptr_t>(out_dev);
    vin.elemsize = vout.elemsize = static_cast<size_t>(n) * sizeof(float);
    buffers[0] = &vin;
    buffers[1] = &vout;
  }
  InferenceCodelet codelet;
  int32_t rc = 0;
  try {
    codelet.get_codelet()->cuda_funcs[0](buffers, &params);
  } catch (const StarPUCodeletException &e) {
    fail(e);
    rc = 1;
  }
  rec[0] = rec_dev;
  rec[1] = rec_worker;
  rec[2] = static_cast<int64_t>(rec_on);
  rec[3] = t0.time_since_epoch().count() != 0;
  rec[4] = t1.time_since_epoch().count() != 0;
  rec[5] = worker_inflight_gauge(worker_id);
  set_worker_context_hooks(WorkerContextHooks{});
  return rc;
}

}  // extern "C"
