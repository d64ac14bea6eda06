// TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
//
// The reference's CPU codelet body, run through the reference's OWN TensorBuilder: this file is linked with
// /root/reference/src/core/tensor_builder.cpp, compiled from where it lies (oracle/Makefile, target _ref), against
// the wheel's LibTorch and oracle/ref_stub/starpu.h (StarPU itself is absent from the image).  What this driver
// adds is the glue StarPU would otherwise provide, restated from
//   run_inference            /root/reference/src/core/starpu_setup.cpp:594-624
//   cpu_inference_func       /root/reference/src/core/starpu_setup.cpp:784-801   (c10::InferenceMode + copy lambda)
//   buffer_byte_size         /root/reference/src/core/starpu_setup.cpp:515-542   (variable interface: elemsize)
//   extract_tensors_from_output / append_ivalue  starpu_setup.cpp:545-587
// around hand-declared starpu_variable_interface structs, as the reference's tests do
// (/root/reference/tests/common/test_utils.hpp:88-135).  Used as bench.py's `cpu_baseline` / `--impl reference`
// (kind "reference") and by tests/test_ref_cpu_codelet.py to pin oracle/bert_oracle.py against LibTorch.
#include <ATen/Parallel.h>
#include <torch/script.h>

#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "tensor_builder.hpp"  // the reference's header (-I /root/reference/src/core)

namespace {
thread_local std::string g_err;

void append_ivalue(const c10::IValue &v, std::vector<at::Tensor> &out) {
  if (v.isTensor()) out.push_back(v.toTensor());
  else if (v.isTuple())
    for (const auto &e : v.toTuple()->elements()) append_ivalue(e, out);
  else if (v.isList())
    for (const c10::IValue &e : v.toList()) append_ivalue(e, out);
  else if (v.isTensorList())
    for (const at::Tensor &t : v.toTensorList()) out.push_back(t);
}

struct Codelet {
  torch::jit::script::Module module;
};
}  // namespace

extern "C" {

const char *ref_cpu_codelet_last_error(void) { return g_err.c_str(); }

// torch::jit::load of the TorchScript artifact (inference_runner.cpp:244-249: load_model -> eval)
int ref_cpu_codelet_open(const char *path, int intra_op_threads, void **handle) {
  try {
    if (intra_op_threads > 0) at::set_num_threads(intra_op_threads);
    auto c = std::make_unique<Codelet>();
    c->module = torch::jit::load(path, torch::kCPU);
    c->module.eval();
    *handle = c.release();
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

void ref_cpu_codelet_close(void *handle) { delete static_cast<Codelet *>(handle); }

// One task = cpu_inference_func over n_in input buffers + ONE output buffer.
//   in_ptrs[i]  host buffer of input i, dtype in_dtype[i] (0 = fp32, 3 = int64: include/spu_host.h SPU_DT_*),
//   dims        the inputs' dims back to back, ndims[i] each (dims[.][0] = effective batch)
//   out         host output buffer of out_bytes
// Runs `warmup` untimed + `iters` timed tasks; ms[k] = wall milliseconds of task k.  0 ok / 1 exception.
int ref_cpu_codelet_run(void *handle, int n_in, const void *const *in_ptrs, const int *in_dtype, const long long *dims,
                        const int *ndims, void *out, unsigned long long out_bytes, int warmup, int iters, double *ms) {
  try {
    auto *c = static_cast<Codelet *>(handle);
    using namespace starpu_server;
    InferenceParams params;
    params.models.model_cpu = &c->module;
    params.num_inputs = static_cast<size_t>(n_in);
    params.num_outputs = 1;
    params.limits.max_inputs = 16;
    params.limits.max_dims = 8;
    std::vector<starpu_variable_interface> ifaces(static_cast<size_t>(n_in) + 1);
    std::vector<StarpuBufferPtr> bufs;
    const long long *d = dims;
    for (int i = 0; i < n_in; ++i) {
      std::vector<int64_t> dv(d, d + ndims[i]);
      d += ndims[i];
      size_t numel = 1;
      for (int64_t x : dv) numel *= static_cast<size_t>(x);
      const at::ScalarType t = in_dtype[i] == 3 ? at::kLong : at::kFloat;
      params.layout.dims.push_back(dv);
      params.layout.num_dims.push_back(static_cast<int64_t>(dv.size()));
      params.layout.input_types.push_back(t);
      ifaces[static_cast<size_t>(i)] = starpu_variable_interface{STARPU_VARIABLE_INTERFACE_ID,
                                                                 reinterpret_cast<uintptr_t>(in_ptrs[i]), 0, 0,
                                                                 numel * c10::elementSize(t)};
      if (i == 0 && !dv.empty()) params.batch_size = dv[0];
    }
    ifaces[static_cast<size_t>(n_in)] =
        starpu_variable_interface{STARPU_VARIABLE_INTERFACE_ID, reinterpret_cast<uintptr_t>(out), 0, 0,
                                  static_cast<size_t>(out_bytes)};
    for (auto &f : ifaces) bufs.push_back(&f);
    const StarpuBufferSpan span(bufs.data(), bufs.size());

    const c10::InferenceMode no_autograd;  // cpu_inference_func
    for (int k = 0; k < warmup + iters; ++k) {
      const auto t0 = std::chrono::steady_clock::now();
      // run_inference: wrap -> forward -> copy every output into its StarPU buffer
      const auto &ivalues = TensorBuilder::prepare_input_ivalues(&params, span, torch::kCPU);
      auto result = params.models.model_cpu->forward(ivalues);
      std::vector<at::Tensor> outputs;
      append_ivalue(result, outputs);
      TORCH_CHECK(outputs.size() == params.num_outputs, "Mismatch between model outputs and StarPU buffers");
      for (size_t i = 0; i < params.num_outputs; ++i) {
        auto *iface = span[params.num_inputs + i];
        std::span<std::byte> buffer(reinterpret_cast<std::byte *>(iface->ptr), iface->elemsize);
        TensorBuilder::copy_output_to_buffer(outputs[i], buffer, outputs[i].numel(), outputs[i].scalar_type());
      }
      const auto t1 = std::chrono::steady_clock::now();
      if (k >= warmup && ms) ms[k - warmup] = std::chrono::duration<double, std::milli>(t1 - t0).count();
    }
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return 1;
  }
}

}  // extern "C"
