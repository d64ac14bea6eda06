// libspu_loader.so: TorchScript state -> replicas of libspu_b200.so (start-up only).
// Restates load_model / clone_model_to_gpus (/root/reference/src/core/inference_runner.cpp:243-275):
// the module is loaded and put in eval mode exactly as there; what differs is the per-device
// step, which uploads bf16 / fp32 device buffers through the C-ABI instead of cloning a
// torch::jit::Module onto the GPU.  The only translation unit of the repo that includes torch.
#include <torch/csrc/jit/ir/constants.h>
#include <torch/csrc/jit/passes/inliner.h>
#include <torch/script.h>

#include <algorithm>
#include <cmath>
#include <functional>
#include <cstdio>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../include/spu_loader.h"

namespace {

thread_local std::string g_error;

struct Named {
  std::string name;
  torch::Tensor t;  // fp32, contiguous, CPU
};

std::vector<std::string> bert_names(int layers) {
  std::vector<std::string> n = {"embeddings.word_embeddings.weight", "embeddings.position_embeddings.weight",
                                "embeddings.token_type_embeddings.weight", "embeddings.LayerNorm.weight",
                                "embeddings.LayerNorm.bias"};
  static const char *per_layer[] = {"attention.self.query.weight",  "attention.self.query.bias",
                                    "attention.self.key.weight",    "attention.self.key.bias",
                                    "attention.self.value.weight",  "attention.self.value.bias",
                                    "attention.output.dense.weight", "attention.output.dense.bias",
                                    "attention.output.LayerNorm.weight", "attention.output.LayerNorm.bias",
                                    "intermediate.dense.weight",    "intermediate.dense.bias",
                                    "output.dense.weight",          "output.dense.bias",
                                    "output.LayerNorm.weight",      "output.LayerNorm.bias"};
  for (int l = 0; l < layers; ++l)
    for (const char *s : per_layer) n.push_back("encoder.layer." + std::to_string(l) + "." + s);
  return n;
}

std::vector<std::string> vit_names(int layers) {
  std::vector<std::string> n = {"conv_proj.weight", "conv_proj.bias", "class_token", "encoder.pos_embedding"};
  static const char *per_layer[] = {"ln_1.weight", "ln_1.bias", "self_attention.in_proj_weight",
                                    "self_attention.in_proj_bias", "self_attention.out_proj.weight",
                                    "self_attention.out_proj.bias", "ln_2.weight", "ln_2.bias", "mlp.0.weight",
                                    "mlp.0.bias", "mlp.3.weight", "mlp.3.bias"};
  for (int l = 0; l < layers; ++l)
    for (const char *s : per_layer) n.push_back("encoder.layers.encoder_layer_" + std::to_string(l) + "." + s);
  for (const char *s : {"encoder.ln.weight", "encoder.ln.bias", "heads.head.weight", "heads.head.bias"}) n.push_back(s);
  return n;
}

// ---- what the tensor shapes do not say: LayerNorm epsilon and the number of attention heads ------------------
// Both are read from the artifact itself, never guessed (hidden / 64 heads is right for BERT-base / ViT-B/L and
// silently wrong for e.g. ViT-H/14, 1280 = 16 x 80): the inlined forward graph carries eps as the constant fifth
// input of every aten::layer_norm, a traced BertModel carries (heads, head_dim) as the two trailing constants of
// the [B, S, heads, head_dim] view of q/k/v, and a scripted torchvision ViT keeps `num_heads` / `eps` as module
// attributes.  A caller that knows better (models/*.yml) passes them to spu_loader_open_ex.
void walk(torch::jit::Block *b, const std::function<void(torch::jit::Node *)> &f) {
  for (torch::jit::Node *n : b->nodes()) {
    f(n);
    for (torch::jit::Block *sub : n->blocks()) walk(sub, f);
  }
}

struct GraphFacts {
  std::vector<double> ln_eps;                   // distinct eps constants of aten::layer_norm nodes
  std::vector<std::pair<int64_t, int64_t>> hd;  // distinct trailing (heads, head_dim) constants of 4-d views
};

GraphFacts read_graph(torch::jit::script::Module &module) {
  GraphFacts facts;
  auto method = module.find_method("forward");
  if (!method) return facts;
  std::shared_ptr<torch::jit::Graph> g = method->graph()->copy();
  torch::jit::Inline(*g);
  auto const_of = [](torch::jit::Value *v) { return torch::jit::toIValue(v); };
  walk(g->block(), [&](torch::jit::Node *n) {
    if (n->kind() == c10::aten::layer_norm && n->inputs().size() >= 5) {
      if (auto iv = const_of(n->input(4)); iv && iv->isDouble()) {
        const double e = iv->toDouble();
        if (std::find(facts.ln_eps.begin(), facts.ln_eps.end(), e) == facts.ln_eps.end()) facts.ln_eps.push_back(e);
      }
    } else if ((n->kind() == c10::aten::view || n->kind() == c10::aten::reshape) && n->inputs().size() == 2) {
      torch::jit::Node *shape = n->input(1)->node();
      if (shape->kind() != c10::prim::ListConstruct || shape->inputs().size() != 4) return;
      auto h = const_of(shape->input(2)), d = const_of(shape->input(3));
      if (!h || !d || !h->isInt() || !d->isInt()) return;
      const std::pair<int64_t, int64_t> hd{h->toInt(), d->toInt()};
      if (std::find(facts.hd.begin(), facts.hd.end(), hd) == facts.hd.end()) facts.hd.push_back(hd);
    }
  });
  return facts;
}

// attribute `name` of the submodule at dotted `path` (tried under each wrapper prefix), if both exist
c10::optional<c10::IValue> find_attr(const torch::jit::script::Module &root, const std::string &path, const std::string &name) {
  for (const char *prefix : {"", "bert.", "model.", "module."}) {
    torch::jit::script::Module m = root;
    bool ok = true;
    std::string rest = std::string(prefix) + path;
    while (ok && !rest.empty()) {
      const size_t dot = rest.find('.');
      const std::string head = rest.substr(0, dot);
      rest = dot == std::string::npos ? "" : rest.substr(dot + 1);
      if (head.empty()) continue;
      if (m.hasattr(head) && m.attr(head).isModule()) m = m.attr(head).toModule();
      else ok = false;
    }
    if (ok && m.hasattr(name)) return m.attr(name);
  }
  return c10::nullopt;
}

float resolve_eps(float given, const GraphFacts &facts, const c10::optional<c10::IValue> &attr) {
  if (given > 0.f) return given;
  if (facts.ln_eps.size() == 1) return static_cast<float>(facts.ln_eps[0]);
  if (facts.ln_eps.size() > 1)
    throw std::runtime_error("the module's LayerNorms use several epsilons; this path supports one (pass it to spu_loader_open_ex to override)");
  if (attr && attr->isDouble()) return static_cast<float>(attr->toDouble());
  throw std::runtime_error("cannot read the LayerNorm epsilon from the TorchScript module (no constant aten::layer_norm eps, no `eps` attribute): pass it to spu_loader_open_ex");
}

int resolve_heads(int given, int hidden, const GraphFacts &facts, const c10::optional<c10::IValue> &attr) {
  int heads = given;
  if (heads <= 0 && attr && attr->isInt()) heads = static_cast<int>(attr->toInt());
  if (heads <= 0) {
    std::vector<int64_t> cands;
    for (auto [h, d] : facts.hd) {
      if (h == -1 && d > 0 && hidden % d == 0) h = hidden / d;  // view(B, S, -1, head_dim), as recent HF BertModel writes it
      if (h > 0 && d > 0 && h * d == hidden && std::find(cands.begin(), cands.end(), h) == cands.end()) cands.push_back(h);
    }
    if (cands.size() == 1) heads = static_cast<int>(cands[0]);
    else if (cands.size() > 1)
      throw std::runtime_error("the forward graph splits the hidden size into heads in more than one way: pass the head count to spu_loader_open_ex");
  }
  if (heads <= 0)
    throw std::runtime_error("cannot read the number of attention heads from the TorchScript module (no `num_heads` attribute, no [B, S, heads, head_dim] view with constant trailing dims): pass it to spu_loader_open_ex");
  if (hidden % heads != 0) throw std::runtime_error("hidden size " + std::to_string(hidden) + " is not divisible by " + std::to_string(heads) + " heads");
  if (hidden / heads != 64)
    throw std::runtime_error("model has " + std::to_string(heads) + " heads of " + std::to_string(hidden / heads) +
                             " (hidden " + std::to_string(hidden) + "); the attention kernels of libspu_b200 are built for head_dim 64 only");
  return heads;
}

}  // namespace

struct spu_module {
  int kind = SPU_LOADER_BERT;
  std::vector<Named> tensors;  // in replica-create order
  spu_bert_desc bert{};
  spu_vit_desc vit{};
};

extern "C" {

const char *spu_loader_last_error(void) { return g_error.c_str(); }

int32_t spu_loader_open(const char *path, int32_t kind, spu_module **out) {
  return spu_loader_open_ex(path, kind, 0, 0.f, out);
}

int32_t spu_loader_open_ex(const char *path, int32_t kind, int32_t heads, float ln_eps, spu_module **out) {
  if (!path || !out || (kind != SPU_LOADER_BERT && kind != SPU_LOADER_VIT)) {
    g_error = "loader_open: bad argument";
    return 1;
  }
  *out = nullptr;
  try {
    // load_model (inference_runner.cpp:243-249)
    torch::jit::script::Module module = torch::jit::load(path, torch::kCPU);
    module.eval();
    const GraphFacts facts = read_graph(module);
    std::map<std::string, torch::Tensor> state;
    auto strip = [&](std::string k) {
      // the exporter wraps BertModel in a module whose attribute is `bert` (import_bert-base-uncased.py:8-25)
      for (const char *prefix : {"bert.", "model.", "module."})
        if (k.rfind(prefix, 0) == 0) return k.substr(std::string(prefix).size());
      return k;
    };
    for (const auto &p : module.named_parameters(/*recurse=*/true)) state[strip(p.name)] = p.value;
    for (const auto &b : module.named_buffers(/*recurse=*/true)) state.emplace(strip(b.name), b.value);
    auto m = std::make_unique<spu_module>();
    m->kind = kind;
    auto has = [&](const std::string &k) { return state.find(k) != state.end(); };
    auto need = [&](const std::string &k) -> const torch::Tensor & {
      auto it = state.find(k);
      if (it == state.end()) throw std::runtime_error("TorchScript module has no parameter '" + k + "'");
      return it->second;
    };
    std::vector<std::string> names;
    if (kind == SPU_LOADER_BERT) {
      int layers = 0;
      while (has("encoder.layer." + std::to_string(layers) + ".output.dense.weight")) ++layers;
      if (layers == 0) throw std::runtime_error("no encoder.layer.N.* parameters: not a BertModel TorchScript");
      const auto &word = need("embeddings.word_embeddings.weight");
      spu_bert_desc &d = m->bert;
      d.layers = layers;
      d.hidden = static_cast<int32_t>(word.size(1));
      d.heads = resolve_heads(heads, d.hidden, facts, find_attr(module, "encoder.layer.0.attention.self", "num_attention_heads"));
      d.ffn = static_cast<int32_t>(need("encoder.layer.0.intermediate.dense.weight").size(0));
      d.vocab = static_cast<int32_t>(word.size(0));
      d.max_pos = static_cast<int32_t>(need("embeddings.position_embeddings.weight").size(0));
      d.type_vocab = static_cast<int32_t>(need("embeddings.token_type_embeddings.weight").size(0));
      d.ln_eps = resolve_eps(ln_eps, facts, find_attr(module, "embeddings.LayerNorm", "eps"));
      names = bert_names(layers);
    } else {
      int layers = 0;
      while (has("encoder.layers.encoder_layer_" + std::to_string(layers) + ".ln_1.weight")) ++layers;
      if (layers == 0) throw std::runtime_error("no encoder.layers.encoder_layer_N.* parameters: not a torchvision ViT");
      const auto &conv = need("conv_proj.weight");
      const auto &pos = need("encoder.pos_embedding");
      spu_vit_desc &d = m->vit;
      d.layers = layers;
      d.hidden = static_cast<int32_t>(conv.size(0));
      d.heads = resolve_heads(heads, d.hidden, facts, find_attr(module, "encoder.layers.encoder_layer_0.self_attention", "num_heads"));
      d.mlp = static_cast<int32_t>(need("encoder.layers.encoder_layer_0.mlp.0.weight").size(0));
      d.patch = static_cast<int32_t>(conv.size(2));
      d.channels = static_cast<int32_t>(conv.size(1));
      const int side = static_cast<int>(std::lround(std::sqrt(static_cast<double>(pos.size(1) - 1))));
      if (side * side + 1 != pos.size(1)) throw std::runtime_error("pos_embedding is not a square grid + class token");
      d.image = side * d.patch;
      d.classes = static_cast<int32_t>(need("heads.head.weight").size(0));
      d.ln_eps = resolve_eps(ln_eps, facts, find_attr(module, "encoder.layers.encoder_layer_0.ln_1", "eps"));
      names = vit_names(layers);
    }
    for (const std::string &n : names) {
      torch::Tensor t = need(n).detach().to(torch::kCPU, torch::kFloat32).contiguous();
      m->tensors.push_back({n, t});
    }
    *out = m.release();
    return 0;
  } catch (const std::exception &e) {
    g_error = std::string("loader_open(") + path + "): " + e.what();
    return 1;
  }
}

void spu_loader_close(spu_module *m) { delete m; }
int32_t spu_loader_kind(const spu_module *m) { return m ? m->kind : -1; }
int32_t spu_loader_num_tensors(const spu_module *m) { return m ? static_cast<int32_t>(m->tensors.size()) : 0; }

int32_t spu_loader_tensor(const spu_module *m, int32_t i, const float **data, uint64_t *numel, char *name,
                          size_t name_cap) {
  if (!m || i < 0 || i >= static_cast<int32_t>(m->tensors.size())) {
    g_error = "loader_tensor: index out of range";
    return 1;
  }
  const Named &n = m->tensors[static_cast<size_t>(i)];
  if (data) *data = n.t.data_ptr<float>();
  if (numel) *numel = static_cast<uint64_t>(n.t.numel());
  if (name && name_cap) std::snprintf(name, name_cap, "%s", n.name.c_str());
  return 0;
}

int32_t spu_loader_bert_desc(const spu_module *m, spu_bert_desc *desc) {
  if (!m || !desc || m->kind != SPU_LOADER_BERT) {
    g_error = "loader_bert_desc: not a BERT module";
    return 1;
  }
  *desc = m->bert;
  return 0;
}

int32_t spu_loader_vit_desc(const spu_module *m, spu_vit_desc *desc) {
  if (!m || !desc || m->kind != SPU_LOADER_VIT) {
    g_error = "loader_vit_desc: not a ViT module";
    return 1;
  }
  *desc = m->vit;
  return 0;
}

int32_t spu_loader_create_replica(const spu_module *m, int32_t device, int32_t max_batch, int32_t max_seq,
                                  int32_t n_workspaces, uint32_t flags, spu_replica **out) {
  if (!m || !out) {
    g_error = "loader_create_replica: bad argument";
    return SPU_ERR_INVALID;
  }
  std::vector<const float *> ptrs;
  ptrs.reserve(m->tensors.size());
  for (const Named &n : m->tensors) ptrs.push_back(n.t.data_ptr<float>());
  // clone_model_to_gpus (inference_runner.cpp:251-275): one device copy of the weights per assignment
  const spu_status st =
      m->kind == SPU_LOADER_BERT
          ? spu_bert_replica_create(&m->bert, ptrs.data(), static_cast<int32_t>(ptrs.size()), device, max_batch,
                                    max_seq, n_workspaces, flags, out)
          : spu_vit_replica_create(&m->vit, ptrs.data(), static_cast<int32_t>(ptrs.size()), device, max_batch,
                                   n_workspaces, flags, out);
  if (st != SPU_OK) g_error = std::string("replica create failed: ") + spu_last_error();
  return static_cast<int32_t>(st);
}

}  // extern "C"
