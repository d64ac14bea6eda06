/*
 * spu_loader.h -- C-ABI of libspu_loader.so: start-up-only weight loading from the
 * TorchScript artifacts the reference serves (models/bert.yml `model: bert_libtorch.pt`,
 * models/vit_l_16.yml), into replicas of libspu_b200.so.
 *
 * Replaces load_model + clone_model_to_gpus of the reference
 * (src/core/inference_runner.cpp:243-275): `torch::jit::load(path)` + `eval()` once on the
 * host, then -- instead of `model_cpu.clone().to(cuda:id)` per device -- the module's
 * named parameters are handed, in the order include/spu_b200.h documents, to
 * spu_bert_replica_create / spu_vit_replica_create for every device of
 * build_gpu_replica_assignments (src/core/inference_runner.cpp:128-183).
 *
 * This is the ONLY library of the repo that links LibTorch, and nothing on the forward
 * path links it: libspu_b200.so and libspu_host.so stay torch-free.  After the replicas
 * exist the module can be closed; no torch code runs per request.
 */
#ifndef SPU_LOADER_H
#define SPU_LOADER_H

#include <stddef.h>
#include <stdint.h>

#include "spu_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

#define SPU_LOADER_BERT 0 /* HF BertModel traced like models/import_bert-base-uncased.py (keys may carry "bert.") */
#define SPU_LOADER_VIT 1  /* torchvision vit_{b,l}_16 scripted / traced like models/import_vit.py */

typedef struct spu_module spu_module;

/* torch::jit::load(path, CPU) + eval(); collects the fp32 host copy of every tensor the
 * replica needs, in SPU_BERT_NUM_TENSORS / SPU_VIT_NUM_TENSORS order, and infers the
 * architecture from their shapes (heads and the LayerNorm epsilon are read from the artifact, see spu_loader_open_ex).  Returns 0, or non-zero with
 * spu_loader_last_error() set (missing file, missing parameter, unsupported shape). */
int32_t spu_loader_open(const char *path, int32_t kind, spu_module **out);
/* The same with the two facts the tensor shapes do not carry supplied by the caller (models/*.yml): heads > 0 and
 * ln_eps > 0 override what the loader reads from the artifact; 0 = read it.  The loader reads the LayerNorm epsilon
 * from the constant eps input of the forward graph's aten::layer_norm nodes (else the `eps` attribute of the first
 * LayerNorm) and the head count from the `num_heads` / `num_attention_heads` attribute of the first attention
 * module (else the constant trailing dims of the [B, S, heads, head_dim] view in the forward graph).  It never
 * guesses: when neither source has them, or when head_dim is not the 64 the attention kernels are built for
 * (e.g. ViT-H/14: 16 heads of 80), the call fails with the reason in spu_loader_last_error(). */
int32_t spu_loader_open_ex(const char *path, int32_t kind, int32_t heads, float ln_eps, spu_module **out);
void spu_loader_close(spu_module *m);

int32_t spu_loader_kind(const spu_module *m);
int32_t spu_loader_num_tensors(const spu_module *m);
/* i-th tensor: fp32, contiguous, host; valid until spu_loader_close.  name may be NULL. */
int32_t spu_loader_tensor(const spu_module *m, int32_t i, const float **data, uint64_t *numel,
                          char *name, size_t name_cap);
/* ln_eps is not recoverable from the weights: 1e-12 (HF BertConfig default) / 1e-6 (torchvision). */
int32_t spu_loader_bert_desc(const spu_module *m, spu_bert_desc *desc);
int32_t spu_loader_vit_desc(const spu_module *m, spu_vit_desc *desc);

/* One replica on `device` from the opened module (call once per entry of
 * build_gpu_replica_assignments).  max_seq is ignored for ViT.  Returns the spu_status of
 * the underlying create call (message: spu_last_error()). */
int32_t spu_loader_create_replica(const spu_module *m, int32_t device, int32_t max_batch,
                                  int32_t max_seq, int32_t n_workspaces, uint32_t flags,
                                  spu_replica **out);

const char *spu_loader_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* SPU_LOADER_H */
