// ViT encoder forward (placeholder until the kernels land in this round).
#include "common.cuh"

extern "C" spu_status spu_vit_replica_create(const spu_vit_desc *, const float *const *, int32_t, int32_t, int32_t,
                                             int32_t, uint32_t, spu_replica **out) {
  if (out) *out = nullptr;
  spu::set_error("vit: not built yet");
  return SPU_ERR_UNSUPPORTED;
}
extern "C" spu_status spu_vit_forward_async(spu_replica *, int32_t, const float *, int32_t, float *, void *) {
  spu::set_error("vit: not built yet");
  return SPU_ERR_UNSUPPORTED;
}
extern "C" uint64_t spu_vit_input_bytes(const spu_replica *, int32_t) { return 0; }
extern "C" uint64_t spu_vit_output_bytes(const spu_replica *, int32_t) { return 0; }
