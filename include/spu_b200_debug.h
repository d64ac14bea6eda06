/*
 * spu_b200_debug.h -- measurement hooks of libspu_b200.so, present ONLY in a library built with
 * -DSPU_DEBUG_HOOKS (CMake: -DSPU_DEBUG_HOOKS=ON; in-tree: SPU_DEBUG_HOOKS=1 python -m
 * starpu_inference_server_b200.build).  The product build does not compile them: they are process-wide
 * switches read by every GEMM launch of every worker thread, and mode 1 / 2 make every GEMM of the process
 * return garbage.  Used by tools/gemm_trace.py and tools/gemm_rates.py on a dedicated process.
 */
#ifndef SPU_B200_DEBUG_H
#define SPU_B200_DEBUG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* CTA `cta` (| local tile << 16) of every following GEMM launch writes u64 %globaltimer stamps of its
 * pipeline to device_buf (128 words); NULL switches it off. */
void spu_debug_gemm_trace(void *device_buf, int32_t cta);
/* 0 = normal, 1 = GEMM launches skip their TMA operand loads, 2 = skip their MMAs, 3 = twice the MMAs per
 * barrier round; bits 8..11 = use fewer ring stages.  Results are garbage for mode != 0. */
void spu_debug_gemm_mode(int32_t mode);

#ifdef __cplusplus
}
#endif
#endif /* SPU_B200_DEBUG_H */
