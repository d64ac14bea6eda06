// tcgen05/TMEM GEMM with TMA-fed operands and fused epilogues (sm_100a).
//   C[M,N] = A[M,K] * W[N,K]^T  (+bias, then GELU | +fp32 residual)
// Both operands are K-major (activations row-major, nn.Linear weights [out,in]),
// which is the native TN form of tcgen05.mma with SWIZZLE_128B smem tiles.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace spu {

enum class GemmIn : int { BF16 = 0, FP8_E4M3 = 1 };

struct GemmEpilogueArgs {
  const float *bias;   // [N] fp32 (may be null)
  const float *resid;  // [M,N] fp32 (EPI_BIAS_RESID)
  void *out;           // bf16 [M,N] (BIAS, BIAS_GELU) or fp32 [M,N] (BIAS_RESID)
  void *out_fp8;       // optional e4m3 [M,N] copy of the bf16 result (fp8 pipelines)
  float acc_scale;     // multiplies the accumulator (fp8: a_scale*w_scale; bf16: 1)
  float out_fp8_inv_scale;  // 1/scale used when writing out_fp8
};

constexpr int kNumBlockN = 4;
constexpr int kBlockNCands[kNumBlockN] = {256, 192, 128, 64};

struct GemmPlan {
  CUtensorMap tmap_a;
  CUtensorMap tmap_b[kNumBlockN];  // one per candidate BLOCK_N that divides N
  bool bn_ok[kNumBlockN];
  int M, N, K;                     // M = row capacity of the A buffer
  int block_n_override;            // 0 = choose per launch from M
  int epilogue;
  GemmIn in;
};

// Pick BLOCK_N: must divide N; prefer the choice that minimises the number of
// persistent rounds times tile cost over 148 SMs.
int choose_block_n(int M, int N);

// Encode the two operand tensor maps.  rows_a is the number of rows the A buffer
// really holds (>= M; tail tiles read but never store rows >= M).
spu_status gemm_make_plan(GemmPlan *plan, const void *A, int rows_a, const void *W, int M, int N,
                          int K, int epilogue, GemmIn in, int block_n_override = 0);

// Enqueue.  Only (M, epilogue pointers) may change between calls with one plan.
spu_status gemm_launch(const GemmPlan &plan, int M, const GemmEpilogueArgs &args,
                       cudaStream_t stream);

}  // namespace spu
