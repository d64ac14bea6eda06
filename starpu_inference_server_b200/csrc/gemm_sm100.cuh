// tcgen05/TMEM GEMM with TMA-fed operands and fused, TMA-stored epilogues (sm_100a).
//   C[M,N] = A[M,K] * W[N,K]^T  (+bias, then GELU | +fp32 residual)
// Both operands are K-major (activations row-major, nn.Linear weights [out,in]),
// which is the native TN form of tcgen05.mma with SWIZZLE_128B smem tiles.
#pragma once
#include <cuda.h>

#include <vector>

#include "common.cuh"

namespace spu {

enum class GemmIn : int { BF16 = 0, FP8_E4M3 = 1 };

// Internal epilogue kinds (the first three are the public SPU_EPI_* values).
constexpr int kEpiBias = SPU_EPI_BIAS;             // bf16 out
constexpr int kEpiBiasGelu = SPU_EPI_BIAS_GELU;    // bf16 out
constexpr int kEpiBiasResid = SPU_EPI_BIAS_RESID;  // fp32 out = acc + bias + fp32 residual
constexpr int kEpiBiasGeluE4m3 = 3;   // e4m3 out = gelu(acc*s + bias) * out_inv_scale
constexpr int kEpiBiasAccum = 4;  // fp32 out += acc*s + bias by TMA reduce-add: BIAS_RESID when resid == out
// y = LayerNorm(acc*s + bias + resid) * gamma + beta over the N columns of a row, fused: the CTAs (or CTA
// pairs) that hold the column blocks of one row block form a thread-block cluster and exchange per-row
// (mean, M2) over distributed shared memory.  Writes y as fp32 (out) and / or bf16 / e4m3 (out2).
constexpr int kEpiBiasResidLN = 5;

struct GemmEpilogueArgs {
  const float *bias;    // [N] fp32 (may be null)
  const float *resid;   // [M,N] fp32 (kEpiBiasResid; may alias out)
  void *out;            // bf16 / fp32 / e4m3 [M,N] according to the epilogue kind
  float acc_scale;      // multiplies the accumulator (fp8: a_scale*w_scale; bf16: 1)
  float out_inv_scale;  // kEpiBiasGeluE4m3 / LN e4m3 output: 1/scale of the quantised output
  // kEpiBiasResidLN only
  const float *gamma = nullptr, *beta = nullptr;  // [N] fp32
  float ln_eps = 0.f;
  void *out2 = nullptr;  // bf16 (out2_e4m3 == 0) or e4m3 [M,N] copy of y; null = none
  int out2_e4m3 = 0;
  // out may be null (no fp32 y) when out2 is given
};

// Can the LayerNorm-fused epilogue serve an output width of N columns (cluster of <= 8 CTAs spanning N)?
bool gemm_ln_supported(int N);

// Operand-B TMA boxes come in multiples of 32 rows: BLOCK_N / cta_group in {32..256}.
constexpr int kNumBBoxes = 8;

struct GemmConfig {
  int cta_group;  // 1: 128 x BLOCK_N tile per CTA; 2: 256 x BLOCK_N per CTA pair
  int block_n;
  int half = 0;   // CTA pairs only: half-footprint CTAs (<= 113 KB smem, 256 TMEM columns), two resident per SM
};

// Tiles measured at warm-up (spu_bert_autotune): the winner per (N, K, epilogue, operand format, M) replaces the
// cost model's choice for exactly that problem.  Written at start-up only, read by every launch.
struct TunedTable {
  struct Entry {
    int N, K, epilogue, in, M;
    GemmConfig cfg;
    float us;  // measured microseconds per launch of the winner
  };
  std::vector<Entry> entries;
  const Entry *find(int N, int K, int epilogue, int in, int M) const {
    for (const Entry &e : entries)
      if (e.M == M && e.N == N && e.K == K && e.epilogue == epilogue && e.in == in) return &e;
    return nullptr;
  }
};

// Activation-side tensor maps of one launch (they carry the live row count M and the epilogue pointers).  A plan
// remembers the last few it encoded: a worker re-launching the same problem eagerly (before its graph exists, or
// with graphs off) pays cuTensorMapEncodeTiled once, not three times per launch.
struct GemmLaunchMaps {
  CUtensorMap a, out, resid, out2;
};
struct GemmMapCacheEntry {
  int M = 0;
  const void *out = nullptr, *resid = nullptr, *out2 = nullptr;
  int out2_e4m3 = 0;
  GemmLaunchMaps maps;
};

struct GemmPlan {
  CUtensorMap tmap_b[kNumBBoxes];  // index = box_rows/32 - 1
  bool b_ok[kNumBBoxes];
  const void *A;                   // activation buffer the plan was made for
  int rows_a;                      // row capacity of A (and of out / resid)
  int N, K;
  int epilogue;
  GemmIn in;
  int force_cg, force_bn;          // 0 = choose per launch from M; force_cg 3 = CTA pair with half-footprint CTAs
  const TunedTable *tuned = nullptr;  // consulted before the cost model when no tile is forced
  // single-threaded like the workspace the plan belongs to
  mutable GemmMapCacheEntry map_cache[4];
  mutable unsigned map_cache_next = 0;
};

// Every (cta_group, BLOCK_N) the kernel is instantiated for that divides N (cta_group 3 = half-footprint pair).
int gemm_candidates(int N, int epilogue, GemmConfig *out, int cap);

// Pick (cta_group, BLOCK_N) for an [M,N,K] problem: BLOCK_N must divide N; minimise
// persistent rounds x per-tile cost (MMA issue, L2->smem feed, epilogue) over 148 SMs.
GemmConfig choose_config(int M, int N, int K, int esize);

// Encode the weight tensor maps.  rows_a is the number of rows the A / out / resid
// buffers really hold (>= any M later passed to gemm_launch).
spu_status gemm_make_plan(GemmPlan *plan, const void *A, int rows_a, const void *W, int N, int K,
                          int epilogue, GemmIn in, int force_cg = 0, int force_bn = 0);

// Enqueue.  Only (M, epilogue pointers) may change between calls with one plan.
spu_status gemm_launch(const GemmPlan &plan, int M, const GemmEpilogueArgs &args,
                       cudaStream_t stream);

// Debug: have CTA `cta` of every following launch write its pipeline time stamps (65 x u64,
// %globaltimer ns) to device_buf; nullptr switches tracing off.  Not thread-safe, tools only.
void gemm_set_trace(unsigned long long *device_buf, int cta);
// Debug: 0 = normal; 1 = no TMA operand loads; 2 = no MMAs (results are garbage; rate experiments only).
void gemm_set_debug_mode(int mode);

}  // namespace spu
