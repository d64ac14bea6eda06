// tcgen05/TMEM fused masked-softmax attention for head_dim 64 (BERT-base / ViT-L heads), sm_100a.
//   ctx[b, q, h, :] = softmax_k( q.k / 8 + mask_bias[b, k] ) . v
// One CTA = 128 query rows of one (batch, head).  Q, K and V tiles arrive by TMA
// (3-D map over qkv viewed as [B, S, 3H]: rows past the end of a sequence are zero-filled, so
// ragged S needs no special load path) into 128B-swizzled smem.  S = Q K^T is ONE 128x128x64
// tcgen05.mma chain into TMEM; a thread then owns (half of) one query row (tcgen05.ld 32x32b), so the
// row max / row sum need no shuffles at all; P goes back to smem as a K-major bf16 operand and
// O = P V runs as a second tcgen05.mma chain whose B operand is V as loaded (MN-major
// descriptor, no transpose pass).  Scores and probabilities never touch HBM.
//   SINGLE (S <= 128): one thread per query row (128 threads); P overwrites the dead Q|K tiles and O reuses
//          the TMEM columns of S: 48 KB smem + 128 TMEM columns + <=128 registers per CTA -> 4 CTAs per
//          SM, all (b, h) of a b<=49 batch resident at once (the kernel is latency-bound there).
//   MULTI  (S  > 128): KV loop with online softmax, TWO threads per query row (256 threads, each owning
//          64 of a tile's 128 key columns and 32 output columns; max / sum exchanged through smem), K
//          double-buffered, S_(j+1) issued under pass 2 of tile j into the second TMEM buffer, O_(j-1)
//          collected one tile late; V single-buffered so that a CTA needs 97 KB -> 2 CTAs per SM.
// Chunks / KV tiles behind the last live key of a hard-masked (padded) sequence are skipped: they add
// exactly 0 to every row sum and to O.
// PACKED mode (cu_seqlens != nullptr): qkv holds the sequences back to back ([total_tokens, 3H], no padding rows),
// sequence b owns rows cu_seqlens[b] .. cu_seqlens[b+1]; a CTA reads its own length, tiles that start behind it exit
// at once, a tile that runs past it sees the next sequence's rows as keys -- masked with -inf like any key past the
// end -- and as query rows that are computed and never stored.
//
// Restates the attention block of HF BertSelfAttention / torchvision MultiheadAttention that
// LibTorch executes inside Module::forward (/root/reference/src/core/starpu_setup.cpp:610);
// oracle: oracle/bert_oracle.py:bert_forward (scores/softmax/ctx lines).
#include <cuda.h>

#include <mutex>

#include "replica.cuh"

namespace spu {
namespace {

constexpr int kHeadDim = 64;
constexpr int kTile = 128;            // query rows per CTA == keys per KV tile
constexpr int kTileBytes = kTile * kHeadDim * 2;  // 16 KB: 128 rows x 128 B, SWIZZLE_128B
constexpr float kLog2e = 1.4426950408889634f;

template <bool MULTI>
struct AttnCfg {
  static constexpr int kBufs = MULTI ? 2 : 1;
  static constexpr int kOffQ = 0;
  static constexpr int kOffK = kTileBytes;
  static constexpr int kOffV = kOffK + kBufs * kTileBytes;
  // SINGLE: P (2 x 16 KB, keys 0-63 | 64-127) overwrites Q|K once S has been read out of TMEM
  // V is single-buffered in both variants (two CTAs of 97 KB per SM): V_j is fetched while pass 2 of tile j
  // runs, right after O_{j-1} has been collected
  static constexpr int kOffP = MULTI ? kOffV + kTileBytes : 0;
  static constexpr int kTiles = MULTI ? kOffP + 2 * kTileBytes : kOffV + kTileBytes;
  static constexpr int kOffBias = kTiles;                       // kBufs x 128 fp32
  static constexpr int kOffRed = kOffBias + kBufs * kTile * 4;   // 2 x 128 fp32: max / sum exchange (MULTI)
  static constexpr int kOffBars = kOffRed + 2 * kTile * 4;       // 6 mbarriers + tmem slot
  static constexpr int kSmemBytes = kOffBars + 96 + 1024;        // + 1024-alignment slack
  // S buffers of 128 fp32 columns (MULTI: two, tile j in buffer j&1); O of tile j overwrites
  // columns [0,64) of its own S buffer once every row owner has read S out
  static constexpr uint32_t kTmemCols = MULTI ? 256 : 128;
  // SINGLE: one thread per query row (128 threads, <=128 registers -> 4 CTAs/SM).
  // MULTI : two threads per row, each owning 64 of a tile's 128 key columns and 32 of the 64
  //         output columns (256 threads x 2 CTAs/SM = the same 16 warps/SM with half the
  //         per-thread softmax chain and register footprint).
  static constexpr int kTpr = MULTI ? 2 : 1;
  static constexpr int kThreads = kTile * kTpr;
  static constexpr int kOCols = kHeadDim / kTpr;
};

__device__ __forceinline__ void tma_load_3d(void *smem_dst, const void *tmap, uint64_t *bar, int32_t c0,
                                            int32_t c1, int32_t c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint16_t pack_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}

// byte offset of 16-byte chunk `chunk` of row `row` inside a 128-row x 128 B SWIZZLE_128B tile
__device__ __forceinline__ uint32_t sw128_off(int row, int chunk) {
  return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4));
}

// Instruction descriptor: bf16 x bf16 -> fp32, M x N, B operand MN-major when b_mn (bit 16).
__host__ __device__ constexpr uint32_t attn_idesc(uint32_t m, uint32_t n, bool b_mn) {
  return make_idesc(1u, m, n) | (b_mn ? (1u << 16) : 0u);
}

// One thread = one query row of S in TMEM (up to 128 fp32 columns), read in up to four 32-column chunks with the
// next chunk's tcgen05.ld in flight while the current one is consumed.
// PIPELINED = false (MULTI, 128-register budget with the running O resident): one chunk at a time.
template <bool PIPELINED, typename F>
__device__ __forceinline__ void for_each_chunk(uint32_t tmem_row, int nch, F &&f) {
  if (nch <= 0) return;  // warp-uniform
  if constexpr (!PIPELINED) {
    for (int c = 0; c < nch; ++c) {
      uint32_t v[32];
      tmem_ld_32x32b_x32(tmem_row + c * 32, v);
      tmem_ld_wait();
      f(c, v);
    }
    return;
  }
  uint32_t va[32], vb[32];
  tmem_ld_32x32b_x32(tmem_row, va);
  tmem_ld_wait();
  if (nch > 1) tmem_ld_32x32b_x32(tmem_row + 32, vb);
  f(0, va);
  if (nch > 1) {
    tmem_ld_wait();
    if (nch > 2) tmem_ld_32x32b_x32(tmem_row + 64, va);
    f(1, vb);
    if (nch > 2) {
      tmem_ld_wait();
      if (nch > 3) tmem_ld_32x32b_x32(tmem_row + 96, vb);
      f(2, va);
      if (nch > 3) {
        tmem_ld_wait();
        f(3, vb);
      }
    }
  }
}

// max_k (s*scale + bias_k); four independent accumulators keep the FMNMX chain short
template <bool BIAS, bool PIPELINED>
__device__ __forceinline__ float row_max(uint32_t tmem_row, int nch, const float *bias, float scale) {
  float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
  for_each_chunk<PIPELINED>(tmem_row, nch, [&](int c, const uint32_t(&v)[32]) {
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      if constexpr (BIAS) {
        const float4 bz = *reinterpret_cast<const float4 *>(bias + c * 32 + i);
        m0 = fmaxf(m0, fmaf(__uint_as_float(v[i + 0]), scale, bz.x));
        m1 = fmaxf(m1, fmaf(__uint_as_float(v[i + 1]), scale, bz.y));
        m2 = fmaxf(m2, fmaf(__uint_as_float(v[i + 2]), scale, bz.z));
        m3 = fmaxf(m3, fmaf(__uint_as_float(v[i + 3]), scale, bz.w));
      } else {
        m0 = fmaxf(m0, __uint_as_float(v[i + 0]));
        m1 = fmaxf(m1, __uint_as_float(v[i + 1]));
        m2 = fmaxf(m2, __uint_as_float(v[i + 2]));
        m3 = fmaxf(m3, __uint_as_float(v[i + 3]));
      }
    }
  });
  const float m = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
  return BIAS ? m : m * scale;  // scale > 0
}

// p_k = 2^(s*scale + bias_k - m) for chunks [c0, c0+nch) -> bf16 into the two 64-key K-major SWIZZLE_128B tiles at s_p;
// returns the fp32 row sum
template <bool BIAS, bool PIPELINED>
__device__ __forceinline__ float row_exp(uint32_t tmem_row, int nch, int c0, const float *bias, float scale, float m,
                                         uint8_t *s_p, int row) {
  float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
  const float neg_m = -m;
  for_each_chunk<PIPELINED>(tmem_row, nch, [&](int cl, const uint32_t(&v)[32]) {
    const int c = c0 + cl;  // chunk index inside the 128-key tile; `bias` and `tmem_row` are already offset by c0
    uint8_t *p_half = s_p + (c >> 1) * kTileBytes;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float p[8];
#pragma unroll
      for (int e = 0; e < 8; e += 4) {
        const int i = q * 8 + e;
        if constexpr (BIAS) {
          const float4 bz = *reinterpret_cast<const float4 *>(bias + cl * 32 + i);
          p[e + 0] = ex2_approx(fmaf(__uint_as_float(v[i + 0]), scale, bz.x) - m);
          p[e + 1] = ex2_approx(fmaf(__uint_as_float(v[i + 1]), scale, bz.y) - m);
          p[e + 2] = ex2_approx(fmaf(__uint_as_float(v[i + 2]), scale, bz.z) - m);
          p[e + 3] = ex2_approx(fmaf(__uint_as_float(v[i + 3]), scale, bz.w) - m);
        } else {
          p[e + 0] = ex2_approx(fmaf(__uint_as_float(v[i + 0]), scale, neg_m));
          p[e + 1] = ex2_approx(fmaf(__uint_as_float(v[i + 1]), scale, neg_m));
          p[e + 2] = ex2_approx(fmaf(__uint_as_float(v[i + 2]), scale, neg_m));
          p[e + 3] = ex2_approx(fmaf(__uint_as_float(v[i + 3]), scale, neg_m));
        }
        l0 += p[e + 0];
        l1 += p[e + 1];
        l2 += p[e + 2];
        l3 += p[e + 3];
      }
      *reinterpret_cast<uint4 *>(p_half + sw128_off(row, (c & 1) * 4 + q)) =
          make_uint4(pack_bf16(p[0], p[1]), pack_bf16(p[2], p[3]), pack_bf16(p[4], p[5]), pack_bf16(p[6], p[7]));
    }
  });
  return (l0 + l1) + (l2 + l3);
}

template <bool MULTI, bool OUT_E4M3>
__global__ void __launch_bounds__(AttnCfg<MULTI>::kThreads, MULTI ? 2 : 4)
attention_tc_kernel(const __grid_constant__ CUtensorMap tm_qkv, const float *__restrict__ mask_bias,
                    void *__restrict__ ctx_out, int seq, int heads, float q_inv_scale,
                    const int32_t *__restrict__ cu_seqlens) {
  using Cfg = AttnCfg<MULTI>;
  constexpr int kOCols = Cfg::kOCols;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t *s_q = smem + Cfg::kOffQ;
  uint8_t *s_k = smem + Cfg::kOffK;
  uint8_t *s_v = smem + Cfg::kOffV;
  uint8_t *s_p = smem + Cfg::kOffP;
  float *s_bias = reinterpret_cast<float *>(smem + Cfg::kOffBias);
  float *s_red = reinterpret_cast<float *>(smem + Cfg::kOffRed);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + Cfg::kOffBars);
  uint64_t *bar_k = bars, *bar_v = bars + 2, *bar_s = bars + 4, *bar_o = bars + 5;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 6);
  int *s_live = reinterpret_cast<int *>(bars + 7);  // one per warp (<= 8)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row = tid & (kTile - 1);          // query row of the tile this thread works on
  const int half = MULTI ? (tid >> 7) : 0;    // which half of the key / output columns (MULTI)
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int hidden = heads * kHeadDim;
  // PACKED: this sequence's rows start at row0 of the [total_tokens, 3H] array (TMA batch coordinate 0) and there are
  // `seq` of them; cu_seqlens was written by an earlier kernel of the chain (stream order, no PDL overlap with it)
  int row0 = 0, bz = b;
  if (cu_seqlens != nullptr) {
    pdl_wait_prior_grid();
    row0 = __ldg(cu_seqlens + b);
    seq = __ldg(cu_seqlens + b + 1) - row0;
    bz = 0;
    if (qt * kTile >= seq) {  // block-uniform: nothing of this tile is live
      pdl_launch_dependents();
      return;
    }
  }

  // thread 0 gets the first loads in flight before anything else; warp 1 allocates TMEM meanwhile
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 6; ++i) mbar_init(bars + i, 1);
    fence_barrier_init();
    pdl_wait_prior_grid();  // qkv is written by the preceding GEMM
    mbar_arrive_expect_tx(&bar_k[0], 2 * kTileBytes);
    tma_load_3d(s_q, &tm_qkv, &bar_k[0], h * kHeadDim, row0 + qt * kTile, bz);
    tma_load_3d(s_k, &tm_qkv, &bar_k[0], hidden + h * kHeadDim, row0, bz);
    mbar_arrive_expect_tx(&bar_v[0], kTileBytes);
    tma_load_3d(s_v, &tm_qkv, &bar_v[0], 2 * hidden + h * kHeadDim, row0, bz);
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::kTmemCols);
    tmem_relinquish();
  }
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  // Hard-masked tail (padding): keys after the last live one add exactly 0 to every row sum and to
  // O, so whole 32-key chunks / KV tiles behind it are skipped.  "Live" = bias above -1e29 (the
  // mask builder writes -1e30); softer additive masks are never skipped.  No live key at all
  // (fully padded sequence -> uniform softmax, as HF) keeps the full length.
  {
    int last = -1;
    if (mask_bias != nullptr)
      for (int k = tid; k < seq; k += Cfg::kThreads)
        if (__ldg(mask_bias + static_cast<size_t>(b) * seq + k) > -1e29f) last = k;
    last = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0) s_live[warp] = last;
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  int live_end = -1;
#pragma unroll
  for (int w = 0; w < Cfg::kThreads / 32; ++w) live_end = max(live_end, s_live[w]);
  live_end = (mask_bias == nullptr || live_end < 0) ? seq : live_end + 1;
  const int n_kv = MULTI ? (live_end + kTile - 1) / kTile : 1;
  const uint32_t tmem_base = *tmem_slot;
  // a warp may only touch TMEM lanes 32*(warp%4) .. +31: warps w and w+4 share the rows, split the columns
  const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>((warp & 3) * 32) << 16);

  constexpr uint32_t kIdescQK = attn_idesc(kTile, kTile, false);
  constexpr uint32_t kIdescPV = attn_idesc(kTile, kHeadDim, true);
  const float scale_log2 = 0.125f * kLog2e;  // 1/sqrt(64) folded into exp2
  float m_run = -INFINITY, l_run = 0.f;      // l_run: this thread's share of the row sum
  float o_run[kOCols];
#pragma unroll
  for (int i = 0; i < kOCols; ++i) o_run[i] = 0.f;

  // MMA issue helpers (thread 0 only)
  auto issue_qk = [&](int j) {  // S_j = Q K_j^T : 128 x 128 x 64 as four K=16 steps (32 B inside the swizzle atom)
    const int kb = MULTI ? (j & 1) : 0;
    mbar_wait(&bar_k[kb], (j >> 1) & 1);
    tcgen05_fence_after();
    const uint64_t dq = make_sw128_kmajor_desc(smem_u32(s_q));
    const uint64_t dk = make_sw128_kmajor_desc(smem_u32(s_k + kb * kTileBytes));
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_base + kb * kTile, dq + 2 * k, dk + 2 * k, kIdescQK, k > 0);
    umma_commit(bar_s);
  };
  // o_run = o_run * corr + this thread's columns of O_tile (tile in S buffer sb)
  auto load_o = [&](int sb, float corr) {
    if constexpr (MULTI) {
      uint32_t va[32];
      tmem_ld_32x32b_x32(tmem_row + sb * kTile + half * 32, va);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) o_run[i] = fmaf(o_run[i], corr, __uint_as_float(va[i]));
    } else {
      uint32_t va[32], vb[32];
      tmem_ld_32x32b_x32(tmem_row, va);
      tmem_ld_32x32b_x32(tmem_row + 32, vb);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        o_run[i] = __uint_as_float(va[i]);
        o_run[32 + i] = __uint_as_float(vb[i]);
      }
    }
  };

  if (tid == 0) {
    if (MULTI && n_kv > 1) {  // second K tile: in flight before the first softmax starts
      mbar_arrive_expect_tx(&bar_k[1], kTileBytes);
      tma_load_3d(s_k + kTileBytes, &tm_qkv, &bar_k[1], hidden + h * kHeadDim, row0 + kTile, bz);
    }
    issue_qk(0);
  }

  // Software pipeline over KV tiles (MULTI): while the row owners run pass 2 of tile j the tensor
  // core computes S_{j+1}; O_{j-1} = P V of the previous tile is collected one tile late, after
  // pass 1 of tile j, so neither MMA round trip is exposed.
  // keys past the sequence end never contribute (-inf); masked keys get the additive bias
  // the mask builder wrote (HF: (1-mask)*finfo.min), pre-multiplied by log2(e)
  auto load_bias = [&](int j) -> float {
    if (tid >= kTile) return 0.f;
    const int key = j * kTile + tid;
    return key < seq ? (mask_bias != nullptr ? __ldg(mask_bias + static_cast<size_t>(b) * seq + key) * kLog2e : 0.f)
                     : -INFINITY;
  };
  float next_bias = load_bias(0);
  float corr_prev = 0.f;
  for (int j = 0; j < n_kv; ++j) {
    const int buf = MULTI ? (j & 1) : 0;
    float *bias = s_bias + buf * kTile;
    const float my_bias = next_bias;
    if (tid < kTile) bias[tid] = my_bias;
    if (MULTI && j + 1 < n_kv) next_bias = load_bias(j + 1);  // L2 latency hidden behind this tile's work
    const int valid = min(kTile, live_end - j * kTile);
    const int nch = (valid + 31) >> 5;  // 32-key chunks that hold any live key (ragged last tile)
    const int c0 = MULTI ? half * 2 : 0;               // first chunk of this thread's key columns
    const int my_nch = MULTI ? min(2, nch - c0) : nch;  // may be <= 0 for the upper half of a short tail
    const uint32_t s_row = tmem_row + buf * kTile + c0 * 32;
    // bias visible to every row owner; a tile whose processed chunks carry no bias (no padding, no
    // tail inside them) takes the cheaper softmax path below (block-uniform branch)
    const bool has_bias = __syncthreads_or(my_bias != 0.f && tid < nch * 32) != 0;
    mbar_wait_warp(bar_s, j & 1);
    tcgen05_fence_after();
    if (MULTI && tid == 0 && j + 2 < n_kv) {  // S_j done: its K buffer takes tile j+2
      mbar_arrive_expect_tx(&bar_k[buf], kTileBytes);
      tma_load_3d(s_k + buf * kTileBytes, &tm_qkv, &bar_k[buf], hidden + h * kHeadDim, row0 + (j + 2) * kTile, bz);
    }

    // ---- pass 1: max of s' = s*scale + bias over this thread's columns of query row `row` ----
    float mx = has_bias ? row_max<true, !MULTI>(s_row, my_nch, bias + c0 * 32, scale_log2)
                        : row_max<false, !MULTI>(s_row, my_nch, bias + c0 * 32, scale_log2);
    if (MULTI) {
      if (j > 0) {  // collect O_{j-1} (long complete): frees its S buffer, its V buffer and the P tile
        mbar_wait_warp(bar_o, (j - 1) & 1);
        tcgen05_fence_after();
        load_o(buf ^ 1, corr_prev);
      }
      s_red[half * kTile + row] = mx;
      tcgen05_fence_before();
      __syncthreads();
      mx = fmaxf(mx, s_red[(half ^ 1) * kTile + row]);
      if (tid == 0) {
        if (j > 0) {  // P V of tile j-1 has completed: the V buffer takes tile j
          mbar_arrive_expect_tx(&bar_v[0], kTileBytes);
          tma_load_3d(s_v, &tm_qkv, &bar_v[0], 2 * hidden + h * kHeadDim, row0 + j * kTile, bz);
        }
        if (j + 1 < n_kv) issue_qk(j + 1);  // runs under pass 2 of tile j
      }
    }
    const float m_new = fmaxf(m_run, mx);
    const float m_use = m_new == -INFINITY ? 0.f : m_new;
    const float corr = ex2_approx(m_run - m_use);  // first tile: 2^-inf = 0
    m_run = m_new;
    corr_prev = corr;

    // ---- pass 2: p = 2^(s' - m), row sum, P -> smem as the K-major bf16 A operand of P V ----
    const float l_tile = has_bias ? row_exp<true, !MULTI>(s_row, my_nch, c0, bias + c0 * 32, scale_log2, m_use, s_p, row)
                                  : row_exp<false, !MULTI>(s_row, my_nch, c0, bias + c0 * 32, scale_log2, m_use, s_p, row);
    l_run = fmaf(l_run, corr, l_tile);
    fence_proxy_async_smem();  // generic-proxy P stores -> visible to the tensor core's async proxy
    tcgen05_fence_before();
    __syncthreads();           // all of S_j read out of TMEM, all of P_j in smem

    // ---- O_j = P V : 128 x 64 x (32*nch), V consumed as loaded (MN-major B, 16 keys = 2 KB per step) ----
    if (tid == 0) {
      tcgen05_fence_after();
      mbar_wait(&bar_v[0], j & 1);
      const uint32_t v_addr = smem_u32(s_v);
      for (int k = 0; k < 2 * nch; ++k) {
        const uint64_t dp = make_sw128_kmajor_desc(smem_u32(s_p + (k >> 2) * kTileBytes)) + 2 * (k & 3);
        const uint64_t dv = make_sw128_kmajor_desc(v_addr + k * 2048);
        umma_f16_ss(tmem_base + buf * kTile, dp, dv, kIdescPV, k > 0);
      }
      umma_commit(bar_o);
    }
  }
  mbar_wait_warp(bar_o, (n_kv - 1) & 1);
  tcgen05_fence_after();
  load_o(MULTI ? ((n_kv - 1) & 1) : 0, corr_prev);
  if (MULTI) {  // the two column owners of a row add their partial sums
    s_red[half * kTile + row] = l_run;
    __syncthreads();
    l_run += s_red[(half ^ 1) * kTile + row];
  }

  // ---- normalise and store: this thread holds kOCols columns of query row `row` ----
  const float inv = 1.f / l_run;
  const int q_row = qt * kTile + row;
  const size_t tok0 = cu_seqlens != nullptr ? static_cast<size_t>(row0) : static_cast<size_t>(b) * seq;
  if constexpr (OUT_E4M3) {
    // per-tensor scaled e4m3 operand of the attention-output GEMM: 64 contiguous bytes per row
    if (q_row < seq) {
      uint8_t *dst = static_cast<uint8_t *>(ctx_out) + (tok0 + q_row) * hidden + h * kHeadDim + half * kOCols;
      const float s = inv * q_inv_scale;
#pragma unroll
      for (int c = 0; c < kOCols / 16; ++c) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int i = c * 16 + e * 4;
          const uint32_t lo = pack_e4m3x2(o_run[i] * s, o_run[i + 1] * s);
          const uint32_t hi = pack_e4m3x2(o_run[i + 2] * s, o_run[i + 3] * s);
          w[e] = lo | (hi << 16);
        }
        *reinterpret_cast<uint4 *>(dst + c * 16) = make_uint4(w[0], w[1], w[2], w[3]);
      }
    }
  } else {
    // stage the rows in the (dead) first P tile, then 128-bit coalesced stores: 4 full 128-byte
    // rows per warp instruction
    uint8_t *stage = s_p;
#pragma unroll
    for (int c = 0; c < kOCols / 8; ++c) {
      const uint4 w = make_uint4(pack_bf16(o_run[c * 8 + 0] * inv, o_run[c * 8 + 1] * inv),
                                 pack_bf16(o_run[c * 8 + 2] * inv, o_run[c * 8 + 3] * inv),
                                 pack_bf16(o_run[c * 8 + 4] * inv, o_run[c * 8 + 5] * inv),
                                 pack_bf16(o_run[c * 8 + 6] * inv, o_run[c * 8 + 7] * inv));
      *reinterpret_cast<uint4 *>(stage + sw128_off(row, half * (kOCols / 8) + c)) = w;
    }
    __nv_bfloat16 *ctx = static_cast<__nv_bfloat16 *>(ctx_out);
    if (MULTI) __syncthreads();  // a row's two halves come from different warps
    else __syncwarp();           // a warp stores exactly the 32 rows it staged
    constexpr int kIters = kTile * 8 / Cfg::kThreads;
#pragma unroll
    for (int it = 0; it < kIters; ++it) {
      const int idx = MULTI ? tid + it * Cfg::kThreads : warp * 256 + lane + it * 32;
      const int r = idx >> 3, c = idx & 7;
      const int grow = qt * kTile + r;
      if (grow < seq)
        *reinterpret_cast<uint4 *>(ctx + (tok0 + grow) * hidden + h * kHeadDim + c * 8) =
            *reinterpret_cast<const uint4 *>(stage + sw128_off(r, c));
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---- host side -----------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

EncodeTiledFn attn_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

template <bool MULTI, bool OUT_E4M3>
spu_status launch_tc(const CUtensorMap &map, const float *mask_bias, void *ctx, int batch, int seq, int heads,
                     cudaStream_t stream, float q_inv_scale, const int32_t *cu_seqlens) {
  using Cfg = AttnCfg<MULTI>;
  auto kern = attention_tc_kernel<MULTI, OUT_E4M3>;
  SPU_CUDA_TRY((ensure_dynamic_smem<attention_tc_kernel<MULTI, OUT_E4M3>>(Cfg::kSmemBytes)));
  dim3 grid((seq + kTile - 1) / kTile, heads, batch);
  SPU_CUDA_TRY(launch_pdl(kern, grid, dim3(Cfg::kThreads), Cfg::kSmemBytes, stream, map, mask_bias, ctx, seq, heads,
                          q_inv_scale, cu_seqlens));
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

}  // namespace

// qkv: bf16 [batch*seq, 3*heads*64] (Q | K | V column blocks), ctx: bf16 or e4m3 [batch*seq, heads*64].
// PACKED (cu_seqlens != nullptr, DEVICE int32 [batch + 1]): qkv / ctx hold total_tokens rows, `seq` is the longest
// sequence (it sizes the grid and picks the kernel variant), mask_bias must be null.
spu_status attention_tc_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq, int heads,
                               cudaStream_t stream, bool out_e4m3, float q_inv_scale, const int32_t *cu_seqlens,
                               int total_tokens) {
  SPU_REQUIRE(qkv && ctx, SPU_ERR_INVALID, "attention: null argument");  // mask_bias may be null (no mask)
  SPU_REQUIRE(cu_seqlens == nullptr || (mask_bias == nullptr && total_tokens > 0), SPU_ERR_INVALID,
              "attention: packed sequences carry no mask and need their total row count");
  SPU_REQUIRE(batch > 0 && seq > 0 && heads > 0, SPU_ERR_INVALID, "attention: empty shape");
  SPU_REQUIRE(batch <= 65535 && heads <= 65535, SPU_ERR_CAPACITY, "attention: grid too large");
  SPU_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0, SPU_ERR_INVALID, "attention: qkv must be 16-byte aligned");
  EncodeTiledFn enc = attn_encode_fn();
  SPU_REQUIRE(enc != nullptr, SPU_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  // qkv viewed as [batch, seq, 3H]: a 128-row box that runs past a sequence end is zero-filled
  const cuuint64_t ld = static_cast<cuuint64_t>(3) * heads * kHeadDim;
  // packed: ONE "batch" of total_tokens rows (rows past the end are zero-filled, rows of the next sequence masked)
  const cuuint64_t rows = cu_seqlens != nullptr ? static_cast<cuuint64_t>(total_tokens) : static_cast<cuuint64_t>(seq);
  cuuint64_t gdim[3] = {ld, rows, cu_seqlens != nullptr ? 1u : static_cast<cuuint64_t>(batch)};
  cuuint64_t gstride[2] = {ld * 2, ld * 2 * rows};
  cuuint32_t box[3] = {kHeadDim, kTile, 1};
  cuuint32_t estride[3] = {1, 1, 1};
  CUtensorMap map;
  const CUresult r = enc(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void *>(qkv), gdim, gstride, box,
                         estride, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("attention: cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
    return SPU_ERR_CUDA;
  }
  if (seq <= kTile)
    return out_e4m3 ? launch_tc<false, true>(map, mask_bias, ctx, batch, seq, heads, stream, q_inv_scale, cu_seqlens)
                    : launch_tc<false, false>(map, mask_bias, ctx, batch, seq, heads, stream, 1.f, cu_seqlens);
  return out_e4m3 ? launch_tc<true, true>(map, mask_bias, ctx, batch, seq, heads, stream, q_inv_scale, cu_seqlens)
                  : launch_tc<true, false>(map, mask_bias, ctx, batch, seq, heads, stream, 1.f, cu_seqlens);
}

}  // namespace spu
