// Fused masked-softmax attention for head_dim 64 (BERT-base / ViT-L heads).
//   ctx[b, q, h, :] = softmax_k( q.k / 8 + mask_bias[b, k] ) . v
// One CTA = 64 query rows of one (batch, head); 4 warps x 16 rows.  K/V stream
// through a double-buffered cp.async smem ring in 64-key tiles with an online
// (running max / running sum) softmax, so scores and probabilities never touch
// HBM.  QK^T and PV use bf16 mma.sync m16n8k16 with fp32 accumulation; row
// reductions are quad warp-shuffles; global traffic is 128-bit coalesced.
//
// Restates the attention block of HF BertSelfAttention that LibTorch executes
// inside Module::forward (/root/reference/src/core/starpu_setup.cpp:610); oracle:
// oracle/bert_oracle.py:bert_forward (scores/softmax/ctx lines).
#include "replica.cuh"

namespace spu {
namespace {

constexpr int kHeadDim = 64;
constexpr int kQTile = 64;
constexpr int kKTile = 64;
constexpr int kThreads = 128;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gsrc, bool valid) {
  const int src_size = valid ? 16 : 0;  // src_size 0 -> 16 bytes of zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(src_size)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2,
                                            uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t &r0, uint32_t &r1,
                                                  uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, "
      "{%8, %9}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Tile of 64 rows x 64 bf16 (128 B per row), 16-byte chunks XOR-swizzled by row so
// ldmatrix row fetches hit 8 distinct bank groups.
__device__ __forceinline__ uint32_t tile_off(int row, int chunk) {
  return static_cast<uint32_t>(row * 128 + ((chunk ^ (row & 7)) << 4));
}

// Copy rows [row0, row0+64) x 64 columns at column offset `col0` of a [*, ld] bf16 matrix.
__device__ __forceinline__ void load_tile_async(uint8_t *smem_tile, const __nv_bfloat16 *g, int ld,
                                                int row0, int rows_valid_end, int col0) {
#pragma unroll
  for (int i = 0; i < (64 * 8) / kThreads; ++i) {
    const int idx = threadIdx.x + i * kThreads;
    const int r = idx >> 3, c = idx & 7;
    const bool ok = (row0 + r) < rows_valid_end;
    const __nv_bfloat16 *src = g + static_cast<size_t>(ok ? row0 + r : row0) * ld + col0 + c * 8;
    cp_async_16(smem_tile + tile_off(r, c), src, ok);
  }
}

__device__ __forceinline__ uint16_t pack_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}

template <bool OUT_E4M3>
__global__ void __launch_bounds__(kThreads)
attention_kernel(const __nv_bfloat16 *__restrict__ qkv, const float *__restrict__ mask_bias,
                 void *__restrict__ ctx_out, int seq, int heads, float q_inv_scale) {
  __shared__ __align__(128) uint8_t s_q[kQTile * 128];
  __shared__ __align__(128) uint8_t s_k[2][kKTile * 128];
  __shared__ __align__(128) uint8_t s_v[2][kKTile * 128];
  __shared__ float s_bias[2][kKTile];

  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int qt = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int hidden = heads * kHeadDim;
  const int ld = 3 * hidden;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const size_t tok0 = static_cast<size_t>(b) * seq;
  const __nv_bfloat16 *base = qkv + tok0 * ld;
  const int n_kv = (seq + kKTile - 1) / kKTile;

  auto load_kv = [&](int buf, int j) {
    load_tile_async(s_k[buf], base, ld, j * kKTile, seq, hidden + h * kHeadDim);
    load_tile_async(s_v[buf], base, ld, j * kKTile, seq, 2 * hidden + h * kHeadDim);
    if (threadIdx.x < kKTile) {
      const int key = j * kKTile + threadIdx.x;
      // keys past the sequence end never contribute (-inf); masked keys get the
      // additive bias the mask builder wrote (HF: (1-mask)*finfo.min)
      s_bias[buf][threadIdx.x] =
          key < seq ? (mask_bias != nullptr ? mask_bias[static_cast<size_t>(b) * seq + key] * kLog2e : 0.f)
                    : -INFINITY;
    }
  };

  load_tile_async(s_q, base, ld, qt * kQTile, seq, h * kHeadDim);
  load_kv(0, 0);
  cp_async_commit();

  uint32_t qf[4][4];
  float o[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const float scale_log2 = 0.125f * kLog2e;  // 1/sqrt(64) folded into exp2

  for (int j = 0; j < n_kv; ++j) {
    const int buf = j & 1;
    if (j + 1 < n_kv) {
      load_kv(buf ^ 1, j + 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (j == 0) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const int row = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int chunk = ks * 2 + (lane >> 4);
        ldmatrix_x4(smem_u32(s_q + tile_off(row, chunk)), qf[ks][0], qf[ks][1], qf[ks][2],
                    qf[ks][3]);
      }
    }
    // ---- S = Q K^T for 16 rows x 64 keys ----
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        const int key = np * 16 + (lane & 7) + (lane >> 4) * 8;
        const int chunk = ks * 2 + ((lane >> 3) & 1);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(smem_u32(s_k[buf] + tile_off(key, chunk)), b0, b1, b2, b3);
        mma_bf16_16816(s[2 * np], qf[ks], b0, b1);
        mma_bf16_16816(s[2 * np + 1], qf[ks], b2, b3);
      }
    }
    // ---- scale + mask + online softmax ----
    float m_new[2] = {m_run[0], m_run[1]};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float bz0 = s_bias[buf][i * 8 + 2 * t], bz1 = s_bias[buf][i * 8 + 2 * t + 1];
      s[i][0] = fmaf(s[i][0], scale_log2, bz0);
      s[i][1] = fmaf(s[i][1], scale_log2, bz1);
      s[i][2] = fmaf(s[i][2], scale_log2, bz0);
      s[i][3] = fmaf(s[i][3], scale_log2, bz1);
      m_new[0] = fmaxf(m_new[0], fmaxf(s[i][0], s[i][1]));
      m_new[1] = fmaxf(m_new[1], fmaxf(s[i][2], s[i][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      m_new[r] = fmaxf(m_new[r], __shfl_xor_sync(0xffffffffu, m_new[r], 1));
      m_new[r] = fmaxf(m_new[r], __shfl_xor_sync(0xffffffffu, m_new[r], 2));
    }
    const float corr0 = exp2f(m_run[0] - m_new[0]);
    const float corr1 = exp2f(m_run[1] - m_new[1]);
    m_run[0] = m_new[0];
    m_run[1] = m_new[1];
    l_run[0] *= corr0;
    l_run[1] *= corr1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      o[i][0] *= corr0;
      o[i][1] *= corr0;
      o[i][2] *= corr1;
      o[i][3] *= corr1;
    }
    uint32_t pf[4][4];  // P as bf16 A-fragments: 4 k-steps of 16 keys
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float p0 = exp2f(s[i][0] - m_new[0]);
      const float p1 = exp2f(s[i][1] - m_new[0]);
      const float p2 = exp2f(s[i][2] - m_new[1]);
      const float p3 = exp2f(s[i][3] - m_new[1]);
      l_run[0] += p0 + p1;
      l_run[1] += p2 + p3;
      pf[i >> 1][(i & 1) * 2 + 0] = pack_bf16(p0, p1);
      pf[i >> 1][(i & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
    // ---- O += P V ----
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int np = 0; np < 4; ++np) {
        const int key = ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        const int chunk = np * 2 + (lane >> 4);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(smem_u32(s_v[buf] + tile_off(key, chunk)), b0, b1, b2, b3);
        mma_bf16_16816(o[2 * np], pf[ks], b0, b1);
        mma_bf16_16816(o[2 * np + 1], pf[ks], b2, b3);
      }
    }
    __syncthreads();  // everyone done with buf before it is refilled
  }

  // ---- normalise, stage through this warp's 16 rows of s_q, 128-bit stores ----
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  const float inv0 = 1.f / l_run[0], inv1 = 1.f / l_run[1];
  if constexpr (OUT_E4M3) {
    // per-tensor scaled e4m3 operand of the attention-output GEMM: 64 rows x 64 bytes staged
    // in this warp's part of s_q, then one 16-byte store per row quarter
    uint8_t *ctx = static_cast<uint8_t *>(ctx_out);
    const float s0 = inv0 * q_inv_scale, s1 = inv1 * q_inv_scale;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r0 = warp * 16 + g, r1 = r0 + 8;
      *reinterpret_cast<uint16_t *>(s_q + r0 * 64 + i * 8 + t * 2) = pack_e4m3x2(o[i][0] * s0, o[i][1] * s0);
      *reinterpret_cast<uint16_t *>(s_q + r1 * 64 + i * 8 + t * 2) = pack_e4m3x2(o[i][2] * s1, o[i][3] * s1);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = lane + i * 32;
      const int r = warp * 16 + (idx >> 2), c = idx & 3;
      const int q_row = qt * kQTile + r;
      if (q_row < seq)
        *reinterpret_cast<uint4 *>(ctx + (tok0 + q_row) * hidden + h * kHeadDim + c * 16) =
            *reinterpret_cast<const uint4 *>(s_q + r * 64 + c * 16);
    }
  } else {
    __nv_bfloat16 *ctx = static_cast<__nv_bfloat16 *>(ctx_out);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r0 = warp * 16 + g, r1 = r0 + 8;
      *reinterpret_cast<uint32_t *>(s_q + tile_off(r0, i) + t * 4) =
          pack_bf16(o[i][0] * inv0, o[i][1] * inv0);
      *reinterpret_cast<uint32_t *>(s_q + tile_off(r1, i) + t * 4) =
          pack_bf16(o[i][2] * inv1, o[i][3] * inv1);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = lane + i * 32;
      const int r = warp * 16 + (idx >> 3), c = idx & 7;
      const int q_row = qt * kQTile + r;
      if (q_row < seq) {
        const uint4 v = *reinterpret_cast<const uint4 *>(s_q + tile_off(r, c));
        *reinterpret_cast<uint4 *>(ctx + (tok0 + q_row) * hidden + h * kHeadDim + c * 8) = v;
      }
    }
  }
}

}  // namespace

spu_status attention_mma_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq,
                                int heads, cudaStream_t stream, bool out_e4m3, float q_inv_scale) {
  SPU_REQUIRE(qkv && ctx, SPU_ERR_INVALID, "attention: null argument");  // mask_bias may be null (no mask)
  SPU_REQUIRE(batch > 0 && seq > 0 && heads > 0, SPU_ERR_INVALID, "attention: empty shape");
  SPU_REQUIRE(batch <= 65535 && heads <= 65535, SPU_ERR_CAPACITY, "attention: grid too large");
  dim3 grid((seq + kQTile - 1) / kQTile, heads, batch);
  const __nv_bfloat16 *q16 = static_cast<const __nv_bfloat16 *>(qkv);
  if (out_e4m3)
    SPU_CUDA_TRY(launch_pdl(attention_kernel<true>, grid, dim3(kThreads), 0, stream, q16, mask_bias, ctx, seq, heads,
                            q_inv_scale));
  else
    SPU_CUDA_TRY(launch_pdl(attention_kernel<false>, grid, dim3(kThreads), 0, stream, q16, mask_bias, ctx, seq, heads,
                            1.f));
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

// Default: the tcgen05/TMEM kernel (attention_sm100.cu), faster than this file's mma.sync kernel at
// every shape of the path on B200 (profiles/r06_attention.md: 1.4-2.2x from S=128 to S=512 and on
// ViT's ragged S=197).  SPU_ATTN=mma keeps the mma.sync kernel selectable as the measured baseline.
spu_status attention_launch(const void *qkv, const float *mask_bias, void *ctx, int batch, int seq,
                            int heads, cudaStream_t stream, bool out_e4m3, float q_inv_scale) {
  static const int forced = [] {
    const char *e = std::getenv("SPU_ATTN");
    if (e == nullptr) return 0;
    return std::string(e) == "mma" ? 2 : (std::string(e) == "tc" ? 1 : 0);
  }();
  const bool tc = forced != 2;
  return tc ? attention_tc_launch(qkv, mask_bias, ctx, batch, seq, heads, stream, out_e4m3, q_inv_scale)
            : attention_mma_launch(qkv, mask_bias, ctx, batch, seq, heads, stream, out_e4m3, q_inv_scale);
}

}  // namespace spu
