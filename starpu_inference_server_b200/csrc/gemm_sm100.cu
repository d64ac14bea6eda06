// tcgen05/TMEM GEMM, TMA operand staging, warp-specialised persistent kernel.
//
// Roles (256 threads, 1 CTA per SM, grid = min(#tiles, #SMs)):
//   warp 0 lane 0 : TMA producer  (global -> SWIZZLE_128B smem ring, mbarrier tx-count)
//   warp 1 lane 0 : MMA issuer    (tcgen05.mma 128 x BLOCK_N x 16, fp32 accum in TMEM,
//                                  tcgen05.commit releases smem stages / publishes tiles)
//   warp 2        : TMEM allocator (2 accumulator buffers -> epilogue overlaps next tile)
//   warps 4..7    : epilogue (tcgen05.ld 32 lanes x 32 cols per warp; thread == output row)
//
// Replaces the cuBLAS/ATen GEMMs LibTorch picks inside Module::forward
// (/root/reference/src/core/starpu_setup.cpp:610) for the QKV, attention-output and
// FFN projections of HF BertModel (oracle/bert_oracle.py).
#include "gemm_sm100.cuh"

#include <mutex>

namespace spu {

namespace {

constexpr int kBlockM = 128;
constexpr int kTileKBytes = 128;  // one SWIZZLE_128B row: 64 bf16 or 128 e4m3
constexpr int kUmmaKBytes = 32;   // K=16 bf16 / K=32 e4m3 per tcgen05.mma
constexpr int kThreads = 256;
constexpr int kEpiWarp0 = 4;
constexpr int kABytes = kBlockM * kTileKBytes;  // 16 KiB
constexpr int kSmemBudget = 227 * 1024 - 2048;  // leave room for alignment + barriers

template <int BLOCK_N>
struct TileCfg {
  static constexpr int kBBytes = BLOCK_N * kTileKBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = kSmemBudget / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTmemCols = (2 * BLOCK_N <= 32)    ? 32
                                   : (2 * BLOCK_N <= 64)  ? 64
                                   : (2 * BLOCK_N <= 128) ? 128
                                   : (2 * BLOCK_N <= 256) ? 256
                                                          : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
};

struct KernelArgs {
  int M, N, K;
  GemmEpilogueArgs epi;
};

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const uint32_t (&v)[32], const KernelArgs &a,
                                               int row, int col0, bool row_ok) {
  // 32 consecutive output columns of one row live in this thread.
  float f[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * a.epi.acc_scale;
  if (a.epi.bias != nullptr) {
    const float4 *b4 = reinterpret_cast<const float4 *>(a.epi.bias + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 b = __ldg(b4 + j);
      f[4 * j + 0] += b.x;
      f[4 * j + 1] += b.y;
      f[4 * j + 2] += b.z;
      f[4 * j + 3] += b.w;
    }
  }
  if constexpr (EPI == SPU_EPI_BIAS_GELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
  }
  if (!row_ok) return;
  const size_t off = static_cast<size_t>(row) * a.N + col0;
  if constexpr (EPI == SPU_EPI_BIAS_RESID) {
    const float4 *r4 = reinterpret_cast<const float4 *>(a.epi.resid + off);
    float4 *o4 = reinterpret_cast<float4 *>(static_cast<float *>(a.epi.out) + off);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 r = r4[j];
      o4[j] = make_float4(f[4 * j] + r.x, f[4 * j + 1] + r.y, f[4 * j + 2] + r.z,
                          f[4 * j + 3] + r.w);
    }
  } else {
    uint4 *o4 = reinterpret_cast<uint4 *>(static_cast<__nv_bfloat16 *>(a.epi.out) + off);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      o4[j] = make_uint4(pack_bf16(f[8 * j], f[8 * j + 1]), pack_bf16(f[8 * j + 2], f[8 * j + 3]),
                         pack_bf16(f[8 * j + 4], f[8 * j + 5]),
                         pack_bf16(f[8 * j + 6], f[8 * j + 7]));
    }
    if (a.epi.out_fp8 != nullptr) {
      // per-tensor scaled e4m3 copy for a following FP8 GEMM
      uint32_t w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float s = a.epi.out_fp8_inv_scale;
        uint16_t lo, hi;
        asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(f[4 * j + 1] * s), "f"(f[4 * j] * s));
        asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(f[4 * j + 3] * s), "f"(f[4 * j + 2] * s));
        w[j] = static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
      }
      uint4 *q4 = reinterpret_cast<uint4 *>(static_cast<uint8_t *>(a.epi.out_fp8) + off);
      q4[0] = make_uint4(w[0], w[1], w[2], w[3]);
      q4[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
  }
}

template <int BLOCK_N, int EPI, int IN_FP8>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b, const KernelArgs args) {
  using Cfg = TileCfg<BLOCK_N>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kStageBytes = Cfg::kStageBytes;
  constexpr uint32_t kIdesc = make_idesc(IN_FP8 ? 0u : 1u, kBlockM, BLOCK_N);
  constexpr int kElemsPerKBlock = IN_FP8 ? 128 : 64;

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte aligned bases.
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + kStages * kStageBytes);
  uint64_t *empty_bar = full_bar + kStages;
  uint64_t *tmem_full_bar = empty_bar + kStages;
  uint64_t *tmem_empty_bar = tmem_full_bar + 2;
  uint32_t *tmem_ptr_smem = reinterpret_cast<uint32_t *>(tmem_empty_bar + 2);

  const int warp_idx = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int n_blocks = args.N / BLOCK_N;
  const int m_blocks = (args.M + kBlockM - 1) / kBlockM;
  const int num_tiles = n_blocks * m_blocks;
  const int num_k_blocks = args.K / kElemsPerKBlock;

  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
  }
  if (warp_idx == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp_idx == 2) {
    tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp_idx == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m_blk = tile / n_blocks;
        const int n_blk = tile % n_blocks;
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t *sa = smem + stage * kStageBytes;
          uint8_t *sb = sa + kABytes;
          mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
          tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kElemsPerKBlock, m_blk * kBlockM);
          tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * kElemsPerKBlock, n_blk * BLOCK_N);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
    __syncwarp();  // reconverge before the CTA-wide barrier below
  } else if (warp_idx == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int local_tile = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
        const int acc = local_tile & 1;
        const uint32_t acc_phase = (local_tile >> 1) & 1u;
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BLOCK_N);
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * kStageBytes);
          const uint64_t desc_a = make_sw128_kmajor_desc(sa);
          const uint64_t desc_b = make_sw128_kmajor_desc(sa + kABytes);
#pragma unroll
          for (int k = 0; k < kTileKBytes / kUmmaKBytes; ++k) {
            // advancing 32 bytes along K inside the 128-byte swizzle row = +2 in the
            // (addr >> 4) start-address field
            const uint64_t adv = static_cast<uint64_t>((k * kUmmaKBytes) >> 4);
            if constexpr (IN_FP8)
              umma_f8_ss(tmem_d, desc_a + adv, desc_b + adv, kIdesc, (kb | k) != 0);
            else
              umma_f16_ss(tmem_d, desc_a + adv, desc_b + adv, kIdesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);  // smem stage reusable once these MMAs retire
          if (kb == num_k_blocks - 1) umma_commit(&tmem_full_bar[acc]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp_idx >= kEpiWarp0) {
    // ===================== epilogue =====================
    const int q = warp_idx - kEpiWarp0;  // == warp_idx % 4: TMEM lane quarter this warp may read
    int local_tile = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local_tile) {
      const int m_blk = tile / n_blocks;
      const int n_blk = tile % n_blocks;
      const int acc = local_tile & 1;
      const uint32_t acc_phase = (local_tile >> 1) & 1u;
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tcgen05_fence_after();
      const int row = m_blk * kBlockM + q * 32 + lane;
      const bool row_ok = row < args.M;
      const uint32_t taddr =
          tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * BLOCK_N);
#pragma unroll 1
      for (int c = 0; c < BLOCK_N / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 32), v);
        tmem_ld_wait();
        epilogue_chunk<EPI>(v, args, row, n_blk * BLOCK_N + c * 32, row_ok);
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp_idx == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---- host side -----------------------------------------------------------------
using EncodeTiledFn = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                   const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                   const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

spu_status make_kmajor_tmap(CUtensorMap *map, const void *base, int rows, int k_elems,
                            int box_rows, GemmIn in) {
  EncodeTiledFn enc = get_encode_fn();
  SPU_REQUIRE(enc != nullptr, SPU_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  const int esize = in == GemmIn::FP8_E4M3 ? 1 : 2;
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(k_elems), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(k_elems) * esize};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kTileKBytes / esize),
                       static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estride[2] = {1, 1};
  const CUresult r = enc(map,
                         in == GemmIn::FP8_E4M3 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                                : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                         2, const_cast<void *>(base), gdim, gstride, box, estride,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
    return SPU_ERR_CUDA;
  }
  return SPU_OK;
}

int bn_index(int bn) {
  for (int i = 0; i < kNumBlockN; ++i)
    if (kBlockNCands[i] == bn) return i;
  return -1;
}

template <int BLOCK_N, int EPI, int IN_FP8>
spu_status launch_cfg(const GemmPlan &plan, int M, const GemmEpilogueArgs &epi,
                      cudaStream_t stream) {
  using Cfg = TileCfg<BLOCK_N>;
  auto kern = gemm_tcgen05_kernel<BLOCK_N, EPI, IN_FP8>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::kSmemBytes);
  });
  SPU_CUDA_TRY(attr_err);
  KernelArgs a{M, plan.N, plan.K, epi};
  const int tiles = (plan.N / BLOCK_N) * ((M + kBlockM - 1) / kBlockM);
  int dev = 0, sms = kNumSMsB200;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = tiles < sms ? tiles : sms;
  kern<<<grid, kThreads, Cfg::kSmemBytes, stream>>>(plan.tmap_a, plan.tmap_b[bn_index(BLOCK_N)], a);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

template <int BLOCK_N, int IN_FP8>
spu_status launch_epi(const GemmPlan &plan, int M, const GemmEpilogueArgs &epi,
                      cudaStream_t stream) {
  switch (plan.epilogue) {
    case SPU_EPI_BIAS: return launch_cfg<BLOCK_N, SPU_EPI_BIAS, IN_FP8>(plan, M, epi, stream);
    case SPU_EPI_BIAS_GELU: return launch_cfg<BLOCK_N, SPU_EPI_BIAS_GELU, IN_FP8>(plan, M, epi, stream);
    case SPU_EPI_BIAS_RESID: return launch_cfg<BLOCK_N, SPU_EPI_BIAS_RESID, IN_FP8>(plan, M, epi, stream);
    default: set_error("unknown epilogue"); return SPU_ERR_INVALID;
  }
}

template <int IN_FP8>
spu_status launch_bn(const GemmPlan &plan, int block_n, int M, const GemmEpilogueArgs &epi,
                     cudaStream_t stream) {
  switch (block_n) {
    case 64: return launch_epi<64, IN_FP8>(plan, M, epi, stream);
    case 128: return launch_epi<128, IN_FP8>(plan, M, epi, stream);
    case 192: return launch_epi<192, IN_FP8>(plan, M, epi, stream);
    case 256: return launch_epi<256, IN_FP8>(plan, M, epi, stream);
    default: set_error("unsupported BLOCK_N"); return SPU_ERR_INVALID;
  }
}

}  // namespace

int choose_block_n(int M, int N) {
  const int m_blocks = (M + kBlockM - 1) / kBlockM;
  int best = 0;
  double best_cost = 1e30;
  for (int bn : kBlockNCands) {
    if (N % bn != 0) continue;
    const int tiles = m_blocks * (N / bn);
    const int rounds = (tiles + kNumSMsB200 - 1) / kNumSMsB200;
    // tile cost ~ MMA cycles (proportional to bn) + fixed per-tile overhead
    const double cost = rounds * (bn + 24.0);
    if (cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

spu_status gemm_make_plan(GemmPlan *plan, const void *A, int rows_a, const void *W, int M, int N,
                          int K, int epilogue, GemmIn in, int block_n_override) {
  SPU_REQUIRE(plan && A && W, SPU_ERR_INVALID, "gemm: null argument");
  const int kelems = in == GemmIn::FP8_E4M3 ? 128 : 64;
  SPU_REQUIRE(M > 0 && N > 0 && K > 0, SPU_ERR_INVALID, "gemm: empty shape");
  SPU_REQUIRE(K % kelems == 0, SPU_ERR_UNSUPPORTED, "gemm: K must be a multiple of the 128-byte K tile");
  SPU_REQUIRE(N % 64 == 0, SPU_ERR_UNSUPPORTED, "gemm: N must be a multiple of 64");
  SPU_REQUIRE(rows_a >= M, SPU_ERR_INVALID, "gemm: A buffer smaller than M");
  plan->M = rows_a;
  plan->N = N;
  plan->K = K;
  plan->epilogue = epilogue;
  plan->in = in;
  plan->block_n_override = block_n_override;
  if (block_n_override)
    SPU_REQUIRE(bn_index(block_n_override) >= 0 && N % block_n_override == 0, SPU_ERR_UNSUPPORTED,
                "gemm: BLOCK_N override does not divide N");
  spu_status st = make_kmajor_tmap(&plan->tmap_a, A, rows_a, K, kBlockM, in);
  if (st != SPU_OK) return st;
  for (int i = 0; i < kNumBlockN; ++i) {
    plan->bn_ok[i] = (N % kBlockNCands[i] == 0);
    if (!plan->bn_ok[i]) continue;
    st = make_kmajor_tmap(&plan->tmap_b[i], W, N, K, kBlockNCands[i], in);
    if (st != SPU_OK) return st;
  }
  return SPU_OK;
}

spu_status gemm_launch(const GemmPlan &plan, int M, const GemmEpilogueArgs &args,
                       cudaStream_t stream) {
  SPU_REQUIRE(M > 0 && M <= plan.M, SPU_ERR_CAPACITY, "gemm: M exceeds plan capacity");
  SPU_REQUIRE(args.out != nullptr, SPU_ERR_INVALID, "gemm: null output");
  if (plan.epilogue == SPU_EPI_BIAS_RESID)
    SPU_REQUIRE(args.resid != nullptr, SPU_ERR_INVALID, "gemm: residual epilogue without residual");
  const int bn = plan.block_n_override ? plan.block_n_override : choose_block_n(M, plan.N);
  SPU_REQUIRE(bn > 0, SPU_ERR_UNSUPPORTED, "gemm: no BLOCK_N divides N");
  return plan.in == GemmIn::FP8_E4M3 ? launch_bn<1>(plan, bn, M, args, stream)
                                     : launch_bn<0>(plan, bn, M, args, stream);
}

}  // namespace spu
