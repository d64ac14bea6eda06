// tcgen05/TMEM GEMM, TMA operand staging, warp-specialised persistent kernel with a
// smem-staged, TMA-stored epilogue.  One kernel template covers two tile shapes:
//   cta_group::1  : one CTA  owns a 128 x BLOCK_N tile  (small M)
//   cta_group::2  : a CTA pair (cluster of 2, one TPC) owns a 256 x BLOCK_N tile; each
//                   CTA stages its own 128 A rows and HALF of the B rows, so L2->smem
//                   traffic per FLOP drops by a third against the 128-row tile.
//
// Roles (384 threads, 1 CTA per SM, grid = cta_group x min(#tiles, #SMs / cta_group)):
//   warp 0 lane 0 : TMA producer  (global -> SWIZZLE_128B smem ring, mbarrier tx-count;
//                   in a pair both CTAs' loads credit the LEADER CTA's full barrier)
//   warp 1 lane 0 : MMA issuer    (leader CTA only: tcgen05.mma (128|256) x BLOCK_N x 16,
//                   fp32 accumulators in TMEM; tcgen05.commit frees smem stages in both
//                   CTAs and publishes finished accumulators)
//   warp 2        : TMEM allocator (2 accumulator buffers -> epilogue overlaps next tile)
//   warps 4..11   : epilogue.  Warp w reads TMEM lane quarter w%4; warps 4-7 take the low
//                   half of the tile's columns, 8-11 the high half.  Per 32-column chunk:
//                   tcgen05.ld 32x32b.x32 -> bias / GELU / residual in registers ->
//                   swizzled st.shared (conflict-free) -> one TMA store per 32x32 chunk
//                   (coalesced, clips the M tail).  The fp32 residual chunk is TMA-loaded
//                   into the same staging slot two chunks ahead and updated in place.
//
// Replaces the cuBLAS/ATen GEMMs LibTorch picks inside Module::forward
// (/root/reference/src/core/starpu_setup.cpp:610) for the QKV, attention-output and
// FFN projections of HF BertModel (oracle/bert_oracle.py).
#include "gemm_sm100.cuh"

#include <atomic>
#include <cstdlib>
#include <mutex>

namespace spu {

namespace {

constexpr int kRowsPerCta = 128;
constexpr int kTileKBytes = 128;  // one SWIZZLE_128B row: 64 bf16 or 128 e4m3
constexpr int kUmmaKBytes = 32;   // K=16 bf16 / K=32 e4m3 per tcgen05.mma
constexpr int kEpiWarp0 = 4;
constexpr int kMaxEpiWarps = 8;
constexpr int kABytes = kRowsPerCta * kTileKBytes;          // 16 KiB

// Footprint of one CTA.  HF = 0: the whole SM (227 KB of shared memory, all 512 TMEM columns, 8 epilogue
// warps).  HF = 1 ("half footprint"): at most half of everything (113 KB, 256 TMEM columns, 4 epilogue
// warps, <= 128 registers x 256 threads), so that TWO CTAs are resident per SM -- the second one is a CTA of
// the kernel the other CUDA worker's stream runs (or this stream's next kernel under programmatic dependent
// launch): its main loop keeps the tensor pipe busy through this CTA's pipeline fill, epilogue and store drain.
template <int HF>
struct Foot {
  static constexpr int kNumEpiWarps = HF ? 4 : 8;
  static constexpr int kThreads = (kEpiWarp0 + kNumEpiWarps) * 32;  // 256 | 384
  static constexpr int kSmemMax = HF ? 113 * 1024 : 227 * 1024;
  static constexpr int kTmemBudget = HF ? 256 : 512;
  static constexpr int kMinCtas = HF ? 2 : 1;
};
constexpr int kBarrierBytes = 512;
constexpr int kMaxStages = 8;

constexpr int kLnMaxNBlocks = 4;  // LayerNorm-fused epilogue: column blocks (CTAs or CTA pairs) per row block
constexpr int kLnChunk2Bytes = 32 * 32 * 2;  // second (bf16 / e4m3) output chunk of the LayerNorm epilogue

template <int EPI>
struct EpiTraits {
  // bytes of one staged 32x32 output chunk
  static constexpr int kElemBytes =
      (EPI == kEpiBiasResid || EPI == kEpiBiasAccum || EPI == kEpiBiasResidLN) ? 4
                                                                               : (EPI == kEpiBiasGeluE4m3 ? 1 : 2);
  static constexpr int kChunkBytes = 32 * 32 * kElemBytes;
  // kSlots staging slots per warp (2 = double-buffered); the LayerNorm epilogue adds two slots for its second
  // (bf16 / e4m3) output
  template <int kSlots>
  static constexpr int warp_bytes() { return kSlots * kChunkBytes + (EPI == kEpiBiasResidLN ? 2 * kLnChunk2Bytes : 0); }
  // LayerNorm epilogue: per-row (mean, M2) contributions of every column half of every column block,
  // double-buffered across tiles: [2][2 * kLnMaxNBlocks][128 rows] float2
  static constexpr int kLnStatsOnly = 2 * 2 * kLnMaxNBlocks * kRowsPerCta * 8;
  // this CTA's column slices of gamma | beta (with the bias too, <192, 2> would drop from 4 to 3 ring stages)
  static constexpr int kLnVecBytes = 2 * 192 * 4;
  static constexpr int kStatsBytes = EPI == kEpiBiasResidLN ? kLnStatsOnly + kLnVecBytes : 0;
};

template <int BLOCK_N, int CG, int EPI, int HF = 0>
struct TileCfg {
  using F = Foot<HF>;
  static constexpr int kNumEpiWarps = F::kNumEpiWarps;
  static constexpr int kColSplit = kNumEpiWarps / 4;  // epilogue warps sharing one TMEM lane quarter split the columns
  static constexpr int kBRows = BLOCK_N / CG;  // B rows staged by one CTA
  static constexpr int kBBytes = kBRows * kTileKBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStatsBytes = EpiTraits<EPI>::kStatsBytes;
  template <int kS>
  static constexpr int stages_with() {
    return (F::kSmemMax - 1024 - kNumEpiWarps * EpiTraits<EPI>::template warp_bytes<kS>() - kStatsBytes - kBarrierBytes) /
           kStageBytes;
  }
  // operand-ring depth comes first: a half-footprint CTA gives up the second staging slot when that buys a stage
  static constexpr int kSlots = (HF && stages_with<1>() > stages_with<2>() && stages_with<2>() < kMaxStages) ? 1 : 2;
  static constexpr int kWarpBytes = EpiTraits<EPI>::template warp_bytes<kSlots>();
  static constexpr int kStagingBytes = kNumEpiWarps * kWarpBytes;
  static constexpr int kStagesRaw = stages_with<kSlots>();
  static constexpr int kStages = kStagesRaw > kMaxStages ? kMaxStages : kStagesRaw;
  // two accumulator buffers (the epilogue of tile i overlaps the main loop of tile i+1) when the TMEM budget allows
  static constexpr int kAccBufs = 2 * BLOCK_N <= F::kTmemBudget ? 2 : 1;
  static constexpr int kAccCols = kAccBufs * BLOCK_N;
  static constexpr int kTmemCols = kAccCols <= 32 ? 32 : kAccCols <= 64 ? 64 : kAccCols <= 128 ? 128 : kAccCols <= 256 ? 256 : 512;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kStagingBytes + kStatsBytes + kBarrierBytes;
  // does the tile fit the footprint with a ring deep enough to cover the TMA latency?
  static constexpr bool kFits = kStages >= 3 && kTmemCols <= F::kTmemBudget && kSmemBytes <= F::kSmemMax;
  static_assert((BLOCK_N / kColSplit) % 32 == 0, "each epilogue warp must own whole 32-column chunks");
  static_assert(!(HF && EPI == kEpiBiasResidLN), "the LayerNorm-fused epilogue is a full-footprint kernel");
};

struct KernelArgs {
  int M, N, K;
  const float *bias;
  float acc_scale;
  float out_inv_scale;
  // Debug trace (tools/gemm_trace.py): when non-null, CTA `trace_cta` writes %globaltimer
  // stamps: [0] kernel entry, [1] after griddepcontrol.wait (MMA thread), [2+kb] full
  // barrier of k-block kb of its first tile observed by the MMA thread (kb < 60),
  // [62] accumulator of the first tile ready (epilogue warp 4), [63] that warp's last store
  // of the first tile issued, [64] kernel exit, [66 + 8c + i] chunk c < 4 of that warp:
  // i = 0 loop top, 1 tcgen05.ld + bias loads issued, 2 staging slot free, 3 tcgen05.ld
  // landed, 4 math + st.shared issued, 5 proxy fence done, 6 TMA store issued, 7 iteration end.
  unsigned long long *trace;
  int trace_cta;
  // Experiments only (spu_debug_gemm_mode): 1 = producer skips the TMA loads (pure MMA +
  // barrier rate on whatever smem holds), 2 = issuer skips the MMAs (pure operand feed rate).
  int dbg;
  // LayerNorm-fused epilogue (kEpiBiasResidLN): y = LN(acc*s + bias + resid) * gamma + beta over the N columns
  const float *gamma, *beta;
  float ln_eps;
  int ln_write_f32;  // y -> tmap_out (fp32)
  int ln_out2;       // 0 none, 1 bf16 y -> tmap_out2, 2 e4m3 (y * out_inv_scale) -> tmap_out2
};

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                             uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(b), "f"(a));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(d), "f"(c));
  return static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
}

template <int BLOCK_N, int CG, int EPI, int IN_FP8, int HF>
__global__ void __launch_bounds__(Foot<HF>::kThreads, Foot<HF>::kMinCtas)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_out,
                    const __grid_constant__ CUtensorMap tmap_resid,
                    const __grid_constant__ CUtensorMap tmap_out2, const KernelArgs args) {
  using Cfg = TileCfg<BLOCK_N, CG, EPI, HF>;
  using Epi = EpiTraits<EPI>;
  static_assert(Cfg::kFits, "tile exceeds the CTA footprint or leaves too shallow a pipeline");
  constexpr int kNumEpiWarps = Cfg::kNumEpiWarps;
  constexpr int kSlots = Cfg::kSlots;
  constexpr int kAccBufs = Cfg::kAccBufs;
  constexpr int kStages = Cfg::kStages;
  constexpr int kStageBytes = Cfg::kStageBytes;
  constexpr uint32_t kIdesc = make_idesc(IN_FP8 ? 0u : 1u, kRowsPerCta * CG, BLOCK_N);
  constexpr int kElemsPerKBlock = IN_FP8 ? 128 : 64;
  constexpr int kTileM = kRowsPerCta * CG;
  constexpr int kColsPerWarp = BLOCK_N / Cfg::kColSplit;
  constexpr int kChunksPerWarp = kColsPerWarp / 32;
  // LayerNorm-fused epilogue: the CG * (N / BLOCK_N) CTAs holding one row block's columns form a cluster
  constexpr bool kLN = EPI == kEpiBiasResidLN;
  constexpr bool kClustered = CG == 2 || kLN;

  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles need 1024-byte aligned bases.
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t *smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t *staging = smem + kStages * kStageBytes;
  uint8_t *stats_base = staging + Cfg::kStagingBytes;
  uint64_t *full_bar = reinterpret_cast<uint64_t *>(stats_base + Cfg::kStatsBytes);
  uint64_t *empty_bar = full_bar + kMaxStages;
  uint64_t *tmem_full_bar = empty_bar + kMaxStages;
  uint64_t *tmem_empty_bar = tmem_full_bar + 2;
  uint64_t *resid_bar = tmem_empty_bar + 2;  // [kMaxEpiWarps][2]
  uint64_t *stats_bar = resid_bar + 2 * kMaxEpiWarps;  // [2], LayerNorm epilogue only
  uint32_t *tmem_ptr_smem = reinterpret_cast<uint32_t *>(stats_bar + 2);

  // warp-uniform by construction AND provably so for the compiler (shuffle from lane 0)
  const int warp_idx = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const int n_blocks = args.N / BLOCK_N;
  const int m_blocks = (args.M + kTileM - 1) / kTileM;
  const uint32_t crank = kClustered ? cluster_ctarank() : 0u;
  const uint32_t cta_rank = crank & static_cast<uint32_t>(CG - 1);      // rank inside the CTA pair
  const uint32_t pair_base = crank & ~static_cast<uint32_t>(CG - 1);    // cluster rank of the pair's leader
  // A "unit" owns tiles: a CTA (pair) striding over all (m, n) tiles, or -- LayerNorm epilogue -- a whole
  // cluster striding over row blocks while each of its CTAs (pairs) keeps ONE column block.
  const int unit_ctas = kLN ? CG * n_blocks : CG;
  const int unit = static_cast<int>(blockIdx.x) / unit_ctas;
  const int num_units = static_cast<int>(gridDim.x) / unit_ctas;
  const int ln_n_blk = static_cast<int>(crank) / CG;
  const int num_tiles = kLN ? m_blocks : n_blocks * m_blocks;
  const int num_k_blocks = args.K / kElemsPerKBlock;
  auto tile_m = [&](int tile) { return kLN ? tile : tile / n_blocks; };
  auto tile_n = [&](int tile) { return kLN ? ln_n_blk : tile % n_blocks; };

  // experiments (spu_debug_gemm_mode bits 8..11): use fewer ring stages than fit
  const int n_stages = (args.dbg >> 8) > 0 && (args.dbg >> 8) < kStages ? (args.dbg >> 8) : kStages;
  const int dbg_mode = args.dbg & 0xff;
  // trace_cta = cta index | (local tile whose epilogue is stamped << 16)
  const bool tracing = args.trace != nullptr && static_cast<int>(blockIdx.x) == (args.trace_cta & 0xffff);
  const int kTraceTile = args.trace_cta >> 16;
  if (tracing && threadIdx.x == 0) args.trace[0] = global_timer_ns();
  if (warp_idx == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    tma_prefetch_desc(&tmap_out);
    if constexpr (EPI == kEpiBiasResid || kLN) tma_prefetch_desc(&tmap_resid);
    if constexpr (kLN) tma_prefetch_desc(&tmap_out2);
  }
  if (warp_idx == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full_bar[s], 1);
      mbar_init(&tmem_empty_bar[s], CG * kNumEpiWarps);  // one arrive per epilogue warp of the pair
    }
    for (int s = 0; s < 2 * kNumEpiWarps; ++s) mbar_init(&resid_bar[s], 1);
    for (int s = 0; s < 2; ++s) mbar_init(&stats_bar[s], 1);
    fence_barrier_init();
  }
  if (warp_idx == 2) {
    if constexpr (CG == 2) {
      tmem_alloc_pair(tmem_ptr_smem, Cfg::kTmemCols);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
      tmem_relinquish();
    }
  }
  tcgen05_fence_before();
  if constexpr (kClustered) cluster_sync_all(); else __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr_smem, 0);
  const uint32_t smem_base_u32 = smem_u32(smem);

  // Programmatic dependent launch: everything above, plus the producer's first WEIGHT-tile
  // loads below, overlaps the previous kernel's tail; memory written by that kernel (the
  // activations, the residual) is only touched after griddepcontrol.wait.  Dependents are
  // released at once: their CTAs start their own prologue as soon as an SM frees up.
  pdl_launch_dependents();

  if (warp_idx == 0) {
    // ===================== TMA producer (both CTAs of a pair) =====================
    // Whole warp in the loop (uniform control flow -> uniform-register TMA operands), one
    // elected lane issues.
    const bool no_tma = dbg_mode == 1 || dbg_mode == 3;
    auto issue_expect = [&](int stage) {
      if (no_tma) {
        if (cta_rank == 0) mbar_arrive(&full_bar[stage]);
        return;
      }
      if constexpr (CG == 2) {
        if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * kStageBytes);
      } else {
        mbar_arrive_expect_tx(&full_bar[stage], kStageBytes);
      }
    };
    auto load_a = [&](int stage, int kb, int a_row) {
      if (no_tma) return;
      uint8_t *sa = smem + stage * kStageBytes;
      if constexpr (CG == 2) tma_load_2d_pair(sa, &tmap_a, &full_bar[stage], kb * kElemsPerKBlock, a_row);
      else tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kElemsPerKBlock, a_row);
    };
    auto load_b = [&](int stage, int kb, int b_row) {
      if (no_tma) return;
      uint8_t *sb = smem + stage * kStageBytes + kABytes;
      if constexpr (CG == 2) tma_load_2d_pair(sb, &tmap_b, &full_bar[stage], kb * kElemsPerKBlock, b_row);
      else tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * kElemsPerKBlock, b_row);
    };
    // weights of the first tile's first stages: independent of the previous kernel
    const int n_pre = unit < num_tiles ? (num_k_blocks < n_stages ? num_k_blocks : n_stages) : 0;
    {
      const int n_blk = tile_n(unit);
      const int b_row = n_blk * BLOCK_N + static_cast<int>(cta_rank) * Cfg::kBRows;
      if (elect_one()) {
        for (int kb = 0; kb < n_pre; ++kb) {
          issue_expect(kb);
          load_b(kb, kb, b_row);
        }
      }
      __syncwarp();
    }
    pdl_wait_prior_grid();
    int stage = 0;
    uint32_t phase = 0;
    bool first = true;
    for (int tile = unit; tile < num_tiles; tile += num_units) {
      const int m_blk = tile_m(tile);
      const int n_blk = tile_n(tile);
      const int a_row = m_blk * kTileM + static_cast<int>(cta_rank) * kRowsPerCta;
      const int b_row = n_blk * BLOCK_N + static_cast<int>(cta_rank) * Cfg::kBRows;
      for (int kb = 0; kb < num_k_blocks; ++kb) {
        if (first && kb < n_pre) {
          if (elect_one()) load_a(stage, kb, a_row);  // barrier armed and B already in flight
        } else {
          mbar_wait_warp(&empty_bar[stage], phase ^ 1u);
          if (elect_one()) {
            issue_expect(stage);
            load_a(stage, kb, a_row);
            load_b(stage, kb, b_row);
          }
        }
        __syncwarp();
        if (++stage == n_stages) {
          stage = 0;
          phase ^= 1u;
        }
      }
      first = false;
    }
  } else if (warp_idx == 1) {
    // ===================== MMA issuer (leader CTA) =====================
    // The WHOLE warp runs the loop (warp-uniform control flow and operands, so ptxas keeps
    // descriptors in uniform registers); one elected lane issues the MMAs and commits.
    pdl_wait_prior_grid();
    if (tracing && lane == 0) args.trace[1] = global_timer_ns();
    if (cta_rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int local_tile = 0;
      const int dbg = dbg_mode;
      for (int tile = unit; tile < num_tiles; tile += num_units, ++local_tile) {
        const int acc = local_tile % kAccBufs;
        const uint32_t acc_phase = (local_tile / kAccBufs) & 1u;
        mbar_wait_warp(&tmem_empty_bar[acc], acc_phase ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BLOCK_N);
        for (int kb = 0; kb < num_k_blocks; ++kb) {
          mbar_wait_warp(&full_bar[stage], phase);
          tcgen05_fence_after();
          if (tracing && lane == 0 && local_tile == 0 && kb < 60) args.trace[2 + kb] = global_timer_ns();
          const uint32_t sa = smem_base_u32 + static_cast<uint32_t>(stage * kStageBytes);
          const uint64_t desc_a = make_sw128_kmajor_desc(sa);
          const uint64_t desc_b = make_sw128_kmajor_desc(sa + kABytes);
          if (elect_one()) {
            auto issue_kblock = [&](uint32_t first_accumulates) {
#pragma unroll
              for (int k = 0; k < kTileKBytes / kUmmaKBytes; ++k) {
                // advancing 32 bytes along K inside the 128-byte swizzle row = +2 in the
                // (addr >> 4) start-address field
                const uint64_t adv = static_cast<uint64_t>((k * kUmmaKBytes) >> 4);
                const uint32_t accumulate = k != 0 ? 1u : first_accumulates;
                if constexpr (CG == 2) {
                  if constexpr (IN_FP8) umma_f8_ss_pair(tmem_d, desc_a + adv, desc_b + adv, kIdesc, accumulate);
                  else umma_f16_ss_pair(tmem_d, desc_a + adv, desc_b + adv, kIdesc, accumulate);
                } else {
                  if constexpr (IN_FP8) umma_f8_ss(tmem_d, desc_a + adv, desc_b + adv, kIdesc, accumulate);
                  else umma_f16_ss(tmem_d, desc_a + adv, desc_b + adv, kIdesc, accumulate);
                }
              }
            };
            if (dbg != 2) issue_kblock(kb != 0);
            if (dbg == 3) issue_kblock(1u);  // experiments: twice the MMAs per barrier round
            // smem stage reusable (in both CTAs) once these MMAs retire
            if constexpr (CG == 2) umma_commit_pair(&empty_bar[stage], pair_base); else umma_commit(&empty_bar[stage]);
            if (kb == num_k_blocks - 1) {
              if constexpr (CG == 2) umma_commit_pair(&tmem_full_bar[acc], pair_base);
              else umma_commit(&tmem_full_bar[acc]);
            }
          }
          __syncwarp();
          if (++stage == n_stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp_idx >= kEpiWarp0) {
    // ============ epilogue with fused residual + LayerNorm (all CTAs of the cluster) ============
    // Thread (warp w, lane l) owns row q*32+l of this CTA's 128-row slab and kVals = BLOCK_N/2 consecutive
    // columns of it.  Pass 1: x = acc*s + bias + resid for those columns, kept in registers (the TMEM buffer is
    // handed back at once); local (mean, M2) of the kVals values.  Exchange: the pair is sent with st.async to
    // every CTA of the cluster that holds the same rows (one per column block), which credits that CTA's
    // statistics mbarrier; Chan's formula merges the 2 * n_blocks contributions.  Pass 2: normalise from
    // registers, stage and TMA-store y as fp32 (next residual) and / or bf16 / e4m3 (next GEMM operand).
    if constexpr (kLN) {
      constexpr int kVals = BLOCK_N / 2;
      constexpr int kChunks = kVals / 32;
      const int e = warp_idx - kEpiWarp0;
      const int q = warp_idx & 3;
      const int half = e >> 2;
      uint8_t *wstage = staging + e * Cfg::kWarpBytes;  // [2][4096] fp32 slots, then [2][2048] second-output slots
      const uint32_t stg = smem_u32(wstage);
      const uint32_t stg2 = stg + 2 * Epi::kChunkBytes;
      uint64_t *my_resid_bar = resid_bar + 2 * e;
      const int n_contrib = 2 * n_blocks;
      const int my_contrib = 2 * ln_n_blk + half;
      const int row_in_cta = q * 32 + lane;
      const uint32_t stats_u32 = smem_u32(stats_base);
      // gamma | beta of this CTA's BLOCK_N columns -> shared memory once (weights: independent of the previous
      // kernel, so this runs under its tail), read back as warp-wide broadcasts in pass 2; the bias slice of this
      // warp is pulled into L1 now and read with broadcast loads in pass 1
      float *vecs = reinterpret_cast<float *>(stats_base + Epi::kLnStatsOnly);
      {
        const int t = static_cast<int>(threadIdx.x) - kEpiWarp0 * 32;
        const int c0 = ln_n_blk * BLOCK_N;
        if (t < BLOCK_N) {
          vecs[t] = __ldg(args.gamma + c0 + t);
          vecs[BLOCK_N + t] = __ldg(args.beta + c0 + t);
        }
        if (lane < kVals / 32)
          asm volatile("prefetch.global.L1 [%0];" ::"l"(args.bias + c0 + half * kVals + lane * 32));
        asm volatile("bar.sync 1, %0;" ::"n"(kNumEpiWarps * 32) : "memory");
      }
      const uint32_t vec_u32 = smem_u32(vecs) + static_cast<uint32_t>(half * kVals * 4);
      pdl_wait_prior_grid();
      // residual slot of chunk c: the two fp32 slots, then (third chunk) the 4 KiB of the second-output slots,
      // which pass 1 does not use -- every residual chunk of the tile is in flight before the main loop ends
      auto resid_slot = [&](int c) { return c < 2 ? c * Epi::kChunkBytes : 2 * Epi::kChunkBytes; };
      static_assert(kChunks <= 3, "three residual slots");
      uint32_t resid_par = 0u;
      int local_tile = 0;
      for (int tile = unit; tile < num_tiles; tile += num_units, ++local_tile) {
        const int acc = local_tile & 1;
        const uint32_t acc_phase = (local_tile >> 1) & 1u;
        const int buf = local_tile & 1;
        const int row0 = tile * kTileM + static_cast<int>(cta_rank) * kRowsPerCta + q * 32;
        const int colw = ln_n_blk * BLOCK_N + half * kVals;
        if (e == 0 && lane == 0)  // this tile's statistics: 8 bytes per row from every contribution
          mbar_arrive_expect_tx(&stats_bar[buf], static_cast<uint32_t>(n_contrib * kRowsPerCta * 8));
        if (lane == 0) {
          tma_store_wait_read<0>();  // the previous tile's stores have read the staging slots
          mbar_arrive_expect_tx(&my_resid_bar[0], kChunks * Epi::kChunkBytes);
#pragma unroll
          for (int c = 0; c < kChunks; ++c)
            tma_load_2d(wstage + resid_slot(c), &tmap_resid, &my_resid_bar[0], colw + c * 32, row0);
        }
        __syncwarp();
        mbar_wait(&tmem_full_bar[acc], acc_phase);
        tcgen05_fence_after();
        const bool tr = tracing && local_tile == kTraceTile && e == 0 && lane == 0;
        if (tr) args.trace[62] = global_timer_ns();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                               static_cast<uint32_t>(acc * BLOCK_N + half * kVals);
        uint32_t v[kChunks][32];
#pragma unroll
        for (int c = 0; c < kChunks; ++c) tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>(c * 32), v[c]);
        tmem_ld_wait();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) {  // accumulator is in registers: the MMA issuer may start the next tile into it
          if constexpr (CG == 2) mbar_arrive_cluster(&tmem_empty_bar[acc], pair_base);
          else mbar_arrive(&tmem_empty_bar[acc]);
        }
        if (tr) args.trace[70] = global_timer_ns();
        // ---- pass 1 ----
        float sum = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
          if (c == 0) {
            mbar_wait(&my_resid_bar[0], resid_par);
            resid_par ^= 1u;
          }
          const uint32_t rowa = stg + static_cast<uint32_t>(resid_slot(c)) + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 bj = __ldg(reinterpret_cast<const float4 *>(args.bias + colw + c * 32) + j);
            const float4 r = ld_shared_f4(rowa + (static_cast<uint32_t>(j ^ (lane & 7)) << 4));
            const float x0 = fmaf(__uint_as_float(v[c][4 * j + 0]), args.acc_scale, bj.x) + r.x;
            const float x1 = fmaf(__uint_as_float(v[c][4 * j + 1]), args.acc_scale, bj.y) + r.y;
            const float x2 = fmaf(__uint_as_float(v[c][4 * j + 2]), args.acc_scale, bj.z) + r.z;
            const float x3 = fmaf(__uint_as_float(v[c][4 * j + 3]), args.acc_scale, bj.w) + r.w;
            v[c][4 * j + 0] = __float_as_uint(x0);
            v[c][4 * j + 1] = __float_as_uint(x1);
            v[c][4 * j + 2] = __float_as_uint(x2);
            v[c][4 * j + 3] = __float_as_uint(x3);
            sum += (x0 + x1) + (x2 + x3);
          }
        }
        const float mean_i = sum * (1.f / static_cast<float>(kVals));
        float m2_i = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c)
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float d = __uint_as_float(v[c][j]) - mean_i;
            m2_i = fmaf(d, d, m2_i);
          }
        if (tr) args.trace[71] = global_timer_ns();
        // ---- exchange ----
        {
          const uint32_t slot_addr =
              stats_u32 + static_cast<uint32_t>(((buf * 2 * kLnMaxNBlocks + my_contrib) * kRowsPerCta + row_in_cta) * 8);
          const uint32_t bar_addr = smem_u32(&stats_bar[buf]);
          for (int nb = 0; nb < n_blocks; ++nb) {
            const uint32_t dst = static_cast<uint32_t>(nb * CG) + cta_rank;
            st_async_f32x2(mapa_u32(slot_addr, dst), mean_i, m2_i, mapa_u32(bar_addr, dst));
          }
        }
        if (tr) args.trace[72] = global_timer_ns();
        mbar_wait_warp(&stats_bar[buf], (local_tile >> 1) & 1u);
        if (tr) args.trace[73] = global_timer_ns();
        float mean = 0.f, rstd;
        {
          const float2 *st = reinterpret_cast<const float2 *>(stats_base) +
                             (buf * 2 * kLnMaxNBlocks) * kRowsPerCta + row_in_cta;
#pragma unroll
          for (int i = 0; i < 2 * kLnMaxNBlocks; ++i)
            if (i < n_contrib) mean += st[i * kRowsPerCta].x;
          mean *= 1.f / static_cast<float>(n_contrib);
          float m2 = 0.f;
#pragma unroll
          for (int i = 0; i < 2 * kLnMaxNBlocks; ++i)
            if (i < n_contrib) {  // re-read from shared memory rather than keep 16 registers live
              const float2 p = st[i * kRowsPerCta];
              const float d = p.x - mean;
              m2 += p.y + static_cast<float>(kVals) * d * d;
            }
          rstd = rsqrtf(m2 / static_cast<float>(args.N) + args.ln_eps);
        }
        // ---- pass 2 ----
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
          const int slot = c & 1;
          const int col0 = colw + c * 32;
          if (c >= 2) {  // the stores issued from this slot pair two chunks ago must have read them
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
          }
          float y[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 g = ld_shared_f4(vec_u32 + static_cast<uint32_t>((c * 32 + 4 * j) * 4));
            const float4 be = ld_shared_f4(vec_u32 + static_cast<uint32_t>((BLOCK_N + c * 32 + 4 * j) * 4));
            y[4 * j + 0] = fmaf((__uint_as_float(v[c][4 * j + 0]) - mean) * rstd, g.x, be.x);
            y[4 * j + 1] = fmaf((__uint_as_float(v[c][4 * j + 1]) - mean) * rstd, g.y, be.y);
            y[4 * j + 2] = fmaf((__uint_as_float(v[c][4 * j + 2]) - mean) * rstd, g.z, be.z);
            y[4 * j + 3] = fmaf((__uint_as_float(v[c][4 * j + 3]) - mean) * rstd, g.w, be.w);
          }
          if (args.ln_write_f32) {  // 32 rows x 128 B, SWIZZLE_128B
            const uint32_t rowa = stg + static_cast<uint32_t>(slot * Epi::kChunkBytes) + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
            for (int j = 0; j < 8; ++j)
              st_shared_v4(rowa + (static_cast<uint32_t>(j ^ (lane & 7)) << 4), __float_as_uint(y[4 * j]),
                           __float_as_uint(y[4 * j + 1]), __float_as_uint(y[4 * j + 2]), __float_as_uint(y[4 * j + 3]));
          }
          if (args.ln_out2 == 1) {  // bf16: 32 rows x 64 B, SWIZZLE_64B
            const uint32_t rowa = stg2 + static_cast<uint32_t>(slot * kLnChunk2Bytes) + static_cast<uint32_t>(lane) * 64u;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_v4(rowa + (static_cast<uint32_t>(j ^ ((lane >> 1) & 3)) << 4), pack_bf16(y[8 * j], y[8 * j + 1]),
                           pack_bf16(y[8 * j + 2], y[8 * j + 3]), pack_bf16(y[8 * j + 4], y[8 * j + 5]),
                           pack_bf16(y[8 * j + 6], y[8 * j + 7]));
          } else if (args.ln_out2 == 2) {  // e4m3: 32 rows x 32 B, no swizzle
            const float sc = args.out_inv_scale;
            const uint32_t rowa = stg2 + static_cast<uint32_t>(slot * kLnChunk2Bytes) + static_cast<uint32_t>(lane) * 32u;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint32_t w[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int o = 16 * j + 4 * i;
                w[i] = pack_e4m3x4(y[o] * sc, y[o + 1] * sc, y[o + 2] * sc, y[o + 3] * sc);
              }
              st_shared_v4(rowa + static_cast<uint32_t>(j << 4), w[0], w[1], w[2], w[3]);
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (args.ln_write_f32) tma_store_2d(&tmap_out, wstage + slot * Epi::kChunkBytes, col0, row0);
            if (args.ln_out2 != 0)
              tma_store_2d(&tmap_out2, wstage + 2 * Epi::kChunkBytes + slot * kLnChunk2Bytes, col0, row0);
            tma_store_commit();
          }
          __syncwarp();
          if (tr) args.trace[74 + c] = global_timer_ns();
        }
        if (tr) args.trace[63] = global_timer_ns();
      }
      if (lane == 0) tma_store_wait_read<0>();
      __syncwarp();
    } else {
    // ===================== epilogue (all CTAs) =====================
    pdl_wait_prior_grid();
    const int e = warp_idx - kEpiWarp0;
    const int q = warp_idx & 3;  // TMEM lane quarter this warp may read
    const int half = e >> 2;     // which part of the tile's columns (8 epilogue warps: halves; 4: all of them)
    const uint32_t stg = smem_u32(staging + e * Cfg::kWarpBytes);
    uint64_t *my_resid_bar = resid_bar + 2 * e;
    uint32_t chunk_ctr = 0;      // live chunks this warp has staged so far (selects slot + parity)
    int local_tile = 0;
    for (int tile = unit; tile < num_tiles; tile += num_units, ++local_tile) {
      const int m_blk = tile / n_blocks;
      const int n_blk = tile % n_blocks;
      const int acc = local_tile % kAccBufs;
      const uint32_t acc_phase = (local_tile / kAccBufs) & 1u;
      const int row0 = m_blk * kTileM + static_cast<int>(cta_rank) * kRowsPerCta + q * 32;
      const int colw = n_blk * BLOCK_N + half * kColsPerWarp;  // first column of this warp
      const bool slab_live = row0 < args.M;  // TMA clips partial slabs; dead ones are skipped
      // Bias: warp-uniform 128-bit loads (one broadcast each).  Chunk c+1's bias is requested
      // into the same registers right after chunk c's FMAs consumed them, so its L2 latency
      // hides behind the staging / fence / store of chunk c.
      auto load_bias = [&](float4 (&dst)[8], int col) {
        if (args.bias != nullptr) {
          const float4 *b4 = reinterpret_cast<const float4 *>(args.bias + col);
#pragma unroll
          for (int j = 0; j < 8; ++j) dst[j] = __ldg(b4 + j);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) dst[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      };
      float4 bv[8];
      load_bias(bv, colw);  // chunk 0, while the main loop still runs
      if constexpr (EPI == kEpiBiasResid) {
        // The first residual chunks of this tile (one per staging slot): fetched while the main loop still runs.
        if (lane == 0 && slab_live) {
          tma_store_wait_read<0>();  // every staging slot is free again
#pragma unroll
          for (int c = 0; c < (kChunksPerWarp < kSlots ? kChunksPerWarp : kSlots); ++c) {
            const uint32_t b = (chunk_ctr + c) % kSlots;
            mbar_arrive_expect_tx(&my_resid_bar[b], Epi::kChunkBytes);
            tma_load_2d(staging + e * Cfg::kWarpBytes + b * Epi::kChunkBytes, &tmap_resid,
                        &my_resid_bar[b], colw + c * 32, row0);
          }
        }
        __syncwarp();
      }
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tcgen05_fence_after();
      if (tracing && local_tile == kTraceTile && e == 0 && lane == 0) args.trace[62] = global_timer_ns();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                             static_cast<uint32_t>(acc * BLOCK_N + half * kColsPerWarp);
      // TMEM -> registers is software-pipelined: chunk c+1's tcgen05.ld is in flight while
      // chunk c is converted, staged and stored (fully unrolled: the two register sets are
      // named statically).
      uint32_t vv[2][32];
      tmem_ld_32x32b_x32(taddr, vv[0]);
#pragma unroll
      for (int c = 0; c < kChunksPerWarp; ++c) {
        uint32_t(&v)[32] = vv[c & 1];
        const int col0 = colw + c * 32;
        const uint32_t b = chunk_ctr % kSlots;
        const uint32_t sbuf = stg + b * Epi::kChunkBytes;
        const bool tr_chunk = tracing && local_tile == kTraceTile && e == 0 && lane == 0 && c < 4;
        if (tr_chunk) args.trace[66 + 8 * c] = global_timer_ns();
        if (tr_chunk) args.trace[67 + 8 * c] = global_timer_ns();
        if constexpr (EPI != kEpiBiasResid) {
          // the store issued from this slot kSlots chunks ago must have drained it
          if (lane == 0) tma_store_wait_read<kSlots - 1>();
          __syncwarp();
        }
        if (tr_chunk) args.trace[68 + 8 * c] = global_timer_ns();
        tmem_ld_wait();
        if (tr_chunk) args.trace[69 + 8 * c] = global_timer_ns();
        if (c + 1 < kChunksPerWarp) {
          tmem_ld_32x32b_x32(taddr + static_cast<uint32_t>((c + 1) * 32), vv[(c + 1) & 1]);
        } else {
          // accumulator fully read: hand the TMEM buffer back to the MMA issuer
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (CG == 2) mbar_arrive_cluster(&tmem_empty_bar[acc], pair_base);
            else mbar_arrive(&tmem_empty_bar[acc]);
          }
        }
        if (slab_live) {  // dead slab (rows >= M): nothing to convert or store
          float f[32];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            f[4 * j + 0] = fmaf(__uint_as_float(v[4 * j + 0]), args.acc_scale, bv[j].x);
            f[4 * j + 1] = fmaf(__uint_as_float(v[4 * j + 1]), args.acc_scale, bv[j].y);
            f[4 * j + 2] = fmaf(__uint_as_float(v[4 * j + 2]), args.acc_scale, bv[j].z);
            f[4 * j + 3] = fmaf(__uint_as_float(v[4 * j + 3]), args.acc_scale, bv[j].w);
          }
          if (c + 1 < kChunksPerWarp) load_bias(bv, col0 + 32);
          if constexpr (EPI == kEpiBiasGelu || EPI == kEpiBiasGeluE4m3) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = gelu_erf(f[j]);
          }
          if constexpr (EPI == kEpiBiasResid) {
            // 32 rows x 128 B, SWIZZLE_128B: 16-byte chunk j of row r sits at j ^ (r & 7)
            mbar_wait(&my_resid_bar[b], (chunk_ctr / kSlots) & 1u);
            const uint32_t rowa = sbuf + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t a = rowa + (static_cast<uint32_t>(j ^ (lane & 7)) << 4);
              const float4 r = ld_shared_f4(a);
              st_shared_v4(a, __float_as_uint(f[4 * j] + r.x), __float_as_uint(f[4 * j + 1] + r.y),
                           __float_as_uint(f[4 * j + 2] + r.z), __float_as_uint(f[4 * j + 3] + r.w));
            }
          } else if constexpr (EPI == kEpiBiasAccum) {
            // 32 rows x 128 B, SWIZZLE_128B; added to the output in place by the TMA reduce below
            const uint32_t rowa = sbuf + static_cast<uint32_t>(lane) * 128u;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t a = rowa + (static_cast<uint32_t>(j ^ (lane & 7)) << 4);
              st_shared_v4(a, __float_as_uint(f[4 * j]), __float_as_uint(f[4 * j + 1]), __float_as_uint(f[4 * j + 2]),
                           __float_as_uint(f[4 * j + 3]));
            }
          } else if constexpr (EPI == kEpiBiasGeluE4m3) {
            // 32 rows x 32 B, no swizzle
            const float sc = args.out_inv_scale;
            const uint32_t rowa = sbuf + static_cast<uint32_t>(lane) * 32u;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              uint32_t w[4];
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const int o = 16 * j + 4 * i;
                w[i] = pack_e4m3x4(f[o] * sc, f[o + 1] * sc, f[o + 2] * sc, f[o + 3] * sc);
              }
              st_shared_v4(rowa + static_cast<uint32_t>(j << 4), w[0], w[1], w[2], w[3]);
            }
          } else {
            // 32 rows x 64 B, SWIZZLE_64B: 16-byte chunk j of row r sits at j ^ ((r >> 1) & 3)
            const uint32_t rowa = sbuf + static_cast<uint32_t>(lane) * 64u;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a = rowa + (static_cast<uint32_t>(j ^ ((lane >> 1) & 3)) << 4);
              st_shared_v4(a, pack_bf16(f[8 * j], f[8 * j + 1]), pack_bf16(f[8 * j + 2], f[8 * j + 3]),
                           pack_bf16(f[8 * j + 4], f[8 * j + 5]), pack_bf16(f[8 * j + 6], f[8 * j + 7]));
            }
          }
          if (tr_chunk) args.trace[70 + 8 * c] = global_timer_ns();
          fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA engine
          __syncwarp();
          if (tr_chunk) args.trace[71 + 8 * c] = global_timer_ns();
          if (lane == 0) {
            if constexpr (EPI == kEpiBiasAccum)
              tma_reduce_add_2d(&tmap_out, staging + e * Cfg::kWarpBytes + b * Epi::kChunkBytes, col0, row0);
            else
              tma_store_2d(&tmap_out, staging + e * Cfg::kWarpBytes + b * Epi::kChunkBytes, col0, row0);
            tma_store_commit();
            if constexpr (EPI == kEpiBiasResid) {
              if (c + kSlots < kChunksPerWarp) {
                // refill this slot with the residual chunk kSlots ahead once the store has read it
                tma_store_wait_read<0>();
                mbar_arrive_expect_tx(&my_resid_bar[b], Epi::kChunkBytes);
                tma_load_2d(staging + e * Cfg::kWarpBytes + b * Epi::kChunkBytes, &tmap_resid,
                            &my_resid_bar[b], col0 + 32 * kSlots, row0);
              }
            }
          }
          __syncwarp();
          if (tr_chunk) args.trace[72 + 8 * c] = global_timer_ns();
          ++chunk_ctr;  // counts live chunks only: slot = ctr % kSlots, residual-barrier parity = ctr / kSlots
        }
      }
      if (tracing && local_tile == kTraceTile && e == 0 && lane == 0) args.trace[63] = global_timer_ns();
    }
    // the staging slots must outlive the bulk stores' smem reads; their global writes are
    // complete (and visible) at grid completion like any other store
    if (lane == 0) tma_store_wait_read<0>();
    __syncwarp();
    }  // !kLN
  }

  tcgen05_fence_before();
  if constexpr (kClustered) cluster_sync_all(); else __syncthreads();
  if (tracing && threadIdx.x == 0) args.trace[64] = global_timer_ns();
  if (warp_idx == 2) {
    tcgen05_fence_after();
    if constexpr (CG == 2) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
    else tmem_dealloc(tmem_base, Cfg::kTmemCols);
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

// Row-major [rows, cols] matrix of `esize`-byte elements, box = box_cols x box_rows.
spu_status make_tmap_2d(CUtensorMap *map, const void *base, int rows, int cols, int esize,
                        CUtensorMapDataType dtype, int box_cols, int box_rows,
                        CUtensorMapSwizzle swizzle) {
  EncodeTiledFn enc = get_encode_fn();
  SPU_REQUIRE(enc != nullptr, SPU_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(cols) * esize};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estride[2] = {1, 1};
  const CUresult r = enc(map, dtype, 2, const_cast<void *>(base), gdim, gstride, box, estride,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
    return SPU_ERR_CUDA;
  }
  return SPU_OK;
}

spu_status make_kmajor_tmap(CUtensorMap *map, const void *base, int rows, int k_elems, int box_rows,
                            GemmIn in) {
  const int esize = in == GemmIn::FP8_E4M3 ? 1 : 2;
  return make_tmap_2d(map, base, rows, k_elems, esize,
                      in == GemmIn::FP8_E4M3 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                             : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                      kTileKBytes / esize, box_rows, CU_TENSOR_MAP_SWIZZLE_128B);
}

// 32x32 output chunk map matching the staging layout of the epilogue kind.
spu_status make_out_tmap(CUtensorMap *map, const void *base, int rows, int cols, int epilogue) {
  switch (epilogue) {
    case kEpiBiasResid:
    case kEpiBiasAccum:
      return make_tmap_2d(map, base, rows, cols, 4, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 32, 32,
                          CU_TENSOR_MAP_SWIZZLE_128B);
    case kEpiBiasGeluE4m3:
      return make_tmap_2d(map, base, rows, cols, 1, CU_TENSOR_MAP_DATA_TYPE_UINT8, 32, 32,
                          CU_TENSOR_MAP_SWIZZLE_NONE);
    default:
      return make_tmap_2d(map, base, rows, cols, 2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 32, 32,
                          CU_TENSOR_MAP_SWIZZLE_64B);
  }
}

using LaunchMaps = GemmLaunchMaps;

template <int BLOCK_N, int CG, int EPI, int IN_FP8, int HF = 0>
spu_status launch_cfg(const GemmPlan &plan, const LaunchMaps &maps, const KernelArgs &a,
                      cudaStream_t stream) {
  using Cfg = TileCfg<BLOCK_N, CG, EPI, HF>;
  if constexpr (!Cfg::kFits) {
    set_error("gemm: this tile / epilogue does not fit the requested CTA footprint");
    return SPU_ERR_UNSUPPORTED;
  } else {
  auto kern = gemm_tcgen05_kernel<BLOCK_N, CG, EPI, IN_FP8, HF>;
  SPU_CUDA_TRY((ensure_dynamic_smem<gemm_tcgen05_kernel<BLOCK_N, CG, EPI, IN_FP8, HF>>(Cfg::kSmemBytes)));
  constexpr bool kLN = EPI == kEpiBiasResidLN;
  const int n_blocks = plan.N / BLOCK_N;
  const int m_blocks = (a.M + kRowsPerCta * CG - 1) / (kRowsPerCta * CG);
  const int cluster = kLN ? CG * n_blocks : CG;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(Foot<HF>::kThreads, 1, 1);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute attrs[2];
  int n_attr = 0;
  static const bool pdl = std::getenv("SPU_GEMM_NO_PDL") == nullptr;  // experiments only
  if (pdl) {
    attrs[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[n_attr].val.programmaticStreamSerializationAllowed = 1;
    ++n_attr;
  }
  if (cluster > 1 || kLN) {
    attrs[n_attr].id = cudaLaunchAttributeClusterDimension;
    attrs[n_attr].val.clusterDim.x = static_cast<unsigned>(cluster);
    attrs[n_attr].val.clusterDim.y = 1;
    attrs[n_attr].val.clusterDim.z = 1;
    ++n_attr;
  }
  cfg.attrs = attrs;
  cfg.numAttrs = n_attr;
  int grid;
  if constexpr (kLN) {
    // one cluster per row block; persistent over row blocks when more exist than clusters fit the device
    static int max_clusters[64][9] = {};  // [device][cluster size], 0 = not asked yet
    int &cached = max_clusters[dev & 63][cluster];
    int fit = __atomic_load_n(&cached, __ATOMIC_RELAXED);
    if (fit == 0) {
      cfg.gridDim = dim3(static_cast<unsigned>(cluster), 1, 1);
      SPU_CUDA_TRY(cudaOccupancyMaxActiveClusters(&fit, kern, &cfg));
      SPU_REQUIRE(fit > 0, SPU_ERR_UNSUPPORTED, "gemm: the LayerNorm-fused cluster does not fit this device");
      if (const char *env = std::getenv("SPU_LN_MAX_CLUSTERS")) {  // experiments: trust the hardware, not the query
        const int forced = std::atoi(env);
        if (forced > 0) fit = forced;
      }
      __atomic_store_n(&cached, fit, __ATOMIC_RELAXED);
    }
    grid = cluster * (m_blocks < fit ? m_blocks : fit);
  } else {
    static int sm_count[64] = {};
    int sms = __atomic_load_n(&sm_count[dev & 63], __ATOMIC_RELAXED);
    if (sms == 0) {
      sms = kNumSMsB200;
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      __atomic_store_n(&sm_count[dev & 63], sms, __ATOMIC_RELAXED);
    }
    const int tiles = n_blocks * m_blocks;
    const int units = sms / CG;
    grid = CG * (tiles < units ? tiles : units);
  }
  cfg.gridDim = dim3(static_cast<unsigned>(grid), 1, 1);
  SPU_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, maps.a, plan.tmap_b[BLOCK_N / CG / 32 - 1], maps.out,
                                  maps.resid, maps.out2, a));
  return SPU_OK;
  }
}

template <int BLOCK_N, int CG, int HF = 0>
spu_status launch_epi(const GemmPlan &plan, const LaunchMaps &maps, const KernelArgs &a,
                      cudaStream_t stream, bool inplace) {
  if (plan.epilogue == kEpiBiasResidLN) {
    // BLOCK_N / 2 values of a row live in registers per thread; a lone CTA stages all of B (3 stages need <= 128)
    if constexpr (HF == 0 && BLOCK_N <= 192 && (CG == 2 || BLOCK_N <= 128)) {
      return plan.in == GemmIn::FP8_E4M3 ? launch_cfg<BLOCK_N, CG, kEpiBiasResidLN, 1>(plan, maps, a, stream)
                                         : launch_cfg<BLOCK_N, CG, kEpiBiasResidLN, 0>(plan, maps, a, stream);
    } else {
      set_error("gemm: the LayerNorm-fused epilogue needs BLOCK_N <= 192 (<= 128 without a CTA pair), full footprint");
      return SPU_ERR_UNSUPPORTED;
    }
  }
  if (plan.in == GemmIn::FP8_E4M3) {
    switch (plan.epilogue) {
      case kEpiBias: return launch_cfg<BLOCK_N, CG, kEpiBias, 1, HF>(plan, maps, a, stream);
      case kEpiBiasResid:
        return inplace ? launch_cfg<BLOCK_N, CG, kEpiBiasAccum, 1, HF>(plan, maps, a, stream)
                       : launch_cfg<BLOCK_N, CG, kEpiBiasResid, 1, HF>(plan, maps, a, stream);
      case kEpiBiasGeluE4m3: return launch_cfg<BLOCK_N, CG, kEpiBiasGeluE4m3, 1, HF>(plan, maps, a, stream);
      case kEpiBiasGelu: return launch_cfg<BLOCK_N, CG, kEpiBiasGelu, 1, HF>(plan, maps, a, stream);
      default: set_error("fp8 gemm: unsupported epilogue"); return SPU_ERR_INVALID;
    }
  }
  switch (plan.epilogue) {
    case kEpiBias: return launch_cfg<BLOCK_N, CG, kEpiBias, 0, HF>(plan, maps, a, stream);
    case kEpiBiasGelu: return launch_cfg<BLOCK_N, CG, kEpiBiasGelu, 0, HF>(plan, maps, a, stream);
    case kEpiBiasGeluE4m3: return launch_cfg<BLOCK_N, CG, kEpiBiasGeluE4m3, 0, HF>(plan, maps, a, stream);
    case kEpiBiasResid:
      return inplace ? launch_cfg<BLOCK_N, CG, kEpiBiasAccum, 0, HF>(plan, maps, a, stream)
                     : launch_cfg<BLOCK_N, CG, kEpiBiasResid, 0, HF>(plan, maps, a, stream);
    default: set_error("bf16 gemm: unsupported epilogue"); return SPU_ERR_INVALID;
  }
}

template <int CG>
spu_status launch_bn(const GemmPlan &plan, int block_n, const LaunchMaps &maps, const KernelArgs &a,
                     cudaStream_t stream, bool inplace) {
  switch (block_n) {
    case 64: return launch_epi<64, CG>(plan, maps, a, stream, inplace);
    case 128: return launch_epi<128, CG>(plan, maps, a, stream, inplace);
    case 192: return launch_epi<192, CG>(plan, maps, a, stream, inplace);
    case 256: return launch_epi<256, CG>(plan, maps, a, stream, inplace);
    default: set_error("unsupported BLOCK_N"); return SPU_ERR_INVALID;
  }
}

// Half-footprint CTAs exist for CTA pairs only (a lone CTA staging all of B would be left with two ring stages).
spu_status launch_bn_half(const GemmPlan &plan, int block_n, const LaunchMaps &maps, const KernelArgs &a,
                          cudaStream_t stream, bool inplace) {
  switch (block_n) {
    case 128: return launch_epi<128, 2, 1>(plan, maps, a, stream, inplace);
    case 192: return launch_epi<192, 2, 1>(plan, maps, a, stream, inplace);
    case 256: return launch_epi<256, 2, 1>(plan, maps, a, stream, inplace);
    default: set_error("gemm: half-footprint tiles are 128, 192 or 256 columns wide"); return SPU_ERR_INVALID;
  }
}

constexpr int kBlockNCands[4] = {256, 192, 128, 64};

#ifdef SPU_DEBUG_HOOKS
// Measurement builds only (include/spu_b200_debug.h): process-wide switches every launch reads.
std::atomic<unsigned long long *> g_trace_buf{nullptr};  // see KernelArgs::trace
std::atomic<int> g_trace_cta{0};
std::atomic<int> g_dbg_mode{0};
#else
// Product build: the hooks do not exist; the kernel's trace / debug branches are fed constants.
constexpr unsigned long long *g_trace_buf = nullptr;
constexpr int g_trace_cta = 0;
constexpr int g_dbg_mode = 0;
#endif

}  // namespace

#ifdef SPU_DEBUG_HOOKS
void gemm_set_trace(unsigned long long *device_buf, int cta) {
  g_trace_buf = device_buf;
  g_trace_cta = cta;
}
void gemm_set_debug_mode(int mode) { g_dbg_mode = mode; }
#endif

// LayerNorm-fused epilogue: the widest column block whose count per row stays within the cluster.
static GemmConfig choose_ln_config(int M, int N) {
  for (int bn : {192, 128, 64})
    if (N % bn == 0 && N / bn <= kLnMaxNBlocks) return GemmConfig{(M > kRowsPerCta || bn > 128) ? 2 : 1, bn};
  return GemmConfig{0, 0};
}

bool gemm_ln_supported(int N) { return choose_ln_config(1024, N).block_n != 0; }

GemmConfig choose_config(int M, int N, int K, int esize) {
  // Cost model fitted to in-situ measurements on B200 (profiles/r05_gemm_issue_path.md):
  // one persistent round costs k_blocks x t_kb(cta_group, BLOCK_N); the last tile's epilogue
  // and the launch / pipeline-fill cost are paid once.  A CTA pair halves the B bytes each SM
  // stages and reads per MMA, and shared-memory bandwidth (TMA fill + MMA operand reads) is
  // what bounds the main loop, so pairs win at every size measured.
  static const double t_kb_ns[2][4] = {{560.0, 480.0, 390.0, 300.0},   // cta_group::1, BLOCK_N 256/192/128/64
                                       {350.0, 290.0, 270.0, 260.0}};  // cta_group::2
  GemmConfig best{0, 0};
  double best_cost = 1e300;
  const int k_blocks = static_cast<int>(static_cast<long long>(K) * esize / kTileKBytes);
  for (int cg = 2; cg >= 1; --cg) {
    const int tile_m = kRowsPerCta * cg;
    const int units = kNumSMsB200 / cg;
    const int m_blocks = (M + tile_m - 1) / tile_m;
    for (int i = 0; i < 4; ++i) {
      const int bn = kBlockNCands[i];
      if (N % bn != 0) continue;
      const int tiles = m_blocks * (N / bn);
      const int rounds = (tiles + units - 1) / units;
      const double tail_ns = 600.0 + 10.0 * bn;
      const double cost = rounds * k_blocks * t_kb_ns[cg - 1][i] + tail_ns + 2500.0;
      if (cost < best_cost) {
        best_cost = cost;
        best = GemmConfig{cg, bn};
      }
    }
  }
  return best;
}

int gemm_candidates(int N, int epilogue, GemmConfig *out, int cap) {
  int n = 0;
  if (epilogue == kEpiBiasResidLN) return 0;
  for (int cg = 2; cg >= 1; --cg)
    for (int bn : kBlockNCands)
      if (N % bn == 0 && n < cap) out[n++] = GemmConfig{cg, bn, 0};
  return n;
}

spu_status gemm_make_plan(GemmPlan *plan, const void *A, int rows_a, const void *W, int N, int K,
                          int epilogue, GemmIn in, int force_cg, int force_bn) {
  SPU_REQUIRE(plan && A && W, SPU_ERR_INVALID, "gemm: null argument");
  const int kelems = in == GemmIn::FP8_E4M3 ? 128 : 64;
  SPU_REQUIRE(rows_a > 0 && N > 0 && K > 0, SPU_ERR_INVALID, "gemm: empty shape");
  SPU_REQUIRE(K % kelems == 0, SPU_ERR_UNSUPPORTED, "gemm: K must be a multiple of the 128-byte K tile");
  SPU_REQUIRE(N % 64 == 0, SPU_ERR_UNSUPPORTED, "gemm: N must be a multiple of 64");
  SPU_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15u) == 0 && (reinterpret_cast<uintptr_t>(W) & 15u) == 0,
              SPU_ERR_INVALID, "gemm: operands must be 16-byte aligned");
  plan->A = A;
  plan->rows_a = rows_a;
  plan->N = N;
  plan->K = K;
  plan->epilogue = epilogue;
  plan->in = in;
  plan->force_cg = force_cg;
  plan->force_bn = force_bn;
  plan->tuned = nullptr;
  if (const char *env = std::getenv("SPU_GEMM_CFG")) {  // "cg:bn", experiments only
    int cg = 0, bn = 0;
    if (std::sscanf(env, "%d:%d", &cg, &bn) == 2 && cg >= 1 && cg <= 3 && bn > 0 && N % bn == 0) {
      plan->force_cg = cg;
      plan->force_bn = bn;
    }
  }
  if (plan->force_bn)
    SPU_REQUIRE(plan->force_cg >= 1 && plan->force_cg <= 3 && N % plan->force_bn == 0 &&
                    plan->force_bn % 64 == 0 && plan->force_bn <= 256 && (plan->force_cg != 3 || plan->force_bn >= 128),
                SPU_ERR_UNSUPPORTED, "gemm: forced tile does not fit N");
  if (epilogue == kEpiBiasResidLN) {
    SPU_REQUIRE(gemm_ln_supported(N), SPU_ERR_UNSUPPORTED,
                "gemm: no LayerNorm-fused tiling spans N with a cluster of <= 8 CTAs");
    if (plan->force_bn)
      SPU_REQUIRE(plan->force_bn <= 192 && N / plan->force_bn <= kLnMaxNBlocks, SPU_ERR_UNSUPPORTED,
                  "gemm: forced tile does not fit the LayerNorm-fused epilogue");
  }
  for (int i = 0; i < kNumBBoxes; ++i) {
    const int box_rows = 32 * (i + 1);
    plan->b_ok[i] = false;
    bool wanted = false;
    for (int bn : kBlockNCands)
      if (N % bn == 0 && (bn == box_rows || bn / 2 == box_rows)) wanted = true;  // lone CTA | CTA pair
    if (!wanted) continue;
    spu_status st = make_kmajor_tmap(&plan->tmap_b[i], W, N, K, box_rows, in);
    if (st != SPU_OK) return st;
    plan->b_ok[i] = true;
  }
  return SPU_OK;
}

spu_status gemm_launch(const GemmPlan &plan, int M, const GemmEpilogueArgs &args,
                       cudaStream_t stream) {
  SPU_REQUIRE(M > 0 && M <= plan.rows_a, SPU_ERR_CAPACITY, "gemm: M exceeds plan capacity");
  SPU_REQUIRE((reinterpret_cast<uintptr_t>(args.out) & 15u) == 0, SPU_ERR_INVALID,
              "gemm: output must be 16-byte aligned");
  const bool ln = plan.epilogue == kEpiBiasResidLN;
  SPU_REQUIRE(args.out != nullptr || (ln && args.out2 != nullptr), SPU_ERR_INVALID, "gemm: null output");
  if (plan.epilogue == kEpiBiasResid || ln)
    SPU_REQUIRE(args.resid != nullptr && (reinterpret_cast<uintptr_t>(args.resid) & 15u) == 0,
                SPU_ERR_INVALID, "gemm: residual epilogue needs a 16-byte aligned residual");
  if (ln)
    SPU_REQUIRE(args.gamma != nullptr && args.beta != nullptr && args.bias != nullptr &&
                    (reinterpret_cast<uintptr_t>(args.out2) & 15u) == 0,
                SPU_ERR_INVALID, "gemm: the LayerNorm-fused epilogue needs bias, gamma, beta and aligned outputs");
  const int esize = plan.in == GemmIn::FP8_E4M3 ? 1 : 2;
  const TunedTable::Entry *tuned =
      (!plan.force_bn && !ln && plan.tuned != nullptr)
          ? plan.tuned->find(plan.N, plan.K, plan.epilogue, static_cast<int>(plan.in), M)
          : nullptr;
  GemmConfig cfg = plan.force_bn ? GemmConfig{plan.force_cg == 3 ? 2 : plan.force_cg, plan.force_bn, plan.force_cg == 3 ? 1 : 0}
                   : ln          ? choose_ln_config(M, plan.N)
                   : tuned       ? tuned->cfg
                                 : choose_config(M, plan.N, plan.K, esize);
  SPU_REQUIRE(cfg.block_n > 0, SPU_ERR_UNSUPPORTED, "gemm: no BLOCK_N divides N");
  if (!plan.force_bn && !ln && !tuned) {
    static const int hf_env = [] { const char *e = std::getenv("SPU_GEMM_HF"); return e ? std::atoi(e) : 0; }();
    const int tiles_hf = ((M + 255) / 256) * (plan.N / cfg.block_n);
    const bool one_round = tiles_hf <= kNumSMsB200 / 2;  // every CTA pair owns a single tile
    if (cfg.cta_group == 2 && cfg.block_n >= 128 && (cfg.block_n < 256 || plan.epilogue != kEpiBiasResid) &&
        (hf_env == 1 || (hf_env == 2 && one_round) || (hf_env == 3 && one_round && plan.K * esize <= 2048)))
      cfg.half = 1;
  }
  SPU_REQUIRE(plan.b_ok[cfg.block_n / cfg.cta_group / 32 - 1], SPU_ERR_UNSUPPORTED,
              "gemm: no weight tensor map for this tile");
  // Activation-side maps carry the live row count: TMA zero-fills A rows >= M and
  // clips output rows >= M, so a dynamic batch never reads or writes stale rows.
  const LaunchMaps *cached = nullptr;
  for (const GemmMapCacheEntry &c : plan.map_cache)
    if (c.M == M && c.out == args.out && c.resid == args.resid && c.out2 == args.out2 && c.out2_e4m3 == args.out2_e4m3)
      cached = &c.maps;
  LaunchMaps fresh;
  if (cached == nullptr) {
    LaunchMaps &maps = fresh;
    spu_status st = make_kmajor_tmap(&maps.a, plan.A, M, plan.K, kRowsPerCta, plan.in);
    if (st != SPU_OK) return st;
    if (ln) {
      st = make_out_tmap(&maps.resid, args.resid, M, plan.N, kEpiBiasResid);
      if (st != SPU_OK) return st;
      if (args.out != nullptr) {
        st = make_out_tmap(&maps.out, args.out, M, plan.N, kEpiBiasResid);
        if (st != SPU_OK) return st;
      } else {
        maps.out = maps.resid;
      }
      if (args.out2 != nullptr) {
        st = make_out_tmap(&maps.out2, args.out2, M, plan.N, args.out2_e4m3 ? kEpiBiasGeluE4m3 : kEpiBias);
        if (st != SPU_OK) return st;
      } else {
        maps.out2 = maps.out;
      }
    } else {
      st = make_out_tmap(&maps.out, args.out, M, plan.N, plan.epilogue);
      if (st != SPU_OK) return st;
      if (plan.epilogue == kEpiBiasResid) {
        st = make_out_tmap(&maps.resid, args.resid, M, plan.N, plan.epilogue);
        if (st != SPU_OK) return st;
      } else {
        maps.resid = maps.out;
      }
      maps.out2 = maps.out;
    }
    GemmMapCacheEntry &slot = plan.map_cache[plan.map_cache_next++ % 4];
    slot.M = M;
    slot.out = args.out;
    slot.resid = args.resid;
    slot.out2 = args.out2;
    slot.out2_e4m3 = args.out2_e4m3;
    slot.maps = fresh;
    cached = &slot.maps;
  }
  const LaunchMaps &maps = *cached;
  KernelArgs a{M, plan.N, plan.K, args.bias, args.acc_scale, args.out_inv_scale,
               static_cast<unsigned long long *>(g_trace_buf), static_cast<int>(g_trace_cta), static_cast<int>(g_dbg_mode),
               args.gamma, args.beta, args.ln_eps, args.out != nullptr ? 1 : 0,
               args.out2 == nullptr ? 0 : (args.out2_e4m3 ? 2 : 1)};
  // residual == output: accumulate in place with a TMA reduce-add instead of load + add + store
  const bool inplace = plan.epilogue == kEpiBiasResid && args.resid == args.out;
  if (cfg.half) return launch_bn_half(plan, cfg.block_n, maps, a, stream, inplace);
  return cfg.cta_group == 2 ? launch_bn<2>(plan, cfg.block_n, maps, a, stream, inplace)
                            : launch_bn<1>(plan, cfg.block_n, maps, a, stream, inplace);
}

}  // namespace spu
