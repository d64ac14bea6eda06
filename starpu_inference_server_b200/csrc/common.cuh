// Shared device/host helpers for libspu_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <string>

#include "../../include/spu_b200.h"

namespace spu {

// ---- host-side error plumbing ------------------------------------------------
void set_error(const std::string &msg);

#define SPU_CUDA_TRY(expr)                                                         \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      ::spu::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));        \
      return SPU_ERR_CUDA;                                                         \
    }                                                                              \
  } while (0)

#define SPU_REQUIRE(cond, status, msg)                                             \
  do {                                                                             \
    if (!(cond)) {                                                                 \
      ::spu::set_error(std::string(msg) + " (" #cond ")");                         \
      return (status);                                                             \
    }                                                                              \
  } while (0)

static inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr int kNumSMsB200 = 148;

#ifdef __CUDACC__
// Launch helper for the row / attention kernels.  With SPU_ROW_PDL=1 they are launched with
// programmatic stream serialisation (they execute griddepcontrol.wait first thing, so that is
// safe); measured on B200 it LOSES 1-3 % at every batch size (profiles/r04_pdl_row_kernels.md:
// early-resident secondaries take registers from the multi-wave primaries), so the default
// is a plain launch and only the GEMMs use PDL.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  static const bool on = std::getenv("SPU_ROW_PDL") != nullptr;  // experiments only
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

#ifdef __CUDACC__
// Opt a kernel into > 48 KB of dynamic shared memory.  The attribute is PER DEVICE, and one process may drive
// several GPUs (spu_server with several device_ids): remember it per (kernel instantiation, device ordinal).
template <auto Kern>
inline cudaError_t ensure_dynamic_smem(int bytes) {
  static unsigned long long done = 0;  // bit d = set on device d (ordinals < 64)
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (__atomic_load_n(&done, __ATOMIC_ACQUIRE) & bit) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(Kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  // largest shared-memory carve-out: kernels sized for two CTAs per SM must not be cut to one by a smaller split
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(Kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (e == cudaSuccess) __atomic_fetch_or(&done, bit, __ATOMIC_RELEASE);
  return e;
}
#endif

#ifdef __CUDACC__
// ---- small device utilities ---------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t *>(&v);
}

// erf GELU as HF "gelu": x * Phi(x), written as max(x,0) - |x| * Phi(-|x|) with
// Phi(-u) = 2^p(u), p a degree-5 weighted least-squares fit of log2(Phi(-u)) on [0,8]
// whose negative leading coefficient keeps 2^p(u) -> 0 beyond the fit range.
// |gelu_fast - gelu_erf| <= 4.6e-6 over all x (checked against scipy.special.erf on
// [-1000,1000]; tests/test_gpu_stages.py pins it through the GEMM epilogue), i.e.
// far below one bf16 ulp of the result.  5 FFMA + 1 MUFU.EX2 + 2 ALU ops per
// element: a 128x256 epilogue tile stays well under that tile's MMA time.
__device__ __forceinline__ float gelu_erf(float x) {
  const float u = fabsf(x);
  float p = fmaf(-2.708492393e-04f, u, 5.526491906e-03f);
  p = fmaf(p, u, -4.766750708e-02f);
  p = fmaf(p, u, -4.647003412e-01f);
  p = fmaf(p, u, -1.148618102e+00f);
  p = fmaf(p, u, -1.000340343e+00f);
  float q;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(q) : "f"(p));
  return fmaf(-u, q, fmaxf(x, 0.f));
}

// ---- mbarrier -------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (-> launch error the host reports) instead of
// hanging the GPU.  The bound is wall-clock (2 s of %globaltimer), checked every 1024
// failed probes so the common path stays a bare try_wait loop.
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 1023u) == 0) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 2000000000ull) __trap();
    }
  }
}

// Warp-collective wait: lane 0 polls, the warp reconverges on __syncwarp.  One poller keeps
// the barrier's smem word free of 32-fold try_wait traffic (measured: the wake-up of a
// 32-lane poll is slower), and callers keep warp-uniform control flow after it.
__device__ __forceinline__ bool mbar_test_wait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_warp(uint64_t *bar, uint32_t parity) {
  if ((threadIdx.x & 31u) == 0) {
    // non-suspending probe: the two pipeline-critical pollers (TMA producer, MMA issuer)
    // must see a phase flip at once; try_wait may park the thread for a system time slice
    if (!mbar_test_wait(bar, parity)) {
      uint32_t spins = 0;
      uint64_t t0 = 0;
      while (!mbar_test_wait(bar, parity)) {
        if ((++spins & 4095u) == 0) {
          const uint64_t now = global_timer_ns();
          if (t0 == 0) t0 = now;
          else if (now - t0 > 2000000000ull) __trap();
        }
      }
    }
  }
  __syncwarp();
}

// ---- TMA ------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const void *tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tmap)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const void *tmap, uint64_t *bar,
                                            int32_t c0, int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void *tmap, const void *smem_src, int32_t c0,
                                             int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(tmap)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
// out[tile] += smem tile (fp32), performed by the L2 atomics unit: one add per element
__device__ __forceinline__ void tma_reduce_add_2d(const void *tmap, const void *smem_src, int32_t c0,
                                                  int32_t c1) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(tmap)),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- tcgen05 / TMEM ---------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// D[tmem] (+)= A[smem] * B[smem]; kind::f16 covers bf16 inputs with fp32 accumulate.
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// kind::f8f6f4: e4m3 x e4m3 -> fp32 (K = 32 per instruction).
__device__ __forceinline__ void umma_f8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread retire.
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                   smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp gets lane (base+i).
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
        "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
        "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
        "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

// ---- clusters / CTA pairs (cta_group::2) -------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Arrive on the mbarrier at the same smem offset in CTA `cta` of this cluster.  Default
// semantics (.release at CTA scope): a cluster-scope release costs a full memory fence
// (measured ~1.3 us per use in the epilogue) and is not needed -- what the arrive orders
// (tcgen05.ld results) is already in registers after tcgen05.wait::ld.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t *bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
// Address of the same shared-memory offset in CTA `cta` of this cluster.
__device__ __forceinline__ uint32_t mapa_u32(uint32_t smem_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta));
  return r;
}
// Store 8 bytes into a peer CTA's shared memory and credit them to that CTA's mbarrier (both addresses
// are shared::cluster addresses from mapa_u32).  No fence: the data is visible to whoever observes the
// barrier's phase completion.
__device__ __forceinline__ void st_async_f32x2(uint32_t remote_addr, float a, float b, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(
                   remote_addr),
               "f"(a), "f"(b), "r"(remote_bar)
               : "memory");
}
// TMA load issued by either CTA of a pair; the transaction bytes are credited to the
// barrier at the same smem offset in the pair's EVEN (leader) CTA.
__device__ __forceinline__ void tma_load_2d_pair(void *smem_dst, const void *tmap, uint64_t *bar,
                                                 int32_t c0, int32_t c1) {
  asm volatile(
      "{\n\t.reg .b32 rb, rk;\n\t"
      "mov.u32 rk, %%cluster_ctarank;\n\t"
      "and.b32 rk, rk, 0xfffffffe;\n\t"
      "mapa.shared::cluster.u32 rb, %2, rk;\n\t"
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [rb];\n\t}" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t *smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
               : "memory");
}
// Pair MMA (issued by one thread of the leader CTA): D rows 0-127 land in the leader's
// TMEM, rows 128-255 in the peer's; each CTA supplies its own A rows and half of B.
__device__ __forceinline__ void umma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                                 uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_f8_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Commit of the pair's MMAs, arriving on the barrier at this offset in BOTH CTAs.
__device__ __forceinline__ void umma_commit_pair(uint64_t *bar, uint32_t pair_base = 0) {
  const uint16_t mask = static_cast<uint16_t>(0x3u << pair_base);  // the pair's two CTAs inside the cluster
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
      "[%0], %1;" ::"r"(smem_u32(bar)),
      "h"(mask)
      : "memory");
}
// Programmatic dependent launch: wait for the previous kernel in the stream to finish
// (and its memory to be visible) / allow the next one to start its prologue.
__device__ __forceinline__ void pdl_wait_prior_grid() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Shared-memory matrix descriptor for a K-major operand tile laid out by TMA with
// SWIZZLE_128B: rows of 128 bytes, 8-row groups of 1024 bytes (SBO), LBO unused.
// Field layout: start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) | version=1 [46,48)
// | layout_type [61,64) with SWIZZLE_128B = 2 (PTX ISA "tcgen05 shared memory
// descriptor"; same encoding CUTLASS uses in cute/arch/mma_sm100_desc.hpp).
__device__ __forceinline__ uint64_t make_sw128_kmajor_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(0) << 16;               // LBO (ignored for swizzled K-major)
  d |= static_cast<uint64_t>(1024u >> 4) << 32;      // SBO = 8 rows * 128 B
  d |= static_cast<uint64_t>(1) << 46;               // descriptor version (Blackwell)
  d |= static_cast<uint64_t>(2) << 61;               // SWIZZLE_128B
  return d;
}

// Instruction descriptor (upper 32 bits of the "idesc" operand):
// c_format [4,6) F32=1 | a_format [7,10) | b_format [10,13) | a_major bit15 | b_major bit16
// | N>>3 [17,23) | M>>4 [24,29).   kind::f16: BF16 = 1.  kind::f8f6f4: E4M3 = 0.
__host__ __device__ constexpr uint32_t make_idesc(uint32_t ab_format, uint32_t umma_m,
                                                 uint32_t umma_n) {
  return (1u << 4) | (ab_format << 7) | (ab_format << 10) | (0u << 15) | (0u << 16) |
         ((umma_n >> 3) << 17) | ((umma_m >> 4) << 24);
}
#endif  // __CUDACC__

}  // namespace spu
