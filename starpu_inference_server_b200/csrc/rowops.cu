// HBM-bound row kernels of the hot path: embedding+LayerNorm+mask builder,
// residual LayerNorm, ragged gather/pad/mask, segmented gather/scatter copies, and
// the toy affine "model".  All are one-pass, 128-bit coalesced, warp-shuffle
// reductions, no shared-memory staging needed (no reuse across rows).
#include "replica.cuh"

namespace spu {
namespace {

constexpr int kRowThreads = 256;            // 8 warps = 8 rows per CTA
constexpr int kMaxChunksPerLane = 8;        // hidden <= 1024 (float4 chunks, 32 lanes)

struct RowStats {
  float mean, rstd;
};

// v holds this lane's float4 chunks of one row (chunk c -> lane c % 32).
__device__ __forceinline__ RowStats row_stats(const float4 (&v)[kMaxChunksPerLane], int nchunk_lane,
                                              int hidden, float eps) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(s) / static_cast<float>(hidden);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
  const float var = warp_sum(q) / static_cast<float>(hidden);
  return {mean, rsqrtf(var + eps)};
}

__device__ __forceinline__ uint32_t pack_e4m3x4(float a, float b, float c, float d) {
  uint16_t lo, hi;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(b), "f"(a));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(d), "f"(c));
  return static_cast<uint32_t>(lo) | (static_cast<uint32_t>(hi) << 16);
}

__device__ __forceinline__ void row_normalise_store(const float4 (&v)[kMaxChunksPerLane],
                                                    int nchunk_lane, RowStats st,
                                                    const float *__restrict__ gamma,
                                                    const float *__restrict__ beta,
                                                    float *out_f32, __nv_bfloat16 *out_bf16,
                                                    size_t row_off, uint8_t *out_q = nullptr,
                                                    float q_inv_scale = 1.f) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i) {
    if (i >= nchunk_lane) break;
    const int c = lane + i * 32;
    const float4 g = __ldg(reinterpret_cast<const float4 *>(gamma) + c);
    const float4 b = __ldg(reinterpret_cast<const float4 *>(beta) + c);
    float4 y;
    y.x = (v[i].x - st.mean) * st.rstd * g.x + b.x;
    y.y = (v[i].y - st.mean) * st.rstd * g.y + b.y;
    y.z = (v[i].z - st.mean) * st.rstd * g.z + b.z;
    y.w = (v[i].w - st.mean) * st.rstd * g.w + b.w;
    if (out_f32 != nullptr) reinterpret_cast<float4 *>(out_f32 + row_off)[c] = y;
    if (out_bf16 != nullptr)
      reinterpret_cast<uint2 *>(out_bf16 + row_off)[c] =
          make_uint2(pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
    if (out_q != nullptr)  // per-tensor scaled FP8 E4M3 operand of the next GEMM
      reinterpret_cast<uint32_t *>(out_q + row_off)[c] =
          pack_e4m3x4(y.x * q_inv_scale, y.y * q_inv_scale, y.z * q_inv_scale, y.w * q_inv_scale);
  }
}

// x = LN(word[ids] + pos[t] + type[0]); mask_bias = mask ? 0 : -1e30.
// Restates BertEmbeddings.forward + the additive mask HF builds (oracle/bert_oracle.py).
__global__ void __launch_bounds__(kRowThreads)
bert_embed_ln_kernel(const int64_t *__restrict__ ids, const int64_t *__restrict__ mask,
                     const float *__restrict__ word, const float *__restrict__ pos,
                     const float *__restrict__ type0, const float *__restrict__ gamma,
                     const float *__restrict__ beta, float eps, int vocab, float *x_f32,
                     __nv_bfloat16 *x_bf16, float *mask_bias, int rows, int seq, int hidden,
                     uint8_t *x_q, float q_inv_scale) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int row = blockIdx.x * (kRowThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunk = hidden >> 2;
  const int nchunk_lane = (nchunk - lane + 31) >> 5;
  int64_t id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);  // memory safety; torch would raise
  const int t = row % seq;
  const float4 *w4 = reinterpret_cast<const float4 *>(word + static_cast<size_t>(id) * hidden);
  const float4 *p4 = reinterpret_cast<const float4 *>(pos + static_cast<size_t>(t) * hidden);
  const float4 *t4 = reinterpret_cast<const float4 *>(type0);
  float4 v[kMaxChunksPerLane];
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i) {
    if (i < nchunk_lane) {
      const int c = lane + i * 32;
      const float4 a = __ldg(w4 + c), b = __ldg(p4 + c), d = __ldg(t4 + c);
      v[i] = make_float4(a.x + b.x + d.x, a.y + b.y + d.y, a.z + b.z + d.z, a.w + b.w + d.w);
    }
  }
  const RowStats st = row_stats(v, nchunk_lane, hidden, eps);
  row_normalise_store(v, nchunk_lane, st, gamma, beta, x_f32, x_bf16,
                      static_cast<size_t>(row) * hidden, x_q, q_inv_scale);
  if (lane == 0) mask_bias[row] = mask[row] != 0 ? 0.f : -1e30f;
}

// ---- packed (unpadded) sequences: rows of all sequences back to back, cu[b] = first row of sequence b ----
struct CuSeqlensArg {
  int32_t v[kMaxPackedBatch + 1];
};

// cu_seqlens travels as a kernel parameter (1 KB): no staging buffer whose reuse a pipelined worker could race
__global__ void write_cu_seqlens_kernel(const CuSeqlensArg cu, int n, int32_t *out) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = cu.v[i];
}

// sequence that owns packed row `row`: last b with cu[b] <= row
__device__ __forceinline__ int seq_of_row(const int32_t *__restrict__ cu, int batch, int row) {
  int lo = 0, hi = batch - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (__ldg(cu + mid) <= row) lo = mid;
    else hi = mid - 1;
  }
  return lo;
}

// Packed embedding: row r of the output is token t = r - cu[b] of sequence b; its id sits at ids[b*seq + t] of the
// padded [batch, seq] array the host gather produced.  No mask bias: packed attention needs none.
__global__ void __launch_bounds__(kRowThreads)
bert_embed_packed_kernel(const int64_t *__restrict__ ids, const int32_t *__restrict__ cu, int batch, int seq,
                         const float *__restrict__ word, const float *__restrict__ pos,
                         const float *__restrict__ type0, const float *__restrict__ gamma,
                         const float *__restrict__ beta, float eps, int vocab, float *x_f32, __nv_bfloat16 *x_bf16,
                         int rows, int hidden) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int row = blockIdx.x * (kRowThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunk = hidden >> 2;
  const int nchunk_lane = (nchunk - lane + 31) >> 5;
  const int b = seq_of_row(cu, batch, row);
  const int t = row - __ldg(cu + b);
  int64_t id = ids[static_cast<size_t>(b) * seq + t];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float4 *w4 = reinterpret_cast<const float4 *>(word + static_cast<size_t>(id) * hidden);
  const float4 *p4 = reinterpret_cast<const float4 *>(pos + static_cast<size_t>(t) * hidden);
  const float4 *t4 = reinterpret_cast<const float4 *>(type0);
  float4 v[kMaxChunksPerLane];
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i) {
    if (i < nchunk_lane) {
      const int c = lane + i * 32;
      const float4 a = __ldg(w4 + c), q = __ldg(p4 + c), d = __ldg(t4 + c);
      v[i] = make_float4(a.x + q.x + d.x, a.y + q.y + d.y, a.z + q.z + d.z, a.w + q.w + d.w);
    }
  }
  const RowStats st = row_stats(v, nchunk_lane, hidden, eps);
  row_normalise_store(v, nchunk_lane, st, gamma, beta, x_f32, x_bf16, static_cast<size_t>(row) * hidden);
}

// Final LayerNorm of the packed forward: reads packed row r, writes the PADDED output row b*seq + t; one warp per
// padded row, rows behind a sequence's length are zero-filled (the reference leaves whatever the model computes
// for [PAD] tokens there; nothing reads them).
__global__ void __launch_bounds__(kRowThreads)
layernorm_unpad_kernel(const float *__restrict__ x, const int32_t *__restrict__ cu, int batch, int seq,
                       const float *__restrict__ gamma, const float *__restrict__ beta, float eps, float *out_f32,
                       __nv_bfloat16 *out_bf16, int hidden) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int prow = blockIdx.x * (kRowThreads / 32) + (threadIdx.x >> 5);  // padded row
  if (prow >= batch * seq) return;
  const int lane = threadIdx.x & 31;
  const int nchunk = hidden >> 2;
  const int nchunk_lane = (nchunk - lane + 31) >> 5;
  const int b = prow / seq, t = prow - b * seq;
  const int start = __ldg(cu + b), len = __ldg(cu + b + 1) - start;
  const size_t off = static_cast<size_t>(prow) * hidden;
  if (t >= len) {
#pragma unroll
    for (int i = 0; i < kMaxChunksPerLane; ++i) {
      if (i >= nchunk_lane) break;
      const int c = lane + i * 32;
      if (out_f32 != nullptr) reinterpret_cast<float4 *>(out_f32 + off)[c] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (out_bf16 != nullptr) reinterpret_cast<uint2 *>(out_bf16 + off)[c] = make_uint2(0u, 0u);
    }
    return;
  }
  const size_t in_off = static_cast<size_t>(start + t) * hidden;
  float4 v[kMaxChunksPerLane];
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) v[i] = reinterpret_cast<const float4 *>(x + in_off)[lane + i * 32];
  const RowStats st = row_stats(v, nchunk_lane, hidden, eps);
  row_normalise_store(v, nchunk_lane, st, gamma, beta, out_f32, out_bf16, off);
}

__global__ void __launch_bounds__(kRowThreads)
layernorm_kernel(const float *__restrict__ x, const float *__restrict__ gamma,
                 const float *__restrict__ beta, float eps, float *out_f32,
                 __nv_bfloat16 *out_bf16, int rows, int hidden, long long in_row_stride,
                 uint8_t *out_q, float q_inv_scale) {
  pdl_wait_prior_grid();
  pdl_launch_dependents();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const int nchunk = hidden >> 2;
  const int nchunk_lane = (nchunk - lane + 31) >> 5;
  const size_t off = static_cast<size_t>(row) * hidden;
  const size_t in_off = static_cast<size_t>(row) * static_cast<size_t>(in_row_stride);
  float4 v[kMaxChunksPerLane];
#pragma unroll
  for (int i = 0; i < kMaxChunksPerLane; ++i)
    if (i < nchunk_lane) v[i] = reinterpret_cast<const float4 *>(x + in_off)[lane + i * 32];
  const RowStats st = row_stats(v, nchunk_lane, hidden, eps);
  row_normalise_store(v, nchunk_lane, st, gamma, beta, out_f32, out_bf16, off, out_q, q_inv_scale);
}

// out = sat_e4m3(in * inv_scale): weights at replica creation, stage tests.
__global__ void __launch_bounds__(256)
quantize_e4m3_kernel(const float *__restrict__ in_f32, const __nv_bfloat16 *__restrict__ in_bf16,
                     uint8_t *__restrict__ out, uint64_t n4, float inv_scale) {
  for (uint64_t k = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n4;
       k += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
    float4 v;
    if (in_f32 != nullptr) {
      v = reinterpret_cast<const float4 *>(in_f32)[k];
    } else {
      const uint2 u = reinterpret_cast<const uint2 *>(in_bf16)[k];
      const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162 *>(&u.x);
      const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162 *>(&u.y);
      v = make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
    }
    reinterpret_cast<uint32_t *>(out)[k] =
        pack_e4m3x4(v.x * inv_scale, v.y * inv_scale, v.z * inv_scale, v.w * inv_scale);
  }
}

// amax[0] = max(amax[0], max |x|) over a bf16 / fp32 buffer (FP8 calibration).  Values are
// non-negative floats, so the integer max of their bit patterns is the float max.
__global__ void __launch_bounds__(256)
amax_kernel(const float *__restrict__ in_f32, const __nv_bfloat16 *__restrict__ in_bf16, uint64_t n,
            float *amax) {
  float m = 0.f;
  for (uint64_t k = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n;
       k += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    m = fmaxf(m, fabsf(in_f32 != nullptr ? in_f32[k] : __bfloat162float(in_bf16[k])));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int *>(amax), __float_as_int(m));
}

// Ragged gather + pad + mask: blockIdx.y = request; each thread moves TWO consecutive tokens
// per step (two 8-byte loads in flight per table, one 16-byte store per output row when the
// row is 16-byte aligned), grid-striding over the padded row.
__global__ void __launch_bounds__(256)
gather_pad_mask_kernel(const int64_t *const *__restrict__ src_ids,
                       const int64_t *const *__restrict__ src_mask,
                       const int32_t *__restrict__ len, int seq, int64_t pad_id,
                       int64_t *__restrict__ ids_out, int64_t *__restrict__ mask_out, int64_t vocab,
                       int32_t *__restrict__ err) {
  const int r = blockIdx.y;
  const int L = len[r];
  const int64_t *sid = src_ids[r];
  const int64_t *smk = src_mask != nullptr ? src_mask[r] : nullptr;
  const size_t row = static_cast<size_t>(r) * seq;
  const bool vec = ((row | static_cast<size_t>(seq)) & 1u) == 0;  // 16-byte aligned pairs
  for (int t = 2 * (blockIdx.x * blockDim.x + threadIdx.x); t < seq; t += 2 * gridDim.x * blockDim.x) {
    const bool in0 = t < L, in1 = t + 1 < L;
    const int64_t i0 = in0 ? sid[t] : pad_id;
    const int64_t i1 = in1 ? sid[t + 1] : pad_id;
    const int64_t m0 = in0 ? (smk != nullptr ? smk[t] : 1) : 0;
    const int64_t m1 = in1 ? (smk != nullptr ? smk[t + 1] : 1) : 0;
    // err != nullptr: a token id outside the embedding table is REPORTED (row + 1, largest such row wins), where
    // LibTorch's embedding would raise; the embedding kernel still clamps it, so nothing reads out of bounds
    if (err != nullptr && ((in0 && (i0 < 0 || i0 >= vocab)) || (in1 && (i1 < 0 || i1 >= vocab)))) atomicMax(err, r + 1);
    if (vec) {
      *reinterpret_cast<longlong2 *>(ids_out + row + t) = make_longlong2(i0, i1);
      *reinterpret_cast<longlong2 *>(mask_out + row + t) = make_longlong2(m0, m1);
    } else {
      ids_out[row + t] = i0;
      mask_out[row + t] = m0;
      if (t + 1 < seq) {
        ids_out[row + t + 1] = i1;
        mask_out[row + t + 1] = m1;
      }
    }
  }
}

// Segmented byte copy.  Segment i: nbytes[i] bytes from (srcs ? srcs[i] : src_base +
// src_off[i]) to (dsts ? dsts[i] : dst_base + dst_off[i]).  blockIdx.y = segment; block b of
// the gridDim.x blocks of a segment owns one contiguous span of it, and every thread keeps up
// to four independent 16-byte loads in flight before its stores (128-bit path when source,
// destination and size are all 16-byte aligned; byte path otherwise).
__global__ void __launch_bounds__(256)
segcopy_kernel(const void *const *__restrict__ srcs, const uint8_t *src_base,
               const uint64_t *__restrict__ src_off, void *const *__restrict__ dsts,
               uint8_t *dst_base, const uint64_t *__restrict__ dst_off,
               const uint64_t *__restrict__ nbytes) {
  const int i = blockIdx.y;
  const uint8_t *s = srcs != nullptr ? static_cast<const uint8_t *>(srcs[i]) : src_base + src_off[i];
  uint8_t *d = dsts != nullptr ? static_cast<uint8_t *>(dsts[i]) : dst_base + dst_off[i];
  const uint64_t n = nbytes[i];
  if (((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(d) | n) & 15u) == 0) {
    const uint64_t n4 = n >> 4;
    const uint64_t span = (n4 + gridDim.x - 1) / gridDim.x;
    const uint64_t lo = span * blockIdx.x;
    const uint64_t hi = lo + span < n4 ? lo + span : n4;
    const uint4 *s4 = reinterpret_cast<const uint4 *>(s);
    uint4 *d4 = reinterpret_cast<uint4 *>(d);
    for (uint64_t base = lo + threadIdx.x; base < hi; base += 4 * 256) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u * 256 < hi) v[u] = __ldcs(s4 + base + u * 256);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u * 256 < hi) __stcs(d4 + base + u * 256, v[u]);
    }
  } else {
    const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t nthr = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t k = tid; k < n; k += nthr) d[k] = s[k];
  }
}

// Stage 4 on per-request buffers: request i receives rows [i*seq, i*seq + rows_i) of the batched
// [n_req*seq, hidden] fp32 output, rows_i = seq (fixed-shape contract) or len[i] (live rows only).
// blockIdx.y = request, the gridDim.x blocks of a request split its bytes; four independent
// 16-byte streaming loads in flight per thread (the output is read once, written once).
__global__ void __launch_bounds__(256)
scatter_requests_kernel(const float *__restrict__ src, const int32_t *__restrict__ len, int seq, int hidden,
                        int full_rows, float *const *__restrict__ dsts, int32_t *__restrict__ err,
                        int32_t *__restrict__ status) {
  const int i = blockIdx.y;
  // hand the gather kernel's finding (0 = every id inside the table) to this task's status word and re-arm the
  // workspace's error word for the next task on this stream
  if (status != nullptr && blockIdx.x == 0 && i == 0 && threadIdx.x == 0) {
    *status = err != nullptr ? *err : 0;
    if (err != nullptr) *err = 0;
  }
  const int rows = full_rows ? seq : min(max(len[i], 0), seq);
  const float *s = src + static_cast<size_t>(i) * seq * hidden;
  float *d = dsts[i];
  const uint64_t n = static_cast<uint64_t>(rows) * hidden;  // floats
  if (((reinterpret_cast<uintptr_t>(d) | (n * 4)) & 15u) == 0) {
    const uint64_t n4 = n >> 2;
    const uint64_t span = (n4 + gridDim.x - 1) / gridDim.x;
    const uint64_t lo = span * blockIdx.x;
    const uint64_t hi = lo + span < n4 ? lo + span : n4;
    const uint4 *s4 = reinterpret_cast<const uint4 *>(s);
    uint4 *d4 = reinterpret_cast<uint4 *>(d);
    for (uint64_t base = lo + threadIdx.x; base < hi; base += 4 * 256) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u * 256 < hi) v[u] = __ldcs(s4 + base + u * 256);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (base + u * 256 < hi) __stcs(d4 + base + u * 256, v[u]);
    }
  } else {
    const uint64_t tid = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const uint64_t nthr = static_cast<uint64_t>(gridDim.x) * blockDim.x;
    for (uint64_t k = tid; k < n; k += nthr) d[k] = s[k];
  }
}

__global__ void __launch_bounds__(256)
affine_kernel(const float *__restrict__ in, float *__restrict__ out, uint64_t n, float scale,
              float shift) {
  for (uint64_t k = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; k < n;
       k += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    out[k] = in[k] * scale + shift;
}

constexpr int kSegBlocksPerSegment = 32;

}  // namespace

spu_status write_cu_seqlens_launch(const int32_t *lens_host, int batch, int32_t *cu_dev, int *total,
                                   cudaStream_t stream) {
  SPU_REQUIRE(lens_host && cu_dev && total, SPU_ERR_INVALID, "cu_seqlens: null argument");
  SPU_REQUIRE(batch > 0 && batch <= kMaxPackedBatch, SPU_ERR_CAPACITY, "cu_seqlens: batch exceeds the packed-path cap");
  CuSeqlensArg cu;
  cu.v[0] = 0;
  for (int b = 0; b < batch; ++b) {
    SPU_REQUIRE(lens_host[b] >= 0, SPU_ERR_INVALID, "cu_seqlens: negative length");
    cu.v[b + 1] = cu.v[b] + lens_host[b];
  }
  *total = cu.v[batch];
  SPU_CUDA_TRY(launch_pdl(write_cu_seqlens_kernel, dim3(1), dim3(256), 0, stream, cu, batch + 1, cu_dev));
  return SPU_OK;
}

spu_status bert_embed_packed_launch(const int64_t *ids, const int32_t *cu, int batch, int seq, int total,
                                    const float *word, const float *pos, const float *type0, const float *gamma,
                                    const float *beta, float eps, int vocab, float *x_f32, void *x_bf16, int hidden,
                                    cudaStream_t stream) {
  SPU_REQUIRE(ids && cu && word && pos && type0 && gamma && beta, SPU_ERR_INVALID, "embed: null argument");
  SPU_REQUIRE(hidden % 4 == 0 && hidden <= 128 * kMaxChunksPerLane && total > 0, SPU_ERR_UNSUPPORTED, "embed: bad shape");
  const int grid = (total + kRowThreads / 32 - 1) / (kRowThreads / 32);
  SPU_CUDA_TRY(launch_pdl(bert_embed_packed_kernel, dim3(grid), dim3(kRowThreads), 0, stream, ids, cu, batch, seq, word,
                          pos, type0, gamma, beta, eps, vocab, x_f32, static_cast<__nv_bfloat16 *>(x_bf16), total, hidden));
  return SPU_OK;
}

spu_status layernorm_unpad_launch(const float *x, const int32_t *cu, int batch, int seq, const float *gamma,
                                  const float *beta, float eps, float *out_f32, void *out_bf16, int hidden,
                                  cudaStream_t stream) {
  SPU_REQUIRE(x && cu && gamma && beta && (out_f32 || out_bf16), SPU_ERR_INVALID, "layernorm: null argument");
  SPU_REQUIRE(hidden % 4 == 0 && hidden <= 128 * kMaxChunksPerLane, SPU_ERR_UNSUPPORTED, "layernorm: bad hidden size");
  const int rows = batch * seq;
  const int grid = (rows + kRowThreads / 32 - 1) / (kRowThreads / 32);
  SPU_CUDA_TRY(launch_pdl(layernorm_unpad_kernel, dim3(grid), dim3(kRowThreads), 0, stream, x, cu, batch, seq, gamma,
                          beta, eps, out_f32, static_cast<__nv_bfloat16 *>(out_bf16), hidden));
  return SPU_OK;
}

SPU_PIN_KERNEL_ARGS(gather_pad_mask_kernel, kGatherKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(gather_pad_mask_kernel, 0);
SPU_PIN_KERNEL_PTR_ARG(gather_pad_mask_kernel, 1);
SPU_PIN_KERNEL_PTR_ARG(gather_pad_mask_kernel, 2);
SPU_PIN_KERNEL_ARGS(scatter_requests_kernel, kScatterKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(scatter_requests_kernel, kScatterArgLen);
SPU_PIN_KERNEL_PTR_ARG(scatter_requests_kernel, kScatterArgDsts);
SPU_PIN_KERNEL_PTR_ARG(scatter_requests_kernel, kScatterArgStatus);
SPU_PIN_KERNEL_ARGS(bert_embed_ln_kernel, kEmbedKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(bert_embed_ln_kernel, kEmbedArgIds);
SPU_PIN_KERNEL_PTR_ARG(bert_embed_ln_kernel, kEmbedArgMask);
SPU_PIN_KERNEL_ARGS(layernorm_kernel, kLayerNormKernelArgs);
SPU_PIN_KERNEL_PTR_ARG(layernorm_kernel, kLayerNormArgOutF32);
SPU_PIN_KERNEL_PTR_ARG(layernorm_kernel, kLayerNormArgOutBf16);

const void *gather_pad_mask_kernel_entry() { return reinterpret_cast<const void *>(&gather_pad_mask_kernel); }
const void *scatter_requests_kernel_entry() { return reinterpret_cast<const void *>(&scatter_requests_kernel); }
const void *bert_embed_kernel_entry() { return reinterpret_cast<const void *>(&bert_embed_ln_kernel); }
const void *layernorm_kernel_entry() { return reinterpret_cast<const void *>(&layernorm_kernel); }

spu_status bert_embed_launch(const int64_t *ids, const int64_t *mask, const float *word,
                             const float *pos, const float *type0, const float *gamma,
                             const float *beta, float eps, int vocab, float *x_f32, void *x_bf16,
                             float *mask_bias, int batch, int seq, int hidden,
                             cudaStream_t stream, void *x_q, float q_inv_scale) {
  SPU_REQUIRE(ids && mask && word && pos && type0 && gamma && beta && mask_bias, SPU_ERR_INVALID,
              "embed: null argument");
  SPU_REQUIRE(hidden % 4 == 0 && hidden <= 128 * kMaxChunksPerLane, SPU_ERR_UNSUPPORTED,
              "embed: hidden must be a multiple of 4 and <= 1024");
  const int rows = batch * seq;
  SPU_REQUIRE(rows > 0, SPU_ERR_INVALID, "embed: empty batch");
  const int grid = (rows + kRowThreads / 32 - 1) / (kRowThreads / 32);
  SPU_CUDA_TRY(launch_pdl(bert_embed_ln_kernel, dim3(grid), dim3(kRowThreads), 0, stream, ids, mask, word, pos,
                          type0, gamma, beta, eps, vocab, x_f32, static_cast<__nv_bfloat16 *>(x_bf16), mask_bias,
                          rows, seq, hidden, static_cast<uint8_t *>(x_q), q_inv_scale));
  return SPU_OK;
}

spu_status layernorm_launch(const float *x, const float *gamma, const float *beta, float eps,
                            float *out_f32, void *out_bf16, int rows, int hidden,
                            cudaStream_t stream, long long in_row_stride, void *out_q,
                            float q_inv_scale) {
  SPU_REQUIRE(x && gamma && beta, SPU_ERR_INVALID, "layernorm: null argument");
  if (in_row_stride == 0) in_row_stride = hidden;
  SPU_REQUIRE(in_row_stride % 4 == 0, SPU_ERR_INVALID, "layernorm: row stride must keep 16-byte alignment");
  SPU_REQUIRE(hidden % 4 == 0 && hidden <= 128 * kMaxChunksPerLane, SPU_ERR_UNSUPPORTED,
              "layernorm: hidden must be a multiple of 4 and <= 1024");
  SPU_REQUIRE(rows > 0, SPU_ERR_INVALID, "layernorm: empty input");
  // rows per CTA: SPU_LN_THREADS (64..256, experiments) / 32; default 8 rows
  static const int threads = [] {
    const char *e = std::getenv("SPU_LN_THREADS");
    const int t = e ? std::atoi(e) : 0;
    return (t == 64 || t == 128 || t == 256) ? t : kRowThreads;
  }();
  const int grid = (rows + threads / 32 - 1) / (threads / 32);
  SPU_CUDA_TRY(launch_pdl(layernorm_kernel, dim3(grid), dim3(threads), 0, stream, x, gamma, beta, eps, out_f32,
                          static_cast<__nv_bfloat16 *>(out_bf16), rows, hidden, in_row_stride,
                          static_cast<uint8_t *>(out_q), q_inv_scale));
  return SPU_OK;
}

spu_status quantize_e4m3_launch(const float *in_f32, const void *in_bf16, void *out, uint64_t n,
                                float inv_scale, cudaStream_t stream) {
  SPU_REQUIRE((in_f32 != nullptr) != (in_bf16 != nullptr) && out != nullptr, SPU_ERR_INVALID,
              "quantize: exactly one input and an output are required");
  SPU_REQUIRE(n % 4 == 0, SPU_ERR_INVALID, "quantize: element count must be a multiple of 4");
  if (n == 0) return SPU_OK;
  const uint64_t blocks = (n / 4 + 255) / 256;
  const int grid = static_cast<int>(blocks > 148ull * 8 ? 148ull * 8 : blocks);
  quantize_e4m3_kernel<<<grid, 256, 0, stream>>>(in_f32, static_cast<const __nv_bfloat16 *>(in_bf16),
                                                 static_cast<uint8_t *>(out), n / 4, inv_scale);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

spu_status scatter_requests_launch(const float *src, const int32_t *len, int n_req, int seq, int hidden,
                                   bool full_rows, float *const *dsts, cudaStream_t stream, int32_t *err,
                                   int32_t *status) {
  SPU_REQUIRE(src && len && dsts, SPU_ERR_INVALID, "scatter_requests: null argument");
  SPU_REQUIRE(n_req > 0 && n_req <= 65535 && seq > 0 && hidden > 0 && hidden % 4 == 0, SPU_ERR_INVALID,
              "scatter_requests: bad shape");
  // ~64 CTAs in all, not 32 per request: the destinations are usually pinned HOST buffers, the stores drain at the
  // PCIe rate (a few hundred KB in flight saturate it) and every resident CTA of this kernel holds thread slots
  // the other worker's GEMM CTAs are waiting for; 64 CTAs x 16 KB in flight still move 12.6 MB into HBM in ~6 us
  static const int total = [] {
    const char *e = std::getenv("SPU_SCATTER_CTAS");
    const int t = e ? std::atoi(e) : 0;
    return t > 0 ? t : 64;
  }();
  const int per_req = total / n_req > 0 ? (total / n_req > kSegBlocksPerSegment ? kSegBlocksPerSegment : total / n_req) : 1;
  dim3 grid(per_req, n_req);
  scatter_requests_kernel<<<grid, 256, 0, stream>>>(src, len, seq, hidden, full_rows ? 1 : 0, dsts, err, status);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

spu_status amax_launch(const float *in_f32, const void *in_bf16, uint64_t n, float *amax,
                       cudaStream_t stream) {
  SPU_REQUIRE((in_f32 != nullptr) != (in_bf16 != nullptr) && amax != nullptr, SPU_ERR_INVALID,
              "amax: exactly one input and an output are required");
  if (n == 0) return SPU_OK;
  const uint64_t blocks = (n + 255) / 256;
  const int grid = static_cast<int>(blocks > 148ull * 4 ? 148ull * 4 : blocks);
  amax_kernel<<<grid, 256, 0, stream>>>(in_f32, static_cast<const __nv_bfloat16 *>(in_bf16), n, amax);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

}  // namespace spu

using namespace spu;

namespace spu {
spu_status gather_pad_mask_launch(const int64_t *const *src_ids, const int64_t *const *src_mask, const int32_t *len,
                                  int n_req, int seq, int64_t pad_id, int64_t *ids_out, int64_t *mask_out,
                                  int64_t vocab, int32_t *err, cudaStream_t stream) {
  if (n_req == 0 || seq == 0) return SPU_OK;  // empty batch: nothing to move
  SPU_REQUIRE(src_ids && len && ids_out && mask_out, SPU_ERR_INVALID, "gather: null argument");
  SPU_REQUIRE(n_req > 0 && seq > 0 && n_req <= 65535, SPU_ERR_INVALID, "gather: bad shape");
  dim3 grid((seq + 511) / 512 > 8 ? 8 : (seq + 511) / 512, n_req);
  gather_pad_mask_kernel<<<grid, 256, 0, stream>>>(src_ids, src_mask, len, seq, pad_id, ids_out, mask_out, vocab, err);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}
}  // namespace spu

extern "C" spu_status spu_gather_pad_mask_async(const int64_t *const *src_ids,
                                                const int64_t *const *src_mask,
                                                const int32_t *len, int32_t n_req, int32_t seq,
                                                int64_t pad_id, int64_t *ids_out,
                                                int64_t *mask_out, void *stream) {
  return gather_pad_mask_launch(src_ids, src_mask, len, n_req, seq, pad_id, ids_out, mask_out, 0, nullptr,
                                as_stream(stream));
}

extern "C" spu_status spu_gather_segments_async(const void *const *srcs, const uint64_t *dst_off,
                                                const uint64_t *nbytes, int32_t n_seg, void *dst,
                                                void *stream) {
  if (n_seg == 0) return SPU_OK;
  SPU_REQUIRE(srcs && dst_off && nbytes && dst, SPU_ERR_INVALID, "gather_segments: null argument");
  SPU_REQUIRE(n_seg > 0 && n_seg <= 65535, SPU_ERR_INVALID, "gather_segments: bad segment count");
  dim3 grid(kSegBlocksPerSegment, n_seg);
  segcopy_kernel<<<grid, 256, 0, as_stream(stream)>>>(srcs, nullptr, nullptr, nullptr,
                                                      static_cast<uint8_t *>(dst), dst_off, nbytes);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

extern "C" spu_status spu_scatter_segments_async(const void *src, const uint64_t *src_off,
                                                 const uint64_t *nbytes, void *const *dsts,
                                                 int32_t n_seg, void *stream) {
  if (n_seg == 0) return SPU_OK;
  SPU_REQUIRE(src && src_off && nbytes && dsts, SPU_ERR_INVALID, "scatter_segments: null argument");
  SPU_REQUIRE(n_seg > 0 && n_seg <= 65535, SPU_ERR_INVALID, "scatter_segments: bad segment count");
  dim3 grid(kSegBlocksPerSegment, n_seg);
  segcopy_kernel<<<grid, 256, 0, as_stream(stream)>>>(nullptr, static_cast<const uint8_t *>(src),
                                                      src_off, dsts, nullptr, nullptr, nbytes);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

extern "C" spu_status spu_affine_f32_async(const float *in, float *out, uint64_t n, float scale,
                                           float shift, void *stream) {
  if (n == 0) return SPU_OK;
  SPU_REQUIRE(in && out, SPU_ERR_INVALID, "affine: null argument");
  const uint64_t blocks = (n + 255) / 256;
  const int grid = static_cast<int>(blocks > 148ull * 8 ? 148ull * 8 : blocks);
  affine_kernel<<<grid, 256, 0, as_stream(stream)>>>(in, out, n, scale, shift);
  SPU_CUDA_TRY(cudaGetLastError());
  return SPU_OK;
}

extern "C" spu_status spu_layernorm_async(const float *x, const float *gamma, const float *beta,
                                          float eps, float *out_f32, void *out_bf16, int32_t rows,
                                          int32_t hidden, void *stream) {
  return layernorm_launch(x, gamma, beta, eps, out_f32, out_bf16, rows, hidden, as_stream(stream), 0, nullptr,
                          1.f);
}

extern "C" spu_status spu_bert_embed_async(const int64_t *ids, const int64_t *mask,
                                           const float *word, const float *pos,
                                           const float *type0, const float *gamma,
                                           const float *beta, float eps, int32_t vocab,
                                           float *x_f32, void *x_bf16, float *mask_bias,
                                           int32_t batch, int32_t seq, int32_t hidden,
                                           void *stream) {
  return bert_embed_launch(ids, mask, word, pos, type0, gamma, beta, eps, vocab, x_f32, x_bf16,
                           mask_bias, batch, seq, hidden, as_stream(stream), nullptr, 1.f);
}

extern "C" spu_status spu_quantize_e4m3_async(const float *in_f32, void *out, uint64_t n, float inv_scale,
                                              void *stream) {
  return quantize_e4m3_launch(in_f32, nullptr, out, n, inv_scale, as_stream(stream));
}
