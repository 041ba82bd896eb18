// refine.cu — exact L2 re-rank of candidate rows (gather-bound).
//
// Replaces nvdb::cuda_l2_topk_batch's kernels (reference src/cuda_refine.cu:405-838). The reference gives one
// THREAD a whole candidate row (32 lanes -> 32 different rows, uncoalesced) and merges K-lists on thread 0
// (58 % of its kernel time, Performance_CUDA.md:267). Here:
//   * one CTA per query, 8 / 16 / 32 LANES per candidate row (whichever keeps every lane busy for the row length): a row
//     is read as consecutive 16-byte pieces by consecutive lanes, four rows in flight per lane group;
//   * the query lives in shared memory; per-lane partial sums use 4 independent chains, then a fixed
//     shuffle tree — deterministic, and closer to the double-precision CPU refine
//     (apps/nvdb_ivf_eval.cpp:232-240) than the reference kernel's 4 x 192-long chains;
//   * selection = warp-local register sorts of 32-key chunks, then rank counting over each chunk's best K (K <= 32; a
//     block bitonic sort of all R keys above that):
//     K smallest written in ascending order (ties: id ascending). No serial merge.
// Semantics kept: ids >= N (incl. 0xFFFFFFFF) are skipped; unfilled output slots are 0xFFFFFFFF / 1e30f
// (src/cuda_refine.cu:435-437, 891-894).
#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

constexpr int kRefineThreads = 256;
constexpr int kRefineWarps = kRefineThreads / 32;

template <int DT>
__device__ __forceinline__ void l2_piece(const uint4 v, const float* __restrict__ q, float scale, float (&acc)[4]) {
  if constexpr (DT == kDtF16) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    const float4 qa = *reinterpret_cast<const float4*>(q);
    const float4 qb = *reinterpret_cast<const float4*>(q + 4);
    const float qv[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      const float d0 = __fsub_rn(qv[2 * i], f.x), d1 = __fsub_rn(qv[2 * i + 1], f.y);
      acc[i] = __fmaf_rn(d0, d0, acc[i]);
      acc[i] = __fmaf_rn(d1, d1, acc[i]);
    }
  } else if constexpr (DT == kDtF32) {
    const float4 qa = *reinterpret_cast<const float4*>(q);
    const float x[4] = {__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)};
    const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float d = __fsub_rn(qv[i], x[i]);
      acc[i] = __fmaf_rn(d, d, acc[i]);
    }
  } else {
    // int8 rows are dequantised as float(x) * scale (one FP32 multiply), include/nvdb/to_f32_row.h:26-31
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 qa = *reinterpret_cast<const float4*>(q + 4 * i);
      const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float x = __fmul_rn((float)(int8_t)(w[i] >> (8 * b)), scale);
        const float d = __fsub_rn(qv[b], x);
        acc[b] = __fmaf_rn(d, d, acc[b]);
      }
    }
  }
}

// LPR lanes share a candidate row (8, 16 or 32: the host picks the one that keeps every lane busy for the row length:
// a 768-byte row is 48 pieces = 16 lanes x 3, a 1536-byte row 32 lanes x 3, a 128-byte row 8 lanes x 1), and every group
// keeps 2 or 4 rows in flight, so the bytes a CTA has outstanding depend little on the shape.
// Loads in flight per thread = rows in flight (U) x pieces per lane (unrolled up to 4): U = 2 for long rows (3+ pieces per
// lane), 4 for short ones — enough to cover the HBM latency at 5-6 resident CTAs per SM without the register count
// cutting the occupancy (4 rows x 3 pieces measured 0.73 of the copy peak at D = 768, 2 x 3 0.89).
// K <= 16: a tournament instead of the counting step of the selection. floor(32 / K) chunks' best K fit one warp sort;
// their best K become one chunk of the next round (written to the other buffer, so a round needs one barrier). R = 500,
// K = 10: 16 chunks -> 6 -> 2 -> 1 in three warp sorts instead of 160 x 160 key compares (16 K cycles of a 100 K-cycle CTA).
// Not inlined: inlined, the two unrolled sorts changed the code of the distance loop (85 K -> 101 K cycles per CTA).
__device__ __noinline__ void select_tournament(uint64_t* skeys, uint64_t* sbuf, uint32_t nchunks, uint32_t K, uint32_t* out_ids,
                                               float* out_dist) {
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t G = 32u / K;
  uint64_t* src = skeys;
  uint64_t* dst = sbuf;
  uint32_t nch = nchunks;
  while (nch > 1u) {
    __syncthreads();
    const uint32_t ngr = (nch + G - 1u) / G;
    for (uint32_t gr = warp; gr < ngr; gr += kRefineThreads / 32) {
      const uint32_t c = gr * G + lane / K;
      uint64_t key = (lane < G * K && c < nch) ? src[c * 32u + lane % K] : 0ull;
#pragma unroll
      for (uint32_t size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
          const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)key, stride);
          const uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(key >> 32), stride);
          const uint64_t o = ((uint64_t)hi << 32) | lo;
          const bool lower = (lane & stride) == 0, desc = (lane & size) == 0 || size == 32;
          key = (lower == desc) ? (key > o ? key : o) : (key < o ? key : o);
        }
      }
      if (lane < K) dst[gr * 32u + lane] = key;
    }
    uint64_t* t = src;
    src = dst;
    dst = t;
    nch = ngr;
  }
  __syncthreads();
  for (uint32_t j = tid; j < K; j += kRefineThreads) {
    const uint64_t sk = src[j];
    const bool ok = sk != 0ull;
    const uint64_t key = ~sk;
    out_ids[j] = ok ? (uint32_t)key : 0xFFFFFFFFu;
    if (out_dist) out_dist[j] = ok ? __uint_as_float((uint32_t)(key >> 32)) : 1e30f;
  }
}

// Selection of the K best of the R keys in skeys (descending; 0 = empty) and their output. Not inlined into the kernel:
// inlined, its unrolled warp sorts changed the code of the distance loop (85 K -> 101 K cycles per CTA at K5).
__device__ __noinline__ void select_and_write(uint64_t* skeys, uint64_t* sbuf, uint32_t R, uint32_t K, uint32_t sortn,
                                              uint32_t* out_ids, float* out_dist) {
  const uint32_t tid = threadIdx.x;
  if (K <= 32u) {
    // Selection in two barrier-light steps instead of a full block sort (36+ barrier-separated stages for 256 keys):
    //   1. every warp sorts chunks of 32 keys in registers (bitonic over shuffles) and leaves each chunk's best
    //      min(K, 32) at the front of the chunk — the final K best are among them;
    //   2. those few keys are ranked by counting (keys are unique per (distance, id); equal keys — the same id listed
    //      twice — take consecutive ranks by position) and a key of rank < K goes straight to its output slot.
    const uint32_t lane = tid & 31u, warp = tid >> 5;
    const uint32_t nchunks = (R + 31u) >> 5;
    for (uint32_t ch = warp; ch < nchunks; ch += kRefineThreads / 32) {
      const uint32_t pos = ch * 32u + lane;
      uint64_t key = pos < R ? skeys[pos] : 0ull;
#pragma unroll
      for (uint32_t size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
          const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)key, stride);
          const uint32_t hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(key >> 32), stride);
          const uint64_t o = ((uint64_t)hi << 32) | lo;
          const bool lower = (lane & stride) == 0, desc = (lane & size) == 0 || size == 32;
          key = (lower == desc) ? (key > o ? key : o) : (key < o ? key : o);
        }
      }
      skeys[pos] = key;   // sortn >= 32 * nchunks (the host rounds it up to a power of two >= 32)
    }
    if (K <= 16u) {
      select_tournament(skeys, sbuf, nchunks, K, out_ids, out_dist);
      return;
    }
    for (uint32_t j = tid; j < K; j += kRefineThreads) {
      out_ids[j] = 0xFFFFFFFFu;
      if (out_dist) out_dist[j] = 1e30f;
    }
    __syncthreads();
    const uint32_t ncand = nchunks * K;
    for (uint32_t i = tid; i < ncand; i += kRefineThreads) {
      const uint32_t pi = (i / K) * 32u + (i % K);
      const uint64_t mine = skeys[pi];
      if (mine == 0ull) continue;
      uint32_t rank = 0;
      for (uint32_t m = 0; m < ncand; ++m) {
        const uint32_t pm = (m / K) * 32u + (m % K);
        const uint64_t o = skeys[pm];
        rank += (o > mine || (o == mine && pm < pi)) ? 1u : 0u;
      }
      if (rank < K) {
        const uint64_t key = ~mine;
        out_ids[rank] = (uint32_t)key;
        if (out_dist) out_dist[rank] = __uint_as_float((uint32_t)(key >> 32));
      }
    }
    return;
  }
  auto sync = [] { __syncthreads(); };
  bitonic_sort_desc(skeys, sortn, tid, kRefineThreads, sync);
  for (uint32_t j = tid; j < K; j += kRefineThreads) {
    const uint64_t sk = (j < sortn) ? skeys[j] : 0ull;
    const bool ok = sk != 0ull;
    const uint64_t key = ~sk;
    out_ids[j] = ok ? (uint32_t)key : 0xFFFFFFFFu;
    if (out_dist) out_dist[j] = ok ? __uint_as_float((uint32_t)(key >> 32)) : 1e30f;
  }
}

template <int DT, int LPR, int kRowsInFlight, int PPL>
__global__ void __launch_bounds__(kRefineThreads)
refine_l2_kernel(const uint8_t* __restrict__ base, const float* __restrict__ scales, uint64_t n, uint32_t dim,
                 const float* __restrict__ queries, const uint32_t* __restrict__ cand, uint32_t R, uint32_t K,
                 uint32_t sortn, uint32_t* __restrict__ out_ids, float* __restrict__ out_dist,
                 unsigned long long* __restrict__ dbg, uint32_t dbg_q) {
  extern __shared__ uint64_t smem64[];
  // optional in-kernel phase sampling (the reference's CUDA_DBG_TIMING, src/cuda_refine.cu:863-872,1116-1144): thread 0
  // of the first dbg_q queries records clock64 at the phase boundaries -> dbg[q*3 + {0: query staging, 1: gather +
  // distances, 2: selection + output}] cycles
  const bool sample = dbg != nullptr && blockIdx.x < dbg_q && threadIdx.x == 0;
  long long tq0 = 0, tq1 = 0, tq2 = 0;
  if (sample) tq0 = clock64();
  uint64_t* skeys = smem64;                                  // [sortn]
  float* sq = reinterpret_cast<float*>(smem64 + sortn);      // [dim]
  uint64_t* sbuf = smem64 + sortn + (dim + 1u) / 2u;         // [sortn / 2]: the other side of the selection tournament
  constexpr int ES = (DT == kDtF32) ? 4 : (DT == kDtF16 ? 2 : 1);
  constexpr int P16 = 16 / ES;
  constexpr uint32_t kGroups = kRefineThreads / LPR;
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  const uint32_t grp = tid / LPR, gl = tid % LPR;
  const size_t row_bytes = (size_t)dim * ES;
  const uint32_t npieces = (uint32_t)(row_bytes >> 4);       // dim * ES is a multiple of 16 (checked by the host)

  for (uint32_t i = tid; i < dim; i += kRefineThreads) sq[i] = queries[(size_t)q * dim + i];
  for (uint32_t i = R + tid; i < sortn; i += kRefineThreads) skeys[i] = 0ull;
  __syncthreads();
  if (sample) tq1 = clock64();

  const uint32_t* cq = cand + (size_t)q * R;
  // the loop bound is the same for every thread (rows past R are skipped inside): the lane groups of one warp must
  // reach the full-mask shuffles below together
  for (uint32_t rb = 0; rb < R; rb += kGroups * kRowsInFlight) {
    const uint32_t r0 = rb + grp * kRowsInFlight;
    uint32_t id[kRowsInFlight];
    bool ok[kRowsInFlight];
    const uint8_t* row[kRowsInFlight];
    float sc[kRowsInFlight], acc[kRowsInFlight][4];
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      id[u] = (r0 + u < R) ? cq[r0 + u] : 0xFFFFFFFFu;
      ok[u] = (uint64_t)id[u] < n;
      row[u] = base + (size_t)(ok[u] ? id[u] : 0) * row_bytes;
      sc[u] = 1.f;
      if constexpr (DT == kDtI8) sc[u] = ok[u] ? __ldg(scales + id[u]) : 0.f;
#pragma unroll
      for (int a4 = 0; a4 < 4; ++a4) acc[u][a4] = 0.f;
    }
    if constexpr (PPL > 0) {
      // exactly PPL pieces per lane (npieces == PPL * LPR): every load of the kRowsInFlight rows is issued before the
      // first use — PPL * kRowsInFlight 16-byte loads in flight per thread
      uint4 v[PPL][kRowsInFlight];
#pragma unroll
      for (int i = 0; i < PPL; ++i)
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u)
          v[i][u] = ok[u] ? ldg128_stream(row[u] + (size_t)(gl + i * LPR) * 16) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int i = 0; i < PPL; ++i)
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) l2_piece<DT>(v[i][u], sq + (gl + i * LPR) * P16, sc[u], acc[u]);
    } else {
#pragma unroll 2
      for (uint32_t c = gl; c < npieces; c += LPR) {
        uint4 v[kRowsInFlight];
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) v[u] = ok[u] ? ldg128_stream(row[u] + (size_t)c * 16) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int u = 0; u < kRowsInFlight; ++u) l2_piece<DT>(v[u], sq + c * P16, sc[u], acc[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < kRowsInFlight; ++u) {
      float d = __fadd_rn(__fadd_rn(acc[u][0], acc[u][1]), __fadd_rn(acc[u][2], acc[u][3]));
#pragma unroll
      for (int off = LPR / 2; off > 0; off >>= 1) d = __fadd_rn(d, __shfl_xor_sync(0xffffffffu, d, off));
      // distance >= +0, so its bit pattern orders like the value; smaller (dist, id) -> larger stored key
      if (gl == 0 && r0 + u < R) skeys[r0 + u] = ok[u] ? ~(((uint64_t)__float_as_uint(d) << 32) | id[u]) : 0ull;
    }
  }
  __syncthreads();
  if (sample) {
    tq2 = clock64();
    dbg[(size_t)q * 3 + 0] = (unsigned long long)(tq1 - tq0);
    dbg[(size_t)q * 3 + 1] = (unsigned long long)(tq2 - tq1);
  }
  select_and_write(skeys, sbuf, R, K, sortn, out_ids + (size_t)q * K, out_dist ? out_dist + (size_t)q * K : nullptr);
  if (dbg != nullptr && blockIdx.x < dbg_q) {  // the selection phase ends when the slowest thread is done
    __syncthreads();
    if (sample) dbg[(size_t)q * 3 + 2] = (unsigned long long)(clock64() - tq2);
  }
}

}  // namespace

cudaError_t refine_l2_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim,
                             const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                             uint32_t* out_ids, float* out_dist, cudaStream_t stream, unsigned long long* dbg,
                             uint32_t dbg_q, uint32_t* smem_bytes_out) {
  if (Q == 0 || K == 0) return cudaSuccess;
  const uint32_t sortn = next_pow2(R < 32 ? 32 : R);
  const size_t smem = (size_t)sortn * 12u + (size_t)((dim + 1u) / 2u) * 8u;  // keys + query + the tournament's second buffer (sortn / 2)
  if (smem > 220u * 1024u) return cudaErrorInvalidValue;
  if (smem_bytes_out) *smem_bytes_out = (uint32_t)smem;
  auto go = [&](auto kern) -> cudaError_t {
    if (smem > 48u * 1024u) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    kern<<<Q, kRefineThreads, smem, stream>>>(reinterpret_cast<const uint8_t*>(base), scales, n, dim, queries, cand, R,
                                              K, sortn, out_ids, out_dist, dbg, dbg_q);
    count_launch();
    return cudaGetLastError();
  };
  // lanes per row: the largest of 32 / 16 / 8 that leaves no lane without a piece in the last pass over the row
  const uint32_t es = dtype == kDtF32 ? 4u : dtype == kDtF16 ? 2u : 1u;
  const uint32_t npieces = (uint32_t)(((size_t)dim * es) >> 4);
  const int lpr = (npieces % 32u == 0u) ? 32 : (npieces % 16u == 0u) ? 16 : (npieces % 8u == 0u || npieces < 16u) ? 8 : (npieces < 48u ? 16 : 32);
  // pieces per lane: exactly 3 (768-byte rows on 16 lanes, 1536-byte rows on 32: the shapes of the reference's tables
  // and of BASELINE.json) gets the fully unrolled form; otherwise a loop, 2 rows in flight for long rows and 4 for short
  const uint32_t ppl = npieces / (uint32_t)lpr;
  const int variant = (npieces == 3u * (uint32_t)lpr) ? 0 : (ppl >= 3u ? 1 : 2);
#define NVDB_REFINE_PICK(DT)                                                                                         \
  switch (variant * 100 + lpr) {                                                                                     \
    case 32: return go(refine_l2_kernel<DT, 32, 2, 3>);                                                              \
    case 16: return go(refine_l2_kernel<DT, 16, 2, 3>);                                                              \
    case 8: return go(refine_l2_kernel<DT, 8, 2, 3>);                                                                \
    case 132: return go(refine_l2_kernel<DT, 32, 2, 0>);                                                             \
    case 116: return go(refine_l2_kernel<DT, 16, 2, 0>);                                                             \
    case 108: return go(refine_l2_kernel<DT, 8, 2, 0>);                                                              \
    case 232: return go(refine_l2_kernel<DT, 32, 4, 0>);                                                             \
    case 216: return go(refine_l2_kernel<DT, 16, 4, 0>);                                                             \
    default: return go(refine_l2_kernel<DT, 8, 4, 0>);                                                               \
  }
  switch (dtype) {
    case kDtF32: NVDB_REFINE_PICK(kDtF32)
    case kDtF16: NVDB_REFINE_PICK(kDtF16)
    case kDtI8: NVDB_REFINE_PICK(kDtI8)
    default: return cudaErrorInvalidValue;
  }
#undef NVDB_REFINE_PICK
}

}  // namespace nvdb_b200
