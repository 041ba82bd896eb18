// topk_merge.cu — reductions of per-CTA / per-shard top-k lists.
//
// Replaces the serial merges of the reference (src/flat_index_omp.cpp:67-75, src/flat_index_pool.cpp:212-214,
// include/nvdb/topK.h:41-49): one CTA per query; candidates are streamed through a shared-memory buffer,
// block-bitonic-sorted, and the best k are carried to the next round.
#include <math.h>

#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

constexpr int kMergeThreads = 512;
constexpr uint32_t kSelectMinKeys = 2048;  // below this one bitonic sort of everything is cheaper than 8 histogram passes

// partial: [nq][nlists][k] ordered keys (0 = empty). out: ids = row_offset + row, scores; padded with
// UINT64_MAX / -inf when fewer than k candidates exist (k > n).
__global__ void __launch_bounds__(kMergeThreads)
merge_partials_kernel(const uint64_t* __restrict__ partial, uint32_t nlists, uint32_t k, uint32_t bufn,
                      uint64_t row_offset, uint64_t* __restrict__ out_ids, float* __restrict__ out_scores,
                      uint64_t* __restrict__ out_keys, size_t key_stride, float* __restrict__ out_thr) {
  extern __shared__ uint64_t buf[];
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  const uint64_t* in = partial + (size_t)q * nlists * k;
  const uint32_t total = nlists * k;
  auto sync = [] { __syncthreads(); };
  uint32_t kept = 0, pos = 0;
  while (pos < total) {
    const uint32_t take = min(bufn - kept, total - pos);
    for (uint32_t i = tid; i < bufn - kept; i += kMergeThreads) buf[kept + i] = (i < take) ? in[pos + i] : 0ull;
    __syncthreads();
    bitonic_sort_desc(buf, bufn, tid, kMergeThreads, sync);
    kept = k;
    pos += take;
  }
  if (out_keys) {  // seed mode: keys + threshold for a later pass
    for (uint32_t j = tid; j < k; j += kMergeThreads) out_keys[(size_t)q * key_stride + j] = (total > 0) ? buf[j] : 0ull;
    if (tid == 0) {
      const uint64_t kth = (total > 0) ? buf[k - 1] : 0ull;
      out_thr[q] = kth != 0ull ? key_score(kth) : -INFINITY;
    }
    return;
  }
  for (uint32_t j = tid; j < k; j += kMergeThreads) {
    const uint64_t key = (total > 0) ? buf[j] : 0ull;
    const bool ok = key != 0ull;
    out_ids[(size_t)q * k + j] = ok ? row_offset + (uint64_t)key_row(key) : ~0ull;
    out_scores[(size_t)q * k + j] = ok ? key_score(key) : -INFINITY;
  }
}

// Large reductions (k = 100 over 148 lists = 14,800 keys): sorting everything costs ~105 bitonic stages of 16 K keys
// (~0.15 ms on one CTA, a fifth of a batch-1 scan), and the "k-th entry of the luckiest list" floor removes only a
// quarter of the keys when k is large. Instead: all keys into shared memory once, an MSB-first radix select (8 passes of
// 8 bits) finds the k-th largest 64-bit key exactly (keys are unique: score | ~row), the k survivors are compacted and
// only they are sorted. Same outputs as merge_partials_kernel.
__global__ void __launch_bounds__(kMergeThreads)
merge_select_kernel(const uint64_t* __restrict__ partial, uint32_t total, uint32_t k, uint32_t kpow2,
                    uint64_t row_offset, uint64_t* __restrict__ out_ids, float* __restrict__ out_scores,
                    uint64_t* __restrict__ out_keys, size_t key_stride, float* __restrict__ out_thr) {
  extern __shared__ uint64_t buf[];  // [total] keys, then [kpow2] survivors
  __shared__ uint32_t hist[256];
  __shared__ uint64_t s_prefix;
  __shared__ uint32_t s_need, s_cnt;
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  const uint64_t* in = partial + (size_t)q * total;
  uint64_t* sv = buf + total;
  auto sync = [] { __syncthreads(); };
  for (uint32_t i = tid; i < total; i += kMergeThreads) buf[i] = in[i];
  if (tid == 0) {
    s_prefix = 0ull;
    s_need = k;
    s_cnt = 0u;
  }
  for (int pass = 0; pass < 8; ++pass) {
    const int shift = 56 - 8 * pass;
    for (uint32_t b = tid; b < 256u; b += kMergeThreads) hist[b] = 0u;
    __syncthreads();
    const uint64_t prefix = s_prefix;
    for (uint32_t i = tid; i < total; i += kMergeThreads) {
      const uint64_t key = buf[i];
      if (pass == 0 || (key >> (shift + 8)) == (prefix >> (shift + 8))) atomicAdd(&hist[(uint32_t)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32u) {
      // lane l owns buckets 255-8l .. 248-8l (descending); find the bucket where the count from the top reaches `need`
      uint32_t c[8], sum = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        c[j] = hist[255u - 8u * tid - (uint32_t)j];
        sum += c[j];
      }
      uint32_t incl = sum;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, off);
        if (tid >= (uint32_t)off) incl += o;
      }
      const uint32_t need = s_need, before = incl - sum;
      if (before < need && need <= incl) {  // exactly one lane (all keys of the prefix are counted, need <= their number)
        uint32_t cum = before;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (cum < need && need <= cum + c[j]) {
            s_prefix = prefix | ((uint64_t)(255u - 8u * tid - (uint32_t)j) << shift);
            s_need = need - cum;
          }
          cum += c[j];
        }
      }
    }
    __syncthreads();
  }
  // total >= k, so the k-th largest key exists; it is 0 (empty) when fewer than k real candidates exist
  const uint64_t kth = s_prefix;
  for (uint32_t i = tid; i < total; i += kMergeThreads) {
    const uint64_t key = buf[i];
    if (key != 0ull && key >= kth) {  // <= k survivors when keys are unique; duplicates (untrusted input) are dropped, not written past sv
      const uint32_t slot = atomicAdd(&s_cnt, 1u);
      if (slot < kpow2) sv[slot] = key;
    }
  }
  __syncthreads();
  const uint32_t cnt = s_cnt < kpow2 ? s_cnt : kpow2;  // <= k when keys are unique
  for (uint32_t i = cnt + tid; i < kpow2; i += kMergeThreads) sv[i] = 0ull;
  __syncthreads();
  bitonic_sort_desc(sv, kpow2, tid, kMergeThreads, sync);
  if (out_keys) {
    for (uint32_t j = tid; j < k; j += kMergeThreads) out_keys[(size_t)q * key_stride + j] = sv[j];
    if (tid == 0) out_thr[q] = sv[k - 1] != 0ull ? key_score(sv[k - 1]) : -INFINITY;
    return;
  }
  for (uint32_t j = tid; j < k; j += kMergeThreads) {
    const uint64_t key = sv[j];
    const bool ok = key != 0ull;
    out_ids[(size_t)q * k + j] = ok ? row_offset + (uint64_t)key_row(key) : ~0ull;
    out_scores[(size_t)q * k + j] = ok ? key_score(key) : -INFINITY;
  }
}

// Cross-shard merge: ids are 64-bit global ids, so rank by pairwise comparison instead of packed keys.
// L = nlists*k entries per query; entry e lands at position rank(e) = #{f better than e}.
__device__ __forceinline__ bool better(uint32_t sa, uint64_t ia, uint32_t pa, uint32_t sb, uint64_t ib, uint32_t pb) {
  if (sa != sb) return sa > sb;
  if (ia != ib) return ia < ib;
  return pa < pb;
}

__global__ void __launch_bounds__(256)
merge_lists_kernel(const uint64_t* __restrict__ ids, const float* __restrict__ scores, size_t id_list_stride,
                   size_t score_list_stride, uint32_t nlists, uint32_t nq, uint32_t k, uint64_t* __restrict__ out_ids,
                   float* __restrict__ out_scores) {
  extern __shared__ uint64_t sm[];
  const uint32_t L = nlists * k;
  uint64_t* sid = sm;                                      // [L]
  uint32_t* ssc = reinterpret_cast<uint32_t*>(sm + L);     // [L] ordered scores (0 = empty)
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  for (uint32_t e = tid; e < L; e += blockDim.x) {
    const uint32_t l = e / k, j = e - l * k;
    const size_t src = (size_t)q * k + j;
    const uint64_t id = ids[(size_t)l * id_list_stride + src];
    sid[e] = id;
    ssc[e] = (id == ~0ull) ? 0u : float_to_ordered(scores[(size_t)l * score_list_stride + src]);
  }
  for (uint32_t j = tid; j < k; j += blockDim.x) {
    out_ids[(size_t)q * k + j] = ~0ull;
    out_scores[(size_t)q * k + j] = -INFINITY;
  }
  __syncthreads();
  for (uint32_t e = tid; e < L; e += blockDim.x) {
    const uint32_t s = ssc[e];
    if (s == 0u) continue;
    const uint64_t id = sid[e];
    uint32_t rank = 0;
    for (uint32_t f = 0; f < L && rank < k; ++f) rank += better(ssc[f], sid[f], f, s, id, e) ? 1u : 0u;
    if (rank < k) {
      out_ids[(size_t)q * k + rank] = id;
      out_scores[(size_t)q * k + rank] = ordered_to_float(s);
    }
  }
}

}  // namespace

static cudaError_t merge_partials_any(const uint64_t* partial, uint32_t nlists, uint32_t nq, uint32_t k,
                                      uint64_t row_offset, uint64_t* out_ids, float* out_scores, uint64_t* out_keys,
                                      size_t key_stride, float* out_thr, cudaStream_t stream) {
  if (nq == 0 || k == 0) return cudaSuccess;
  // many keys per query: radix select + sort of the k survivors (merge_select_kernel) when everything fits shared memory
  {
    const uint32_t total = nlists * k, kpow2 = next_pow2(k < 2u ? 2u : k);
    const size_t smem = ((size_t)total + kpow2) * 8u;
    if (total > kSelectMinKeys && total >= k && smem <= 200u * 1024u) {
      static bool attr_done = false;  // per process; the attribute is per function and device-wide once set high enough
      if (!attr_done || smem > 48u * 1024u) {
        cudaError_t e = cudaFuncSetAttribute(merge_select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_done = true;
      }
      merge_select_kernel<<<nq, kMergeThreads, smem, stream>>>(partial, total, k, kpow2, row_offset, out_ids, out_scores,
                                                              out_keys, key_stride, out_thr);
      count_launch();
      return cudaGetLastError();
    }
  }
  // buffer: at least 2k (so every round makes progress), ideally everything in one round, capped at 16384 keys
  uint32_t bufn = next_pow2(nlists * k);
  if (bufn > 16384u) bufn = 16384u;
  const uint32_t min_buf = next_pow2(2u * k);
  if (bufn < min_buf) bufn = min_buf;
  if (bufn < 2u) bufn = 2u;
  const size_t smem = (size_t)bufn * 8u;
  if (smem > 48u * 1024u) {
    cudaError_t e = cudaFuncSetAttribute(merge_partials_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  merge_partials_kernel<<<nq, kMergeThreads, smem, stream>>>(partial, nlists, k, bufn, row_offset, out_ids, out_scores,
                                                             out_keys, key_stride, out_thr);
  count_launch();
  return cudaGetLastError();
}

cudaError_t merge_partials_launch(const uint64_t* partial, uint32_t nlists, uint32_t nq, uint32_t k,
                                  uint64_t row_offset, uint64_t* out_ids, float* out_scores, cudaStream_t stream) {
  return merge_partials_any(partial, nlists, nq, k, row_offset, out_ids, out_scores, nullptr, 0, nullptr, stream);
}

cudaError_t merge_partials_seed_launch(const uint64_t* partial, uint32_t nlists, uint32_t nq, uint32_t k,
                                       uint64_t* out_keys, size_t key_stride, float* out_thr, cudaStream_t stream) {
  return merge_partials_any(partial, nlists, nq, k, 0, nullptr, nullptr, out_keys, key_stride, out_thr, stream);
}

cudaError_t merge_lists_launch(const uint64_t* ids, const float* scores, size_t id_list_stride,
                               size_t score_list_stride, uint32_t nlists, uint32_t nq, uint32_t k, uint64_t* out_ids,
                               float* out_scores, cudaStream_t stream) {
  if (nq == 0 || k == 0) return cudaSuccess;
  const size_t smem = (size_t)nlists * k * 12u + 8u;
  if (smem > 200u * 1024u) return cudaErrorInvalidValue;
  if (smem > 48u * 1024u) {
    cudaError_t e = cudaFuncSetAttribute(merge_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  merge_lists_kernel<<<nq, 256, smem, stream>>>(ids, scores, id_list_stride, score_list_stride, nlists, nq, k, out_ids,
                                                out_scores);
  count_launch();
  return cudaGetLastError();
}

}  // namespace nvdb_b200
