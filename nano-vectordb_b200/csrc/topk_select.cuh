// topk_select.cuh — per-CTA streaming top-k selection shared by the scan kernels.
//
// Replaces the reference's per-thread heaps / TopKBuffer (src/flat_index.cpp:28-38, include/nvdb/topK.h:23-63):
// every scoring thread compares its score with the CTA's current k-th best ("threshold") and, only when it
// passes, appends an ordered key to a shared-memory candidate list. When a list may overflow on the next tile
// it is compacted: block bitonic sort, keep the best k, raise the threshold. After the first few tiles
// P(append) ~ k / rows_seen, so selection costs nothing next to the HBM stream.
#pragma once
#include "common.cuh"

namespace nvdb_b200 {

template <int NQ>
struct TopKSelect {
  uint64_t* keys;   // [NQ][cap]
  uint32_t* cnt;    // [NQ]
  float* thr;       // [NQ]
  uint32_t cap;     // power of two, >= k + rows_per_tile
  uint32_t k;
  uint32_t pending_bound;  // uniform upper bound on max_q(cnt[q]) - (entries kept by the last compaction)

  __device__ __forceinline__ void init(uint64_t* keys_, uint32_t* cnt_, float* thr_, uint32_t cap_, uint32_t k_,
                                       uint32_t tid) {
    keys = keys_;
    cnt = cnt_;
    thr = thr_;
    cap = cap_;
    k = k_;
    pending_bound = 0;
    if (tid < NQ) {
      cnt[tid] = 0;
      thr[tid] = -INFINITY;
    }
  }

  // fast path: one compare against a broadcast shared-memory word
  __device__ __forceinline__ bool offer(int qi, float score, uint32_t local_row) {
    if (score >= thr[qi]) {
      const uint32_t pos = atomicAdd(&cnt[qi], 1u);
      keys[(uint32_t)qi * cap + pos] = make_key(score, local_row);
      return true;
    }
    return false;
  }

  // Called by all `nthreads` selecting threads after a tile's offers. `pushed_threads` must be uniform
  // (bar.red.popc of "this thread appended something"), `rows_per_tile` the max appends per query per tile.
  template <typename SyncFn>
  __device__ __forceinline__ void end_tile(uint32_t pushed_threads, uint32_t rows_per_tile, uint32_t tid,
                                           uint32_t nthreads, SyncFn sync) {
    pending_bound += pushed_threads;
    if (k + pending_bound + rows_per_tile > cap) compact(tid, nthreads, sync);
  }

  // Keep the best k of every list in descending order, publish the new thresholds. Ends with a barrier.
  // Lists of at most 2 x `nthreads` entries (nearly every compaction of the streaming scan: a tile appends at most one
  // key per thread) are ranked by counting - a thread counts the keys above its own (two) and stores them at their
  // ranks: one barrier per list
  // instead of the 45 barrier-separated stages of a 512-key bitonic sort (5.6 us per compaction, during which the TMA
  // ring fills up and the HBM stream stalls). Keys of one list are distinct (one per row). Longer lists: bitonic sort.
  template <typename SyncFn>
  __device__ void compact(uint32_t tid, uint32_t nthreads, SyncFn sync) {
    bool small = true;
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi) small = small && cnt[qi] <= 2u * nthreads;
    if (small) {
#pragma unroll 1
      for (int qi = 0; qi < NQ; ++qi) {
        const uint32_t c = cnt[qi];
        uint64_t* kq = keys + (uint32_t)qi * cap;
        const uint64_t m0 = tid < c ? kq[tid] : 0ull, m1 = tid + nthreads < c ? kq[tid + nthreads] : 0ull;
        uint32_t r0 = 0, r1 = 0;
        if (c <= nthreads) {  // one key per thread (the usual case)
          if (tid < c)
            for (uint32_t m = 0; m < c; ++m) r0 += kq[m] > m0 ? 1u : 0u;
        } else {
          for (uint32_t m = 0; m < c; ++m) {
            const uint64_t o = kq[m];
            r0 += o > m0 ? 1u : 0u;
            r1 += o > m1 ? 1u : 0u;
          }
        }
        sync();  // every key of this list has been read
        if (tid < c && r0 < k) kq[r0] = m0;
        if (tid + nthreads < c && r1 < k) kq[r1] = m1;
        for (uint32_t j = c + tid; j < k; j += nthreads) kq[j] = 0ull;  // zero padding of a list with fewer than k keys
      }
      sync();
      if (tid < NQ) {
        const uint32_t c = cnt[tid] < k ? cnt[tid] : k;
        cnt[tid] = c;
        thr[tid] = (c == k) ? key_score(keys[(uint32_t)tid * cap + k - 1]) : -INFINITY;
      }
      pending_bound = 0;
      sync();
      return;
    }
#pragma unroll 1
    for (int qi = 0; qi < NQ; ++qi) {
      const uint32_t c = cnt[qi];
      for (uint32_t i = c + tid; i < cap; i += nthreads) keys[(uint32_t)qi * cap + i] = 0ull;
    }
    sync();
    const uint32_t half = cap >> 1;
    for (uint32_t size = 2; size <= cap; size <<= 1) {
      for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
#pragma unroll 1
        for (int qi = 0; qi < NQ; ++qi) {
          uint64_t* kq = keys + (uint32_t)qi * cap;
          for (uint32_t t = tid; t < half; t += nthreads) {
            const uint32_t lo = 2 * t - (t & (stride - 1));
            const uint32_t hi = lo + stride;
            const bool desc = ((lo & size) == 0);
            const uint64_t a = kq[lo], b = kq[hi];
            if ((a < b) == desc) {
              kq[lo] = b;
              kq[hi] = a;
            }
          }
        }
        sync();
      }
    }
    if (tid < NQ) {
      const uint32_t c = cnt[tid] < k ? cnt[tid] : k;
      cnt[tid] = c;
      // threshold = score of the current k-th best; admission is `>=` so that equal-score rows reach the
      // exact (score, id) ordering in the sort instead of being dropped by the float compare.
      thr[tid] = (c == k) ? key_score(keys[(uint32_t)tid * cap + k - 1]) : -INFINITY;
    }
    pending_bound = 0;
    sync();
  }

  // after a final compact(): list qi holds its best min(cnt,k) keys in descending order, zero padded.
  __device__ __forceinline__ void store(uint64_t* out /*[NQ][stride]*/, size_t q_stride, uint32_t nq_valid,
                                        uint32_t tid, uint32_t nthreads) const {
    for (uint32_t i = tid; i < nq_valid * k; i += nthreads) {
      const uint32_t qi = i / k, j = i - qi * k;
      out[(size_t)qi * q_stride + j] = keys[qi * cap + j];
    }
  }
};

}  // namespace nvdb_b200
