// exchange.cuh — device side of the NVLink peer-memory exchange, shared by the stand-alone exchange kernel
// (exchange.cu) and the tail of the streaming scan's last CTA (scan_gemv.cu): flag publish / wait at system scope and
// the cross-shard merge under (score desc, global id asc) — the rule of merge_lists_kernel (topk_merge.cu).
#pragma once
#include <math.h>

#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Threads tid < world each watch one flag of the OWN window until it reaches `epoch` (monotonic epochs, wrap-safe
// comparison). Gives up after timeout_ns (sticky *status = 1) so that a missing peer cannot hang the GPU. All `nthreads`
// threads of `sync` must call; returns false when any flag did not arrive.
template <class Sync>
__device__ __forceinline__ bool xchg_wait(const uint32_t* flags, uint32_t world, uint32_t epoch, uint32_t* status,
                                          unsigned long long timeout_ns, uint32_t tid, uint32_t* s_fail, Sync sync) {
  if (tid == 0) *s_fail = 0u;
  sync();
  if (tid < world) {
    const uint32_t* flag = flags + tid;
    const unsigned long long t0 = globaltimer_ns();
    uint32_t spins = 0;
    while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
      if ((++spins & 63u) == 0u) {
        __nanosleep(64);
        if (*(volatile uint32_t*)status != 0u) {  // somebody already gave up
          *s_fail = 1u;
          break;
        }
        if (globaltimer_ns() - t0 > timeout_ns) {
          atomicExch(status, 1u);
          *s_fail = 1u;
          break;
        }
      }
    }
  }
  sync();
  return *s_fail == 0u;
}

__device__ __forceinline__ bool xchg_better(uint32_t sa, uint64_t ia, uint32_t pa, uint32_t sb, uint64_t ib, uint32_t pb) {
  if (sa != sb) return sa > sb;
  if (ia != ib) return ia < ib;
  return pa < pb;
}

// Merge the world*k candidates of query q (list l = slot l of the own window, packed [nk u64 ids][nk f32 scores]) into
// out_ids / out_scores [k] (already offset to the query). Peers wrote the slots, so they are read through L2
// (ld.global.cg). sid / ssc: shared scratch of world*k u64 + world*k u32. When !ok only the padding is written.
template <class Sync>
__device__ __forceinline__ void xchg_merge_query(const unsigned char* slots, size_t slot_bytes, size_t nk, uint32_t world,
                                                 uint32_t k, uint32_t q, uint64_t* sid, uint32_t* ssc,
                                                 uint64_t* out_ids, float* out_scores, bool ok, uint32_t tid,
                                                 uint32_t nthreads, Sync sync) {
  const uint32_t L = world * k;
  for (uint32_t j = tid; j < k; j += nthreads) {
    out_ids[j] = ~0ull;
    out_scores[j] = -INFINITY;
  }
  if (!ok) return;
  for (uint32_t e = tid; e < L; e += nthreads) {
    const uint32_t l = e / k, j = e - l * k;
    const uint64_t* ids = reinterpret_cast<const uint64_t*>(slots + (size_t)l * slot_bytes);
    const float* scores = reinterpret_cast<const float*>(ids + nk);
    const size_t src = (size_t)q * k + j;
    const uint64_t id = __ldcg(ids + src);
    sid[e] = id;
    ssc[e] = (id == ~0ull) ? 0u : float_to_ordered(__ldcg(scores + src));
  }
  sync();
  for (uint32_t e = tid; e < L; e += nthreads) {
    const uint32_t s = ssc[e];
    if (s == 0u) continue;
    const uint64_t id = sid[e];
    uint32_t rank = 0;
    for (uint32_t f = 0; f < L && rank < k; ++f) rank += xchg_better(ssc[f], sid[f], f, s, id, e) ? 1u : 0u;
    if (rank < k) {
      out_ids[rank] = id;
      out_scores[rank] = ordered_to_float(s);
    }
  }
  sync();
}

}  // namespace nvdb_b200
