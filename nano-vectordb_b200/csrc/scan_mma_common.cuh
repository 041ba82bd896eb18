// scan_mma_common.cuh — pieces shared by the tensor-core scan kernels (scan_mma.cu: one CTA per tile, cta_group::1;
// scan_mma2.cu: CTA pairs, cta_group::2): geometry constants and the per-thread top-k containers of the epilogue.
#pragma once
#include "common.cuh"

namespace nvdb_b200 {
namespace {

constexpr int kMmaThreads = 192;
constexpr int kBlockM = 128;
constexpr uint32_t kTmemCols = 512;
constexpr int kMmaRegK = 16;  // k <= 16: per-thread top-k lists in registers, else in shared memory
constexpr uint32_t kAccColsPerBuf = 256;
constexpr uint64_t kSentinelKey = ((uint64_t)0x007FFFFFu << 32);  // (ordered(-inf), row 0xFFFFFFFF): below any real key


// Shared-memory variant of the per-thread container: a binary MIN-heap of the k best keys, entry i at list[i * 128]
// (the 128 epilogue threads interleave their lists, so a warp's accesses are conflict-free). The root is the k-th
// best: a key that does not beat it is rejected with one load; an accepted key replaces the root and sifts down in
// <= log2(k) steps (an insertion-sorted list paid k/2 shifts per insert, which dominated the scan at k = 100).
// Unfilled slots hold the sentinel (smaller than any real key), so the heap needs no size counter.
// Returns the score of the new k-th best (-inf while the heap still holds sentinels).
// Lists are addressed as 32-bit shared-window addresses (explicit ld/st.shared: the kernels' 1024-byte-aligned
// dynamic-smem base has lost its address space, and generic LD.E/ST.E pay the window check on every heap step).
__device__ __forceinline__ uint64_t lds_u64(uint32_t addr) {
  uint64_t v;
  asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u64(uint32_t addr, uint64_t v) {
  asm volatile("st.shared.u64 [%0], %1;" ::"r"(addr), "l"(v) : "memory");
}
constexpr uint32_t kHeapStride = kBlockM * 8;  // bytes between consecutive entries of one thread's list

__device__ __noinline__ float heap_insert(uint32_t list, uint32_t k, uint64_t key, float thr) {
  if (key <= lds_u64(list)) return thr;
  uint32_t i = 0;
  for (;;) {
    const uint32_t l = 2 * i + 1;
    if (l >= k) break;
    uint32_t c = l;
    uint64_t cv = lds_u64(list + l * kHeapStride);
    if (l + 1 < k) {
      const uint64_t cr = lds_u64(list + (l + 1) * kHeapStride);
      if (cr < cv) {
        cv = cr;
        c = l + 1;
      }
    }
    if (cv >= key) break;
    sts_u64(list + i * kHeapStride, cv);
    i = c;
  }
  sts_u64(list + i * kHeapStride, key);
  return key_score(lds_u64(list));
}

// The per-thread top-k container. KREG == 0: min-heap in shared memory (any k that fits; stored UNSORTED — the merge
// kernel sorts). KREG > 0 (k <= KREG): the list
// lives in registers — an insert is ~KREG compare/select pairs with no memory latency, which matters because a warp
// runs the insert path whenever ANY of its 32 queries admits a row.
template <int KREG>
struct ThreadTopK {
  // right-aligned: the best entry is e[KREG - k], the k-th best is ALWAYS e[KREG - 1] (a fixed register); the unused
  // leading entries hold UINT64_MAX so nothing ever shifts into them. All indices below are compile-time constants.
  uint64_t e[KREG];
  uint32_t k;
  __device__ __forceinline__ void init(uint64_t*, uint32_t k_) {
    k = k_;
#pragma unroll
    for (int u = 0; u < KREG; ++u) e[u] = ((uint32_t)u + k >= (uint32_t)KREG) ? kSentinelKey : ~0ull;
  }
  __device__ __forceinline__ float insert(uint64_t key, float thr) {
    if (key <= e[KREG - 1]) return thr;
#pragma unroll
    for (int u = KREG - 1; u >= 1; --u) e[u] = (e[u - 1] < key) ? e[u - 1] : ((e[u] < key) ? key : e[u]);
    if (e[0] < key) e[0] = key;
    return key_score(e[KREG - 1]);
  }
  // out[jj] = jj-th best, sentinel (row 0xFFFFFFFF) -> 0 (empty)
  __device__ __forceinline__ void store(uint64_t* out) const {
#pragma unroll
    for (int u = 0; u < KREG; ++u)
      if ((uint32_t)u + k >= (uint32_t)KREG) out[(uint32_t)u + k - (uint32_t)KREG] = ((uint32_t)e[u] == 0u) ? 0ull : e[u];
  }
};
template <>
struct ThreadTopK<0> {
  uint32_t list;  // shared-window address of this thread's entry 0
  uint32_t k;
  __device__ __forceinline__ void init(uint64_t* l, uint32_t k_) {
    list = smem_u32(l);
    k = k_;
    for (uint32_t jj = 0; jj < k; ++jj) sts_u64(list + jj * kHeapStride, kSentinelKey);
  }
  __device__ __forceinline__ float insert(uint64_t key, float thr) { return heap_insert(list, k, key, thr); }
  __device__ __forceinline__ void store(uint64_t* out) const {
    for (uint32_t jj = 0; jj < k; ++jj) {
      const uint64_t key = lds_u64(list + jj * kHeapStride);
      out[jj] = ((uint32_t)key == 0u) ? 0ull : key;
    }
  }
};

}  // namespace
}  // namespace nvdb_b200
