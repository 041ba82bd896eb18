// flat_index_cuda.h — C++ host mirror of the reference's index API over the C ABI (include/nvdb_b200.h).
//
//   nvdb::FlatIndexCuda            ~ nvdb::FlatIndex / FlatIndexOMP / FlatIndexPool
//                                    (include/nvdb/flat_index.h:9-20, flat_index_pool.h:13-25): constructed from a
//                                    non-owning `const VectorDataset*`, `search_topk_dot(q, k)` returns
//                                    `std::vector<SearchResult>` sorted by score descending, length min(k, N);
//                                    "Empty base" / "Null query" are std::runtime_error, k == 0 -> {}.
//   search_topk_dot_batch          ~ the bench-side batched scan (apps/nvdb_bench.cpp:47-159) as an index method.
//   nvdb::cuda_l2_topk_batch(...)  ~ include/nvdb/cuda_refine.h:25-38, same argument list and output conventions
//                                    (ascending L2^2, 0xFFFFFFFF / 1e30f padding, CudaRefineTiming), but errors are
//                                    exceptions instead of exit() and the device copy of the base is an explicit,
//                                    cached index handle keyed by pointer AND shape (the reference's cache ignores
//                                    the pointer, src/cuda_refine.cu:188-203).
// No compute happens here: every call lands in libnvdb_b200.so.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "dataset.h"
#include "nvdb_b200.h"

namespace nvdb {

struct SearchResult {  // include/nvdb/topK.h:9-12
  uint64_t id;
  float score;
};

struct CudaRefineTiming {  // include/nvdb/cuda_refine.h:7-22, field for field
  float h2d_ms = 0, kernel_ms = 0, d2h_ms = 0, total_ms = 0;
  uint32_t threads = 0;
  uint32_t nwarps = 0;
  uint32_t K = 0;
  uint32_t R = 0;
  size_t shmem_bytes = 0;
  // debug timing (clock64 sample; CUDA_DBG_TIMING=1, CUDA_DBG_Q=<n>). Phases of this kernel: "write" = the query into
  // shared memory, "dist" = gather + distances, "merge" = selection + output (include/nvdb_b200.h).
  uint32_t dbg_q = 0;
  double dbg_dist_cycles_avg = 0.0;
  double dbg_write_cycles_avg = 0.0;
  double dbg_merge_cycles_avg = 0.0;
  double dbg_dist_pct = 0.0;
  double dbg_write_pct = 0.0;
  double dbg_merge_pct = 0.0;
};

namespace detail {
inline void check(int status, const nvdb_b200_index* h) {
  if (status != NVDB_B200_OK) {
    const char* m = nvdb_b200_last_error(h);
    throw std::runtime_error((m && *m) ? std::string(m) : std::string(nvdb_b200_status_string(status)));
  }
}
}  // namespace detail

class FlatIndexCuda {
 public:
  // Copies the dataset (typically the mmap of a vecbin file) to `device`. The dataset is not referenced afterwards.
  explicit FlatIndexCuda(const VectorDataset* base, int device = 0, uint64_t row_offset = 0) {
    if (!base || base->count() == 0) throw std::runtime_error("Empty base");
    detail::check(nvdb_b200_index_create(base->data_ptr(), base->scales_ptr(), base->dtype(), base->count(), base->dim(),
                                         device, row_offset, &h_),
                  nullptr);
    n_ = base->count();
    dim_ = base->dim();
  }
  // Raw host rows (used by cuda_l2_topk_batch).
  FlatIndexCuda(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim, int device = 0) {
    detail::check(nvdb_b200_index_create(payload, scales, dtype, n, dim, device, 0, &h_), nullptr);
    n_ = n;
    dim_ = dim;
  }
  ~FlatIndexCuda() { nvdb_b200_index_destroy(h_); }
  FlatIndexCuda(const FlatIndexCuda&) = delete;
  FlatIndexCuda& operator=(const FlatIndexCuda&) = delete;

  uint64_t count() const { return n_; }
  uint32_t dim() const { return dim_; }
  nvdb_b200_index* handle() const { return h_; }
  void set_path(int path) { detail::check(nvdb_b200_index_set_path(h_, path), h_); }

  std::vector<SearchResult> search_topk_dot(const float* q, uint32_t k) const {
    auto r = search_topk_dot_batch(q, 1, k);
    return r.empty() ? std::vector<SearchResult>{} : std::move(r[0]);
  }

  // queries: nq * dim floats. One result list per query, each of length min(k, N).
  std::vector<std::vector<SearchResult>> search_topk_dot_batch(const float* queries, uint32_t nq, uint32_t k,
                                                               nvdb_b200_timing* timing = nullptr) const {
    if (!queries) throw std::runtime_error("Null query");
    std::vector<std::vector<SearchResult>> out(nq);
    if (k == 0 || nq == 0) return out;
    std::vector<uint64_t> ids((size_t)nq * k);
    std::vector<float> sc((size_t)nq * k);
    detail::check(nvdb_b200_search(h_, queries, nq, k, ids.data(), sc.data(), timing), h_);
    const uint32_t kk = (uint64_t)k < n_ ? k : (uint32_t)n_;  // src/flat_index.cpp:24
    for (uint32_t qi = 0; qi < nq; ++qi) {
      out[qi].resize(kk);
      for (uint32_t j = 0; j < kk; ++j) out[qi][j] = SearchResult{ids[(size_t)qi * k + j], sc[(size_t)qi * k + j]};
    }
    return out;
  }

  void refine_l2(const float* queries, const uint32_t* cand_ids, uint32_t Q, uint32_t R, uint32_t K,
                 std::vector<uint32_t>& out_ids, std::vector<float>& out_dist, nvdb_b200_timing* timing = nullptr,
                 bool want_dist = true) const {
    out_ids.assign((size_t)Q * K, 0xFFFFFFFFu);  // src/cuda_refine.cu:891-894
    out_dist.assign((size_t)Q * K, 1e30f);
    if (Q == 0 || R == 0 || K == 0) return;
    detail::check(nvdb_b200_refine_l2(h_, queries, cand_ids, Q, R, K, out_ids.data(), want_dist ? out_dist.data() : nullptr,
                                      timing),
                  h_);
  }

 private:
  nvdb_b200_index* h_ = nullptr;
  uint64_t n_ = 0;
  uint32_t dim_ = 0;
};

// Same signature as the reference's refine entry point. base_dtype: 1 = f32, 2 = f16 (3 = i8 is not expressible through
// this signature because it carries no scales; use FlatIndexCuda::refine_l2). The reference's environment switches are
// honoured (src/cuda_refine.cu:863-876): CUDA_RETURN_DIST=0 -> ids only (out_dist is cleared), CUDA_PINNED=1 -> the
// caller's query / candidate buffers are page-locked (once per buffer) and read by the copy engine in place,
// CUDA_DBG_TIMING=1 (+ CUDA_DBG_Q) -> in-kernel phase sampling into the timing struct.
inline void cuda_l2_topk_batch(const void* base_ptr, uint32_t base_dtype, uint64_t N, uint32_t D,
                               const float* queries_f32, const uint32_t* cand_ids, uint32_t Q, uint32_t R, uint32_t K,
                               std::vector<uint32_t>& out_ids, std::vector<float>& out_dist,
                               CudaRefineTiming* timing = nullptr) {
  struct Cache {
    std::mutex mu;
    const void* ptr = nullptr;
    uint32_t dtype = 0, dim = 0;
    uint64_t n = 0;
    std::unique_ptr<FlatIndexCuda> idx;
    const void *reg_q = nullptr, *reg_c = nullptr;  // buffers registered for CUDA_PINNED=1
    size_t reg_q_bytes = 0, reg_c_bytes = 0;
    ~Cache() {
      if (reg_q) nvdb_b200_host_unregister(const_cast<void*>(reg_q));
      if (reg_c) nvdb_b200_host_unregister(const_cast<void*>(reg_c));
    }
  };
  static Cache c;
  if (timing) *timing = CudaRefineTiming{};
  if (base_dtype != 1 && base_dtype != 2) throw std::runtime_error("Unsupported base dtype (Float32/Float16 only)");
  auto env_int = [](const char* k, int defv) {
    const char* v = std::getenv(k);
    return v ? std::atoi(v) : defv;
  };
  const bool want_dist = env_int("CUDA_RETURN_DIST", 1) != 0;
  const bool pinned = env_int("CUDA_PINNED", 0) != 0;
  std::lock_guard<std::mutex> g(c.mu);
  if (K == 0 || Q == 0 || R == 0) {  // src/cuda_refine.cu:853-857
    out_ids.clear();
    out_dist.clear();
    return;
  }
  if (!c.idx || c.ptr != base_ptr || c.dtype != base_dtype || c.n != N || c.dim != D) {
    c.idx.reset();
    c.idx = std::make_unique<FlatIndexCuda>(base_ptr, nullptr, base_dtype, N, D, 0);
    c.ptr = base_ptr;
    c.dtype = base_dtype;
    c.n = N;
    c.dim = D;
  }
  if (pinned) {
    auto pin = [](const void*& have, size_t& have_bytes, const void* want, size_t bytes) {
      if (have == want && have_bytes >= bytes) return;
      if (have) nvdb_b200_host_unregister(const_cast<void*>(have));
      have = nullptr;
      if (nvdb_b200_host_register(const_cast<void*>(want), bytes) == NVDB_B200_OK) {
        have = want;
        have_bytes = bytes;
      }
    };
    pin(c.reg_q, c.reg_q_bytes, queries_f32, (size_t)Q * D * sizeof(float));
    pin(c.reg_c, c.reg_c_bytes, cand_ids, (size_t)Q * R * sizeof(uint32_t));
  }
  nvdb_b200_timing t{};
  c.idx->refine_l2(queries_f32, cand_ids, Q, R, K, out_ids, out_dist, &t, want_dist);
  if (!want_dist) out_dist.clear();  // src/cuda_refine.cu:893-894
  if (timing) {
    timing->h2d_ms = t.h2d_ms;
    timing->kernel_ms = t.kernel_ms;
    timing->d2h_ms = t.d2h_ms;
    timing->total_ms = t.total_ms;
    timing->threads = t.threads;
    timing->nwarps = t.threads / 32;
    timing->K = K;
    timing->R = R;
    timing->shmem_bytes = t.shmem_bytes;
    timing->dbg_q = t.dbg_q;
    timing->dbg_dist_cycles_avg = t.dbg_dist_cycles_avg;
    timing->dbg_write_cycles_avg = t.dbg_write_cycles_avg;
    timing->dbg_merge_cycles_avg = t.dbg_merge_cycles_avg;
    const double tot = timing->dbg_dist_cycles_avg + timing->dbg_write_cycles_avg + timing->dbg_merge_cycles_avg;
    if (tot > 0.0) {
      timing->dbg_dist_pct = timing->dbg_dist_cycles_avg / tot;
      timing->dbg_write_pct = timing->dbg_write_cycles_avg / tot;
      timing->dbg_merge_pct = timing->dbg_merge_cycles_avg / tot;
    }
  }
}

}  // namespace nvdb
