// oracle/ref_shim.cpp — TEST INFRASTRUCTURE ONLY.
//
// A C-ABI door into the UNMODIFIED reference (compiled from /root/reference/src/*.cpp by oracle/Makefile `ref`).
// It contains no algorithm of its own: every function forwards to a reference symbol.
//   ref_dot_*        -> nvdb::dot_f32 / dot_f32_f16base / dot_f32_i8base   (reference src/simd_dot.cpp:52,127,202)
//   ref_dataset_*    -> nvdb::VectorDataset::load + accessors             (reference src/vector_dataset.cpp:24-157)
//   ref_search       -> FlatIndex / FlatIndexOMP / FlatIndexAsync / FlatIndexPool ::search_topk_dot
//                       (reference src/flat_index.cpp:16, flat_index_omp.cpp:16, flat_index_async.cpp:22,
//                        flat_index_pool.cpp:194)
// Used by tests/ (to pin the C restatement) and by bench.py's cpu_baseline / --impl reference legs.

#include "nvdb/flat_index.h"
#include "nvdb/flat_index_async.h"
#include "nvdb/flat_index_omp.h"
#include "nvdb/flat_index_pool.h"
#include "nvdb/simd_dot.h"
#include "nvdb/vector_dataset.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {
struct RefHandle {
  nvdb::VectorDataset ds;
  std::unique_ptr<nvdb::FlatIndexPool> pool;
  int pool_threads = 0;
};
thread_local std::string g_err;
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }

float ref_dot_f32(const float* a, const float* b, uint32_t dim) { return nvdb::dot_f32(a, b, dim); }
float ref_dot_f16(const float* q, const uint16_t* x, uint32_t dim) { return nvdb::dot_f32_f16base(q, x, dim); }
float ref_dot_i8(const float* q, const int8_t* x, uint32_t dim, float scale) {
  return nvdb::dot_f32_i8base(q, x, dim, scale);
}
void ref_set_force_scalar(int v) { nvdb::set_force_scalar(v != 0); }

void* ref_dataset_load(const char* path) {
  try {
    auto* h = new RefHandle();
    h->ds.load(path);
    return h;
  } catch (const std::exception& e) {
    g_err = e.what();
    return nullptr;
  }
}
void ref_dataset_free(void* h) { delete static_cast<RefHandle*>(h); }
uint64_t ref_dataset_count(void* h) { return static_cast<RefHandle*>(h)->ds.count(); }
uint32_t ref_dataset_dim(void* h) { return static_cast<RefHandle*>(h)->ds.dim(); }
uint32_t ref_dataset_dtype(void* h) { return static_cast<RefHandle*>(h)->ds.dtype(); }

// mode: 0=st 1=omp 2=async 3=pool. Returns number of results written (<= k), or -1 on exception.
int64_t ref_search(void* hv, int mode, int threads, const float* q, uint32_t k, uint64_t* out_ids,
                   float* out_scores) {
  auto* h = static_cast<RefHandle*>(hv);
  try {
    std::vector<nvdb::SearchResult> r;
    if (mode == 0) {
      r = nvdb::FlatIndex(&h->ds).search_topk_dot(q, k);
    } else if (mode == 1) {
      r = nvdb::FlatIndexOMP(&h->ds).search_topk_dot(q, k);
    } else if (mode == 2) {
      r = nvdb::FlatIndexAsync(&h->ds).search_topk_dot(q, k, threads);
    } else if (mode == 3) {
      if (!h->pool || h->pool_threads != threads) {
        h->pool.reset();
        h->pool = std::make_unique<nvdb::FlatIndexPool>(&h->ds, threads);
        h->pool_threads = threads;
      }
      r = h->pool->search_topk_dot(q, k);
    } else {
      g_err = "bad mode";
      return -1;
    }
    for (size_t i = 0; i < r.size(); ++i) {
      out_ids[i] = r[i].id;
      out_scores[i] = r[i].score;
    }
    return (int64_t)r.size();
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

}  // extern "C"
