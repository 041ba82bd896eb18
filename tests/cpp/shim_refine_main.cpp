// Test binary for the drop-in symbol nvdb::cuda_l2_topk_batch (nano-vectordb_b200/host/flat_index_cuda.h): called with
// the exact argument list of the reference's call site (apps/nvdb_ivf_eval.cpp:542-549) on inputs read from files; the
// outputs go back to files for the Python test, which compares them with the reference's own kernel.
//   shim_refine_main <base.f16> <queries.f32> <cand.u32> <N> <D> <Q> <R> <K> <out_ids.u32> <out_dist.f32>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <vector>

#include "flat_index_cuda.h"

template <class T>
static std::vector<T> slurp(const char* path, size_t count) {
  std::vector<T> v(count);
  std::ifstream in(path, std::ios::binary);
  in.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(count * sizeof(T)));
  if (!in) {
    std::cerr << "short read: " << path << "\n";
    std::exit(2);
  }
  return v;
}

int main(int argc, char** argv) {
  if (argc < 11) return 1;
  const uint64_t N = std::strtoull(argv[4], nullptr, 10);
  const uint32_t D = (uint32_t)std::atoi(argv[5]), Q = (uint32_t)std::atoi(argv[6]), R = (uint32_t)std::atoi(argv[7]),
                 K = (uint32_t)std::atoi(argv[8]);
  auto base = slurp<uint16_t>(argv[1], (size_t)N * D);
  auto queries = slurp<float>(argv[2], (size_t)Q * D);
  auto cand = slurp<uint32_t>(argv[3], (size_t)Q * R);
  try {
    std::vector<uint32_t> topk_ids;
    std::vector<float> topk_dist;
    nvdb::CudaRefineTiming t{};
    for (int rep = 0; rep < 2; ++rep)  // second call: the cached index handle and the registered buffers are reused
      nvdb::cuda_l2_topk_batch(base.data(), /*base_dtype=*/2, N, D, queries.data(), cand.data(), Q, R, K, topk_ids, topk_dist, &t);
    std::ofstream(argv[9], std::ios::binary).write(reinterpret_cast<const char*>(topk_ids.data()), (std::streamsize)(topk_ids.size() * 4));
    std::ofstream(argv[10], std::ios::binary).write(reinterpret_cast<const char*>(topk_dist.data()), (std::streamsize)(topk_dist.size() * 4));
    std::printf("ids=%zu dist=%zu h2d_ms=%.4f kernel_ms=%.4f d2h_ms=%.4f total_ms=%.4f threads=%u nwarps=%u K=%u R=%u shmem=%zu dbg_q=%u "
                "dbg_dist=%.1f dbg_write=%.1f dbg_merge=%.1f\n",
                topk_ids.size(), topk_dist.size(), t.h2d_ms, t.kernel_ms, t.d2h_ms, t.total_ms, t.threads, t.nwarps, t.K, t.R,
                t.shmem_bytes, t.dbg_q, t.dbg_dist_cycles_avg, t.dbg_write_cycles_avg, t.dbg_merge_cycles_avg);
    // K*Q*R == 0 clears the outputs (src/cuda_refine.cu:853-857)
    nvdb::cuda_l2_topk_batch(base.data(), 2, N, D, queries.data(), cand.data(), 0, R, K, topk_ids, topk_dist, nullptr);
    if (!topk_ids.empty() || !topk_dist.empty()) return 3;
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return 4;
  }
}
