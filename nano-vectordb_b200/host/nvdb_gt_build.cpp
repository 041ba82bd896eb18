// nvdb_gt_build (B200) — exact ground-truth ids for a query set, byte-compatible with the reference tool.
//
//   nvdb_gt_build <base.vecbin> <query.vecbin> <k> <out.gtbin>
//
// Same argv, messages, exit codes and output file as apps/nvdb_gt_build.cpp:22-129 of the reference: a 64-byte
// GtBinHeader ("NVDBGT01", version 1, metric 1 = DotEquivalentL2, k, dim, Q, N; include/nvdb/gtbin_format.h:18-35)
// followed by uint32 ids[Q * k] in rank order. The reference computes each query with FlatIndex / FlatIndexOMP on the
// host cores (GT_MODE=st|omp); here the whole query set goes through nvdb::FlatIndexCuda in batches (the exact
// tensor-core path for f16 / i8 bases, the streaming scan for f32), whose ids are identical to the reference's except
// inside exact-score tie groups, where the order is canonical (id ascending). Env: GT_BATCH (default 1024), GT_GPU (0).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "dataset.h"
#include "flat_index_cuda.h"

namespace {
#pragma pack(push, 1)
struct GtBinHeader {
  uint64_t magic;
  uint32_t version;
  uint32_t metric;
  uint32_t k;
  uint32_t dim;
  uint64_t Q;
  uint64_t N;
  uint8_t reserved[64 - (8 + 4 + 4 + 4 + 4 + 8 + 8)];
};
#pragma pack(pop)
static_assert(sizeof(GtBinHeader) == 64, "GtBinHeader must be 64 bytes");
constexpr uint64_t kGtMagic = 0x4E56444247543031ULL;  // "NVDBGT01"

int getenv_int(const char* name, int def) {
  const char* v = std::getenv(name);
  return v ? std::atoi(v) : def;
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 5) {
    std::cerr << "Usage: nvdb_gt_build <base.vecbin> <query.vecbin> <k> <out.gtbin>\n"
              << "Notes:\n"
              << "  - Produces GT ids for L2 ranking by computing exact top-k dot products on the GPU.\n"
              << "    This is valid because embeddings are assumed L2-normalized: L2^2 = 2 - 2*dot.\n"
              << "Env:\n"
              << "  GT_BATCH=1024    queries per GPU batch\n"
              << "  GT_GPU=0         device ordinal\n";
    return 1;
  }
  try {
    const std::string base_path = argv[1], query_path = argv[2], out_path = argv[4];
    const uint32_t k = static_cast<uint32_t>(std::stoul(argv[3]));
    const uint32_t batch = (uint32_t)std::max(1, getenv_int("GT_BATCH", 1024));
    const int gpu = getenv_int("GT_GPU", 0);

    nvdb::VectorDataset base, query;
    base.load(base_path);
    query.load(query_path);
    if (query.dtype() != 1) {
      std::cerr << "Query must be float32 vecbin (got dtype=" << query.dtype() << ")\n";
      return 2;
    }
    if (base.dim() != query.dim()) {
      std::cerr << "Dim mismatch: base.dim=" << base.dim() << " query.dim=" << query.dim() << "\n";
      return 3;
    }
    if (k == 0) {
      std::cerr << "k must be > 0\n";
      return 4;
    }
    const uint64_t N = base.count(), Q = query.count();
    const uint32_t d = base.dim();
    std::cout << "GT build: N=" << N << " Q=" << Q << " d=" << d << " k=" << k << " mode=cuda\n";

    nvdb::FlatIndexCuda index(&base, gpu);
    std::vector<uint32_t> gt_ids(static_cast<size_t>(Q) * k, 0);
    for (uint64_t q0 = 0; q0 < Q; q0 += batch) {
      const uint32_t nq = (uint32_t)std::min<uint64_t>(batch, Q - q0);
      auto res = index.search_topk_dot_batch(query.vector_ptr_f32(q0), nq, k);
      for (uint32_t qi = 0; qi < nq; ++qi) {
        if (res[qi].size() != k) {  // base smaller than k; the reference is strict about it too
          std::cerr << "GT size mismatch at qi=" << (q0 + qi) << " got=" << res[qi].size() << " expected=" << k << "\n";
          return 5;
        }
        for (uint32_t j = 0; j < k; ++j) gt_ids[(q0 + qi) * k + j] = (uint32_t)res[qi][j].id;
      }
      if ((q0 + nq) % 200 == 0 || q0 + nq == Q) std::cout << "GT " << (q0 + nq) << "/" << Q << "\n";
    }

    GtBinHeader h{};
    h.magic = kGtMagic;
    h.version = 1;
    h.metric = 1;
    h.k = k;
    h.dim = d;
    h.Q = Q;
    h.N = N;
    std::ofstream out(out_path, std::ios::binary);
    if (!out) {
      std::cerr << "Failed to open output: " << out_path << "\n";
      return 6;
    }
    out.write(reinterpret_cast<const char*>(&h), sizeof(h));
    out.write(reinterpret_cast<const char*>(gt_ids.data()), static_cast<std::streamsize>(gt_ids.size() * sizeof(uint32_t)));
    out.close();
    std::cout << "Wrote GT: " << out_path << " (header=64B, payload=" << (gt_ids.size() * sizeof(uint32_t)) << " bytes)\n";
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return 10;
  }
}
