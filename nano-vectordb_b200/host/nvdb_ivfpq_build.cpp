// nvdb_ivfpq_build (B200) — IVF-PQ index construction on the GPU with the reference tool's command line.
//
//   nvdb_ivfpq_build <base.vecbin> <out.nvdbivf>
//   Env: IVF_NLIST=4096  IVF_TRAIN=200000  PQ_M=48  PQ_BITS=8  USE_OPQ=0  OPQ_NITER=25  (+ IVF_NITER=25, IVF_SEED=1234, IVF_GPU=0)
//
// Same argv, environment variables, messages and exit codes as apps/nvdb_ivfpq_build.cpp:24-64 (1 usage, 2 "PQ_M must
// divide dim"). The reference builds faiss::IndexIVFPQ on the host cores and writes FAISS' own file format; here the
// base goes to the device once and nvdb_b200_ivfpq_build trains + encodes there (kmeans.cu / ivfpq.cu); the index is
// written in this repo's own container (ivfpq_file.h). USE_OPQ=1 trains the OPQ rotation on the device as well
// (OPQ_NITER rounds, the reference's default 25: apps/nvdb_ivfpq_build.cpp:84).
#include <chrono>
#include <cstdlib>
#include <iostream>
#include <string>

#include "dataset.h"
#include "flat_index_cuda.h"
#include "ivfpq_file.h"

static int getenv_int(const char* k, int defv) {
  const char* v = std::getenv(k);
  return v ? std::atoi(v) : defv;
}

int main(int argc, char** argv) {
  if (argc < 3) {
    std::cerr << "Usage: nvdb_ivfpq_build <base.vecbin> <out.nvdbivf>\n"
              << "Env:\n"
              << "  IVF_NLIST=4096\n"
              << "  IVF_TRAIN=200000\n"
              << "  PQ_M=48            (must divide dim)\n"
              << "  PQ_BITS=8          (8 only)\n"
              << "  USE_OPQ=0          (0/1; optional)\n"
              << "  OPQ_NITER=25\n"
              << "  IVF_NITER=25  IVF_SEED=1234  IVF_GPU=0\n";
    return 1;
  }
  const std::string base_path = argv[1], out_path = argv[2];
  const int nlist = getenv_int("IVF_NLIST", 4096);
  const int ntrain = getenv_int("IVF_TRAIN", 200000);
  const int pq_m = getenv_int("PQ_M", 48);
  const int pq_bits = getenv_int("PQ_BITS", 8);
  const int use_opq = getenv_int("USE_OPQ", 0);
  try {
    nvdb::VectorDataset base;
    base.load(base_path);
    const uint32_t d = base.dim();
    const uint64_t N = base.count();
    const uint64_t train_n = std::min<uint64_t>(N, (uint64_t)std::max(ntrain, 0));
    if (pq_m <= 0 || (int)d % pq_m != 0) {
      std::cerr << "PQ_M must divide dim. dim=" << d << " PQ_M=" << pq_m << "\n";
      return 2;
    }
    std::cout << "Base: N=" << N << " d=" << d << " dtype=" << base.dtype() << " nlist=" << nlist << " ntrain=" << train_n
              << " PQ_M=" << pq_m << " PQ_BITS=" << pq_bits << " USE_OPQ=" << use_opq << "\n";
    nvdb::FlatIndexCuda flat(&base, getenv_int("IVF_GPU", 0));
    nvdb_b200_ivfpq* ivf = nullptr;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = nvdb_b200_ivfpq_build_opq(flat.handle(), (uint32_t)nlist, (uint32_t)pq_m, (uint32_t)pq_bits, train_n,
                                             getenv_int("IVF_NITER", 25), use_opq ? getenv_int("OPQ_NITER", 25) : 0,
                                             (uint64_t)getenv_int("IVF_SEED", 1234), &ivf);
    if (rc != NVDB_B200_OK) {
      std::cerr << "Error: " << nvdb_b200_ivfpq_last_error(nullptr) << "\n";
      return 10;
    }
    const auto t1 = std::chrono::steady_clock::now();
    std::cout << "Train+Add: " << std::chrono::duration<double>(t1 - t0).count() << " s\n";
    nvdb::ivfpq_save(ivf, out_path);
    nvdb_b200_ivfpq_destroy(ivf);
    std::cout << "Saved IVF-" << (use_opq ? "OPQ-PQ" : "PQ") << " index: " << out_path << "\n";
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return 10;
  }
}
