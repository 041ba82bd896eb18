// nvdb_bench (mode=cuda) — the reference's benchmark harness for the B200 path.
//
// Same positional arguments and the same stdout keys as the reference harness (apps/nvdb_bench.cpp:254-432):
//   nvdb_bench <base.vecbin> <query.vecbin> <k> [mode=cuda] [gpus=1] [warmup=5] [batch_q=1] [path=auto]
// `threads` of the reference becomes `gpus` (row shards, one per device); `tile_vecs` / `prefetch_dist` have no
// meaning on the GPU and are replaced by `path` = auto | gemv | mma | mma_fast | mma_exact. Per-query (or per-batch) latencies are
// std::chrono::steady_clock around the host-buffer search call, i.e. end to end incl. H2D of the queries and D2H of
// the results; percentiles use the reference's linear interpolation (:370-377); bytes_per_query /
// payload_equiv_bandwidth_GBps as in :414-418. Extra keys: gt_match (when GT_PATH names a gtbin written by the
// reference's nvdb_gt_build), kernel_ms_avg, path_used.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <memory>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "dataset.h"
#include "flat_index_cuda.h"

namespace {

double ms_since(std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
  return std::chrono::duration<double, std::milli>(b - a).count();
}

// Row shards on `gpus` devices of this process; one host thread per device issues its shard's search, and the shards
// exchange and merge their lists on the devices over NVLink peer memory (csrc/exchange.cu; the reference merges its
// row blocks serially on the calling thread, src/flat_index_pool.cpp:212-214) — the host only receives shard 0's copy of
// the final top-k. One shard == plain FlatIndexCuda.
class ShardedIndex {
 public:
  // shard g lives on device g % ndev (ndev < gpus only with NVDB_BENCH_OVERSUBSCRIBE=1: several shards share a device)
  ShardedIndex(const nvdb::VectorDataset& base, int gpus, int path, int ndev) {
    const uint64_t n = base.count(), per = (n + gpus - 1) / gpus;
    const size_t row_bytes = (size_t)base.dim() * nvdb::bytes_per_elem(base.dtype());
    for (int g = 0; g < gpus; ++g) {
      const uint64_t r0 = std::min<uint64_t>(n, (uint64_t)g * per), nr = std::min<uint64_t>(n, r0 + per) - r0;
      if (nr == 0) continue;
      nvdb_b200_index* h = nullptr;
      const char* payload = static_cast<const char*>(base.data_ptr()) + r0 * row_bytes;
      const float* scales = base.scales_ptr() ? base.scales_ptr() + r0 : nullptr;
      nvdb::detail::check(nvdb_b200_index_create(payload, scales, base.dtype(), nr, base.dim(), g % ndev, r0, &h), nullptr);
      nvdb::detail::check(nvdb_b200_index_set_path(h, path), h);
      shards_.push_back(h);
      devices_.push_back(g % ndev);
    }
    shared_device_ = (int)shards_.size() > ndev;
    n_ = n;
  }
  ~ShardedIndex() {
    drop_exchange();
    for (auto* h : shards_) nvdb_b200_index_destroy(h);
  }
  // ids/scores: nq*k, padded like nvdb_b200_search
  void search(const float* q, uint32_t nq, uint32_t k, std::vector<uint64_t>& ids, std::vector<float>& sc,
              nvdb_b200_timing* t0) {
    const size_t S = shards_.size(), nk = (size_t)nq * k;
    ids.resize(nk);
    sc.resize(nk);
    if (S == 1) {
      nvdb::detail::check(nvdb_b200_search(shards_[0], q, nq, k, ids.data(), sc.data(), t0), shards_[0]);
      return;
    }
    ensure_exchange(8 * (nk + (nk + 1) / 2));
    // A shard's first call of a shape grows its workspaces, and cudaMalloc / cudaFree wait for the whole device —
    // including, when shards share one device (NVDB_BENCH_OVERSUBSCRIBE), the other shard's kernel, which is spinning
    // for this shard's result (and with peer access enabled a free may wait on the peers as well). Grow everything with
    // one plain search per shard first (no exchange, one shard at a time): the library sizes every workspace of a
    // (nq, k) shape for its worst case on that first call (capi.cu: reserve_exact_ws), so no later search of the shape
    // allocates (NVDB_B200_XCHG_NO_ALLOC=1 turns that into a checked error).
    if (warmed_.insert({nq, k}).second) {
      std::vector<uint64_t> wi(nk);
      std::vector<float> wsc(nk);
      for (auto* h : shards_) nvdb::detail::check(nvdb_b200_search(h, q, nq, k, wi.data(), wsc.data(), nullptr), h);
    }
    if (shared_device_) {
      // Shards sharing a device (NVDB_BENCH_OVERSUBSCRIBE): kernels that wait for one another must not be separate
      // launches on one GPU, so every shard scans + pushes first and only then every shard waits + merges.
      for (size_t s = 0; s < S; ++s)
        nvdb::detail::check(nvdb_b200_search_exchange_split(shards_[s], xchg_[s], q, nq, k, nullptr, nullptr, 1), shards_[s]);
      for (size_t s = 0; s < S; ++s)
        nvdb::detail::check(nvdb_b200_search_exchange_split(shards_[s], xchg_[s], q, nq, k, s == 0 ? ids.data() : nullptr,
                                                            s == 0 ? sc.data() : nullptr, 2), shards_[s]);
      if (t0) *t0 = nvdb_b200_timing{};
      return;
    }
    std::vector<std::thread> th;
    std::vector<int> rc(S, 0);
    std::vector<nvdb_b200_timing> tt(S);
    for (size_t s = 0; s < S; ++s)
      th.emplace_back([&, s] {
        rc[s] = nvdb_b200_search_exchange(shards_[s], xchg_[s], q, nq, k, s == 0 ? ids.data() : nullptr,
                                          s == 0 ? sc.data() : nullptr, &tt[s]);
      });
    for (auto& x : th) x.join();
    for (size_t s = 0; s < S; ++s) nvdb::detail::check(rc[s], shards_[s]);
    if (t0) *t0 = tt[0];
  }
  size_t shards() const { return shards_.size(); }

 private:
  void drop_exchange() {
    for (auto* x : xchg_) nvdb_b200_exchange_destroy(x);
    xchg_.clear();
    slot_bytes_ = 0;
  }
  // windows of all shards, mapped into each other by plain device pointers (one process)
  void ensure_exchange(size_t need) {
    if (need <= slot_bytes_) return;
    drop_exchange();
    size_t slot = 1 << 16;
    while (slot < need) slot <<= 1;
    const uint32_t S = (uint32_t)shards_.size();
    std::vector<void*> wins(S, nullptr);
    for (uint32_t s = 0; s < S; ++s) {
      nvdb_b200_exchange* x = nullptr;
      if (nvdb_b200_exchange_create(devices_[s], S, s, slot, &x) != NVDB_B200_OK)
        throw std::runtime_error(std::string("exchange_create: ") + nvdb_b200_exchange_last_error(nullptr));
      xchg_.push_back(x);
      nvdb_b200_exchange_window(x, &wins[s], nullptr);
    }
    for (uint32_t s = 0; s < S; ++s)
      if (nvdb_b200_exchange_connect(xchg_[s], nullptr, wins.data()) != NVDB_B200_OK)
        throw std::runtime_error(std::string("exchange_connect: ") + nvdb_b200_exchange_last_error(xchg_[s]));
    slot_bytes_ = slot;
  }

  std::vector<nvdb_b200_index*> shards_;
  std::vector<int> devices_;
  std::vector<nvdb_b200_exchange*> xchg_;
  size_t slot_bytes_ = 0;
  uint64_t n_ = 0;
  bool shared_device_ = false;
  std::set<std::pair<uint32_t, uint32_t>> warmed_;  // (nq, k) shapes whose workspaces have been grown
};

// gtbin (include/nvdb/gtbin_format.h:18-35): 64-byte header {u64 magic "NVDBGT01", u32 version, u32 metric, u32 k,
// u32 dim, u64 Q, u64 N, pad} + u32 ids[Q*k]
bool load_gtbin(const std::string& path, uint32_t* k, uint64_t* Q, std::vector<uint32_t>* ids) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  unsigned char h[64];
  f.read(reinterpret_cast<char*>(h), 64);
  if (!f) return false;
  uint64_t magic;
  std::memcpy(&magic, h, 8);
  if (magic != 0x4E56444247543031ULL) return false;
  std::memcpy(k, h + 16, 4);
  std::memcpy(Q, h + 24, 8);
  ids->resize((size_t)*Q * *k);
  f.read(reinterpret_cast<char*>(ids->data()), (std::streamsize)ids->size() * 4);
  return (bool)f;
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 4) {
    std::cerr << "Usage: nvdb_bench <base.vecbin> <query.vecbin> <k> [mode=cuda] [gpus=1] [warmup=5] [batch_q=1] "
                 "[path=auto|gemv|mma|mma_fast|mma_exact]\n";
    return 1;
  }
  try {
    const std::string base_path = argv[1], query_path = argv[2];
    const uint32_t k = (uint32_t)std::stoul(argv[3]);
    const std::string mode = argc >= 5 ? argv[4] : "cuda";
    int gpus = argc >= 6 ? std::stoi(argv[5]) : 1;
    const int warmup = argc >= 7 ? std::stoi(argv[6]) : 5;
    const int batch_q = std::max(1, argc >= 8 ? std::stoi(argv[7]) : 1);
    const std::string path_s = argc >= 9 ? argv[8] : "auto";
    if (mode != "cuda") {
      std::cerr << "mode=" << mode << " is one of the reference's CPU scans; this harness drives mode=cuda only\n";
      return 1;
    }
    int path = NVDB_B200_PATH_AUTO;
    if (path_s == "gemv") path = NVDB_B200_PATH_GEMV;
    else if (path_s == "mma") path = NVDB_B200_PATH_MMA;
    else if (path_s == "mma_fast") path = NVDB_B200_PATH_MMA_FAST;
    else if (path_s == "mma_exact") path = NVDB_B200_PATH_MMA_EXACT;
    else if (path_s != "auto") throw std::runtime_error("Unknown path: " + path_s);
    nvdb::VectorDataset base, query;
    base.load(base_path);
    query.load(query_path);
    if (base.dim() != query.dim()) {
      std::cerr << "Dim mismatch: base.dim=" << base.dim() << ", query.dim=" << query.dim() << "\n";
      return 2;
    }
    if (query.dtype() != 1) throw std::runtime_error("Dataset is not float32");  // queries must be f32 (raw12 / vecbin f32)
    int ndev = 0;
    nvdb::detail::check(nvdb_b200_device_count(&ndev), nullptr);
    if (gpus <= 0) gpus = ndev;
    const char* over = std::getenv("NVDB_BENCH_OVERSUBSCRIBE");
    if (gpus > ndev && !(over && over[0] == '1')) throw std::runtime_error("gpus=" + std::to_string(gpus) + " but only " + std::to_string(ndev) + " CUDA devices");
    std::cout << "mode=" << mode << " gpus=" << gpus << "\n";
    const auto tu0 = std::chrono::steady_clock::now();
    ShardedIndex index(base, gpus, path, ndev);
    const double upload_ms = ms_since(tu0, std::chrono::steady_clock::now());
    std::cout << "Base count=" << base.count() << " dim=" << base.dim() << " | Query count=" << query.count()
              << " | k=" << k << " | warmup=" << warmup << "\n";

    std::vector<uint64_t> ids;
    std::vector<float> sc;
    for (int i = 0; i < warmup; ++i) index.search(query.vector_ptr(0), 1, k, ids, sc, nullptr);

    const uint64_t Q = query.count();
    std::vector<double> lat_ms;
    std::vector<uint64_t> all_ids((size_t)Q * k);
    volatile float sink = 0.0f;
    double kernel_ms_sum = 0;
    uint32_t path_used = 0;
    const auto t_all0 = std::chrono::steady_clock::now();
    for (uint64_t qb = 0; qb < Q; qb += (uint64_t)batch_q) {
      const uint32_t b = (uint32_t)std::min<uint64_t>((uint64_t)batch_q, Q - qb);
      nvdb_b200_timing t{};
      const auto t0 = std::chrono::steady_clock::now();
      index.search(query.vector_ptr_f32(qb), b, k, ids, sc, &t);  // query rows are contiguous in the file
      lat_ms.push_back(ms_since(t0, std::chrono::steady_clock::now()));
      kernel_ms_sum += t.kernel_ms;
      path_used = t.path;
      for (uint32_t qi = 0; qi < b; ++qi) {
        sink = sink + sc[(size_t)qi * k];
        std::copy(ids.begin() + (size_t)qi * k, ids.begin() + (size_t)(qi + 1) * k, all_ids.begin() + (size_t)(qb + qi) * k);
      }
    }
    const double total_ms = ms_since(t_all0, std::chrono::steady_clock::now());

    std::sort(lat_ms.begin(), lat_ms.end());
    auto pct = [&](double p) {
      if (lat_ms.empty()) return 0.0;
      const double pos = (p / 100.0) * (double)(lat_ms.size() - 1);
      const size_t i0 = (size_t)std::floor(pos), i1 = std::min(i0 + 1, lat_ms.size() - 1);
      const double frac = pos - (double)i0;
      return lat_ms[i0] * (1.0 - frac) + lat_ms[i1] * frac;
    };
    if (batch_q > 1) std::cout << "batch_samples=" << lat_ms.size() << "\n";
    const double avg_query = total_ms / (double)Q, qps = (double)Q * 1000.0 / total_ms;
    std::cout << std::fixed << std::setprecision(3);
    std::cout << "Total:     " << total_ms << " ms\n";
    std::cout << "Avg_query: " << avg_query << " ms/query  (" << qps << " QPS)\n";
    if (batch_q > 1) {
      const double nb = (double)lat_ms.size();
      std::cout << "Avg_batch: " << total_ms / nb << " ms/batch  (" << nb * 1000.0 / total_ms << " batches/s)\n";
      std::cout << "batch_p50: " << pct(50) << " ms\n" << "batch_p95: " << pct(95) << " ms\n" << "batch_p99: " << pct(99) << " ms\n";
    } else {
      std::cout << "p50:       " << pct(50) << " ms\n" << "p95:       " << pct(95) << " ms\n" << "p99:       " << pct(99) << " ms\n";
    }
    std::cout << "sink=" << sink << "\n";
    const double bytes_per_query = (double)nvdb::bytes_for_payload_and_aux(base.count(), base.dim(), base.dtype());
    std::cout << std::setprecision(0) << "bytes_per_query=" << bytes_per_query << "\n" << std::setprecision(3);
    std::cout << "payload_equiv_bandwidth_GBps=" << (avg_query > 0 ? bytes_per_query * 1e-6 / avg_query : 0.0) << "\n";
    if (batch_q > 1) std::cout << "(note) payload_equiv_bandwidth_GBps may exceed DRAM peak due to cache reuse\n";
    std::cout << "batch_q=" << batch_q << " path=" << path_s << " path_used=" << path_used << " shards=" << index.shards() << "\n";
    std::cout << "kernel_ms_avg=" << kernel_ms_sum / (double)lat_ms.size() << " upload_ms=" << upload_ms
              << " upload_GBps=" << (double)nvdb::bytes_for_payload_and_aux(base.count(), base.dim(), base.dtype()) * 1e-6 / upload_ms << "\n";

    if (const char* gt = std::getenv("GT_PATH")) {
      uint32_t gk = 0;
      uint64_t gq = 0;
      std::vector<uint32_t> gids;
      if (!load_gtbin(gt, &gk, &gq, &gids) || gq != Q || gk < std::min<uint64_t>(k, base.count())) {
        std::cerr << "GT_PATH: unreadable gtbin or shape mismatch\n";
        return 3;
      }
      const uint32_t kk = (uint32_t)std::min<uint64_t>(k, base.count());
      uint64_t same = 0;
      for (uint64_t qi = 0; qi < Q; ++qi)
        for (uint32_t j = 0; j < kk; ++j) same += (all_ids[(size_t)qi * k + j] == (uint64_t)gids[(size_t)qi * gk + j]);
      std::cout << std::setprecision(6) << "gt_match=" << (double)same / (double)(Q * kk) << " gt_positions=" << Q * kk << "\n";
    }
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "error: " << e.what() << "\n";
    return 4;
  }
}
