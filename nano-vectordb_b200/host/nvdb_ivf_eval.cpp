// nvdb_ivf_eval (B200) — ANN evaluation harness with the reference tool's command line and report, both stages on the GPU.
//
//   nvdb_ivf_eval <base.vecbin> <ivf.nvdbivf> <query.vecbin> <k>
//   Env: IVF_NPROBE=16  WARMUP=5  REFINE_K=0  EVAL_MODE=full|ann_only  GT_PATH=<gtbin>  PIPELINE=staged|interleaved
//        EXACT_METRIC=L2|DOT  CUDA_DBG_TIMING=1 CUDA_DBG_Q=<n> CUDA_DBG_DIR=<dir>  (+ IVF_GPU=0)
//
// Follows apps/nvdb_ivf_eval.cpp: argv / exit codes 1-3, 7-10 (:309-380), the staged protocol (:470-662: stage A one
// query per call with its own wall-clock latency, stage B ONE batched refine whose per-query share is added to every
// ANN latency), the recall definition (|GT ∩ pred| / k averaged over queries, :640-652) and the printed report
// (:708-779: Recall@k, ANN / TOTAL Avg p50 p95 p99, one RESULT line of key=value pairs). Differences: the index file is
// this repo's container (ivfpq_file.h), stage A is nvdb_b200_ivfpq_search instead of faiss::Index::search, stage B is
// nvdb_b200_refine_l2 (always the CUDA refine; CUDA_REFINE=0's host loop does not exist here). Without GT_PATH the
// reference computes exact neighbours on the host cores (EXACT_METRIC=L2: exact_topk_l2_ids, DOT: FlatIndexOMP,
// apps/nvdb_ivf_eval.cpp:353,384-385,431-435); here EXACT_METRIC=DOT is the exact flat scan on the GPU and EXACT_METRIC=L2
// (the default, as in the reference) re-ranks the flat scan's top max(4k, 64) by exact L2 with the refine kernel — the L2
// ranking itself whenever the true L2 top-k lies inside that dot-product shortlist, which holds for the L2-normalised
// embeddings the tool chain assumes (apps/nvdb_gt_build.cpp:33-35), where the two rankings coincide.
// PIPELINE=interleaved (apps/nvdb_ivf_eval.cpp:465,664-693): ANN + refine per query, one latency sample for both.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <string>
#include <vector>

#include "dataset.h"
#include "flat_index_cuda.h"
#include "ivfpq_file.h"

namespace {
int getenv_int(const char* k, int defv) {
  const char* v = std::getenv(k);
  return v ? std::atoi(v) : defv;
}
std::string getenv_str(const char* k, const char* defv) {
  const char* v = std::getenv(k);
  return v ? std::string(v) : std::string(defv);
}
struct LatStats {
  double avg_ms = 0, p50_ms = 0, p95_ms = 0, p99_ms = 0, qps = 0;
};
LatStats lat_stats(std::vector<double> v) {
  LatStats s;
  if (v.empty()) return s;
  double sum = 0;
  for (double x : v) sum += x;
  s.avg_ms = sum / (double)v.size();
  s.qps = s.avg_ms > 0 ? 1000.0 / s.avg_ms : 0;
  std::sort(v.begin(), v.end());
  auto pct = [&](double p) { return v[std::min(v.size() - 1, (size_t)(p * (double)(v.size() - 1) + 0.5))]; };
  s.p50_ms = pct(0.50);
  s.p95_ms = pct(0.95);
  s.p99_ms = pct(0.99);
  return s;
}
#pragma pack(push, 1)
struct GtBinHeader {  // include/nvdb/gtbin_format.h:18-35
  uint64_t magic;
  uint32_t version, metric, k, dim;
  uint64_t Q, N;
  uint8_t reserved[64 - 40];
};
#pragma pack(pop)
constexpr uint64_t kGtMagic = 0x4E56444247543031ULL;

// the reference's clock64 phase dump (apps/nvdb_ivf_eval.cpp:47-126): same file name pattern, columns and DBG_TSV line
void write_dbg_tsv_if_enabled(const nvdb_b200_timing& t, int k, uint64_t Q, uint32_t R, int nprobe, int refine_k) {
  if (getenv_int("CUDA_DBG_TIMING", 0) == 0 || t.dbg_q == 0) return;
  const std::string outdir = getenv_str("CUDA_DBG_DIR", "./results_dbg");
  std::filesystem::create_directories(outdir);
  const int pinned = getenv_int("CUDA_PINNED", 0), ret = getenv_int("CUDA_RETURN_DIST", 1);
  const std::string git = getenv_str("GIT_SHA", "NA");
  const std::string path = outdir + "/dbg_K" + std::to_string(k) + "_Q" + std::to_string(Q) + "_R" + std::to_string(R) + "_th" +
                           std::to_string(t.threads) + "_mode=b200_optin=1_pinned=" + std::to_string(pinned) +
                           "_ret=" + std::to_string(ret) + "_git=" + git + ".tsv";
  std::ofstream ofs(path, std::ios::out | std::ios::trunc);
  if (!ofs) {
    std::cerr << "[WARN] cannot write DBG TSV: " << path << "\n";
    return;
  }
  const double tot = (double)t.dbg_dist_cycles_avg + t.dbg_write_cycles_avg + t.dbg_merge_cycles_avg;
  const double pd = tot > 0 ? t.dbg_dist_cycles_avg / tot : 0, pw = tot > 0 ? t.dbg_write_cycles_avg / tot : 0,
               pm = tot > 0 ? t.dbg_merge_cycles_avg / tot : 0;
  ofs << "k\tQ\tR\tnprobe\trefine_k\tkernel_mode\tcuda_threads\tcuda_nwarps\tcuda_shmem_bytes\tcuda_shmem_optin\t"
         "cuda_pinned\tcuda_return_dist\tgit_rev\tdbg_q\tdbg_dist_cycles_avg\tdbg_write_cycles_avg\tdbg_merge_cycles_avg\t"
         "dbg_dist_pct\tdbg_write_pct\tdbg_merge_pct\n";
  ofs << k << "\t" << Q << "\t" << R << "\t" << nprobe << "\t" << refine_k << "\tb200\t" << t.threads << "\t" << t.threads / 32 << "\t"
      << t.shmem_bytes << "\t1\t" << pinned << "\t" << ret << "\t" << git << "\t" << t.dbg_q << "\t" << std::fixed
      << std::setprecision(3) << t.dbg_dist_cycles_avg << "\t" << t.dbg_write_cycles_avg << "\t" << t.dbg_merge_cycles_avg << "\t"
      << std::setprecision(6) << pd << "\t" << pw << "\t" << pm << "\n";
  std::cout << "DBG_TSV=" << path << " dbg_q=" << t.dbg_q << " dist%=" << std::setprecision(3) << pd * 100.0
            << " write%=" << pw * 100.0 << " merge%=" << pm * 100.0 << "\n";
}
}  // namespace

int main(int argc, char** argv) {
  if (argc < 5) {
    std::cerr << "Usage: nvdb_ivf_eval <base.vecbin> <ivf.nvdbivf> <query.vecbin> <k>\n"
              << "Env:\n"
              << "  IVF_NPROBE=16\n"
              << "  WARMUP=5\n"
              << "  REFINE_K=0\n"
              << "  GT_PATH=<query.gtbin>\n"
              << "  EVAL_MODE=full|ann_only   (default: full)\n"
              << "  PIPELINE=staged|interleaved   (default: staged)\n"
              << "  EXACT_METRIC=L2|DOT   (default: L2)\n";
    return 1;
  }
  const std::string base_path = argv[1], index_path = argv[2], query_path = argv[3];
  const int k = std::atoi(argv[4]);
  const int nprobe = getenv_int("IVF_NPROBE", 16), warmup = getenv_int("WARMUP", 5), refine_k = getenv_int("REFINE_K", 0);
  const int k_search = refine_k > 0 ? std::max(refine_k, k) : k;
  const int gpu = getenv_int("IVF_GPU", 0);
  std::cout << "REFINE_K=" << refine_k << " (k_search=" << k_search << ")\n";
  try {
    nvdb::VectorDataset base, query;
    base.load(base_path);
    query.load(query_path);
    if (query.dtype() != 1) {
      std::cerr << "Query must be float32 vecbin\n";
      return 2;
    }
    if (base.dim() != query.dim()) {
      std::cerr << "Dim mismatch: base.dim=" << base.dim() << " query.dim=" << query.dim() << "\n";
      return 3;
    }
    if (k <= 0) {
      std::cerr << "k must be > 0\n";
      return 4;
    }
    const uint32_t d = base.dim();
    const uint64_t Q = query.count();
    const bool ann_only = getenv_str("EVAL_MODE", "full") == "ann_only";
    const std::string pipeline = getenv_str("PIPELINE", "staged");
    const bool interleaved = pipeline == "interleaved";
    if (!interleaved && pipeline != "staged") {
      std::cerr << "PIPELINE must be staged or interleaved\n";
      return 6;
    }
    const std::string exact_metric = getenv_str("EXACT_METRIC", "L2");
    if (exact_metric != "L2" && exact_metric != "DOT") {
      std::cerr << "EXACT_METRIC must be L2 or DOT\n";
      return 6;
    }
    std::cout << "PIPELINE=" << pipeline << " EXACT_METRIC=" << exact_metric << "\n";
    const std::string gt_path = getenv_str("GT_PATH", "");
    std::vector<uint32_t> gt;  // [Q * k]
    if (!gt_path.empty() && !ann_only) {
      std::ifstream in(gt_path, std::ios::binary | std::ios::ate);
      const std::streamoff size = in ? (std::streamoff)in.tellg() : -1;
      if (size < (std::streamoff)sizeof(GtBinHeader)) {
        std::cerr << "GT file too small: " << gt_path << "\n";
        return 7;
      }
      in.seekg(0);
      GtBinHeader h{};
      in.read(reinterpret_cast<char*>(&h), sizeof(h));
      if (h.magic != kGtMagic || h.version != 1) {
        std::cerr << "Bad GT header: " << gt_path << "\n";
        return 8;
      }
      if (h.k != (uint32_t)k || h.dim != d || h.Q != Q || h.N != base.count()) {
        std::cerr << "GT mismatch: (k/d/Q/N) expected=(" << k << "/" << d << "/" << Q << "/" << base.count() << ") got=("
                  << h.k << "/" << h.dim << "/" << h.Q << "/" << h.N << ")\n";
        return 9;
      }
      const std::streamoff expect = (std::streamoff)(sizeof(GtBinHeader) + Q * (uint64_t)k * 4);
      if (size != expect) {
        std::cerr << "GT size mismatch: expected=" << expect << " got=" << size << "\n";
        return 10;
      }
      gt.resize((size_t)Q * k);
      in.read(reinterpret_cast<char*>(gt.data()), (std::streamsize)(gt.size() * 4));
      std::cout << "GT_PATH=" << gt_path << " (cached ground truth enabled)\n";
    }

    nvdb::FlatIndexCuda flat(&base, gpu);
    nvdb_b200_ivfpq* ivf = nvdb::ivfpq_load(index_path, gpu);
    uint64_t n_idx = 0;
    uint32_t dim_idx = 0, nlist = 0, pq_m = 0;
    nvdb_b200_ivfpq_info(ivf, &n_idx, &dim_idx, &nlist, &pq_m, nullptr);
    if (dim_idx != d || n_idx != base.count()) {
      std::cerr << "Index does not belong to this base (n/d)\n";
      return 5;
    }
    int has_opq = 0;
    nvdb_b200_ivfpq_rotation(ivf, nullptr, &has_opq);
    std::cout << "IVF nlist=" << nlist << " nprobe=" << nprobe << " | d=" << d << " k=" << k << " Q=" << Q << " | pq_m=" << pq_m
              << (has_opq ? " | pretransform=OPQ" : "") << "\n";

    if (gt.empty() && !ann_only) {  // exact neighbours from the flat scan (see the header comment)
      gt.resize((size_t)Q * k);
      const bool l2 = exact_metric == "L2";
      const uint32_t ks = l2 ? (uint32_t)std::min<uint64_t>(base.count(), (uint64_t)std::max(4 * k, 64)) : (uint32_t)k;
      std::vector<uint32_t> shortlist, l2_ids;
      std::vector<float> l2_dist;
      for (uint64_t q0 = 0; q0 < Q; q0 += 1024) {
        const uint32_t nq = (uint32_t)std::min<uint64_t>(1024, Q - q0);
        auto res = flat.search_topk_dot_batch(query.vector_ptr_f32(q0), nq, ks);
        if (!l2) {
          for (uint32_t i = 0; i < nq; ++i)
            for (int j = 0; j < k; ++j) gt[(q0 + i) * k + j] = j < (int)res[i].size() ? (uint32_t)res[i][j].id : 0xFFFFFFFFu;
          continue;
        }
        shortlist.assign((size_t)nq * ks, 0xFFFFFFFFu);
        for (uint32_t i = 0; i < nq; ++i)
          for (uint32_t j = 0; j < ks && j < res[i].size(); ++j) shortlist[(size_t)i * ks + j] = (uint32_t)res[i][j].id;
        flat.refine_l2(query.vector_ptr_f32(q0), shortlist.data(), nq, ks, (uint32_t)k, l2_ids, l2_dist);
        std::memcpy(gt.data() + q0 * k, l2_ids.data(), (size_t)nq * k * 4);
      }
    }

    auto ms_since = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
      return std::chrono::duration<double, std::milli>(b - a).count();
    };
    std::vector<uint32_t> I((size_t)k_search), I_all((size_t)Q * k_search);
    for (int i = 0; i < warmup && Q > 0; ++i)
      nvdb_b200_ivfpq_search(ivf, query.vector_ptr_f32(0), 1, (uint32_t)nprobe, (uint32_t)k_search, I.data(), nullptr);

    std::vector<double> ann_lat_ms, total_lat_ms;
    ann_lat_ms.reserve(Q);
    std::vector<uint32_t> pred_il;  // interleaved pipeline: per-query refine results
    if (interleaved) {
      // ---- interleaved: ANN then refine of the same query, one call each, one latency sample for the pair
      // (apps/nvdb_ivf_eval.cpp:664-693); the candidates never leave the device (nvdb_b200_ivfpq_search_refine)
      const bool refine_il = !ann_only && refine_k > 0;
      pred_il.assign((size_t)Q * k, 0xFFFFFFFFu);
      std::vector<uint32_t> one((size_t)k);
      for (uint64_t qi = 0; qi < Q; ++qi) {
        const auto t0 = std::chrono::steady_clock::now();
        int st;
        if (refine_il)
          st = nvdb_b200_ivfpq_search_refine(ivf, flat.handle(), query.vector_ptr_f32(qi), 1, (uint32_t)nprobe, (uint32_t)k_search,
                                             (uint32_t)k, one.data(), nullptr);
        else
          st = nvdb_b200_ivfpq_search(ivf, query.vector_ptr_f32(qi), 1, (uint32_t)nprobe, (uint32_t)k_search, I.data(), nullptr);
        if (st != NVDB_B200_OK) {
          std::cerr << "Error: " << nvdb_b200_ivfpq_last_error(ivf) << "\n";
          return 11;
        }
        const double ms = ms_since(t0, std::chrono::steady_clock::now());
        ann_lat_ms.push_back(ms);
        total_lat_ms.push_back(ms);
        for (int j = 0; j < k; ++j) pred_il[qi * k + j] = refine_il ? one[j] : I[j];
      }
    }
    // ---- stage A: ANN only, one query per call
    for (uint64_t qi = 0; qi < Q && !interleaved; ++qi) {
      const auto t0 = std::chrono::steady_clock::now();
      if (nvdb_b200_ivfpq_search(ivf, query.vector_ptr_f32(qi), 1, (uint32_t)nprobe, (uint32_t)k_search, I.data(), nullptr) !=
          NVDB_B200_OK) {
        std::cerr << "Error: " << nvdb_b200_ivfpq_last_error(ivf) << "\n";
        return 11;
      }
      ann_lat_ms.push_back(ms_since(t0, std::chrono::steady_clock::now()));
      std::memcpy(I_all.data() + qi * k_search, I.data(), (size_t)k_search * 4);
    }
    // ---- stage B: one batched refine over all candidates
    double refine_ms_total = 0, refine_ms_per_q = 0;
    nvdb_b200_timing t{};
    std::vector<uint32_t> pred((size_t)Q * k, 0xFFFFFFFFu);
    const bool refine = !ann_only && refine_k > 0;
    if (interleaved) {
      pred = pred_il;
    } else if (refine) {
      std::vector<uint32_t> out_ids;
      std::vector<float> out_dist;
      if (getenv_int("CUDA_REFINE_WARMUP", 1) != 0)
        flat.refine_l2(query.vector_ptr_f32(0), I_all.data(), (uint32_t)std::min<uint64_t>(Q, 16), (uint32_t)k_search, (uint32_t)k,
                       out_ids, out_dist);
      const auto t0 = std::chrono::steady_clock::now();
      flat.refine_l2(query.vector_ptr_f32(0), I_all.data(), (uint32_t)Q, (uint32_t)k_search, (uint32_t)k, out_ids, out_dist, &t);
      refine_ms_total = ms_since(t0, std::chrono::steady_clock::now());
      refine_ms_per_q = Q ? refine_ms_total / (double)Q : 0;
      pred = out_ids;
      std::cout << "CUDA_REFINE=1 refine_ms_total=" << refine_ms_total << " refine_ms_per_q=" << refine_ms_per_q
                << " | h2d_ms=" << t.h2d_ms << " kernel_ms=" << t.kernel_ms << " d2h_ms=" << t.d2h_ms << "\n";
      write_dbg_tsv_if_enabled(t, k, Q, (uint32_t)k_search, nprobe, refine_k);
    } else {
      for (uint64_t qi = 0; qi < Q; ++qi)
        for (int j = 0; j < k; ++j) pred[qi * k + j] = I_all[qi * k_search + j];
    }
    if (!interleaved)
      for (double a : ann_lat_ms) total_lat_ms.push_back(a + (refine ? refine_ms_per_q : 0.0));

    double recall_sum = 0;
    if (!ann_only)
      for (uint64_t qi = 0; qi < Q; ++qi) {
        int hit = 0;
        for (int i = 0; i < k; ++i) {
          const uint32_t p = pred[qi * k + i];
          if (p == 0xFFFFFFFFu) continue;
          for (int j = 0; j < k; ++j)
            if (gt[qi * k + j] == p) {
              ++hit;
              break;
            }
        }
        recall_sum += (double)hit / (double)k;
      }

    const LatStats annS = lat_stats(ann_lat_ms), totalS = lat_stats(total_lat_ms);
    if (!ann_only)
      std::cout << "Recall@" << k << ": " << std::setprecision(4) << (Q ? recall_sum / (double)Q : 0.0) << std::setprecision(3)
                << "\n";
    else
      std::cout << "Recall@" << k << ": N/A (EVAL_MODE=ann_only)\n";
    std::cout << "ANN Avg:   " << annS.avg_ms << " ms/query  (" << annS.qps << " QPS)\n";
    std::cout << "ANN p50:   " << annS.p50_ms << " ms\n";
    std::cout << "ANN p95:   " << annS.p95_ms << " ms\n";
    std::cout << "ANN p99:   " << annS.p99_ms << " ms\n";
    std::cout << "TOTAL Avg:   " << totalS.avg_ms << " ms/query  (" << totalS.qps << " QPS)\n";
    std::cout << "TOTAL p50:   " << totalS.p50_ms << " ms\n";
    std::cout << "TOTAL p95:   " << totalS.p95_ms << " ms\n";
    std::cout << "TOTAL p99:   " << totalS.p99_ms << " ms\n";
    std::cout.setf(std::ios::fixed);
    std::cout << std::setprecision(6) << "RESULT"
              << " pipeline=" << pipeline << " exact_metric=" << exact_metric << " refine_k=" << refine_k << " k_search=" << k_search
              << " nprobe=" << nprobe << " Q=" << Q << " k=" << k
              << " cuda_refine=" << (refine ? 1 : 0) << " refine_enabled=" << (refine ? 1 : 0)
              << " refine_backend=" << (refine ? "cuda" : "none") << " recall=" << (ann_only || !Q ? 0.0 : recall_sum / (double)Q)
              << " ann_avg_ms=" << annS.avg_ms << " ann_p99_ms=" << annS.p99_ms << " total_avg_ms=" << totalS.avg_ms
              << " total_p99_ms=" << totalS.p99_ms << " refine_ms_total=" << refine_ms_total
              << " refine_ms_per_q=" << refine_ms_per_q << " refine_h2d_ms=" << t.h2d_ms << " refine_kernel_ms=" << t.kernel_ms
              << " refine_d2h_ms=" << t.d2h_ms << " git_rev=" << getenv_str("GIT_SHA", "NA") << "\n";
    nvdb_b200_ivfpq_destroy(ivf);
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return 12;
  }
}
