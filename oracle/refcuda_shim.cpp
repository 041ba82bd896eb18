// oracle/refcuda_shim.cpp — TEST INFRASTRUCTURE ONLY.
// C-ABI door into the reference's nvdb::cuda_l2_topk_batch (reference include/nvdb/cuda_refine.h:25-38,
// src/cuda_refine.cu:839-1173), compiled unmodified for sm_100a by `make -C oracle refcuda`.
// Used on the GPU box as (a) a second oracle for the refine kernel and (b) the timed GPU baseline.
#include "nvdb/cuda_refine.h"

#include <cstdint>
#include <cstring>
#include <vector>

extern "C" int refcuda_l2_topk_batch(const void* base_ptr, uint32_t base_dtype, uint64_t N, uint32_t D,
                                      const float* queries, const uint32_t* cand_ids, uint32_t Q, uint32_t R,
                                      uint32_t K, uint32_t* out_ids, float* out_dist, float* timing4) {
  std::vector<uint32_t> ids;
  std::vector<float> dist;
  nvdb::CudaRefineTiming t{};
  nvdb::cuda_l2_topk_batch(base_ptr, base_dtype, N, D, queries, cand_ids, Q, R, K, ids, dist, &t);
  if (!ids.empty()) std::memcpy(out_ids, ids.data(), ids.size() * sizeof(uint32_t));
  if (!dist.empty() && out_dist) std::memcpy(out_dist, dist.data(), dist.size() * sizeof(float));
  if (timing4) {
    timing4[0] = t.h2d_ms;
    timing4[1] = t.kernel_ms;
    timing4[2] = t.d2h_ms;
    timing4[3] = t.total_ms;
  }
  return (int)ids.size();
}
