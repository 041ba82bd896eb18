// scan.h — host-visible launch interfaces of the kernels (internal to the library; the public surface is
// include/nvdb_b200.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nvdb_b200 {

constexpr int kScanRowsPerTile = 256;
constexpr int kScanMaxQueriesPerPass = 8;

struct ScanArgs {
  const float* queries;  // [nq][dim] f32, device
  const float* scales;   // [n_rows] (i8) or null
  const void* base;      // payload (generic kernel only)
  uint64_t* partial;     // [nq][grid][k] ordered keys
  uint32_t n_rows;
  uint32_t dim;
  uint32_t k;
  uint32_t cap;        // per-query candidate list capacity (power of two >= k + rows per tile)
  uint32_t nq;         // valid queries in this pass (<= template NQ)
  uint32_t num_tiles;  // ceil(n_rows / 256)
  uint32_t kchunks;    // ceil(dim * elem_size / 128)
  int stages;
};

struct ScanLaunch {
  int dtype;
  bool use_tma;
  const CUtensorMap* tmap;
  int grid;
  cudaStream_t stream;
};

uint32_t scan_gemv_smem_bytes(bool use_tma, int stages, int nq_template, uint32_t dim, uint32_t cap);
cudaError_t scan_gemv_launch(const ScanLaunch& l, const ScanArgs& a, int nq_template);

// partial [nq][nlists][k] keys -> final ids / scores [nq][k]
cudaError_t merge_partials_launch(const uint64_t* partial, uint32_t nlists, uint32_t nq, uint32_t k,
                                  uint64_t row_offset, uint64_t* out_ids, float* out_scores, cudaStream_t stream);
// per-shard results (global ids, scores), list l at ids + l*id_list_stride / scores + l*score_list_stride, each
// [nq][k] -> [nq][k]
cudaError_t merge_lists_launch(const uint64_t* ids, const float* scores, size_t id_list_stride,
                               size_t score_list_stride, uint32_t nlists, uint32_t nq, uint32_t k, uint64_t* out_ids,
                               float* out_scores, cudaStream_t stream);

cudaError_t refine_l2_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim,
                             const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                             uint32_t* out_ids, float* out_dist, cudaStream_t stream);

cudaError_t synth_fill_launch(int dtype, uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row, void* payload,
                              float* scales, cudaStream_t stream);

// ---- tensor-core scan (scan_mma.cu)
struct MmaArgs {
  const float* scales;   // [n_rows] row scales (i8) or null
  const float* qscale;   // [mpad]  f16: 2^-e of the query prescale;  i8: s1
  const float* qscale2;  // [mpad]  i8: s2
  uint64_t* partial;     // [nq][P][k] ordered keys (0 = empty)
  void* dbg;             // optional raw accumulators [NACC][mpad][n_rows] (tests only), else null
  uint32_t n_rows, dim, k, nq;
  uint32_t num_tiles;    // ceil(n_rows / BLOCK_N)
  uint32_t kchunks;      // ceil(dim * elem_size / 128)
  uint32_t G, P;         // M-tiles (128 queries each), CTAs per M-tile; grid = G * P
  uint32_t mpad;         // G * 128
  uint32_t block_n;      // base rows per tile (template selector; the base tensor map's box rows)
  int stages;
  int evict_first_base;  // base tiles are read once (G == 1): keep them out of L2
};

bool scan_mma_plan(int dtype, int limbs, uint32_t k, int max_smem, uint32_t* block_n, int* stages);
cudaError_t scan_mma_prep_launch(int dtype, int limbs, const float* queries, uint32_t nq, uint32_t dim, uint32_t mpad,
                                 void* A, float* qs1, float* qs2, cudaStream_t stream);
cudaError_t scan_mma_launch(int dtype, int limbs, const CUtensorMap* tmap_a, const CUtensorMap* tmap_b, const MmaArgs& a,
                            cudaStream_t stream);

void count_launch(uint32_t n = 1);

}  // namespace nvdb_b200
