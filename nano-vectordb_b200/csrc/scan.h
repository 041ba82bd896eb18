// scan.h — host-visible launch interfaces of the kernels (internal to the library; the public surface is
// include/nvdb_b200.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nvdb_b200 {

constexpr int kScanRowsPerTile = 256;
constexpr int kScanMaxQueriesPerPass = 8;

// Optional tail of the streaming scan's last CTA: exchange the final lists with the peer GPUs and merge (exchange.cuh).
constexpr uint32_t kXchgMaxWorld = 16;
constexpr uint32_t kXchgHeaderBytes = 1024;  // flags[kXchgMaxWorld] (u32) at the start of every window
struct ScanXchg {
  unsigned char* win[kXchgMaxWorld];  // windows of all ranks as mapped in this process (win[rank] = own)
  uint32_t* status;                   // sticky error word (own device): 1 = timed out waiting for a peer
  unsigned long long timeout_ns;
  size_t parity_off;                  // kXchgHeaderBytes + (epoch & 1) * world * slot_bytes
  size_t slot_bytes;
  uint32_t world = 0, rank = 0, epoch = 0;  // world == 0: no exchange
};

struct ScanArgs {
  const float* queries;  // [nq][dim] f32, device
  const float* scales;   // [n_rows] (i8) or null
  const void* base;      // payload (generic kernel only)
  uint64_t* partial;     // [nq][grid][k] ordered keys
  // fused final merge (TMA kernel only): when done_counter != null the last CTA to finish writes the final top-k of this
  // pass's queries to out_ids / out_scores ([nq][k]); needs next_pow2(grid * k) * 8 bytes <= stages * 32 KB
  uint32_t* done_counter;  // zero-initialised; the merging CTA resets it
  uint64_t* out_ids;
  float* out_scores;
  ScanXchg xchg;  // with the fused merge only, single pass: out_* receive the merge over ALL shards
  uint64_t row_offset;
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
// same reduction, but the k best KEYS of query q go to out_keys + q * key_stride (e.g. one more list of a later pass)
// and the score of the k-th best (-inf when fewer than k exist) to out_thr[q]: the seed of a later pass.
cudaError_t merge_partials_seed_launch(const uint64_t* partial, uint32_t nlists, uint32_t nq, uint32_t k,
                                       uint64_t* out_keys, size_t key_stride, float* out_thr, cudaStream_t stream);
// per-shard results (global ids, scores), list l at ids + l*id_list_stride / scores + l*score_list_stride, each
// [nq][k] -> [nq][k]
cudaError_t merge_lists_launch(const uint64_t* ids, const float* scores, size_t id_list_stride,
                               size_t score_list_stride, uint32_t nlists, uint32_t nq, uint32_t k, uint64_t* out_ids,
                               float* out_scores, cudaStream_t stream);

cudaError_t refine_l2_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim,
                             const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                             uint32_t* out_ids, float* out_dist, cudaStream_t stream, unsigned long long* dbg = nullptr,
                             uint32_t dbg_q = 0, uint32_t* smem_bytes_out = nullptr);
constexpr int kRefineThreadsPerCta = 256;

cudaError_t synth_fill_launch(int dtype, uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row, void* payload,
                              float* scales, cudaStream_t stream);

// ---- tensor-core scan (scan_mma.cu)
struct MmaArgs {
  const float* scales;   // [n_rows] row scales (i8) or null
  const float* qscale;   // [mpad]  f16: 2^-e of the query prescale;  i8: s1
  const float* qscale2;  // [mpad]  i8: s2
  uint64_t* partial;     // [nq][lists][k] ordered keys (0 = empty); CTA j of an M-tile writes list j
  const float* thr_init; // [mpad] admission threshold seeds (a lower bound of each query's final k-th best) or null
  void* dbg;             // optional raw accumulators [NACC][mpad][n_rows] (tests only), else null
  uint32_t n_rows, dim, k, nq;
  uint32_t tile0;        // first base tile of this pass
  uint32_t num_tiles;    // one past the last base tile of this pass (<= ceil(n_rows / BLOCK_N))
  uint32_t kchunks;      // ceil(dim * elem_size / 128)
  uint32_t G, P;         // M-tiles (128 queries each), CTAs per M-tile; grid = G * P
  uint32_t lists;        // lists per query in `partial` (>= P)
  uint32_t mpad;         // G * 128
  uint32_t block_n;      // base rows per tile (template selector; the base tensor map's box rows)
  int stages;
  int evict_first_base;  // base tiles are read once (G == 1): keep them out of L2
  int resident_a;        // i8: query limbs resident in shared memory (knob: 0 none, 1 the first, 2 both if they fit — pair kernel)
  int max_smem;          // opt-in shared memory per CTA
};

bool scan_mma_plan(int dtype, int limbs, uint32_t k, int max_smem, uint32_t* block_n, int* stages);
cudaError_t scan_mma_prep_launch(int dtype, int limbs, const float* queries, uint32_t nq, uint32_t dim, uint32_t mpad,
                                 void* A, float* qs1, float* qs2, cudaStream_t stream);
cudaError_t scan_mma_launch(int dtype, int limbs, const CUtensorMap* tmap_a, const CUtensorMap* tmap_b, const MmaArgs& a,
                            cudaStream_t stream);

// ---- tensor-core scan for small batches, base rows on M / queries on N (scan_rows.cu)
struct RowsPlan {
  uint32_t bp;      // queries padded to a multiple of 16 (columns per limb)
  uint32_t nb;      // MMA N: f16 bp, i8 2 * bp
  uint32_t cap;     // candidate-list capacity per query (power of two >= 4 * kp)
  int stages;       // ring depth (16 KB per stage)
  uint32_t smem_bytes;
};
struct RowsArgs {
  const float* queries;   // [nq][dim] f32, 16-byte aligned
  const float* scales;    // [n_rows] row scales (i8) or null
  const float* thr_init;  // [nq] admission threshold seeds or null
  uint64_t* partial;      // [nq][lists][kp] ordered keys (0 = empty); CTA j writes list j
  uint32_t* seed_ws;      // cross-CTA threshold seeding: [0] arrival counter, [16 + cta * bp + q] best ordered score of
                          // CTA cta for query q after its first tile; null = off
  uint32_t seed_target;   // counter value that means "every CTA of THIS launch has published" (launch number * grid)
  uint32_t seed_rounds;   // publication rounds (after tile 0, 8, 64 of every CTA), <= 3; every launch on a workspace uses the same
  uint32_t n_rows, dim, kp, nq;
  uint32_t num_tiles;     // ceil(n_rows / 128)
  uint32_t kchunks;       // ceil(dim * elem_size / 128)
  uint32_t lists;         // lists per query in `partial` (>= grid)
  uint32_t bp, nb, cap;
  int stages;
};
bool scan_rows_plan(int dtype, uint32_t nq, uint32_t kp, uint32_t dim, int max_smem, RowsPlan* out);
// tmap_base128: the base tensor map with 128-row boxes
cudaError_t scan_rows_launch(int dtype, const CUtensorMap* tmap_base128, const RowsArgs& a, int grid, cudaStream_t stream);

// ---- exact re-score + certificate of tensor-core candidates (rescore.cu)
struct RescoreArgs {
  const void* base;          // payload
  const float* scales;       // [n_rows] (i8) or null
  const float* queries;      // [nq][dim] f32
  const uint64_t* cand_ids;  // [nq][kp] LOCAL rows ranked by approximate score (~0 = empty)
  const float* cand_approx;  // [nq][kp] approximate scores
  uint64_t* out_ids;         // [nq][k] row_offset + row
  float* out_scores;         // [nq][k] exact (reference-order FP32) scores
  uint32_t* flags;           // [nq] 1 = certificate failed, the query must be re-run on the exact scan
  uint32_t* nflag;           // [1] number of flagged queries (caller zeroes it)
  uint64_t row_offset;
  uint64_t cand_offset;      // subtracted from cand_ids before use (candidates produced with a row_offset)
  uint32_t n_rows, dim, kp, k;
  float max_l2, max_l1;      // max over rows of |row|_2, |row|_1 (times the row scale for i8), rounded up
  float rel_err;             // bound on |approx - exact| / (|q|_2 * max_l2) excluding the i8 limb residual term
  // fused merge: when `partial` is set the candidates are selected here from the scan's per-CTA lists [nq][nlists][kp]
  // (each sorted descending, 0 = empty) and cand_ids / cand_approx are not read
  const uint64_t* partial;
  uint32_t nlists;
  uint32_t* nflag_next;      // optional: zeroed by the kernel (the flag counter of the next call on the stream)
  uint32_t* done_ctr;        // optional pair: device word (zero between launches) counting finished CTAs, and a
  uint32_t* host_nflag;      // device-visible pinned host word that the LAST CTA stores the final flag count to (no D2H copy)
  uint32_t sel_off;          // set by rescore_exact_launch
  int stage_rows;            // set by rescore_exact_launch: candidate rows staged in shared memory, 8 lanes per candidate
};
cudaError_t rescore_exact_launch(int dtype, const RescoreArgs& a, uint32_t nq, cudaStream_t stream);
cudaError_t row_norm_max_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim, uint32_t* out2,
                                int num_sms, cudaStream_t stream);
// dst[j] = src[list[j]] (rows of `dim` floats) / dst[list[j]] = src[j] (k ids + k scores per query)
cudaError_t gather_queries_launch(const float* src, const uint32_t* list, uint32_t n, uint32_t dim, float* dst,
                                  cudaStream_t stream);
cudaError_t scatter_topk_launch(const uint64_t* src_ids, const float* src_scores, const uint32_t* list, uint32_t n,
                                uint32_t k, uint64_t* dst_ids, float* dst_scores, cudaStream_t stream);

// CTA-pair variant (scan_mma2.cu): a.G even, a.P = clusters per pair column, tmap_b with scan_mma2_half_n(dtype)-row
// boxes (f16: 128 of a 256-row tile; i8: 64 of a 128-row tile). Tile rows = 2 * scan_mma2_half_n(dtype).
uint32_t scan_mma2_half_n(int dtype);
bool scan_mma2_plan(int dtype, int limbs, uint32_t k, int max_smem, int* stages);
cudaError_t scan_mma2_launch(int dtype, int limbs, const CUtensorMap* tmap_a, const CUtensorMap* tmap_b, const MmaArgs& a,
                             cudaStream_t stream);

// reference converters on the device (convert.cu): f32 -> f16 (RNE); f32 | f16 -> i8 + per-row scale
cudaError_t convert_launch(int src_dtype, const void* src, int dst_dtype, uint64_t n, uint32_t dim, void* dst, float* scales,
                           int num_sms, cudaStream_t stream);

void count_launch(uint32_t n = 1);

}  // namespace nvdb_b200
