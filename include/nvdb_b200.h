/*
 * nvdb_b200.h — C ABI of the B200-native flat-scan top-k + exact-refine path.
 *
 * This is the drop-in boundary for ONE hot path of ntnu-yjyjames/nano-vectordb (citations into that repo):
 *   - nvdb::FlatIndex{,OMP,Async,Pool}::search_topk_dot(q, k)       include/nvdb/flat_index.h:13,
 *                                                                   src/flat_index.cpp:16-48 (+_omp/_async/_pool)
 *   - the bench-side batched scan                                   apps/nvdb_bench.cpp:47-159,161-251
 *   - nvdb::cuda_l2_topk_batch(...)                                 include/nvdb/cuda_refine.h:25-38,
 *                                                                   src/cuda_refine.cu:839-1173
 * Plain pointers and sizes only; no C++/torch types; every function returns an int status (0 = OK) and never
 * exits or throws across the boundary (the reference's CUDA library calls exit(); src/cuda_refine.cu:18-23).
 * There is no CPU fallback behind these symbols: without a CUDA device they return NVDB_B200_ERR_CUDA.
 *
 * dtype codes are the vecbin codes (include/nvdb/vecbin_format.h:10-14): 1 = f32, 2 = f16, 3 = i8 (+ one f32
 * scale per row). Result order is canonical: score descending, then id ascending (flat scan); distance
 * ascending, then id ascending (refine). See DESIGN.md "Parity contract".
 *
 * Thread-safety: calls on different index handles are independent; calls on ONE handle serialise on its
 * stream (same contract as FlatIndexPool / cuda_l2_topk_batch: one query batch in flight per index).
 */
#ifndef NVDB_B200_H
#define NVDB_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NVDB_B200_VERSION 101

enum {
  NVDB_B200_OK = 0,
  NVDB_B200_ERR_ARG = 1,      /* null pointer / bad dtype / bad shape */
  NVDB_B200_ERR_EMPTY = 2,    /* "Empty base" (src/flat_index.cpp:17) */
  NVDB_B200_ERR_CUDA = 3,     /* CUDA runtime / driver error; see nvdb_b200_last_error */
  NVDB_B200_ERR_OOM = 4,      /* device allocation failed */
  NVDB_B200_ERR_UNSUPPORTED = 5
};

enum { NVDB_B200_F32 = 1, NVDB_B200_F16 = 2, NVDB_B200_I8 = 3 };

/* which scan kernel family nvdb_b200_search* uses */
enum {
  NVDB_B200_PATH_AUTO = 0, /* GEMV for small batches, tensor-core GEMM above the crossover */
  NVDB_B200_PATH_GEMV = 1, /* bit-exact FP32-FMA streaming scan (all dtypes) */
  NVDB_B200_PATH_MMA = 2,  /* tcgen05 tensor-core scan (f16: query = fp16 hi + lo limbs; i8: two int8 query limbs,
                              exact int32 limb dot products), tolerance-mode parity vs the FP32-FMA reference */
  NVDB_B200_PATH_MMA_FAST = 3, /* f16 only: single fp16 query limb (half the tensor work, ~5e-5 relative error) */
  NVDB_B200_PATH_MMA_EXACT = 4 /* tensor-core candidate scan (f16: one limb, i8: two limbs) keeping k' > k rows per
                                  query, exact FP32 re-score of the candidates in the reference's summation order, and a
                                  per-query error-bound certificate; queries that fail it are re-run on PATH_GEMV. Ids and
                                  scores are bit-identical to PATH_GEMV. AUTO uses this for batches of >= 2 queries (k <= 12; >= 3 otherwise).
                                  Synchronises the stream once per call. */
};

typedef struct nvdb_b200_index nvdb_b200_index;

/* mirrors nvdb::CudaRefineTiming (include/nvdb/cuda_refine.h:7-22): the four event times, the launch shape and the
 * optional in-kernel phase sampling of the refine kernel (clock64 at the phase boundaries of the first dbg_q queries:
 * enabled by the reference's own switches, environment CUDA_DBG_TIMING=1 and CUDA_DBG_Q=<n, default 32>, when a timing
 * struct is passed to nvdb_b200_refine_l2; src/cuda_refine.cu:863-872,1116-1144). Phases here: "write" = the query into
 * shared memory, "dist" = gather + distances + key writes, "merge" = selection + output. */
typedef struct nvdb_b200_timing {
  float h2d_ms;
  float kernel_ms;
  float d2h_ms;
  float total_ms;
  uint32_t kernel_launches; /* kernels of this library launched by the call */
  uint32_t path;            /* NVDB_B200_PATH_* actually used */
  uint32_t threads;         /* refine: threads per CTA */
  uint32_t shmem_bytes;     /* refine: dynamic shared memory per CTA */
  uint32_t dbg_q;           /* queries sampled (0 = sampling off) */
  float dbg_dist_cycles_avg, dbg_write_cycles_avg, dbg_merge_cycles_avg;
} nvdb_b200_timing;

int nvdb_b200_version(void);
const char* nvdb_b200_status_string(int status);
/* last error text; idx may be NULL for errors raised before a handle existed (thread-local). */
const char* nvdb_b200_last_error(const nvdb_b200_index* idx);
int nvdb_b200_device_count(int* out_count);

/* ---- index lifecycle --------------------------------------------------------------------------------------
 * Replaces: `FlatIndex(const VectorDataset* base)` (non-owning view of the mmap; include/nvdb/flat_index.h:11)
 * and the hidden process-global device cache of the refine library (src/cuda_refine.cu:179-204).
 * payload_host: n*dim elements, row-major (the vecbin payload, e.g. the mmap); copied to `device` through a
 * chunked pinned double buffer. scales_host: n floats (i8 only, else NULL). row_offset is added to local row
 * ids in results (row-sharded multi-GPU indices). */
int nvdb_b200_index_create(const void* payload_host, const float* scales_host, uint32_t dtype, uint64_t n,
                           uint32_t dim, int device, uint64_t row_offset, nvdb_b200_index** out);
/* Non-owning view over rows already resident in device memory (16-byte aligned). */
int nvdb_b200_index_create_device(const void* payload_dev, const float* scales_dev, uint32_t dtype, uint64_t n,
                                  uint32_t dim, int device, uint64_t row_offset, nvdb_b200_index** out);
/* Synthetic base generated on the device: global rows [first_row, first_row+n) of the counter-based
 * generator restated in oracle/nvdb_oracle.c (orc_synth_*), converted with the reference converters'
 * semantics (tools/nvdb_convert_f16.cpp, apps/nvdb_quantize_i8.cpp). */
int nvdb_b200_index_create_synthetic(uint32_t dtype, uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row,
                                     int device, uint64_t row_offset, nvdb_b200_index** out);
int nvdb_b200_index_destroy(nvdb_b200_index* idx);
int nvdb_b200_index_info(const nvdb_b200_index* idx, uint64_t* n, uint32_t* dim, uint32_t* dtype, int* device,
                         uint64_t* row_offset, const void** payload_dev, const float** scales_dev);
/* Loader fast path: wall time and bytes of the host -> device upload done by nvdb_b200_index_create
 * (chunked, multi-threaded staging into two pinned buffers overlapping the DMA; the reference does one synchronous
 * pageable cudaMemcpy of the whole base, src/cuda_refine.cu:179-204). 0 / 0 for device-resident and synthetic indexes. */
int nvdb_b200_index_upload_info(const nvdb_b200_index* idx, double* seconds, uint64_t* bytes);
int nvdb_b200_index_set_path(nvdb_b200_index* idx, int path);
/* Copy rows [row0,row0+nrows) of the device-resident payload (and scales) back to the host. */
int nvdb_b200_index_read_rows(const nvdb_b200_index* idx, uint64_t row0, uint64_t nrows, void* payload_host,
                              float* scales_host);

/* ---- flat scan ----------------------------------------------------------------------------------------------
 * Replaces FlatIndex::search_topk_dot (nq == 1) and batched_scan_* (nq > 1). queries: nq*dim f32.
 * out_ids / out_scores: nq*k, score-descending; when k > n the tail is padded with id = UINT64_MAX,
 * score = -inf (the reference clamps k to n, src/flat_index.cpp:24). k == 0 or nq == 0 is a no-op. */
int nvdb_b200_search(nvdb_b200_index* idx, const float* queries_host, uint32_t nq, uint32_t k, uint64_t* out_ids,
                     float* out_scores, nvdb_b200_timing* timing);
/* Same, all buffers in device memory, asynchronous on `stream` (a cudaStream_t; NULL = the index's stream).
 * One search of an index is in flight at a time: its workspaces are per index, so a call that arrives on another stream
 * than the previous one is ordered behind it with an event, never overlapped. NVDB_B200_PATH_MMA_EXACT (what AUTO picks
 * for batches) synchronises `stream` once per call (the certificate verdict comes back to the host). */
int nvdb_b200_search_device(nvdb_b200_index* idx, const float* queries_dev, uint32_t nq, uint32_t k,
                            uint64_t* out_ids_dev, float* out_scores_dev, void* stream);
/* Allocation-free variant of nvdb_b200_search_device. It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it). Same arguments, same result (bit for bit: the streaming scan and its merge kernel), but
 * the scan's scratch - the per-CTA partial lists, nq * lists * k * 8 bytes with lists = 148 or 296 on a B200 - lives in
 * a file-scope static __device__ array of NVDB_B200_STATIC_SCRATCH_BYTES that every GPU which loads the library holds
 * until the process exits. The call does no cudaMalloc / cudaFree / allocating library call, never touches the index's
 * workspaces and always takes the streaming scan (the PATH setting is ignored; meant for 1-8 queries).
 * A request whose scratch exceeds the maximum returns NVDB_B200_ERR_UNSUPPORTED with the sizes in last_error.
 * ALL calls in the process on one GPU share the array (whatever the index or stream): calls that overlap on the GPU race.
 * The array keeps the previous call's contents; every launch rewrites the part it reads, nothing is assumed zero. */
#define NVDB_B200_STATIC_SCRATCH_BYTES (4u << 20)
int nvdb_b200_search_device_static(nvdb_b200_index* idx, const float* queries_dev, uint32_t nq, uint32_t k,
                                   uint64_t* out_ids_dev, float* out_scores_dev, void* stream);
/* Merge `nlists` per-shard results (each [nq][k], laid out list-major: [nlists][nq][k]) into one [nq][k]
 * under (score desc, id asc). Device buffers; asynchronous on `stream`. This is the step after the all-gather
 * of a row-sharded index. */
int nvdb_b200_merge_topk_device(int device, const uint64_t* ids_dev, const float* scores_dev, uint32_t nlists,
                                uint32_t nq, uint32_t k, uint64_t* out_ids_dev, float* out_scores_dev,
                                void* stream);

/* Same merge over the PACKED per-rank layout that one all-gather produces (nano-vectordb_b200/sharded.py):
 * list l starts at packed + l*W 8-byte words, W = nq*k + (nq*k+1)/2, and holds [nq*k u64 ids][nq*k f32 scores]. */
int nvdb_b200_merge_topk_packed_device(int device, const void* packed_dev, uint32_t nlists, uint32_t nq, uint32_t k,
                                       uint64_t* out_ids_dev, float* out_scores_dev, void* stream);

/* ---- exchange over NVLink peer memory (row-sharded index, one process or one handle per GPU) ----------------------
 * The step after the local scans of a row-sharded index (the reference's serial merge of its row
 * blocks, src/flat_index_pool.cpp:212-214) as ONE kernel per rank: push the packed local result (layout of
 * nvdb_b200_merge_topk_packed_device) into every peer's window with NVLink stores, publish a flag, wait for the peers'
 * flags, merge world*k candidates per query under (score desc, id asc). No NCCL call and no second launch.
 *   create   allocates this rank's window (slot_bytes >= 8 * (nq*k + (nq*k+1)/2) of the largest call);
 *   connect  maps the peers: `handles` = world * NVDB_B200_IPC_HANDLE_BYTES bytes gathered from every rank's
 *            _ipc_handle (ranks in other processes), or `windows` = world device pointers from _window (same process);
 *   merge    phases 3 = push + wait + merge in one launch (every rank must call it, all calls of one exchange on one
 *            stream, same order on every rank); phases 1 / 2 split it into two launches (logical shards sharing one
 *            GPU: push all ranks first, then merge all ranks);
 *   status   0 = fine, 1 = a wait timed out (NVDB_B200_XCHG_TIMEOUT_MS, default 30000): the outputs of that call were
 *            filled with the padding values (UINT64_MAX / -inf) and later calls fail the same way. Synchronises. */
#define NVDB_B200_IPC_HANDLE_BYTES 64
typedef struct nvdb_b200_exchange nvdb_b200_exchange;
int nvdb_b200_exchange_create(int device, uint32_t world, uint32_t rank, uint64_t slot_bytes, nvdb_b200_exchange** out);
int nvdb_b200_exchange_ipc_handle(nvdb_b200_exchange* x, unsigned char* handle64);
int nvdb_b200_exchange_window(nvdb_b200_exchange* x, void** window_dev, uint64_t* window_bytes);
int nvdb_b200_exchange_connect(nvdb_b200_exchange* x, const unsigned char* handles, void* const* windows);
int nvdb_b200_exchange_merge(nvdb_b200_exchange* x, const void* local_packed_dev, uint32_t nq, uint32_t k,
                             uint64_t* out_ids_dev, float* out_scores_dev, int phases, void* stream);
int nvdb_b200_exchange_status(nvdb_b200_exchange* x, uint32_t* status);
/* Collective recovery after a timed-out exchange: every rank calls it (no rank inside an exchange call, and a barrier
 * between the resets and the next search); clears the sticky status, the flags of the own window and the epochs. */
int nvdb_b200_exchange_reset(nvdb_b200_exchange* x);
const char* nvdb_b200_exchange_last_error(const nvdb_b200_exchange* x);
int nvdb_b200_exchange_destroy(nvdb_b200_exchange* x);
/* The whole step of one rank of a row-sharded index: local scan of `idx` (NULL = this rank's shard is empty) + exchange
 * + merge; out_* receive the top-k over ALL shards on every rank. Single-query / single-pass streaming scans do it in
 * ONE launch (the scan's last CTA pushes its list to the peers, waits for theirs and merges); other shapes run the
 * scan into the exchange's staging buffer and then the exchange kernel. Same calling rules as nvdb_b200_exchange_merge. */
int nvdb_b200_search_exchange_device(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_dev, uint32_t nq,
                                     uint32_t k, uint64_t* out_ids_dev, float* out_scores_dev, void* stream);
/* Host-buffer form (the shape of nvdb_b200_search; runs on the index's stream and synchronises it): every rank calls it
 * — in one process, one host thread per device (nano-vectordb_b200/host/nvdb_bench.cpp). out_ids / out_scores may both
 * be NULL on ranks that do not need the result. A timed-out exchange is an error (NVDB_B200_ERR_CUDA). */
int nvdb_b200_search_exchange(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_host, uint32_t nq,
                              uint32_t k, uint64_t* out_ids, float* out_scores, nvdb_b200_timing* timing);

/* The same step in two calls, for logical ranks that SHARE one GPU (kernels that wait for one another must not be separate
 * launches on one device — nothing guarantees that they are resident together): phase 1 = local scan + push, called for
 * every rank first; phase 2 = wait (finds every flag) + merge + results, called for every rank afterwards. queries_host
 * is read by phase 1 only. */
int nvdb_b200_search_exchange_split(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_host, uint32_t nq,
                                    uint32_t k, uint64_t* out_ids, float* out_scores, int phase);

/* Kernel-selection knobs for A/B measurements and tests; results never depend on them. Defaults can also be set with
 * the environment variables NVDB_B200_MMA_SEED / NVDB_B200_GEMV_CTAS / NVDB_B200_MMA_2CTA at index creation. */
enum {
  NVDB_B200_TUNE_MMA_SEED = 1,  /* threshold-seeding pass of the tensor-core scan: 0 off, 1 auto (default), 2 always */
  NVDB_B200_TUNE_GEMV_CTAS = 2, /* CTAs per SM of the single-query streaming scan: 1 (default) | 2 */
  NVDB_B200_TUNE_FUSED_MERGE = 4, /* final merge inside the streaming scan's last CTA: 0 never, 1 auto (default), 2 always */
  NVDB_B200_TUNE_MMA_RESIDENT = 5, /* i8 tensor scan, query limbs resident in shared memory: 0 none, 1 the first (default), 2 both when they fit (CTA pairs) */
  NVDB_B200_TUNE_EXACT_TIER2 = 6, /* exact path: queries whose certificate failed get a second tensor-core pass with wider candidate lists before the streaming scan: 0 off, 1 on (default), > 1 that many list entries (A/B measurements); env NVDB_B200_EXACT_TIER2 */
  NVDB_B200_TUNE_ROWS = 7, /* exact path, small batches: row-tile tensor kernel (base rows on M, the whole batch resident in shared memory; scan_rows.cu): 0 never, 1 auto (default: int8 bases whenever the batch and its candidate lists fit, f16 bases up to 32 queries — where it measured faster), 2 whenever it fits; env NVDB_B200_ROWS */
  NVDB_B200_TUNE_ROWS_SEED = 8, /* its cross-CTA threshold seeding: 0 off, 1 auto (default: from 40 queries), 2 always; env NVDB_B200_ROWS_SEED */
  NVDB_B200_TUNE_MMA_2CTA = 3   /* CTA-pair (cta_group::2) tensor-core scan (f16, i8): 0 never, 1 (default) / 2 whenever the batch has an even number of 128-query tiles */
};
int nvdb_b200_index_set_tuning(nvdb_b200_index* idx, int knob, int value);

/* Certificate statistics of NVDB_B200_PATH_MMA_EXACT since creation / the last reset: queries answered through it and
 * how many of them failed the (first) certificate and were re-run. */
int nvdb_b200_index_exact_stats(nvdb_b200_index* idx, uint64_t* queries, uint64_t* fallbacks, int reset);
/* Where the `fallbacks` of nvdb_b200_index_exact_stats went: `tier2` of them were re-run on the tensor-core scan with
 * wider candidate lists, `streamed` were (finally) answered by the bit-exact streaming scan — the ones the second
 * certificate rejected too, or all of them when no wider list is available. Read BEFORE a resetting exact_stats call. */
int nvdb_b200_index_exact_tiers(nvdb_b200_index* idx, uint64_t* tier2, uint64_t* streamed);

/* Two-stage search (coarse scan + exact re-rank, both on the device): `coarse` and `fine` are two indexes over the SAME
 * rows on one device (e.g. the INT8 and the FP16 copy of a base); the coarse scan keeps R >= k candidates per query
 * (R <= 4096), which are re-ranked by their exact reference-order dot products over `fine`. Halves the bytes scanned at
 * FP16-grade scores; approximate (recall < 1 is possible when a true neighbour falls below rank R in the coarse scan). */
int nvdb_b200_search_two_stage(nvdb_b200_index* coarse, nvdb_b200_index* fine, const float* queries_host, uint32_t nq,
                               uint32_t k, uint32_t R, uint64_t* out_ids, float* out_scores, nvdb_b200_timing* timing);

/* Test hook for the tensor-core scan: runs the NVDB_B200_PATH_MMA / _MMA_FAST scan of `nq` host queries (k = 1) and
 * returns the RAW accumulators of every (query, row) pair: f16 base -> out_acc[nq][n] float bit patterns (before
 * the power-of-two rescale); i8 base -> out_acc[2][nq][n] int32 limb dot products I1, I2 (exact integers).
 * Small cases only (nq * n * 8 bytes of device scratch). */
int nvdb_b200_debug_mma_accumulators(nvdb_b200_index* idx, const float* queries_host, uint32_t nq, int path,
                                     uint32_t* out_acc);

/* ---- exact refine -------------------------------------------------------------------------------------------
 * Replaces nvdb::cuda_l2_topk_batch: for each of Q queries, squared-L2 distance to its R candidate rows
 * (ids >= n, incl. 0xFFFFFFFF, are skipped), K smallest in ascending order; unfilled slots are
 * id = 0xFFFFFFFF, dist = 1e30f (src/cuda_refine.cu:891-894). out_dist may be NULL (CUDA_RETURN_DIST=0). */
int nvdb_b200_refine_l2(nvdb_b200_index* idx, const float* queries_host, const uint32_t* cand_ids_host,
                        uint32_t Q, uint32_t R, uint32_t K, uint32_t* out_ids, float* out_dist,
                        nvdb_b200_timing* timing);
int nvdb_b200_refine_l2_device(nvdb_b200_index* idx, const float* queries_dev, const uint32_t* cand_ids_dev,
                               uint32_t Q, uint32_t R, uint32_t K, uint32_t* out_ids_dev, float* out_dist_dev,
                               void* stream);

/* Page-lock caller memory (cudaHostRegister / cudaHostUnregister behind the C ABI, so that host code needs no CUDA
 * headers): host-buffer calls copy straight from / to registered or pinned buffers instead of staging them through the
 * library's own pinned buffers — the analogue of the reference's CUDA_PINNED=1 (src/cuda_refine.cu:875,903-916). */
int nvdb_b200_host_register(void* ptr, size_t bytes);
int nvdb_b200_host_unregister(void* ptr);

/* ---- IVF-PQ candidate generation ---------------------------------------------------------
 * The stage in front of the refine: replaces the FAISS calls of the reference's apps — build:
 * apps/nvdb_ivfpq_build.cpp:78-135 (IndexIVFPQ(IndexFlatL2, d, nlist, PQ_M, PQ_BITS, METRIC_L2), train on the first
 * IVF_TRAIN rows, add all rows); search: apps/nvdb_ivf_eval.cpp:398-408,482 (ivf->nprobe, idx->search(1, q, k_search)).
 * FAISS is not vendored by the reference and not present here: parity is UNPINNED; the algorithm is the
 * published IVFADC with residual encoding (+ optional OPQ rotation), checked against oracle/ivfpq_ref.py.
 * All kernels run on `base`'s device; base rows are read from its device-resident payload (f32 / f16 / i8+scale). */
typedef struct nvdb_b200_ivfpq nvdb_b200_ivfpq;
/* pq_bits must be 8; pq_m must divide dim (exit code 2 of nvdb_ivfpq_build); ntrain = 0 -> all rows (the tool's
 * IVF_TRAIN=200000 default is the caller's); niter <= 0 -> 25 Lloyd iterations. Deterministic for a given seed. */
int nvdb_b200_ivfpq_build(nvdb_b200_index* base, uint32_t nlist, uint32_t pq_m, uint32_t pq_bits, uint64_t ntrain,
                          int niter, uint64_t seed, nvdb_b200_ivfpq** out);
/* Same with an OPQ rotation in front (the reference's USE_OPQ=1: faiss::OPQMatrix + IndexPreTransform,
 * apps/nvdb_ivfpq_build.cpp:78-91, OPQ_NITER default 25): opq_niter rounds of {PQ of the rotated training rows,
 * orthogonal Procrustes update of the rotation}; 0 = no rotation. Queries are rotated inside search; the refine stage
 * works on the original rows. */
int nvdb_b200_ivfpq_build_opq(nvdb_b200_index* base, uint32_t nlist, uint32_t pq_m, uint32_t pq_bits, uint64_t ntrain,
                              int niter, int opq_niter, uint64_t seed, nvdb_b200_ivfpq** out);
/* The rotation R (dim x dim, row-major, x' = x R): *has_rotation = 0/1; rotation may be NULL. set_rotation installs one on
 * an imported index (NULL removes it). */
int nvdb_b200_ivfpq_rotation(const nvdb_b200_ivfpq* ivf, float* rotation, int* has_rotation);
int nvdb_b200_ivfpq_set_rotation(nvdb_b200_ivfpq* ivf, const float* rotation);
int nvdb_b200_ivfpq_destroy(nvdb_b200_ivfpq* ivf);
const char* nvdb_b200_ivfpq_last_error(const nvdb_b200_ivfpq* ivf);
int nvdb_b200_ivfpq_info(const nvdb_b200_ivfpq* ivf, uint64_t* n, uint32_t* dim, uint32_t* nlist, uint32_t* pq_m,
                         uint64_t* index_bytes);
/* Trained parameters and inverted lists to the host (any pointer may be NULL): coarse [nlist*dim], codebooks
 * [pq_m*256*(dim/pq_m)], list_off [nlist+1], ids [n] (row id of every grouped position), codes [n*pq_m] (grouped). */
int nvdb_b200_ivfpq_export(const nvdb_b200_ivfpq* ivf, float* coarse, float* codebooks, uint32_t* list_off, uint32_t* ids,
                           uint8_t* codes);
/* The inverse of export: an index from host arrays (e.g. loaded from a file the caller wrote from an export, or trained
 * elsewhere). Rows are NOT re-encoded; list_off must ascend from 0 to n. This is how an index is saved and re-loaded:
 * the reference's faiss::write_index / read_index file format belongs to FAISS and is not reproduced. */
int nvdb_b200_ivfpq_import(int device, uint32_t dim, uint32_t nlist, uint32_t pq_m, uint64_t n, const float* coarse,
                           const float* codebooks, const uint32_t* list_off, const uint32_t* ids, const uint8_t* codes,
                           nvdb_b200_ivfpq** out);
/* R best candidates per query by ADC distance over the nprobe nearest lists, ordered (distance asc, id asc); unfilled
 * slots id = 0xFFFFFFFF, dist = 1e30f (what the refine stage skips). 1 <= R <= 2048. out_dist may be NULL. */
int nvdb_b200_ivfpq_search(nvdb_b200_ivfpq* ivf, const float* queries_host, uint32_t nq, uint32_t nprobe, uint32_t R,
                           uint32_t* out_ids, float* out_dist);
int nvdb_b200_ivfpq_search_device(nvdb_b200_ivfpq* ivf, const float* queries_dev, uint32_t nq, uint32_t nprobe, uint32_t R,
                                  uint32_t* cand_ids_dev, float* cand_dist_dev, void* stream);
/* candidates -> exact L2 refine against `base` -> k best, with the candidates never leaving the device (the reference
 * moves them host -> device per batch: src/cuda_refine.cu:926-929). Output as nvdb_b200_refine_l2. */
int nvdb_b200_ivfpq_search_refine(nvdb_b200_ivfpq* ivf, nvdb_b200_index* base, const float* queries_host, uint32_t nq,
                                  uint32_t nprobe, uint32_t R, uint32_t k, uint32_t* out_ids, float* out_dist);

/* ---- synthetic data on the device ----------------------------------------------------------------------------
 * Fill caller-owned device buffers with global rows [first_row, first_row+n) of the synthetic generator in
 * `dtype` (payload n*dim elements; scales n floats for i8, else NULL). Asynchronous on `stream`. */
int nvdb_b200_synth_fill_device(int device, uint32_t dtype, uint64_t n, uint32_t dim, uint64_t seed,
                                uint64_t first_row, void* payload_dev, float* scales_dev, void* stream);

/* ---- converters ------------------------------------------------------------------------------------------------
 * The reference's fixture tools on the device, bit-identical outputs (tests/golden): f32 -> f16 is IEEE round-to-
 * nearest-even (tools/nvdb_convert_f16.cpp:20-119); f32 | f16 -> i8 is scale = max|x|/127 (1 for an all-zero row),
 * q = clamp(lrint(x * (1.0f/scale)), -127, 127) in FP32 with the scale stored per row (apps/nvdb_quantize_i8.cpp:57-85).
 * The host-buffer form streams the rows through the device in pinned chunks; dst_scales is required for i8 only. */
int nvdb_b200_convert(int device, uint32_t src_dtype, const void* src_host, uint32_t dst_dtype, uint64_t n, uint32_t dim,
                      void* dst_payload_host, float* dst_scales_host);
int nvdb_b200_convert_device(int device, uint32_t src_dtype, const void* src_dev, uint32_t dst_dtype, uint64_t n,
                             uint32_t dim, void* dst_payload_dev, float* dst_scales_dev, void* stream);

/* ---- introspection for the harness ------------------------------------------------------------------------ */
/* kernels launched by this library in this process so far (bench.py's gpu_launches). */
uint64_t nvdb_b200_launch_count(void);
/* Per-kernel CUDA-event timing of nvdb_b200_search*(): enable, run searches, then read the summed device
 * time of the scan kernels and of the merge kernel over `calls` searches (read synchronises and resets).
 * This is the analogue of the reference's ev1->ev2 kernel timing (src/cuda_refine.cu:1147-1157). */
int nvdb_b200_index_profile(nvdb_b200_index* idx, int enable);
int nvdb_b200_index_profile_read(nvdb_b200_index* idx, double* scan_ms_sum, double* merge_ms_sum, uint32_t* calls);

#ifdef __cplusplus
}
#endif
#endif /* NVDB_B200_H */
