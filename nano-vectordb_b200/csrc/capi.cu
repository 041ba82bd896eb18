// capi.cu — the C ABI (include/nvdb_b200.h): index handle, uploads, search / refine orchestration.
// No CPU compute path lives here: every result is produced by the kernels in this directory.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/nvdb_b200.h"
#include "common.cuh"
#include "exchange.h"
#include "scan.h"

using namespace nvdb_b200;

namespace nvdb_b200 {
static std::atomic<uint64_t> g_launches{0};
void count_launch(uint32_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace nvdb_b200

struct nvdb_b200_index {
  int device = 0;
  uint32_t dtype = 0;
  uint64_t n = 0;
  uint32_t dim = 0;
  uint64_t row_offset = 0;
  void* payload = nullptr;
  float* scales = nullptr;
  bool owns = false;
  int path = NVDB_B200_PATH_AUTO;
  int num_sms = 148;
  int max_smem = 227 * 1024;
  int max_smem_sm = 228 * 1024;  // per SM (all resident CTAs together)
  cudaStream_t stream = nullptr;
  // One stream per index: the workspaces, the streaming scan's done-counter and the query tensor map are per index, so
  // two searches of one index must not overlap. `cur_stream` is the stream of the latest call; a call on another
  // stream is ordered after it with an event (adopt_stream), never run concurrently.
  cudaStream_t cur_stream = nullptr;
  cudaEvent_t ev_switch = nullptr;
  // debug guard (env NVDB_B200_XCHG_NO_ALLOC=1): while an exchange call is in flight a workspace may not grow — ranks
  // sharing one device would dead-lock behind a spinning peer kernel
  bool guard_no_alloc = false, in_exchange = false;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  double upload_s = 0;         // host -> device time of the base (index_create), and the bytes moved
  uint64_t upload_bytes = 0;
  bool tma_ok = false;
  CUtensorMap tmap;
  // tensor-core scan: base tensor map with box rows = scan_mma_block_n(dtype); query (A operand) tensor map cache
  bool mma_ok = false;
  CUtensorMap tmap_mma_b64; // box rows 64: the half tile each CTA of an i8 CTA pair loads
  CUtensorMap tmap_mma_b;   // box rows 128 (the streaming scan's `tmap` has 256-row boxes and serves the 256-row tiles)
  CUtensorMap tmap_mma_a;
  const void* tmap_a_ptr = nullptr;
  uint32_t tmap_a_rows = 0;
  // grow-only workspaces
  void* d_ws[24] = {};
  size_t d_ws_bytes[24] = {};
  void* h_pin[3] = {nullptr, nullptr, nullptr};
  size_t h_pin_bytes[3] = {0, 0, 0};
  // exact tensor-core path (rescore.cu): row-norm maxima of the base (lazy) and certificate statistics
  bool norms_ready = false;
  float max_l2 = 0.f, max_l1 = 0.f;
  uint32_t rows_launches = 0;          // launches of the row-tile kernel that used the seeding workspace (its counter's epoch)
  const void* rows_seed_seen = nullptr;  // WS_ROWS_SEED allocation whose counter has been zeroed
  // adaptive first tier: running rate of first-certificate failures (per k); when most queries of large batches fail the
  // narrow lists anyway (clustered bases at k = 100: 61 %), the wide lists are scanned FIRST and the narrow pass is only
  // probed every 8th call
  float exact_fail_ewma = 0.f;
  uint32_t exact_ewma_k = 0, exact_batches = 0;
  bool exact_wide_first = false;
  uint64_t exact_calls = 0;            // candidate-scan + re-score chains enqueued: parity picks the flag counter slot
  const void* flag_ws_seen = nullptr;  // WS_X_FLAGS allocation whose counter slots have been zeroed
  uint64_t exact_queries = 0, exact_fallbacks = 0;
  uint64_t exact_tier2 = 0, exact_gemv = 0;  // flagged queries re-run with wider lists / on the streaming scan
  // tuning knobs (nvdb_b200_index_set_tuning; defaults may be overridden by environment variables at creation)
  int tune_mma_seed = 1;   // 0 off, 1 auto (>= 256 queries and >= 32 tiles per CTA), 2 whenever there are >= 2 tiles per CTA
  int tune_mma_resident = 1; // i8 tensor scan: query limb 0 resident in shared memory (0: stream both limbs)
  int tune_fused_merge = 1;  // final merge inside the streaming scan's last CTA (0: separate merge kernel)
  int tune_gemv_ctas = 1;  // CTAs per SM of the single-query streaming scan (1 | 2)
  int tune_mma_2cta = 1;   // CTA-pair (cta_group::2) tensor-core scan: 0 never, 1 (default) / 2 whenever the shape allows
  int tune_rows = 1;       // small batches of the exact path on the row-tile tensor kernel (scan_rows.cu): 0 never, 1 auto
                           // (int8 bases: measured faster than the query-tile kernels for every batch that fits, f16: never),
                           // 2 whenever the batch fits
  int tune_rows_seed = 1;  // its cross-CTA threshold seeding: 0 off, 1 auto (>= 40 queries, where it pays), 2 always
  int tune_exact_tier2 = 1; // exact path: flagged queries get a second tensor-core pass with wider candidate lists (0: off;
                            // > 1: that many list entries instead of the default width, for A/B measurements)
  std::mutex mu;
  mutable std::string err;
  // optional per-kernel event timing (bench.py's roofline figure)
  bool profiling = false;
  std::vector<cudaEvent_t> prof_ev;  // triples: before scan passes, after scan passes, after merge
  size_t prof_used = 0;
};

namespace {

thread_local std::string t_err;
enum { WS_QUERIES = 0, WS_PARTIAL = 1, WS_OUT_IDS = 2, WS_OUT_SCORES = 3, WS_CAND = 4, WS_OUT_IDS32 = 5, WS_MMA_A = 6,
       WS_MMA_QS = 7, WS_DBG = 8, WS_X_IDS = 9, WS_X_APX = 10, WS_X_FLAGS = 11, WS_X_Q = 12, WS_X_OUT_IDS = 13,
       WS_X_OUT_SC = 14, WS_SEED = 15, WS_GEMV_CTR = 16, WS_X_LIST = 17, WS_X_Q2 = 18, WS_X_OUT2 = 19, WS_ROWS_SEED = 20 };

int fail(const nvdb_b200_index* idx, int code, const std::string& msg) {
  t_err = msg;
  if (idx) idx->err = msg;
  return code;
}
int fail_cuda(const nvdb_b200_index* idx, cudaError_t e, const char* what) {
  return fail(idx, e == cudaErrorMemoryAllocation ? NVDB_B200_ERR_OOM : NVDB_B200_ERR_CUDA,
              std::string(what) + ": " + cudaGetErrorString(e));
}
#define CK(call, what)                                   \
  do {                                                   \
    cudaError_t _e = (call);                             \
    if (_e != cudaSuccess) return fail_cuda(idx, _e, what); \
  } while (0)

// Staging copies into pinned memory: one core moves ~10 GB/s, which is slower than the H2D copy that follows; large
// blocks are split over a few short-lived threads (>= 2 MB per thread).
void par_memcpy(void* dst, const void* src, size_t bytes, size_t max_threads = 6) {
  const size_t per_thread = (size_t)2 << 20;
  size_t T = bytes / per_thread;
  if (T > max_threads) T = max_threads;
  if (T <= 1) {  // the per-query path: no thread, no look at the machine
    std::memcpy(dst, src, bytes);
    return;
  }
  static const unsigned hw = std::thread::hardware_concurrency();
  if (hw && T > hw) T = hw;
  if (T <= 1) {
    std::memcpy(dst, src, bytes);
    return;
  }
  const size_t slice = ((bytes / T) + 4095) & ~(size_t)4095;
  std::vector<std::thread> th;
  th.reserve(T - 1);
  for (size_t t = 1; t < T; ++t) {
    const size_t off = t * slice;
    if (off >= bytes) break;
    const size_t len = bytes - off < slice ? bytes - off : slice;
    th.emplace_back([=] { std::memcpy((char*)dst + off, (const char*)src + off, len); });
  }
  std::memcpy(dst, src, slice < bytes ? slice : bytes);
  for (auto& x : th) x.join();
}

size_t elem_size(uint32_t dtype) { return dtype == kDtF32 ? 4 : dtype == kDtF16 ? 2 : dtype == kDtI8 ? 1 : 0; }

int ensure_ws(nvdb_b200_index* idx, int which, size_t bytes) {
  if (bytes <= idx->d_ws_bytes[which]) return NVDB_B200_OK;
  if (idx->guard_no_alloc && idx->in_exchange)
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "workspace " + std::to_string(which) + " would grow inside an exchange call (NVDB_B200_XCHG_NO_ALLOC)");
  if (idx->d_ws[which]) {
    CK(cudaStreamSynchronize(idx->stream), "sync before workspace grow");
    if (idx->cur_stream && idx->cur_stream != idx->stream) CK(cudaStreamSynchronize(idx->cur_stream), "sync before workspace grow");
    cudaFree(idx->d_ws[which]);
    idx->d_ws[which] = nullptr;
    idx->d_ws_bytes[which] = 0;
  }
  size_t cap = 4096;
  while (cap < bytes) cap <<= 1;
  CK(cudaMalloc(&idx->d_ws[which], cap), "cudaMalloc workspace");
  idx->d_ws_bytes[which] = cap;
  return NVDB_B200_OK;
}
int ensure_pin(nvdb_b200_index* idx, int which, size_t bytes) {
  if (bytes <= idx->h_pin_bytes[which]) return NVDB_B200_OK;
  if (idx->guard_no_alloc && idx->in_exchange)
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "pinned buffer would grow inside an exchange call (NVDB_B200_XCHG_NO_ALLOC)");
  if (idx->h_pin[which]) {
    CK(cudaStreamSynchronize(idx->stream), "sync before pinned grow");
    if (idx->cur_stream && idx->cur_stream != idx->stream) CK(cudaStreamSynchronize(idx->cur_stream), "sync before pinned grow");
    cudaFreeHost(idx->h_pin[which]);
    idx->h_pin[which] = nullptr;
    idx->h_pin_bytes[which] = 0;
  }
  size_t cap = 4096;
  while (cap < bytes) cap <<= 1;
  CK(cudaMallocHost(&idx->h_pin[which], cap), "cudaMallocHost");
  idx->h_pin_bytes[which] = cap;
  return NVDB_B200_OK;
}

// Calls of one index run on ONE stream at a time (the per-index state above): a call that arrives on another stream is
// ordered behind everything enqueued so far.
int adopt_stream(nvdb_b200_index* idx, cudaStream_t s) {
  if (idx->cur_stream == s) return NVDB_B200_OK;
  if (idx->cur_stream != nullptr || s != idx->stream) {
    cudaStream_t prev = idx->cur_stream ? idx->cur_stream : idx->stream;
    CK(cudaEventRecord(idx->ev_switch, prev), "stream switch record");
    CK(cudaStreamWaitEvent(s, idx->ev_switch, 0), "stream switch wait");
  }
  idx->cur_stream = s;
  return NVDB_B200_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess) return (EncodeTiledFn) nullptr;
    if (qres != cudaDriverEntryPointSuccess) return (EncodeTiledFn) nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// The streaming scan needs: whole 8-chains only (no scalar tail), 16-byte row pitch, 16-byte aligned base.
bool tma_eligible(const nvdb_b200_index* idx) {
  const uint32_t step = idx->dtype == kDtI8 ? 16u : 8u;
  if (idx->dim % step) return false;
  if ((idx->dim * elem_size(idx->dtype)) % 16) return false;
  if (reinterpret_cast<uintptr_t>(idx->payload) % 16) return false;
  if (idx->n > 0x7FFFFF00ull) return false;
  return true;
}

int encode_2d(const nvdb_b200_index* idx, CUtensorMap* out, uint32_t dtype, const void* ptr, uint64_t rows, uint32_t dim,
              uint32_t box_rows) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(idx, NVDB_B200_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  const size_t es = elem_size(dtype);
  const CUtensorMapDataType dt = dtype == kDtF32   ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : dtype == kDtF16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                                   : CU_TENSOR_MAP_DATA_TYPE_UINT8;
  cuuint64_t gdim[2] = {dim, rows};
  cuuint64_t gstride[1] = {(cuuint64_t)dim * es};
  cuuint32_t box[2] = {(cuuint32_t)(128 / es), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, dt, 2, const_cast<void*>(ptr), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(idx, NVDB_B200_ERR_CUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
  return NVDB_B200_OK;
}

int build_tmap(nvdb_b200_index* idx) {
  idx->tma_ok = idx->mma_ok = false;
  if (!tma_eligible(idx)) return NVDB_B200_OK;
  int rc = encode_2d(idx, &idx->tmap, idx->dtype, idx->payload, idx->n, idx->dim, (uint32_t)kScanRowsPerTile);
  if (rc) return rc;
  idx->tma_ok = true;
  if (idx->dtype == kDtF16 || idx->dtype == kDtI8) {
    rc = encode_2d(idx, &idx->tmap_mma_b, idx->dtype, idx->payload, idx->n, idx->dim, 128u);
    if (rc) return rc;
    rc = encode_2d(idx, &idx->tmap_mma_b64, idx->dtype, idx->payload, idx->n, idx->dim, 64u);
    if (rc) return rc;
    idx->mma_ok = true;
  }
  return NVDB_B200_OK;
}

int new_handle(uint32_t dtype, uint64_t n, uint32_t dim, int device, uint64_t row_offset, nvdb_b200_index** out) {
  const nvdb_b200_index* idx = nullptr;
  if (!out) return fail(nullptr, NVDB_B200_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (elem_size(dtype) == 0) return fail(nullptr, NVDB_B200_ERR_ARG, "Unsupported base dtype (Float32/Float16/Int8 only)");
  if (n == 0 || dim == 0) return fail(nullptr, NVDB_B200_ERR_EMPTY, "Empty base");
  if (n > 0xFFFFFFF0ull) return fail(nullptr, NVDB_B200_ERR_UNSUPPORTED, "more than 2^32 rows per device shard");
  CK(cudaSetDevice(device), "cudaSetDevice");
  auto* h = new nvdb_b200_index();
  h->device = device;
  h->dtype = dtype;
  h->n = n;
  h->dim = dim;
  h->row_offset = row_offset;
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e == cudaSuccess) {
    h->num_sms = prop.multiProcessorCount;
    h->max_smem = (int)prop.sharedMemPerBlockOptin;
    h->max_smem_sm = (int)prop.sharedMemPerMultiprocessor;
  }
  if (const char* ev = getenv("NVDB_B200_MMA_SEED")) h->tune_mma_seed = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_GEMV_CTAS")) h->tune_gemv_ctas = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_MMA_2CTA")) h->tune_mma_2cta = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_FUSED_MERGE")) h->tune_fused_merge = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_EXACT_TIER2")) h->tune_exact_tier2 = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_ROWS")) h->tune_rows = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_ROWS_SEED")) h->tune_rows_seed = atoi(ev);
  if (const char* ev = getenv("NVDB_B200_XCHG_NO_ALLOC")) h->guard_no_alloc = atoi(ev) != 0;
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_switch, cudaEventDisableTiming);
  for (int i = 0; i < 4 && e == cudaSuccess; ++i) e = cudaEventCreate(&h->ev[i]);
  if (e != cudaSuccess) {
    delete h;
    return fail_cuda(nullptr, e, "index init");
  }
  *out = h;
  return NVDB_B200_OK;
}

}  // namespace

// =====================================================================================================
// lifecycle
// =====================================================================================================
extern "C" {

int nvdb_b200_version(void) { return NVDB_B200_VERSION; }

const char* nvdb_b200_status_string(int s) {
  switch (s) {
    case NVDB_B200_OK: return "ok";
    case NVDB_B200_ERR_ARG: return "invalid argument";
    case NVDB_B200_ERR_EMPTY: return "Empty base";
    case NVDB_B200_ERR_CUDA: return "CUDA error";
    case NVDB_B200_ERR_OOM: return "out of device memory";
    case NVDB_B200_ERR_UNSUPPORTED: return "unsupported";
    default: return "unknown status";
  }
}

const char* nvdb_b200_last_error(const nvdb_b200_index* idx) { return idx ? idx->err.c_str() : t_err.c_str(); }

int nvdb_b200_device_count(int* out_count) {
  const nvdb_b200_index* idx = nullptr;
  if (!out_count) return fail(nullptr, NVDB_B200_ERR_ARG, "out_count is NULL");
  *out_count = 0;
  CK(cudaGetDeviceCount(out_count), "cudaGetDeviceCount");
  return NVDB_B200_OK;
}

uint64_t nvdb_b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int nvdb_b200_index_destroy(nvdb_b200_index* idx) {
  if (!idx) return NVDB_B200_OK;
  cudaSetDevice(idx->device);
  if (idx->stream) cudaStreamSynchronize(idx->stream);
  if (idx->cur_stream && idx->cur_stream != idx->stream) cudaStreamSynchronize(idx->cur_stream);
  if (idx->owns) {
    if (idx->payload) cudaFree(idx->payload);
    if (idx->scales) cudaFree(idx->scales);
  }
  for (auto& p : idx->d_ws)
    if (p) cudaFree(p);
  for (auto& p : idx->h_pin)
    if (p) cudaFreeHost(p);
  for (auto& e : idx->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : idx->prof_ev) cudaEventDestroy(e);
  if (idx->ev_switch) cudaEventDestroy(idx->ev_switch);
  if (idx->stream) cudaStreamDestroy(idx->stream);
  delete idx;
  return NVDB_B200_OK;
}

static int alloc_base(nvdb_b200_index* idx) {
  const size_t bytes = (size_t)idx->n * idx->dim * elem_size(idx->dtype);
  CK(cudaMalloc(&idx->payload, bytes), "cudaMalloc base payload");
  idx->owns = true;
  if (idx->dtype == kDtI8) CK(cudaMalloc((void**)&idx->scales, (size_t)idx->n * 4), "cudaMalloc scales");
  return NVDB_B200_OK;
}

// Upload path: host (typically the mmap of a vecbin file) -> two pinned staging buffers -> cudaMemcpyAsync,
// the memcpy of chunk i+1 overlapping the DMA of chunk i. The reference does one synchronous pageable
// cudaMemcpy of the whole base (src/cuda_refine.cu:198).
// Process-wide staging buffers of the loader: three 64 MB pinned chunks, allocated on the first large upload and kept
// (page-locking 128 MB cost ~45 ms of a 160 ms upload when every index allocated its own). Uploads serialise on them.
namespace {
constexpr size_t kStageChunk = (size_t)64 << 20;
constexpr int kStageBufs = 3;
std::mutex g_stage_mu;
void* g_stage[kStageBufs] = {nullptr, nullptr, nullptr};
}  // namespace

static int upload(nvdb_b200_index* idx, void* dst, const void* src, size_t bytes) {
  const size_t chunk = kStageChunk;
  const auto t0 = std::chrono::steady_clock::now();
  int rc;
  if (bytes <= chunk) {  // small payloads (tests, scales of small bases): the index's own pinned buffer
    if ((rc = ensure_pin(idx, 0, bytes))) return rc;
    par_memcpy(idx->h_pin[0], src, bytes, 16);
    CK(cudaMemcpyAsync(dst, idx->h_pin[0], bytes, cudaMemcpyHostToDevice, idx->stream), "H2D base");
    CK(cudaStreamSynchronize(idx->stream), "upload sync");
  } else {
    std::lock_guard<std::mutex> g(g_stage_mu);
    for (auto& p : g_stage)
      if (!p) CK(cudaHostAlloc(&p, chunk, cudaHostAllocPortable), "cudaHostAlloc staging");
    cudaEvent_t done[kStageBufs];
    bool used[kStageBufs] = {false, false, false};
    for (auto& e : done) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming), "upload event");
    size_t off = 0;
    int b = 0;
    cudaError_t err = cudaSuccess;
    while (off < bytes && err == cudaSuccess) {
      const size_t take = bytes - off < chunk ? bytes - off : chunk;
      if (used[b]) err = cudaEventSynchronize(done[b]);
      // the staging copy (page faults of a cold mmap included) is the slow side: one core moves ~10 GB/s, the copy
      // engine ~50 GB/s, so each chunk is split over up to 16 threads; with three buffers two copies can be queued
      // behind the one in flight
      par_memcpy(g_stage[b], (const char*)src + off, take, 16);
      if (err == cudaSuccess) err = cudaMemcpyAsync((char*)dst + off, g_stage[b], take, cudaMemcpyHostToDevice, idx->stream);
      if (err == cudaSuccess) err = cudaEventRecord(done[b], idx->stream);
      used[b] = true;
      off += take;
      b = (b + 1) % kStageBufs;
    }
    if (err == cudaSuccess) err = cudaStreamSynchronize(idx->stream);
    for (auto& e : done) cudaEventDestroy(e);
    CK(err, "H2D base chunks");
  }
  idx->upload_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  idx->upload_bytes += bytes;
  return NVDB_B200_OK;
}

int nvdb_b200_index_create(const void* payload_host, const float* scales_host, uint32_t dtype, uint64_t n,
                           uint32_t dim, int device, uint64_t row_offset, nvdb_b200_index** out) {
  if (!payload_host) return fail(nullptr, NVDB_B200_ERR_ARG, "payload is NULL");
  if (dtype == kDtI8 && !scales_host) return fail(nullptr, NVDB_B200_ERR_ARG, "Dataset is not int8 (scales missing)");
  int rc = new_handle(dtype, n, dim, device, row_offset, out);
  if (rc) return rc;
  nvdb_b200_index* idx = *out;
  rc = alloc_base(idx);
  if (!rc) rc = upload(idx, idx->payload, payload_host, (size_t)n * dim * elem_size(dtype));
  if (!rc && dtype == kDtI8) rc = upload(idx, idx->scales, scales_host, (size_t)n * 4);
  if (!rc) rc = build_tmap(idx);
  if (rc) {
    t_err = idx->err;
    nvdb_b200_index_destroy(idx);
    *out = nullptr;
  }
  return rc;
}

int nvdb_b200_index_create_device(const void* payload_dev, const float* scales_dev, uint32_t dtype, uint64_t n,
                                  uint32_t dim, int device, uint64_t row_offset, nvdb_b200_index** out) {
  if (!payload_dev) return fail(nullptr, NVDB_B200_ERR_ARG, "payload is NULL");
  if (dtype == kDtI8 && !scales_dev) return fail(nullptr, NVDB_B200_ERR_ARG, "Dataset is not int8 (scales missing)");
  int rc = new_handle(dtype, n, dim, device, row_offset, out);
  if (rc) return rc;
  nvdb_b200_index* idx = *out;
  idx->payload = const_cast<void*>(payload_dev);
  idx->scales = const_cast<float*>(scales_dev);
  idx->owns = false;
  rc = build_tmap(idx);
  if (rc) {
    t_err = idx->err;
    nvdb_b200_index_destroy(idx);
    *out = nullptr;
  }
  return rc;
}

int nvdb_b200_convert_device(int device, uint32_t src_dtype, const void* src_dev, uint32_t dst_dtype, uint64_t n,
                             uint32_t dim, void* dst_payload_dev, float* dst_scales_dev, void* stream) {
  const nvdb_b200_index* idx = nullptr;
  const bool ok = (src_dtype == kDtF32 && dst_dtype == kDtF16) || ((src_dtype == kDtF32 || src_dtype == kDtF16) && dst_dtype == kDtI8);
  if (!ok) return fail(nullptr, NVDB_B200_ERR_ARG, "convert: supported conversions are f32->f16 and f32|f16->i8");
  if (n == 0 || dim == 0) return NVDB_B200_OK;
  if (!src_dev || !dst_payload_dev || (dst_dtype == kDtI8 && !dst_scales_dev))
    return fail(nullptr, NVDB_B200_ERR_ARG, "convert: NULL buffer");
  CK(cudaSetDevice(device), "cudaSetDevice");
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
  CK(convert_launch((int)src_dtype, src_dev, (int)dst_dtype, n, dim, dst_payload_dev, dst_scales_dev, prop.multiProcessorCount,
                    (cudaStream_t)stream),
     "convert launch");
  return NVDB_B200_OK;
}

// Host-buffer converter (what nvdb_convert_f16 / nvdb_quantize_i8 do on the CPU): rows stream through the device in
// chunks — pinned staging, H2D, convert kernel, D2H — two chunks in flight.
int nvdb_b200_convert(int device, uint32_t src_dtype, const void* src_host, uint32_t dst_dtype, uint64_t n, uint32_t dim,
                      void* dst_payload_host, float* dst_scales_host) {
  const nvdb_b200_index* idx = nullptr;
  const bool ok = (src_dtype == kDtF32 && dst_dtype == kDtF16) || ((src_dtype == kDtF32 || src_dtype == kDtF16) && dst_dtype == kDtI8);
  if (!ok) return fail(nullptr, NVDB_B200_ERR_ARG, "convert: supported conversions are f32->f16 and f32|f16->i8");
  if (n == 0 || dim == 0) return NVDB_B200_OK;
  if (!src_host || !dst_payload_host || (dst_dtype == kDtI8 && !dst_scales_host))
    return fail(nullptr, NVDB_B200_ERR_ARG, "convert: NULL buffer");
  CK(cudaSetDevice(device), "cudaSetDevice");
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
  const size_t ses = elem_size(src_dtype), des = elem_size(dst_dtype);
  const size_t src_row = (size_t)dim * ses, dst_row = (size_t)dim * des;
  uint64_t chunk = (64ull << 20) / src_row;  // ~64 MB of source rows per chunk
  if (chunk == 0) chunk = 1;
  if (chunk > n) chunk = n;
  struct Slot {
    void *h_src = nullptr, *h_dst = nullptr, *d_src = nullptr, *d_dst = nullptr;
    float *h_sc = nullptr, *d_sc = nullptr;
    cudaStream_t s = nullptr;
    uint64_t row0 = 0, rows = 0;
  } slot[2];
  int rc = NVDB_B200_OK;
  auto cleanup = [&] {
    for (auto& sl : slot) {
      if (sl.s) cudaStreamSynchronize(sl.s);
      cudaFreeHost(sl.h_src);
      cudaFreeHost(sl.h_dst);
      cudaFreeHost(sl.h_sc);
      cudaFree(sl.d_src);
      cudaFree(sl.d_dst);
      cudaFree(sl.d_sc);
      if (sl.s) cudaStreamDestroy(sl.s);
    }
  };
#define CKC(call, what)                      \
  do {                                       \
    cudaError_t _e = (call);                 \
    if (_e != cudaSuccess) {                 \
      rc = fail_cuda(idx, _e, what);         \
      cleanup();                             \
      return rc;                             \
    }                                        \
  } while (0)
  for (auto& sl : slot) {
    CKC(cudaStreamCreateWithFlags(&sl.s, cudaStreamNonBlocking), "convert stream");
    CKC(cudaMallocHost(&sl.h_src, chunk * src_row), "convert pinned");
    CKC(cudaMallocHost(&sl.h_dst, chunk * dst_row), "convert pinned");
    CKC(cudaMallocHost((void**)&sl.h_sc, chunk * 4), "convert pinned");
    CKC(cudaMalloc(&sl.d_src, chunk * src_row), "convert device buffer");
    CKC(cudaMalloc(&sl.d_dst, chunk * dst_row), "convert device buffer");
    CKC(cudaMalloc((void**)&sl.d_sc, chunk * 4), "convert device buffer");
  }
  auto drain = [&](Slot& sl) -> cudaError_t {
    if (sl.rows == 0) return cudaSuccess;
    cudaError_t e = cudaStreamSynchronize(sl.s);
    if (e != cudaSuccess) return e;
    std::memcpy((char*)dst_payload_host + sl.row0 * dst_row, sl.h_dst, sl.rows * dst_row);
    if (dst_dtype == kDtI8) std::memcpy(dst_scales_host + sl.row0, sl.h_sc, sl.rows * 4);
    sl.rows = 0;
    return cudaSuccess;
  };
  int cur = 0;
  for (uint64_t r0 = 0; r0 < n; r0 += chunk, cur ^= 1) {
    Slot& sl = slot[cur];
    CKC(drain(sl), "convert drain");
    const uint64_t rows = n - r0 < chunk ? n - r0 : chunk;
    std::memcpy(sl.h_src, (const char*)src_host + r0 * src_row, rows * src_row);
    CKC(cudaMemcpyAsync(sl.d_src, sl.h_src, rows * src_row, cudaMemcpyHostToDevice, sl.s), "convert H2D");
    CKC(convert_launch((int)src_dtype, sl.d_src, (int)dst_dtype, rows, dim, sl.d_dst, sl.d_sc, prop.multiProcessorCount, sl.s),
        "convert launch");
    CKC(cudaMemcpyAsync(sl.h_dst, sl.d_dst, rows * dst_row, cudaMemcpyDeviceToHost, sl.s), "convert D2H");
    if (dst_dtype == kDtI8) CKC(cudaMemcpyAsync(sl.h_sc, sl.d_sc, rows * 4, cudaMemcpyDeviceToHost, sl.s), "convert D2H scales");
    sl.row0 = r0;
    sl.rows = rows;
  }
  CKC(drain(slot[0]), "convert drain");
  CKC(drain(slot[1]), "convert drain");
#undef CKC
  cleanup();
  return NVDB_B200_OK;
}

int nvdb_b200_synth_fill_device(int device, uint32_t dtype, uint64_t n, uint32_t dim, uint64_t seed,
                                uint64_t first_row, void* payload_dev, float* scales_dev, void* stream) {
  const nvdb_b200_index* idx = nullptr;
  if (!payload_dev || elem_size(dtype) == 0 || (dtype == kDtI8 && !scales_dev))
    return fail(nullptr, NVDB_B200_ERR_ARG, "synth_fill: bad arguments");
  CK(cudaSetDevice(device), "cudaSetDevice");
  CK(synth_fill_launch((int)dtype, n, dim, seed, first_row, payload_dev, scales_dev, (cudaStream_t)stream), "synth_fill");
  return NVDB_B200_OK;
}

int nvdb_b200_index_create_synthetic(uint32_t dtype, uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row,
                                     int device, uint64_t row_offset, nvdb_b200_index** out) {
  int rc = new_handle(dtype, n, dim, device, row_offset, out);
  if (rc) return rc;
  nvdb_b200_index* idx = *out;
  rc = alloc_base(idx);
  if (!rc) {
    cudaError_t e = synth_fill_launch((int)dtype, n, dim, seed, first_row, idx->payload, idx->scales, idx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(idx->stream);
    if (e != cudaSuccess) rc = fail_cuda(idx, e, "synthetic fill");
  }
  if (!rc) rc = build_tmap(idx);
  if (rc) {
    t_err = idx->err;
    nvdb_b200_index_destroy(idx);
    *out = nullptr;
  }
  return rc;
}

int nvdb_b200_index_info(const nvdb_b200_index* idx, uint64_t* n, uint32_t* dim, uint32_t* dtype, int* device,
                         uint64_t* row_offset, const void** payload_dev, const float** scales_dev) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (n) *n = idx->n;
  if (dim) *dim = idx->dim;
  if (dtype) *dtype = idx->dtype;
  if (device) *device = idx->device;
  if (row_offset) *row_offset = idx->row_offset;
  if (payload_dev) *payload_dev = idx->payload;
  if (scales_dev) *scales_dev = idx->scales;
  return NVDB_B200_OK;
}

int nvdb_b200_index_upload_info(const nvdb_b200_index* idx, double* seconds, uint64_t* bytes) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (seconds) *seconds = idx->upload_s;
  if (bytes) *bytes = idx->upload_bytes;
  return NVDB_B200_OK;
}

int nvdb_b200_index_set_path(nvdb_b200_index* idx, int path) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (path < NVDB_B200_PATH_AUTO || path > NVDB_B200_PATH_MMA_EXACT) return fail(idx, NVDB_B200_ERR_ARG, "bad path");
  idx->path = path;
  return NVDB_B200_OK;
}

int nvdb_b200_index_read_rows(const nvdb_b200_index* idx, uint64_t row0, uint64_t nrows, void* payload_host,
                              float* scales_host) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (row0 + nrows > idx->n) return fail(idx, NVDB_B200_ERR_ARG, "Index out of range");
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  CK(cudaStreamSynchronize(idx->stream), "sync");
  if (idx->cur_stream && idx->cur_stream != idx->stream) CK(cudaStreamSynchronize(idx->cur_stream), "sync");
  const size_t rb = (size_t)idx->dim * elem_size(idx->dtype);
  if (payload_host && nrows)
    CK(cudaMemcpy(payload_host, (const char*)idx->payload + row0 * rb, nrows * rb, cudaMemcpyDeviceToHost), "D2H rows");
  if (scales_host && idx->scales && nrows)
    CK(cudaMemcpy(scales_host, idx->scales + row0, nrows * 4, cudaMemcpyDeviceToHost), "D2H scales");
  return NVDB_B200_OK;
}

}  // extern "C"

// =====================================================================================================
// flat scan
// =====================================================================================================
namespace {

// done-counter of the streaming scan's fused merge: zeroed once, on the stream of the call that creates it (the merging
// CTA resets it after every launch)
int ensure_gemv_ctr(nvdb_b200_index* idx, cudaStream_t stream) {
  if (idx->d_ws[WS_GEMV_CTR] != nullptr) return NVDB_B200_OK;
  int rc = ensure_ws(idx, WS_GEMV_CTR, 256);
  if (rc) return rc;
  CK(cudaMemsetAsync(idx->d_ws[WS_GEMV_CTR], 0, 256, stream), "memset merge counter");
  return NVDB_B200_OK;
}

struct ScanPlan {
  int nq_template = 1;
  int stages = 4;
  uint32_t cap = 512;
  int grid = 1;
};

// Queries per pass: as many as fit next to >= 3 ring stages (each 32 KB) in shared memory, at most 8.
int gemv_ctas(const nvdb_b200_index* idx) {
  // measured (profiles/r1_gemv_ffma2_ctas.jsonl): one CTA with the deep ring beats two for every dtype
  return idx->tune_gemv_ctas == 2 ? 2 : 1;
}

// NVDB_B200_GEMV_MAX_STAGES=3..6 (0 = unset); read per call: the A/B probes and a test flip it inside one process
static int gemv_max_stages_override() {
  const char* e = getenv("NVDB_B200_GEMV_MAX_STAGES");
  if (!e) return 0;
  const int v = atoi(e);
  return (v >= 3 && v <= 6) ? v : 0;
}

int plan_scan(const nvdb_b200_index* idx, uint32_t nq, uint32_t k, ScanPlan* p) {
  const uint64_t need = (uint64_t)k + kScanRowsPerTile;
  if (need > 8192) return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "k too large for the fused selector (k <= 7936)");
  p->cap = next_pow2((uint32_t)need);
  if (p->cap < 512) p->cap = 512;
  const uint32_t tiles = (uint32_t)((idx->n + kScanRowsPerTile - 1) / kScanRowsPerTile);
  p->grid = (int)(tiles < (uint32_t)idx->num_sms ? tiles : (uint32_t)idx->num_sms);
  int want = nq >= 8 ? 8 : nq >= 4 ? 4 : nq >= 2 ? 2 : 1;
  if (nq == 3) want = 4;
  if (nq > 4 && nq < 8) want = 8;
  for (int t = want; t >= 1; t >>= 1) {
    if (!idx->tma_ok) {
      if (scan_gemv_smem_bytes(false, 0, t, idx->dim, p->cap) <= (uint32_t)idx->max_smem) {
        p->nq_template = t;
        p->stages = 0;
        return NVDB_B200_OK;
      }
      continue;
    }
    const uint32_t fixed = scan_gemv_smem_bytes(true, 0, t, idx->dim, p->cap);
    if (fixed + 3u * 32768u > (uint32_t)idx->max_smem) continue;
    int st = (int)(((uint32_t)idx->max_smem - fixed) / (32768u + 16u));
    // Ring depth: FP16 rows run fastest with 4 x 32 KB in flight per SM (batch 1, 2.9 M x 768: 6 stages 0.68-0.70 ms,
    // 5 0.67, 4 0.64-0.65, 3 0.65-0.67; profiles/r2_ring_depth_sweep.jsonl) - a deeper ring spreads the requests of
    // 148 SMs over more DRAM pages at a time. INT8 (issue-bound) and FP32 do not care / prefer 5-6.
    // NVDB_B200_GEMV_MAX_STAGES=3..6 overrides (A/B).
    int st_cap = idx->dtype == NVDB_B200_F16 ? 4 : 6;
    if (const int v = gemv_max_stages_override()) st_cap = v;
    if (st > st_cap) st = st_cap;
    // single-query passes may run two CTAs per SM (more consumer warps to cover the int8 conversion chain); each CTA
    // then gets half of the shared memory. Tuning knob NVDB_B200_TUNE_GEMV_CTAS (default 1).
    if (t == 1 && gemv_ctas(idx) == 2) {
      const uint32_t half = ((uint32_t)idx->max_smem_sm / 2u) - 1024u;  // 1 KB per-CTA reservation
      int st2 = half > fixed ? (int)((half - fixed) / (32768u + 16u)) : 0;
      if (st2 >= 3 && tiles >= 2u * (uint32_t)idx->num_sms) {
        st = st2 > 6 ? 6 : st2;
        p->grid = 2 * idx->num_sms;
      }
    }
    while (st > 3 && scan_gemv_smem_bytes(true, st, t, idx->dim, p->cap) > (uint32_t)idx->max_smem) --st;
    p->nq_template = t;
    p->stages = st;
    return NVDB_B200_OK;
  }
  return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "dim / k too large for shared memory");
}

int prof_begin(nvdb_b200_index* idx, cudaStream_t stream, cudaEvent_t** pev) {
  *pev = nullptr;
  if (!idx->profiling) return NVDB_B200_OK;
  if (idx->prof_used + 3 > idx->prof_ev.size()) {
    for (int i = 0; i < 3; ++i) {
      cudaEvent_t e;
      CK(cudaEventCreate(&e), "profile event");
      idx->prof_ev.push_back(e);
    }
  }
  *pev = &idx->prof_ev[idx->prof_used];
  idx->prof_used += 3;
  CK(cudaEventRecord((*pev)[0], stream), "profile event record");
  return NVDB_B200_OK;
}

// ---- tensor-core path: prep (limb split) -> scan_mma_kernel -> merge of the per-CTA lists -------------------
// `dbg` (tests only): raw accumulators [NACC][mpad][n]; mpad_out receives the padded query count.
int mma_search_impl(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                    float* d_out_scores, cudaStream_t stream, uint32_t* launches, int limbs, void* dbg,
                    uint64_t row_offset) {
  if (!idx->mma_ok)
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED,
                "tensor-core scan needs an f16 / i8 base with a 16-byte row pitch and whole 8/16-element chains");
  if (idx->dtype == kDtI8) limbs = 2;
  int stages = 0;
  uint32_t block_n = 0;
  if (!scan_mma_plan((int)idx->dtype, limbs, k, idx->max_smem, &block_n, &stages))
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "k too large for the tensor-core scan's shared memory");
  const size_t es = elem_size(idx->dtype);
  const uint32_t max_q = 128u * (uint32_t)idx->num_sms;
  cudaEvent_t* pev = nullptr;
  int rc = prof_begin(idx, stream, &pev);
  if (rc) return rc;
  // every pass handles up to 128 * num_sms queries; G M-tiles x P CTAs (or CTA pairs) each.
  // CTA-pair kernel (scan_mma2.cu): an even number of M-tiles, P = clusters per pair column; f16: 256-row tiles,
  // i8: 128-row tiles (two limb accumulators), each CTA loading half a tile. The pair halves the base-tile bytes every
  // SM writes to and reads from shared memory per MMA step; both single-CTA kernels are bound by that shared-memory
  // traffic, not by the tensor pipe (profiles/README.md). Knob MMA_2CTA: 0 never, 1 / 2 whenever the shape allows.
  int stages2 = 0;
  const bool pair_ok = !dbg && idx->tune_mma_2cta != 0 && scan_mma2_plan((int)idx->dtype, limbs, k, idx->max_smem, &stages2);
  const uint32_t pair_bn = 2u * scan_mma2_half_n((int)idx->dtype);
  struct PassPlan {
    bool pair;
    uint32_t G, P, bn, tiles;
  };
  auto plan_pass = [&](uint32_t nqp) {
    PassPlan pp;
    pp.G = (nqp + 127u) / 128u;
    // an odd number (>= 3) of query tiles: one more, empty, tile buys the CTA-pair kernel. Measured
    // (profiles/r1_ab_pair_pad_odd.jsonl): i8 -7 / -12 / -21 % at 3 / 5 / 7 tiles, f16 top-100 -7 % at 3 tiles, but f16
    // top-10 +15 / +15 / +7 % (its single-CTA kernel with register lists is the better one there). Knob MMA_2CTA = 2
    // forces the padding, 1 (default) follows the measurement.
    const bool pad_default = idx->dtype == kDtI8 || k > 16u;
    if (pair_ok && pp.G >= 3u && (pp.G % 2u) == 1u && pp.G + 1u <= (uint32_t)idx->num_sms &&
        (idx->tune_mma_2cta == 2 || (idx->tune_mma_2cta == 1 && pad_default)))
      pp.G += 1u;
    pp.pair = pair_ok && pp.G >= 2u && (pp.G % 2u) == 0u;
    pp.bn = pp.pair ? pair_bn : block_n;
    pp.tiles = (uint32_t)((idx->n + pp.bn - 1) / pp.bn);
    pp.P = pp.pair ? ((uint32_t)idx->num_sms / 2u) / (pp.G / 2u) : (uint32_t)idx->num_sms / pp.G;
    if (pp.P > pp.tiles) pp.P = pp.tiles;
    return pp;
  };
  uint32_t maxP = 1;
  for (uint32_t q0 = 0; q0 < nq; q0 += max_q) {
    const PassPlan pp = plan_pass(nq - q0 < max_q ? nq - q0 : max_q);
    if (pp.P > maxP) maxP = pp.P;
  }
  if ((rc = ensure_ws(idx, WS_PARTIAL, (size_t)nq * (maxP + 1) * k * 8))) return rc;
  for (uint32_t q0 = 0; q0 < nq; q0 += max_q) {
    const uint32_t nqp = nq - q0 < max_q ? nq - q0 : max_q;
    const PassPlan pp = plan_pass(nqp);
    const uint32_t tiles_all = pp.tiles;
    MmaArgs a{};
    a.G = pp.G;
    a.P = pp.P;
    a.mpad = a.G * 128u;
    if ((rc = ensure_ws(idx, WS_MMA_A, (size_t)limbs * a.mpad * idx->dim * es))) return rc;
    if ((rc = ensure_ws(idx, WS_MMA_QS, (size_t)3 * a.mpad * 4))) return rc;
    void* A = idx->d_ws[WS_MMA_A];
    float* qs1 = (float*)idx->d_ws[WS_MMA_QS];
    float* qs2 = qs1 + a.mpad;
    float* thr_seed = qs2 + a.mpad;
    if (idx->tmap_a_ptr != A || idx->tmap_a_rows != (uint32_t)limbs * a.mpad) {
      if ((rc = encode_2d(idx, &idx->tmap_mma_a, idx->dtype, A, (uint64_t)limbs * a.mpad, idx->dim, 128u))) return rc;
      idx->tmap_a_ptr = A;
      idx->tmap_a_rows = (uint32_t)limbs * a.mpad;
    }
    CK(scan_mma_prep_launch((int)idx->dtype, limbs, d_queries + (size_t)q0 * idx->dim, nqp, idx->dim, a.mpad, A, qs1, qs2,
                            stream),
       "mma query prep launch");
    a.scales = idx->scales;
    a.qscale = qs1;
    a.qscale2 = qs2;
    a.dbg = dbg;
    a.n_rows = (uint32_t)idx->n;
    a.dim = idx->dim;
    a.k = k;
    a.nq = nqp;
    a.kchunks = (uint32_t)(((size_t)idx->dim * es + 127) / 128);
    a.stages = pp.pair ? stages2 : stages;
    a.block_n = pp.bn;
    a.evict_first_base = (a.G == 1);
    a.resident_a = idx->tune_mma_resident;
    a.max_smem = idx->max_smem;
    const CUtensorMap* tmb = pp.pair ? (idx->dtype == kDtI8 ? &idx->tmap_mma_b64 : &idx->tmap_mma_b)
                                     : (pp.bn == 256 ? &idx->tmap : &idx->tmap_mma_b);
    auto launch_scan = [&](const MmaArgs& args) {
      return pp.pair ? scan_mma2_launch((int)idx->dtype, limbs, &idx->tmap_mma_a, tmb, args, stream)
                     : scan_mma_launch((int)idx->dtype, limbs, &idx->tmap_mma_a, tmb, args, stream);
    };
    uint64_t* partial = (uint64_t*)idx->d_ws[WS_PARTIAL];
    // Threshold seeding: a first pass over ~1/16 of the base (the same number of tiles for every CTA) gives each query
    // its k-th best score among those rows — a valid lower bound of its final k-th best. The main pass over the
    // remaining rows starts from that threshold, so its epilogue threads admit ~k*15/P rows instead of
    // ~k*ln(rows per CTA / k) (the per-thread list insert is the expensive, warp-divergent part of the epilogue).
    // Auto mode seeds only multi-tile batches (G >= 2): a single M-tile is HBM-bound and the two extra launches cost more
    // than the inserts they save (profiles/r1_ab_*.jsonl).
    // The seed pass's own top-k joins the final merge as list number P.
    uint32_t seed_tiles = 0;
    if (!dbg && idx->tune_mma_seed && (idx->tune_mma_seed == 2 ? tiles_all >= 2u * a.P : (a.G >= 2u && tiles_all >= 32u * a.P))) {
      seed_tiles = ((tiles_all / 16u + a.P - 1u) / a.P) * a.P;
      if ((rc = ensure_ws(idx, WS_SEED, (size_t)nqp * a.P * k * 8))) return rc;
      MmaArgs s0 = a;
      s0.partial = (uint64_t*)idx->d_ws[WS_SEED];
      s0.lists = a.P;
      s0.tile0 = 0;
      s0.num_tiles = seed_tiles;
      s0.thr_init = nullptr;
      CK(launch_scan(s0), "scan_mma seed launch");
      CK(merge_partials_seed_launch(s0.partial, a.P, nqp, k, partial + (size_t)a.P * k, (size_t)(a.P + 1) * k, thr_seed,
                                    stream),
         "merge_partials seed launch");
      if (launches) *launches += 2;
    }
    a.partial = partial;
    a.lists = a.P + (seed_tiles ? 1u : 0u);
    a.tile0 = seed_tiles;
    a.num_tiles = tiles_all;
    a.thr_init = seed_tiles ? thr_seed : nullptr;
    CK(launch_scan(a), "scan_mma launch");
    if (launches) *launches += 2;
    // merge this pass right away (the partial buffer is reused by the next pass)
    if (pev && q0 + max_q >= nq) CK(cudaEventRecord(pev[1], stream), "profile event record");
    CK(merge_partials_launch(a.partial, a.lists, nqp, k, row_offset, d_out_ids + (size_t)q0 * k,
                             d_out_scores + (size_t)q0 * k, stream),
       "merge_partials launch");
    if (launches) ++*launches;
  }
  if (pev) CK(cudaEventRecord(pev[2], stream), "profile event record");
  return NVDB_B200_OK;
}

// `xchg` (row-sharded index): when the call is ONE pass with the fused merge, the scan's last CTA also runs the exchange
// with the peer GPUs and d_out_* receive the merge over all shards (*tail_used = true); otherwise the call behaves as
// without it and the caller runs the exchange kernel on the local result.
int gemv_search_impl(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                     float* d_out_scores, cudaStream_t stream, uint32_t* launches, nvdb_b200_exchange* xchg = nullptr,
                     bool* tail_used = nullptr) {
  ScanPlan plan;
  int rc = plan_scan(idx, nq, k, &plan);
  if (rc) return rc;
  if ((rc = ensure_ws(idx, WS_PARTIAL, (size_t)nq * plan.grid * k * 8))) return rc;
  uint64_t* partial = (uint64_t*)idx->d_ws[WS_PARTIAL];

  ScanLaunch L;
  L.dtype = (int)idx->dtype;
  L.use_tma = idx->tma_ok;
  L.tmap = &idx->tmap;
  L.grid = plan.grid;
  L.stream = stream;
  ScanArgs a;
  a.scales = idx->scales;
  a.base = idx->payload;
  a.n_rows = (uint32_t)idx->n;
  a.dim = idx->dim;
  a.k = k;
  a.cap = plan.cap;
  a.num_tiles = (uint32_t)((idx->n + kScanRowsPerTile - 1) / kScanRowsPerTile);
  a.kchunks = (uint32_t)(((size_t)idx->dim * elem_size(idx->dtype) + 127) / 128);
  a.stages = plan.stages;
  // fused final merge in the scan's last CTA when the survivors' buffer fits the TMA ring (tuning knob FUSED_MERGE)
  // (auto: up to 4096 keys per query — beyond that one CTA sorting alone is slower than the 512-thread merge kernel:
  // k = 100, 148 lists: 0.95 vs 0.86 ms, profiles/r1_ab_fused_merge.jsonl; 2 = whenever the buffer fits)
  const size_t merge_keys = (size_t)next_pow2((uint32_t)plan.grid * k);
  // (round 2: with the rank-counting merge the last CTA keeps only ~k of the keys, so it also takes k = 100 - 14,800 keys
  // - whenever the keys and the lists' j-th entries fit the ring)
  const bool fused = idx->tma_ok && idx->tune_fused_merge && merge_keys * 8u <= (size_t)plan.stages * 32768u &&
                     (idx->tune_fused_merge == 2 || merge_keys <= 4096u ||
                      ((size_t)plan.grid * k + (size_t)plan.grid) * 8u <= (size_t)plan.stages * 32768u);
  a.done_counter = nullptr;
  a.out_ids = nullptr;
  a.out_scores = nullptr;
  a.row_offset = idx->row_offset;
  a.xchg = ScanXchg{};
  if (tail_used) *tail_used = false;
  if (fused) {
    if ((rc = ensure_gemv_ctr(idx, stream))) return rc;
    a.done_counter = (uint32_t*)idx->d_ws[WS_GEMV_CTR];
  }
  cudaEvent_t* pev = nullptr;
  if ((rc = prof_begin(idx, stream, &pev))) return rc;
  // The epoch of a row-sharded search starts only now, after everything that can fail without a launch: a rank that
  // bumped its epoch and then returned an error would be one epoch ahead of its peers for good.
  const bool tail = xchg && fused && nq <= (uint32_t)plan.nq_template &&
                    (size_t)xchg->world * k * 12u + 8u <= (size_t)plan.stages * 32768u;
  if (tail) {
    exchange_begin_epoch(xchg, &a.xchg);
    if (tail_used) *tail_used = true;
  }
  for (uint32_t q0 = 0; q0 < nq; q0 += (uint32_t)plan.nq_template) {
    const uint32_t left = nq - q0;
    // shrink the template for the remainder so a tail of 1-2 queries does not pay for 8
    int t = plan.nq_template;
    while (t > 1 && (uint32_t)(t / 2) >= left) t >>= 1;
    a.nq = left < (uint32_t)t ? left : (uint32_t)t;
    a.queries = d_queries + (size_t)q0 * idx->dim;
    a.partial = partial + (size_t)q0 * plan.grid * k;
    if (fused) {
      a.out_ids = d_out_ids + (size_t)q0 * k;
      a.out_scores = d_out_scores + (size_t)q0 * k;
    }
    {
      const cudaError_t le = scan_gemv_launch(L, a, t);
      if (le != cudaSuccess) {
        if (tail) xchg->epoch -= 1u;  // nothing was launched for this epoch (a tail call is a single pass)
        return fail_cuda(idx, le, "scan_gemv launch");
      }
    }
    count_launch();
    if (launches) ++*launches;
  }
  if (pev) CK(cudaEventRecord(pev[1], stream), "profile event record");
  if (!fused) {
    CK(merge_partials_launch(partial, (uint32_t)plan.grid, nq, k, idx->row_offset, d_out_ids, d_out_scores, stream),
       "merge_partials launch");
    if (launches) ++*launches;
  }
  if (pev) CK(cudaEventRecord(pev[2], stream), "profile event record");
  return NVDB_B200_OK;
}

// ---- exact tensor-core path: candidate scan (k' > k) -> exact re-score + certificate -> streaming-scan fallback ----
// f16: ONE fp16 query limb (half the tensor work of PATH_MMA); i8: two int8 limbs. Results are bit-identical to
// PATH_GEMV (and therefore to the reference) for every query: certified ones by construction (rescore.cu), the rest
// because they are re-run on the streaming scan. Synchronises `stream` once (the flag count comes back to the host).
uint32_t exact_kprime(uint32_t k) {
  if (k <= 12u) return 16u;  // per-thread register lists of the scan hold 16 entries
  const uint32_t extra = k / 4u > 8u ? k / 4u : 8u;  // k = 100 -> 125: ~25 ranks of slack against the error margin
  return k + extra;
}

int ensure_norms(nvdb_b200_index* idx, cudaStream_t stream) {
  if (idx->norms_ready) return NVDB_B200_OK;
  int rc;
  if ((rc = ensure_ws(idx, WS_X_FLAGS, 4096))) return rc;
  if ((rc = ensure_pin(idx, 2, 4096))) return rc;
  uint32_t* d2 = (uint32_t*)idx->d_ws[WS_X_FLAGS];
  CK(row_norm_max_launch((int)idx->dtype, idx->payload, idx->scales, idx->n, idx->dim, d2, idx->num_sms, stream),
     "row_norm_max launch");
  CK(cudaMemcpyAsync(idx->h_pin[2], d2, 8, cudaMemcpyDeviceToHost, stream), "D2H norms");
  CK(cudaStreamSynchronize(stream), "norms sync");
  std::memcpy(&idx->max_l2, idx->h_pin[2], 4);
  std::memcpy(&idx->max_l1, (char*)idx->h_pin[2] + 4, 4);
  idx->norms_ready = true;
  return NVDB_B200_OK;
}

bool mma_exact_supported(const nvdb_b200_index* idx, uint32_t k) {
  uint32_t bn;
  int st;
  return idx->mma_ok && scan_mma_plan((int)idx->dtype, idx->dtype == kDtF16 ? 1 : 2, exact_kprime(k), idx->max_smem, &bn, &st);
}

int mma_exact_finish(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                     float* d_out_scores, cudaStream_t stream, uint32_t* launches, uint32_t* fallbacks);

// Small batches: candidate scan on the row-tile kernel (scan_rows.cu: base rows on M, the whole batch resident as the B
// operand) — one HBM pass, tensor work proportional to the batch. false = the batch / k' does not fit, use scan_mma*.
bool rows_kernel_fits(const nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t kp, RowsPlan* plan) {
  if (!idx->mma_ok || !idx->tune_rows) return false;
  // measured (profiles/r2_probe_rows_ab_v2.jsonl, interleaved on one box, 4-stage ring): int8 0.36-0.43 ms against
  // 0.51-0.59 ms of the query-tile kernels for 1..32 queries (0.52-0.65 against 0.59-0.67 at 48..128); f16 0.66-0.68
  // against 0.69-0.71 up to 16 queries, 0.76 / 0.78 at 32, and the other way round from 48 on (0.80 against 0.76: the
  // one-limb query tile is cheap to re-stream and 256-row tiles need half the MMA issues)
  if (idx->tune_rows == 1 && idx->dtype != kDtI8 && nq > 32u) return false;
  if ((reinterpret_cast<uintptr_t>(d_queries) & 15u) || (idx->dim % 4u)) return false;  // float4 reads of the queries
  return scan_rows_plan((int)idx->dtype, nq, kp, idx->dim, idx->max_smem, plan);
}

// cand_ids == null: no merge launch — the caller's re-score kernel selects the candidates from the per-CTA lists itself
// (*out_partial / *out_lists describe them).
int rows_candidate_scan(nvdb_b200_index* idx, const RowsPlan& plan, const float* d_queries, uint32_t nq, uint32_t kp,
                        uint64_t* cand_ids, float* cand_apx, cudaStream_t stream, uint32_t* launches,
                        const uint64_t** out_partial = nullptr, uint32_t* out_lists = nullptr) {
  const uint32_t tiles = (uint32_t)((idx->n + 127u) / 128u);
  const int grid = (int)(tiles < (uint32_t)idx->num_sms ? tiles : (uint32_t)idx->num_sms);
  int rc;
  if ((rc = ensure_ws(idx, WS_PARTIAL, (size_t)nq * grid * kp * 8))) return rc;
  cudaEvent_t* pev = nullptr;
  if ((rc = prof_begin(idx, stream, &pev))) return rc;
  RowsArgs a{};
  a.queries = d_queries;
  a.scales = idx->scales;
  a.thr_init = nullptr;
  a.partial = (uint64_t*)idx->d_ws[WS_PARTIAL];
  // cross-CTA threshold seeding workspace: arrival counter (epoch = launch number, so no memset per launch) + one word
  // per (CTA, query)
  // (pays from ~40 queries on: below that its publish / poll / select steps cost more than the appends they save)
  const bool seed = idx->tune_rows_seed == 2 || (idx->tune_rows_seed == 1 && nq >= 40u);
  if (seed && (uint32_t)grid >= kp && grid <= 256) {  // the kernel's own condition: every CTA publishes or none
    // 4 rounds of [grid][bp <= 128] words after the 64-byte header of counters; one size, so it is never re-allocated
    if ((rc = ensure_ws(idx, WS_ROWS_SEED, 64u + 4u * 256u * 128u * 4u))) return rc;
    if (idx->rows_seed_seen != idx->d_ws[WS_ROWS_SEED]) {
      CK(cudaMemsetAsync(idx->d_ws[WS_ROWS_SEED], 0, 64, stream), "memset seeding counter");
      idx->rows_seed_seen = idx->d_ws[WS_ROWS_SEED];
      idx->rows_launches = 0;
    }
    a.seed_ws = (uint32_t*)idx->d_ws[WS_ROWS_SEED];
    a.seed_target = ++idx->rows_launches * (uint32_t)grid;
    // one round (after the first tile): a second and third one (tiles 8, 64) cut the appends further but measured no
    // faster (64 int8 queries 0.461 / 0.465 ms with 1 / 3 rounds, 0.482 without seeding: profiles/r2_rows_startup.jsonl)
    a.seed_rounds = 1;
    if (const char* e = getenv("NVDB_B200_ROWS_ROUNDS")) a.seed_rounds = (uint32_t)atoi(e) > 3u ? 3u : (uint32_t)atoi(e) < 1u ? 1u : (uint32_t)atoi(e);
  }
  a.n_rows = (uint32_t)idx->n;
  a.dim = idx->dim;
  a.kp = kp;
  a.nq = nq;
  a.num_tiles = tiles;
  a.kchunks = (uint32_t)(((size_t)idx->dim * elem_size(idx->dtype) + 127) / 128);
  a.lists = (uint32_t)grid;
  a.bp = plan.bp;
  a.nb = plan.nb;
  a.cap = plan.cap;
  a.stages = plan.stages;
  CK(scan_rows_launch((int)idx->dtype, &idx->tmap_mma_b, a, grid, stream), "scan_rows launch");
  if (pev) CK(cudaEventRecord(pev[1], stream), "profile event record");
  if (launches) ++*launches;
  if (cand_ids) {
    CK(merge_partials_launch(a.partial, a.lists, nq, kp, 0, cand_ids, cand_apx, stream), "merge_partials launch");
    if (launches) ++*launches;
  }
  if (pev) CK(cudaEventRecord(pev[2], stream), "profile event record");
  if (out_partial) *out_partial = a.partial;
  if (out_lists) *out_lists = a.lists;
  return NVDB_B200_OK;
}

// Candidate scan with k' = kp entries per query + exact re-score + certificate, enqueued on `stream`; the number of
// flagged queries lands in the pinned word h_pin[2][0] (flags per query stay on the device, WS_X_FLAGS + 4).
int mma_exact_enqueue(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint32_t kp,
                      uint64_t* d_out_ids, float* d_out_scores, cudaStream_t stream, uint32_t* launches) {
  int rc;
  if ((rc = ensure_ws(idx, WS_X_IDS, (size_t)nq * kp * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_X_APX, (size_t)nq * kp * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_FLAGS, ((size_t)nq + 4) * 4))) return rc;
  uint64_t* cand_ids = (uint64_t*)idx->d_ws[WS_X_IDS];
  float* cand_apx = (float*)idx->d_ws[WS_X_APX];
  // flag counter: two slots used alternately; the re-score kernel of call i zeroes the slot of call i + 1 (stream order),
  // so no memset is launched per search. Slots are zeroed here whenever the workspace is new.
  uint32_t* ws_flags = (uint32_t*)idx->d_ws[WS_X_FLAGS];
  if (idx->flag_ws_seen != ws_flags) {
    CK(cudaMemsetAsync(ws_flags, 0, 16, stream), "memset flag counters");
    idx->flag_ws_seen = ws_flags;
    idx->exact_calls = 0;
  }
  const uint32_t slot = (uint32_t)(idx->exact_calls++ & 1u);
  uint32_t* nflag = ws_flags + slot;
  uint32_t* flags = ws_flags + 4;
  const int limbs = idx->dtype == kDtF16 ? 1 : 2;
  RowsPlan rplan;
  const uint64_t* fused_partial = nullptr;
  uint32_t fused_lists = 0;
  if (rows_kernel_fits(idx, d_queries, nq, kp, &rplan)) {
    // few enough keys per query (<= 4096): the re-score kernel selects the candidates from the per-CTA lists itself
    const uint32_t tiles = (uint32_t)((idx->n + 127u) / 128u);
    const uint32_t lists = tiles < (uint32_t)idx->num_sms ? tiles : (uint32_t)idx->num_sms;
    const bool fuse = (size_t)lists * kp <= 4096u;
    if ((rc = rows_candidate_scan(idx, rplan, d_queries, nq, kp, fuse ? nullptr : cand_ids, cand_apx, stream, launches,
                                  &fused_partial, &fused_lists)))
      return rc;
    if (!fuse) fused_partial = nullptr;
  } else if ((rc = mma_search_impl(idx, d_queries, nq, kp, cand_ids, cand_apx, stream, launches, limbs, nullptr, 0))) {
    return rc;
  }

  RescoreArgs r{};
  r.partial = fused_partial;
  r.nlists = fused_lists;
  r.nflag_next = ws_flags + (slot ^ 1u);
  r.base = idx->payload;
  r.scales = idx->scales;
  r.queries = d_queries;
  r.cand_ids = cand_ids;
  r.cand_approx = cand_apx;
  r.out_ids = d_out_ids;
  r.out_scores = d_out_scores;
  r.flags = flags;
  r.nflag = nflag;
  r.row_offset = idx->row_offset;
  r.n_rows = (uint32_t)idx->n;
  r.dim = idx->dim;
  r.kp = kp;
  r.k = k;
  r.max_l2 = idx->max_l2;
  r.max_l1 = idx->max_l1;
  const double D = (double)idx->dim, u24 = 1.0 / 16777216.0;
  r.rel_err = idx->dtype == kDtF16
                  ? (float)(D * 2.0 * u24 + (D / 8.0 + 4.0) * u24)  // + |delta|_2 of the limb split, added per query
                  : (float)((D / 8.0 + 16.0) * u24);
  // the verdict (number of flagged queries) is stored into the pinned word by the kernel's last CTA
  r.done_ctr = ws_flags + 3;
  r.host_nflag = (uint32_t*)idx->h_pin[2];
  CK(rescore_exact_launch((int)idx->dtype, r, nq, stream), "rescore launch");
  if (launches) ++*launches;
  return NVDB_B200_OK;
}

// Second tier for queries whose first certificate failed: the same one-limb candidate scan with much wider lists
// (k' = 64 for k <= 12, else the widest list the scan's shared memory holds, about 1.5 k). The certificate fails when
// fewer than k' - k rows separate the k-th exact score from (k'-th approximate score + margin); bases whose rows cluster
// (hundreds of near-equal scores around the k-th) do that to 13 % (k = 10) - 61 % (k = 100) of the queries of an FP16
// 2.9 M-row base with 256 clusters, and the streaming scan at 8 queries per pass then costs 10 - 30x the tensor-core
// scan (profiles/r1_clustered_2p9M_before_tier2.jsonl). 0 = no wider list available.
uint32_t exact_kprime2(const nvdb_b200_index* idx, uint32_t k) {
  if (!idx->tune_exact_tier2) return 0u;
  const uint32_t kp = exact_kprime(k);
  const uint32_t want[4] = {idx->tune_exact_tier2 > 1 ? (uint32_t)idx->tune_exact_tier2 : (k <= 12u ? 64u : 4u * k), 2u * k,
                            k + k / 2u + 8u, k + k / 3u + 8u};
  const int limbs = idx->dtype == kDtF16 ? 1 : 2;
  for (uint32_t w : want) {
    uint32_t bn;
    int st;
    if (w > kp && scan_mma_plan((int)idx->dtype, limbs, w, idx->max_smem, &bn, &st) &&
        ((size_t)idx->dim * 4u + 15u) + (size_t)next_pow2(w) * 8u <= 200u * 1024u)
      return w;
  }
  return 0u;
}

// Every workspace the exact path can touch for a batch of nq queries, sized for the worst case (all nq queries flagged by
// the first certificate, all of them again by the second) BEFORE anything is launched. Which queries fail a certificate
// is data dependent, and a workspace that grows in the middle of a row-sharded search (cudaFree / cudaMalloc wait for
// the whole device) dead-locks ranks that share a GPU behind a spinning peer kernel; sized like this, one
// plain search of a (nq, k) shape grows everything any later search of that shape can need.
int reserve_exact_ws(nvdb_b200_index* idx, uint32_t nq, uint32_t k, cudaStream_t stream) {
  const uint32_t kp = exact_kprime(k), kp2 = exact_kprime2(idx, k), kw = kp2 > kp ? kp2 : kp;
  const size_t qb = (size_t)idx->dim * 4, es = elem_size(idx->dtype);
  const int limbs = idx->dtype == kDtF16 ? 1 : 2;
  int rc;
  if ((rc = ensure_ws(idx, WS_X_IDS, (size_t)nq * kw * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_X_APX, (size_t)nq * kw * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_FLAGS, ((size_t)nq + 4) * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_LIST, 2 * (size_t)nq * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_Q, (size_t)nq * qb))) return rc;
  if ((rc = ensure_ws(idx, WS_X_OUT_IDS, (size_t)nq * k * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_X_OUT_SC, (size_t)nq * k * 4))) return rc;
  if (kp2) {
    if ((rc = ensure_ws(idx, WS_X_Q2, (size_t)nq * qb))) return rc;
    if ((rc = ensure_ws(idx, WS_X_OUT2, (size_t)nq * k * 12))) return rc;
  }
  // per-CTA candidate lists: a tensor-core pass of m queries writes m * (P + 1) lists with P <= num_sms / ceil(m / 128),
  // i.e. at most 128 * num_sms + m lists (m * (num_sms + 1) for a single query tile); the streaming-scan fallback writes
  // nq * grid lists of k entries
  const size_t sms = (size_t)idx->num_sms;
  const size_t lists_mma = nq <= 128u * sms ? std::min((size_t)nq * (sms + 1), 128u * sms + nq + 128u) : (size_t)nq * (sms + 1);
  ScanPlan plan;
  if ((rc = plan_scan(idx, nq, k, &plan))) return rc;
  const size_t part = std::max(lists_mma * kw * 8, (size_t)nq * plan.grid * k * 8);
  if ((rc = ensure_ws(idx, WS_PARTIAL, part))) return rc;
  if ((rc = ensure_ws(idx, WS_SEED, lists_mma * kw * 8))) return rc;
  const size_t mpad = (((size_t)(nq < 128u * sms ? nq : 128u * sms) + 127u) / 128u + 1u) * 128u;  // + 1: CTA-pair padding tile
  if ((rc = ensure_ws(idx, WS_MMA_A, (size_t)limbs * mpad * idx->dim * es))) return rc;
  if ((rc = ensure_ws(idx, WS_MMA_QS, 3 * mpad * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_ROWS_SEED, 64u + (size_t)idx->num_sms * 2u * 128u * 4u))) return rc;
  return ensure_gemv_ctr(idx, stream);
}

// `deferred` != null: only enqueue (scan, re-score, D2H of the flag count) and set *deferred; the caller synchronises the
// stream together with its own copies and then calls mma_exact_finish (the host-buffer API does this, so the common
// no-fallback case costs ONE stream sync per search). deferred == null: finish here (device-buffer API).
int mma_exact_search_impl(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                          float* d_out_scores, cudaStream_t stream, uint32_t* launches, bool* deferred) {
  if (!mma_exact_supported(idx, k))
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "exact tensor-core scan: unsupported base layout or k too large");
  int rc;
  if ((rc = ensure_norms(idx, stream))) return rc;
  // pinned words: [0] flag count, [4, 4 + nq) flags, then two index lists of up to nq entries (mma_exact_finish)
  if ((rc = ensure_pin(idx, 2, (3 * (size_t)nq + 8) * 4))) return rc;
  if ((rc = reserve_exact_ws(idx, nq, k, stream))) return rc;
  if (idx->exact_ewma_k != k) {
    idx->exact_ewma_k = k;
    idx->exact_fail_ewma = 0.f;
    idx->exact_batches = 0;
  }
  const uint32_t kp2 = exact_kprime2(idx, k);
  idx->exact_wide_first = kp2 != 0 && nq >= 64u && idx->exact_fail_ewma > 0.35f && (++idx->exact_batches & 7u) != 0u;
  if ((rc = mma_exact_enqueue(idx, d_queries, nq, k, idx->exact_wide_first ? kp2 : exact_kprime(k), d_out_ids, d_out_scores,
                              stream, launches)))
    return rc;
  if (deferred) {
    *deferred = true;
    return NVDB_B200_OK;
  }
  CK(cudaStreamSynchronize(stream), "exact scan sync");
  return mma_exact_finish(idx, d_queries, nq, k, d_out_ids, d_out_scores, stream, launches, nullptr);
}

// After the stream has been synchronised: read the flag count and re-run flagged queries — first on the tensor-core
// scan with wider lists (exact_kprime2), and what that cannot certify either on the streaming scan
// (asynchronous on `stream`; *fallbacks tells the caller whether the outputs were rewritten).
int mma_exact_finish(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                     float* d_out_scores, cudaStream_t stream, uint32_t* launches, uint32_t* fallbacks) {
  int rc;
  uint32_t* h = (uint32_t*)idx->h_pin[2];
  const uint32_t nf = h[0];
  idx->exact_queries += nq;
  idx->exact_fallbacks += nf;
  if (fallbacks) *fallbacks = nf;
  if (!idx->exact_wide_first && nq >= 64u) idx->exact_fail_ewma = 0.5f * idx->exact_fail_ewma + 0.5f * (float)nf / (float)nq;
  if (nf == 0) return NVDB_B200_OK;

  // the wide lists were scanned first (adaptive first tier): what they could not certify goes to the streaming scan
  const uint32_t kp2 = idx->exact_wide_first ? 0u : exact_kprime2(idx, k);
  if (nf == nq && !kp2) {
    idx->exact_gemv += nf;
    return gemv_search_impl(idx, d_queries, nq, k, d_out_ids, d_out_scores, stream, launches);
  }
  // flags -> host -> ascending list of flagged queries (pinned) -> device; returns the list length
  auto fetch_list = [&](uint32_t n, uint32_t* h_list, uint32_t* d_list, uint32_t* count) -> int {
    uint32_t* flags = (uint32_t*)idx->d_ws[WS_X_FLAGS] + 4;
    CK(cudaMemcpyAsync(h + 4, flags, (size_t)n * 4, cudaMemcpyDeviceToHost, stream), "D2H flags");
    CK(cudaStreamSynchronize(stream), "exact scan sync");
    uint32_t m = 0;
    for (uint32_t q = 0; q < n; ++q)
      if (h[4 + q]) h_list[m++] = q;
    if (m) CK(cudaMemcpyAsync(d_list, h_list, (size_t)m * 4, cudaMemcpyHostToDevice, stream), "H2D flagged list");
    *count = m;
    return NVDB_B200_OK;
  };
  const size_t qb = (size_t)idx->dim * 4;
  if ((rc = ensure_ws(idx, WS_X_LIST, 2 * (size_t)nq * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_Q, (size_t)nf * qb))) return rc;
  if ((rc = ensure_ws(idx, WS_X_OUT_IDS, (size_t)nf * k * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_X_OUT_SC, (size_t)nf * k * 4))) return rc;
  uint32_t* d_list1 = (uint32_t*)idx->d_ws[WS_X_LIST];
  uint32_t* d_list2 = d_list1 + nq;
  uint32_t* h_list1 = h + 4 + nq;
  uint32_t* h_list2 = h_list1 + nq;
  float* gq = (float*)idx->d_ws[WS_X_Q];
  uint64_t* fo_ids = (uint64_t*)idx->d_ws[WS_X_OUT_IDS];
  float* fo_sc = (float*)idx->d_ws[WS_X_OUT_SC];
  uint32_t n1 = 0;
  if ((rc = fetch_list(nq, h_list1, d_list1, &n1))) return rc;
  if (n1 != nf) return fail(idx, NVDB_B200_ERR_CUDA, "exact path: flag count and flags disagree");
  CK(gather_queries_launch(d_queries, d_list1, n1, idx->dim, gq, stream), "gather flagged queries");
  if (launches) ++*launches;

  if (!kp2) {  // no wider list: flagged queries go straight to the streaming scan
    idx->exact_gemv += n1;
    if ((rc = gemv_search_impl(idx, gq, n1, k, fo_ids, fo_sc, stream, launches))) return rc;
  } else {
    idx->exact_tier2 += n1;
    if ((rc = mma_exact_enqueue(idx, gq, n1, k, kp2, fo_ids, fo_sc, stream, launches))) return rc;
    CK(cudaStreamSynchronize(stream), "exact tier-2 sync");
    const uint32_t nf2 = h[0];
    if (nf2) {  // still not certified (exact duplicates, pathological ties): the bit-exact streaming scan settles them
      idx->exact_gemv += nf2;
      if ((rc = ensure_ws(idx, WS_X_Q2, (size_t)nf2 * qb))) return rc;
      if ((rc = ensure_ws(idx, WS_X_OUT2, (size_t)nf2 * k * 12))) return rc;
      float* gq2 = (float*)idx->d_ws[WS_X_Q2];
      uint64_t* f2_ids = (uint64_t*)idx->d_ws[WS_X_OUT2];
      float* f2_sc = (float*)((char*)idx->d_ws[WS_X_OUT2] + (size_t)nf2 * k * 8);
      uint32_t n2 = 0;
      if ((rc = fetch_list(n1, h_list2, d_list2, &n2))) return rc;
      if (n2 != nf2) return fail(idx, NVDB_B200_ERR_CUDA, "exact path: tier-2 flag count and flags disagree");
      CK(gather_queries_launch(gq, d_list2, n2, idx->dim, gq2, stream), "gather tier-2 queries");
      if ((rc = gemv_search_impl(idx, gq2, n2, k, f2_ids, f2_sc, stream, launches))) return rc;
      CK(scatter_topk_launch(f2_ids, f2_sc, d_list2, n2, k, fo_ids, fo_sc, stream), "scatter tier-2 results");
      if (launches) *launches += 2;
    }
  }
  CK(scatter_topk_launch(fo_ids, fo_sc, d_list1, n1, k, d_out_ids, d_out_scores, stream), "scatter fallback results");
  if (launches) ++*launches;
  return NVDB_B200_OK;
}

// AUTO: one or two queries take the bit-exact streaming scan; from kMmaCrossover queries on the exact tensor-core path
// wins (measured, profiles/r1_batch_sweep_*.jsonl: the streaming scan serves 2 queries per pass at the HBM roofline
// and is FMA-bound from 4; the tensor-core scan reads the base once per 128 * num_sms queries). Both are bit-identical
// to the reference, so AUTO never changes results.
constexpr uint32_t kMmaCrossover = 3;
// From how many queries AUTO leaves the streaming scan: int8 with the row-tile kernel already at 2 (0.38 ms against the
// streaming scan's 0.50 ms for two queries: 41 instructions per 16 bytes and query make it issue-bound), else 3.
// A host-buffer call (nvdb_b200_search: synchronises anyway) with an int8 base takes the row-tile kernel from ONE query
// on: device time is the same (0.357 against 0.363 ms), but the issue-bound streaming scan runs at the ramping clock of
// a GPU that idled between two calls (0.422 ms per call against 0.372, profiles/r2_bench_n1_v5.json).
static uint32_t auto_crossover(const nvdb_b200_index* idx, uint32_t k, bool exchange, bool host_call = false) {
  if (idx->dtype == kDtI8 && idx->tune_rows) return (host_call && !exchange && k <= 12u) ? 1u : 2u;
  // f16, two queries, small k: the one-limb query-tile kernel streams the base at 6.9 TB/s, the streaming scan with two
  // queries per pass at 6.2 (0.692 against 0.726 ms per step, profiles/r2_probe_rows_ab.jsonl). Row-sharded searches keep
  // the streaming scan there: its last CTA runs the exchange in the same launch.
  if (idx->dtype == kDtF16 && k <= 12u && !exchange) return 2u;
  return kMmaCrossover;
}

int search_device_impl(nvdb_b200_index* idx, const float* d_queries, uint32_t nq, uint32_t k, uint64_t* d_out_ids,
                       float* d_out_scores, cudaStream_t stream, uint32_t* launches, uint32_t* path_used,
                       bool* deferred = nullptr) {
  int path = idx->path;
  if (path == NVDB_B200_PATH_AUTO)
    path = (nq >= auto_crossover(idx, k, false, deferred != nullptr) && mma_exact_supported(idx, k)) ? NVDB_B200_PATH_MMA_EXACT
                                                                                                    : NVDB_B200_PATH_GEMV;
  if (path_used) *path_used = (uint32_t)path;
  if (path == NVDB_B200_PATH_GEMV) return gemv_search_impl(idx, d_queries, nq, k, d_out_ids, d_out_scores, stream, launches);
  if (path == NVDB_B200_PATH_MMA_EXACT)
    return mma_exact_search_impl(idx, d_queries, nq, k, d_out_ids, d_out_scores, stream, launches, deferred);
  return mma_search_impl(idx, d_queries, nq, k, d_out_ids, d_out_scores, stream, launches,
                         path == NVDB_B200_PATH_MMA_FAST ? 1 : 2, nullptr, idx->row_offset);
}

}  // namespace

extern "C" {

int nvdb_b200_search_device(nvdb_b200_index* idx, const float* queries_dev, uint32_t nq, uint32_t k,
                            uint64_t* out_ids_dev, float* out_scores_dev, void* stream) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!queries_dev) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  if (!out_ids_dev || !out_scores_dev) return fail(idx, NVDB_B200_ERR_ARG, "output buffer is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  cudaStream_t s = stream ? (cudaStream_t)stream : idx->stream;
  int rc = adopt_stream(idx, s);
  if (rc) return rc;
  return search_device_impl(idx, queries_dev, nq, k, out_ids_dev, out_scores_dev, s, nullptr, nullptr);
}

// scan + exchange + merge of one rank; idx->mu held by the caller
static int search_exchange_impl(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_dev, uint32_t nq,
                                uint32_t k, uint64_t* out_ids_dev, float* out_scores_dev, cudaStream_t s) {
  const size_t nk = (size_t)nq * k, words = nk + (nk + 1) / 2;
  if (words * 8u > x->slot_bytes) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange: packed result larger than the slot");
  if (!x->connected && x->world > 1) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange: exchange is not connected");
  if (idx->device != x->device) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange: index and exchange live on different devices");
  int path = idx->path;
  if (path == NVDB_B200_PATH_AUTO)
    path = (nq >= auto_crossover(idx, k, true) && mma_exact_supported(idx, k)) ? NVDB_B200_PATH_MMA_EXACT : NVDB_B200_PATH_GEMV;
  uint64_t* l_ids = reinterpret_cast<uint64_t*>(x->staging);
  float* l_scores = reinterpret_cast<float*>(l_ids + nk);
  int rc;
  if (path == NVDB_B200_PATH_GEMV) {
    // one launch when the call is a single pass: the scan's last CTA exchanges and merges (scan_gemv.cu)
    bool tail = false;
    // the tail writes the global result to out_*; without it the local result goes to the staging buffer
    ScanPlan plan;
    if ((rc = plan_scan(idx, nq, k, &plan))) return rc;
    const bool single = nq <= (uint32_t)plan.nq_template;
    rc = gemv_search_impl(idx, queries_dev, nq, k, single ? out_ids_dev : l_ids, single ? out_scores_dev : l_scores, s,
                          nullptr, single ? x : nullptr, &tail);
    if (rc) return rc;
    if (tail) return NVDB_B200_OK;
    if (single) {  // not fused after all (knob off, large k): the local result sits in out_*, stage it
      CK(cudaMemcpyAsync(l_ids, out_ids_dev, nk * 8u, cudaMemcpyDeviceToDevice, s), "stage ids");
      CK(cudaMemcpyAsync(l_scores, out_scores_dev, nk * 4u, cudaMemcpyDeviceToDevice, s), "stage scores");
    }
  } else {
    rc = search_device_impl(idx, queries_dev, nq, k, l_ids, l_scores, s, nullptr, nullptr);
    if (rc) return rc;
  }
  rc = nvdb_b200_exchange_merge(x, x->staging, nq, k, out_ids_dev, out_scores_dev, 3, s);
  if (rc) return fail(idx, rc, nvdb_b200_exchange_last_error(x));
  return NVDB_B200_OK;
}

int nvdb_b200_search_exchange_device(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_dev, uint32_t nq,
                                     uint32_t k, uint64_t* out_ids_dev, float* out_scores_dev, void* stream) {
  if (!x) return fail(idx, NVDB_B200_ERR_ARG, "exchange is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!out_ids_dev || !out_scores_dev) return fail(idx, NVDB_B200_ERR_ARG, "output buffer is NULL");
  if (!idx) {  // an empty shard contributes padding
    int rc = nvdb_b200_exchange_merge(x, nullptr, nq, k, out_ids_dev, out_scores_dev, 3, stream);
    return rc ? fail(nullptr, rc, nvdb_b200_exchange_last_error(x)) : NVDB_B200_OK;
  }
  if (!queries_dev) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  cudaStream_t s = stream ? (cudaStream_t)stream : idx->stream;
  int rc = adopt_stream(idx, s);
  if (rc) return rc;
  idx->in_exchange = true;
  rc = search_exchange_impl(idx, x, queries_dev, nq, k, out_ids_dev, out_scores_dev, s);
  idx->in_exchange = false;
  return rc;
}

// phases: 3 = the whole step (scan + push + wait + merge, results to the host); 1 = local scan + push only; 2 = wait +
// merge + results of the epoch a phase-1 call of this rank started.
static int search_exchange_host(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_host, uint32_t nq,
                                uint32_t k, uint64_t* out_ids, float* out_scores, nvdb_b200_timing* timing, int phases) {
  if (timing) *timing = nvdb_b200_timing{};
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (!x) return fail(idx, NVDB_B200_ERR_ARG, "exchange is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if ((phases & 1) && !queries_host) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  if ((out_ids == nullptr) != (out_scores == nullptr)) return fail(idx, NVDB_B200_ERR_ARG, "give both output buffers or none");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  const size_t qbytes = (size_t)nq * idx->dim * 4, ibytes = (size_t)nq * k * 8, sbytes = (size_t)nq * k * 4;
  int rc;
  if ((rc = ensure_ws(idx, WS_QUERIES, qbytes))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_IDS, ibytes + sbytes))) return rc;
  if ((rc = ensure_pin(idx, 0, qbytes))) return rc;
  if ((rc = ensure_pin(idx, 1, ibytes + sbytes))) return rc;
  cudaStream_t s = idx->stream;
  if ((rc = adopt_stream(idx, s))) return rc;
  uint64_t* d_ids = (uint64_t*)idx->d_ws[WS_OUT_IDS];
  float* d_scores = (float*)((char*)idx->d_ws[WS_OUT_IDS] + ibytes);
  const uint64_t l0 = g_launches.load(std::memory_order_relaxed);
  if (timing) CK(cudaEventRecord(idx->ev[0], s), "event");
  if (phases & 1) {
    std::memcpy(idx->h_pin[0], queries_host, qbytes);
    CK(cudaMemcpyAsync(idx->d_ws[WS_QUERIES], idx->h_pin[0], qbytes, cudaMemcpyHostToDevice, s), "H2D queries");
  }
  if (timing) CK(cudaEventRecord(idx->ev[1], s), "event");
  if (phases == 3) {
    idx->in_exchange = true;
    rc = search_exchange_impl(idx, x, (const float*)idx->d_ws[WS_QUERIES], nq, k, d_ids, d_scores, s);
    idx->in_exchange = false;
    if (rc) return rc;
  } else if (phases == 1) {  // local result into the exchange's staging buffer, then push without waiting
    const size_t nk = (size_t)nq * k, words = nk + (nk + 1) / 2;
    if (words * 8u > x->slot_bytes) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange: packed result larger than the slot");
    if (idx->device != x->device) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange: index and exchange live on different devices");
    uint64_t* l_ids = reinterpret_cast<uint64_t*>(x->staging);
    if ((rc = search_device_impl(idx, (const float*)idx->d_ws[WS_QUERIES], nq, k, l_ids, reinterpret_cast<float*>(l_ids + nk), s,
                                 nullptr, nullptr)))
      return rc;
    if ((rc = nvdb_b200_exchange_merge(x, x->staging, nq, k, nullptr, nullptr, 1, s))) return fail(idx, rc, nvdb_b200_exchange_last_error(x));
  } else {
    if ((rc = nvdb_b200_exchange_merge(x, nullptr, nq, k, d_ids, d_scores, 2, s))) return fail(idx, rc, nvdb_b200_exchange_last_error(x));
  }
  if (timing) CK(cudaEventRecord(idx->ev[2], s), "event");
  char* hp = (char*)idx->h_pin[1];
  if (out_ids && (phases & 2)) CK(cudaMemcpyAsync(hp, d_ids, ibytes + sbytes, cudaMemcpyDeviceToHost, s), "D2H results");
  if (timing) CK(cudaEventRecord(idx->ev[3], s), "event");
  CK(cudaStreamSynchronize(s), "search sync");
  if (phases & 2) {
    uint32_t st = 0;
    if ((rc = nvdb_b200_exchange_status(x, &st))) return fail(idx, rc, nvdb_b200_exchange_last_error(x));
    if (st != 0) return fail(idx, NVDB_B200_ERR_CUDA, nvdb_b200_exchange_last_error(x));
    if (out_ids) {
      std::memcpy(out_ids, hp, ibytes);
      std::memcpy(out_scores, hp + ibytes, sbytes);
    }
  }
  if (timing) {
    cudaEventElapsedTime(&timing->h2d_ms, idx->ev[0], idx->ev[1]);
    cudaEventElapsedTime(&timing->kernel_ms, idx->ev[1], idx->ev[2]);
    cudaEventElapsedTime(&timing->d2h_ms, idx->ev[2], idx->ev[3]);
    cudaEventElapsedTime(&timing->total_ms, idx->ev[0], idx->ev[3]);
    timing->kernel_launches = (uint32_t)(g_launches.load(std::memory_order_relaxed) - l0);  // process-wide: an upper bound
    timing->path = (uint32_t)idx->path;
  }
  return NVDB_B200_OK;
}

int nvdb_b200_search_exchange(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_host, uint32_t nq,
                              uint32_t k, uint64_t* out_ids, float* out_scores, nvdb_b200_timing* timing) {
  return search_exchange_host(idx, x, queries_host, nq, k, out_ids, out_scores, timing, 3);
}

int nvdb_b200_search_exchange_split(nvdb_b200_index* idx, nvdb_b200_exchange* x, const float* queries_host, uint32_t nq,
                                    uint32_t k, uint64_t* out_ids, float* out_scores, int phase) {
  if (phase != 1 && phase != 2) return fail(idx, NVDB_B200_ERR_ARG, "search_exchange_split: phase must be 1 (scan + push) or 2 (merge)");
  return search_exchange_host(idx, x, queries_host, nq, k, out_ids, out_scores, nullptr, phase);
}

int nvdb_b200_search(nvdb_b200_index* idx, const float* queries_host, uint32_t nq, uint32_t k, uint64_t* out_ids,
                     float* out_scores, nvdb_b200_timing* timing) {
  if (timing) *timing = nvdb_b200_timing{};
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!queries_host) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  if (!out_ids || !out_scores) return fail(idx, NVDB_B200_ERR_ARG, "output buffer is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  const size_t qbytes = (size_t)nq * idx->dim * 4, ibytes = (size_t)nq * k * 8, sbytes = (size_t)nq * k * 4;
  int rc;
  if ((rc = ensure_ws(idx, WS_QUERIES, qbytes))) return rc;
  // ids and scores share one device buffer and one pinned buffer: a single D2H copy per call; the four timing events
  // are recorded only when the caller asked for timings (each costs a few microseconds of a 0.7 ms batch-1 call)
  if ((rc = ensure_ws(idx, WS_OUT_IDS, ibytes + sbytes))) return rc;
  if ((rc = ensure_pin(idx, 0, qbytes))) return rc;
  if ((rc = ensure_pin(idx, 1, ibytes + sbytes))) return rc;
  cudaStream_t s = idx->stream;
  if ((rc = adopt_stream(idx, s))) return rc;
  // Small results (a handful of queries) are written by the kernels straight into the pinned host buffer, which the
  // device addresses directly (unified addressing): no device-to-host copy is enqueued, the stream synchronisation
  // below makes the stores visible. Larger results go through device memory and one copy.
  char* hp = (char*)idx->h_pin[1];
  const bool direct_out = ibytes + sbytes <= 16384u;  // up to 128 queries x top-10
  uint64_t* d_ids = direct_out ? (uint64_t*)hp : (uint64_t*)idx->d_ws[WS_OUT_IDS];
  float* d_scores = direct_out ? (float*)(hp + ibytes) : (float*)((char*)idx->d_ws[WS_OUT_IDS] + ibytes);
  par_memcpy(idx->h_pin[0], queries_host, qbytes);
  if (timing) CK(cudaEventRecord(idx->ev[0], s), "event");
  CK(cudaMemcpyAsync(idx->d_ws[WS_QUERIES], idx->h_pin[0], qbytes, cudaMemcpyHostToDevice, s), "H2D queries");
  if (timing) CK(cudaEventRecord(idx->ev[1], s), "event");
  uint32_t launches = 0, path_used = 0;
  bool deferred = false;
  rc = search_device_impl(idx, (const float*)idx->d_ws[WS_QUERIES], nq, k, d_ids, d_scores, s, &launches, &path_used,
                          &deferred);
  if (rc) return rc;
  if (timing) CK(cudaEventRecord(idx->ev[2], s), "event");
  if (!direct_out) CK(cudaMemcpyAsync(hp, d_ids, ibytes + sbytes, cudaMemcpyDeviceToHost, s), "D2H results");
  if (timing) CK(cudaEventRecord(idx->ev[3], s), "event");
  CK(cudaStreamSynchronize(s), "search sync");
  if (deferred) {  // exact tensor-core path: the certificate verdict arrived with the results
    uint32_t nfall = 0;
    rc = mma_exact_finish(idx, (const float*)idx->d_ws[WS_QUERIES], nq, k, d_ids, d_scores, s, &launches, &nfall);
    if (rc) return rc;
    if (nfall) {  // flagged queries were re-run on the streaming scan: fetch the corrected results
      if (timing) CK(cudaEventRecord(idx->ev[2], s), "event");
      if (!direct_out) CK(cudaMemcpyAsync(hp, d_ids, ibytes + sbytes, cudaMemcpyDeviceToHost, s), "D2H results");
      if (timing) CK(cudaEventRecord(idx->ev[3], s), "event");
      CK(cudaStreamSynchronize(s), "search sync");
    }
  }
  std::memcpy(out_ids, hp, ibytes);
  std::memcpy(out_scores, hp + ibytes, sbytes);
  if (timing) {
    cudaEventElapsedTime(&timing->h2d_ms, idx->ev[0], idx->ev[1]);
    cudaEventElapsedTime(&timing->kernel_ms, idx->ev[1], idx->ev[2]);
    cudaEventElapsedTime(&timing->d2h_ms, idx->ev[2], idx->ev[3]);
    cudaEventElapsedTime(&timing->total_ms, idx->ev[0], idx->ev[3]);
    timing->kernel_launches = launches;
    timing->path = path_used;
  }
  return NVDB_B200_OK;
}

int nvdb_b200_index_set_tuning(nvdb_b200_index* idx, int knob, int value) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  switch (knob) {
    case NVDB_B200_TUNE_MMA_SEED: idx->tune_mma_seed = value; break;
    case NVDB_B200_TUNE_GEMV_CTAS: idx->tune_gemv_ctas = value; break;
    case NVDB_B200_TUNE_MMA_2CTA: idx->tune_mma_2cta = value; break;
    case NVDB_B200_TUNE_FUSED_MERGE: idx->tune_fused_merge = value; break;
    case NVDB_B200_TUNE_MMA_RESIDENT: idx->tune_mma_resident = value; break;
    case NVDB_B200_TUNE_EXACT_TIER2: idx->tune_exact_tier2 = value; break;
    case NVDB_B200_TUNE_ROWS: idx->tune_rows = value; break;
    case NVDB_B200_TUNE_ROWS_SEED: idx->tune_rows_seed = value; break;
    default: return fail(idx, NVDB_B200_ERR_ARG, "unknown tuning knob");
  }
  return NVDB_B200_OK;
}

int nvdb_b200_index_exact_stats(nvdb_b200_index* idx, uint64_t* queries, uint64_t* fallbacks, int reset) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  if (queries) *queries = idx->exact_queries;
  if (fallbacks) *fallbacks = idx->exact_fallbacks;
  if (reset) idx->exact_queries = idx->exact_fallbacks = idx->exact_tier2 = idx->exact_gemv = 0;
  return NVDB_B200_OK;
}

int nvdb_b200_index_exact_tiers(nvdb_b200_index* idx, uint64_t* tier2, uint64_t* streamed) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  if (tier2) *tier2 = idx->exact_tier2;
  if (streamed) *streamed = idx->exact_gemv;
  return NVDB_B200_OK;
}

// Two-stage search: coarse scan over a compact copy of the base (typically INT8) keeps R
// candidates per query, which are re-ranked by their EXACT reference-order dot products over the fine copy (FP16 /
// FP32) without leaving the device. Approximate by construction (a true neighbour the coarse scan ranks below R is
// lost); scores of the returned ids are bit-identical to what the fine index's own scan would return for them.
int nvdb_b200_search_two_stage(nvdb_b200_index* coarse, nvdb_b200_index* fine, const float* queries_host, uint32_t nq,
                               uint32_t k, uint32_t R, uint64_t* out_ids, float* out_scores, nvdb_b200_timing* timing) {
  nvdb_b200_index* idx = coarse;
  if (timing) *timing = nvdb_b200_timing{};
  if (!coarse || !fine) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!queries_host) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  if (!out_ids || !out_scores) return fail(idx, NVDB_B200_ERR_ARG, "output buffer is NULL");
  if (coarse == fine || coarse->device != fine->device || coarse->n != fine->n || coarse->dim != fine->dim ||
      coarse->row_offset != fine->row_offset)
    return fail(idx, NVDB_B200_ERR_ARG, "two_stage: coarse and fine must be two copies of the same rows on one device");
  if (R < k) R = k;
  if (R > 4096u) return fail(idx, NVDB_B200_ERR_ARG, "two_stage: R <= 4096");
  std::scoped_lock g(coarse->mu, fine->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  const size_t qbytes = (size_t)nq * idx->dim * 4, ibytes = (size_t)nq * k * 8, sbytes = (size_t)nq * k * 4;
  int rc;
  if ((rc = ensure_ws(idx, WS_QUERIES, qbytes))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_IDS, ibytes))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_SCORES, sbytes))) return rc;
  if ((rc = ensure_ws(idx, WS_X_IDS, (size_t)nq * R * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_X_APX, (size_t)nq * R * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_X_FLAGS, ((size_t)nq + 4) * 4))) return rc;
  if ((rc = ensure_pin(idx, 0, qbytes))) return rc;
  if ((rc = ensure_pin(idx, 1, ibytes + sbytes))) return rc;
  cudaStream_t s = idx->stream;
  if ((rc = adopt_stream(idx, s))) return rc;
  if ((rc = adopt_stream(fine, s))) return rc;
  std::memcpy(idx->h_pin[0], queries_host, qbytes);
  CK(cudaEventRecord(idx->ev[0], s), "event");
  CK(cudaMemcpyAsync(idx->d_ws[WS_QUERIES], idx->h_pin[0], qbytes, cudaMemcpyHostToDevice, s), "H2D queries");
  CK(cudaEventRecord(idx->ev[1], s), "event");
  const float* dq = (const float*)idx->d_ws[WS_QUERIES];
  uint64_t* cand = (uint64_t*)idx->d_ws[WS_X_IDS];
  float* capx = (float*)idx->d_ws[WS_X_APX];
  uint32_t launches = 0;
  // stage 1: coarse top-R (tolerance-mode tensor scan for batches, streaming scan otherwise — no re-score needed here)
  uint32_t bn;
  int st;
  const bool tensor = coarse->mma_ok && nq >= 3 && scan_mma_plan((int)coarse->dtype, 2, R, coarse->max_smem, &bn, &st);
  rc = tensor ? mma_search_impl(coarse, dq, nq, R, cand, capx, s, &launches, coarse->dtype == kDtF16 ? 1 : 2, nullptr,
                                coarse->row_offset)
              : gemv_search_impl(coarse, dq, nq, R, cand, capx, s, &launches);
  if (rc) return rc;
  // stage 2: exact re-rank over the fine copy
  RescoreArgs r{};
  r.base = fine->payload;
  r.scales = fine->scales;
  r.queries = dq;
  r.cand_ids = cand;
  r.cand_approx = capx;
  r.out_ids = (uint64_t*)idx->d_ws[WS_OUT_IDS];
  r.out_scores = (float*)idx->d_ws[WS_OUT_SCORES];
  r.nflag = (uint32_t*)idx->d_ws[WS_X_FLAGS] + 2;  // scratch word (slots 0 / 1 belong to the exact path's certificate)
  r.flags = (uint32_t*)idx->d_ws[WS_X_FLAGS] + 4;
  r.row_offset = fine->row_offset;
  r.cand_offset = coarse->row_offset;
  r.n_rows = (uint32_t)fine->n;
  r.dim = fine->dim;
  r.kp = R;
  r.k = k;
  r.max_l2 = r.max_l1 = INFINITY;  // no certificate: every query is "flagged", the flags are ignored
  r.rel_err = 0.f;
  CK(rescore_exact_launch((int)fine->dtype, r, nq, s), "rescore launch");
  ++launches;
  CK(cudaEventRecord(idx->ev[2], s), "event");
  char* hp = (char*)idx->h_pin[1];
  CK(cudaMemcpyAsync(hp, idx->d_ws[WS_OUT_IDS], ibytes, cudaMemcpyDeviceToHost, s), "D2H ids");
  CK(cudaMemcpyAsync(hp + ibytes, idx->d_ws[WS_OUT_SCORES], sbytes, cudaMemcpyDeviceToHost, s), "D2H scores");
  CK(cudaEventRecord(idx->ev[3], s), "event");
  CK(cudaEventSynchronize(idx->ev[3]), "search sync");
  std::memcpy(out_ids, hp, ibytes);
  std::memcpy(out_scores, hp + ibytes, sbytes);
  if (timing) {
    cudaEventElapsedTime(&timing->h2d_ms, idx->ev[0], idx->ev[1]);
    cudaEventElapsedTime(&timing->kernel_ms, idx->ev[1], idx->ev[2]);
    cudaEventElapsedTime(&timing->d2h_ms, idx->ev[2], idx->ev[3]);
    cudaEventElapsedTime(&timing->total_ms, idx->ev[0], idx->ev[3]);
    timing->kernel_launches = launches;
    timing->path = tensor ? NVDB_B200_PATH_MMA : NVDB_B200_PATH_GEMV;
  }
  return NVDB_B200_OK;
}

int nvdb_b200_debug_mma_accumulators(nvdb_b200_index* idx, const float* queries_host, uint32_t nq, int path,
                                     uint32_t* out_acc) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (!queries_host || !out_acc || nq == 0) return fail(idx, NVDB_B200_ERR_ARG, "debug_mma: bad arguments");
  if (path != NVDB_B200_PATH_MMA && path != NVDB_B200_PATH_MMA_FAST) return fail(idx, NVDB_B200_ERR_ARG, "bad path");
  if (nq > 128u * (uint32_t)idx->num_sms) return fail(idx, NVDB_B200_ERR_ARG, "debug_mma: too many queries");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  const uint32_t nacc = idx->dtype == kDtI8 ? 2u : 1u, mpad = (nq + 127u) / 128u * 128u;
  const size_t qbytes = (size_t)nq * idx->dim * 4, dbytes = (size_t)nacc * mpad * idx->n * 4;
  int rc;
  if ((rc = ensure_ws(idx, WS_QUERIES, qbytes))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_IDS, (size_t)nq * 8))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_SCORES, (size_t)nq * 4))) return rc;
  if ((rc = ensure_ws(idx, WS_DBG, dbytes))) return rc;
  cudaStream_t s = idx->stream;
  if ((rc = adopt_stream(idx, s))) return rc;
  CK(cudaMemsetAsync(idx->d_ws[WS_DBG], 0, dbytes, s), "memset");
  CK(cudaMemcpyAsync(idx->d_ws[WS_QUERIES], queries_host, qbytes, cudaMemcpyHostToDevice, s), "H2D queries");
  rc = mma_search_impl(idx, (const float*)idx->d_ws[WS_QUERIES], nq, 1, (uint64_t*)idx->d_ws[WS_OUT_IDS],
                       (float*)idx->d_ws[WS_OUT_SCORES], s, nullptr, path == NVDB_B200_PATH_MMA_FAST ? 1 : 2,
                       idx->d_ws[WS_DBG], idx->row_offset);
  if (rc) return rc;
  CK(cudaStreamSynchronize(s), "debug_mma sync");
  for (uint32_t a = 0; a < nacc; ++a)
    CK(cudaMemcpy(out_acc + (size_t)a * nq * idx->n, (const uint32_t*)idx->d_ws[WS_DBG] + (size_t)a * mpad * idx->n,
                  (size_t)nq * idx->n * 4, cudaMemcpyDeviceToHost),
       "D2H accumulators");
  return NVDB_B200_OK;
}

int nvdb_b200_merge_topk_device(int device, const uint64_t* ids_dev, const float* scores_dev, uint32_t nlists,
                                uint32_t nq, uint32_t k, uint64_t* out_ids_dev, float* out_scores_dev,
                                void* stream) {
  const nvdb_b200_index* idx = nullptr;
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!ids_dev || !scores_dev || !out_ids_dev || !out_scores_dev || nlists == 0)
    return fail(nullptr, NVDB_B200_ERR_ARG, "merge_topk: bad arguments");
  CK(cudaSetDevice(device), "cudaSetDevice");
  CK(merge_lists_launch(ids_dev, scores_dev, (size_t)nq * k, (size_t)nq * k, nlists, nq, k, out_ids_dev, out_scores_dev,
                        (cudaStream_t)stream),
     "merge_lists launch");
  return NVDB_B200_OK;
}

int nvdb_b200_merge_topk_packed_device(int device, const void* packed_dev, uint32_t nlists, uint32_t nq, uint32_t k,
                                       uint64_t* out_ids_dev, float* out_scores_dev, void* stream) {
  const nvdb_b200_index* idx = nullptr;
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!packed_dev || !out_ids_dev || !out_scores_dev || nlists == 0)
    return fail(nullptr, NVDB_B200_ERR_ARG, "merge_topk_packed: bad arguments");
  CK(cudaSetDevice(device), "cudaSetDevice");
  const size_t nk = (size_t)nq * k, words = nk + (nk + 1) / 2;
  const uint64_t* ids = (const uint64_t*)packed_dev;
  const float* scores = (const float*)(ids + nk);
  CK(merge_lists_launch(ids, scores, words, 2 * words, nlists, nq, k, out_ids_dev, out_scores_dev, (cudaStream_t)stream),
     "merge_lists launch");
  return NVDB_B200_OK;
}

// =====================================================================================================
// refine
// =====================================================================================================
static int refine_check(const nvdb_b200_index* idx) {
  if ((idx->dim * elem_size(idx->dtype)) % 16)
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "refine needs a 16-byte row pitch (dim*elem_size % 16 == 0)");
  return NVDB_B200_OK;
}

int nvdb_b200_refine_l2_device(nvdb_b200_index* idx, const float* queries_dev, const uint32_t* cand_ids_dev,
                               uint32_t Q, uint32_t R, uint32_t K, uint32_t* out_ids_dev, float* out_dist_dev,
                               void* stream) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (Q == 0 || R == 0 || K == 0) return NVDB_B200_OK;  // src/cuda_refine.cu:853-857
  if (!queries_dev || !cand_ids_dev || !out_ids_dev) return fail(idx, NVDB_B200_ERR_ARG, "refine: NULL buffer");
  int rc = refine_check(idx);
  if (rc) return rc;
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  // stateless on the index (no workspace): any stream, no ordering against the index's searches needed
  CK(refine_l2_launch((int)idx->dtype, idx->payload, idx->scales, idx->n, idx->dim, queries_dev, cand_ids_dev, Q, R, K,
                      out_ids_dev, out_dist_dev, stream ? (cudaStream_t)stream : idx->stream),
     "refine launch");
  return NVDB_B200_OK;
}

// true when the driver can DMA from / to `p` directly (cudaMallocHost / cudaHostRegister memory)
static bool host_is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

int nvdb_b200_host_register(void* ptr, size_t bytes) {
  const nvdb_b200_index* idx = nullptr;
  if (!ptr || bytes == 0) return fail(nullptr, NVDB_B200_ERR_ARG, "host_register: NULL / empty buffer");
  cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
  if (e == cudaErrorHostMemoryAlreadyRegistered) {
    cudaGetLastError();
    return NVDB_B200_OK;
  }
  CK(e, "cudaHostRegister");
  return NVDB_B200_OK;
}

int nvdb_b200_host_unregister(void* ptr) {
  const nvdb_b200_index* idx = nullptr;
  if (!ptr) return NVDB_B200_OK;
  cudaError_t e = cudaHostUnregister(ptr);
  if (e == cudaErrorHostMemoryNotRegistered) {
    cudaGetLastError();
    return NVDB_B200_OK;
  }
  CK(e, "cudaHostUnregister");
  return NVDB_B200_OK;
}

int nvdb_b200_refine_l2(nvdb_b200_index* idx, const float* queries_host, const uint32_t* cand_ids_host, uint32_t Q,
                        uint32_t R, uint32_t K, uint32_t* out_ids, float* out_dist, nvdb_b200_timing* timing) {
  if (timing) *timing = nvdb_b200_timing{};
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (Q == 0 || R == 0 || K == 0) return NVDB_B200_OK;
  if (!queries_host || !cand_ids_host || !out_ids) return fail(idx, NVDB_B200_ERR_ARG, "refine: NULL buffer");
  int rc = refine_check(idx);
  if (rc) return rc;
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  const size_t qb = (size_t)Q * idx->dim * 4, cb = (size_t)Q * R * 4, ob = (size_t)Q * K * 4;
  // Registered / pinned caller buffers are used in place (no staging copy: 41 MB at the K5 shape, which was 3 of the
  // call's 3.7 ms); pageable ones go through the library's pinned staging buffers in chunks.
  const bool in_pinned = host_is_pinned(queries_host) && host_is_pinned(cand_ids_host);
  const bool out_pinned = host_is_pinned(out_ids) && (!out_dist || host_is_pinned(out_dist));
  if ((rc = ensure_ws(idx, WS_QUERIES, qb))) return rc;
  if ((rc = ensure_ws(idx, WS_CAND, cb))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_IDS32, ob))) return rc;
  if ((rc = ensure_ws(idx, WS_OUT_SCORES, ob))) return rc;
  if (!in_pinned && (rc = ensure_pin(idx, 0, qb + cb))) return rc;
  if (!out_pinned && (rc = ensure_pin(idx, 1, 2 * ob))) return rc;
  // in-kernel phase sampling (the reference's CUDA_DBG_TIMING / CUDA_DBG_Q switches; only with a timing struct)
  uint32_t dbg_q = 0;
  if (timing) {
    const char* on = getenv("CUDA_DBG_TIMING");
    if (on && atoi(on) != 0) {
      const char* dq = getenv("CUDA_DBG_Q");
      dbg_q = dq ? (uint32_t)atoi(dq) : 32u;
      if (dbg_q > Q) dbg_q = Q;
    }
  }
  unsigned long long* dbg_dev = nullptr;
  if (dbg_q) {
    if ((rc = ensure_ws(idx, WS_DBG, (size_t)dbg_q * 24))) return rc;
    if ((rc = ensure_pin(idx, 2, (size_t)dbg_q * 24 + 64))) return rc;
    dbg_dev = (unsigned long long*)idx->d_ws[WS_DBG];
  }
  cudaStream_t s = idx->stream;
  if ((rc = adopt_stream(idx, s))) return rc;
  if (dbg_dev) CK(cudaMemsetAsync(dbg_dev, 0, (size_t)dbg_q * 24, s), "memset dbg");
  const char* hin_q = in_pinned ? (const char*)queries_host : (const char*)idx->h_pin[0];
  const char* hin_c = in_pinned ? (const char*)cand_ids_host : (const char*)idx->h_pin[0] + qb;
  char* hout_i = out_pinned ? (char*)out_ids : (char*)idx->h_pin[1];
  char* hout_d = out_pinned ? (char*)out_dist : (char*)idx->h_pin[1] + ob;
  // Chunks of queries flow through (staging copy ->) H2D -> kernel -> D2H, so the host's copy of chunk c+1 into pinned
  // memory overlaps the device work on chunk c (the reference moves everything, then computes: src/cuda_refine.cu:
  // 926-1115). With `timing`, or with pinned caller buffers, the call is one chunk: the four events then bracket H2D /
  // kernel / D2H.
  const size_t per_q = (size_t)idx->dim * 4 + (size_t)R * 4;
  uint32_t qc = (timing || in_pinned) ? Q : (uint32_t)(((size_t)8 << 20) / per_q);
  if (qc == 0) qc = 1;
  if (qc > Q) qc = Q;
  if (timing && !in_pinned) {  // staged up front: ev[0] -> ev[1] is the H2D copy alone, as in the reference's timing
    par_memcpy(idx->h_pin[0], queries_host, qb);
    par_memcpy((char*)idx->h_pin[0] + qb, cand_ids_host, cb);
  }
  uint32_t smem_bytes = 0;
  CK(cudaEventRecord(idx->ev[0], s), "event");
  for (uint32_t q0 = 0; q0 < Q; q0 += qc) {
    const uint32_t nqc = Q - q0 < qc ? Q - q0 : qc;
    const size_t qoff = (size_t)q0 * idx->dim * 4, qlen = (size_t)nqc * idx->dim * 4;
    const size_t coff = (size_t)q0 * R * 4, clen = (size_t)nqc * R * 4;
    const size_t ooff = (size_t)q0 * K * 4, olen = (size_t)nqc * K * 4;
    if (!timing && !in_pinned) {
      par_memcpy((char*)idx->h_pin[0] + qoff, (const char*)queries_host + qoff, qlen);
      par_memcpy((char*)idx->h_pin[0] + qb + coff, (const char*)cand_ids_host + coff, clen);
    }
    CK(cudaMemcpyAsync((char*)idx->d_ws[WS_QUERIES] + qoff, hin_q + qoff, qlen, cudaMemcpyHostToDevice, s), "H2D queries");
    CK(cudaMemcpyAsync((char*)idx->d_ws[WS_CAND] + coff, hin_c + coff, clen, cudaMemcpyHostToDevice, s), "H2D candidates");
    if (timing) CK(cudaEventRecord(idx->ev[1], s), "event");
    CK(refine_l2_launch((int)idx->dtype, idx->payload, idx->scales, idx->n, idx->dim,
                        (const float*)((char*)idx->d_ws[WS_QUERIES] + qoff),
                        (const uint32_t*)((char*)idx->d_ws[WS_CAND] + coff), nqc, R, K,
                        (uint32_t*)((char*)idx->d_ws[WS_OUT_IDS32] + ooff), (float*)((char*)idx->d_ws[WS_OUT_SCORES] + ooff), s,
                        q0 == 0 ? dbg_dev : nullptr, dbg_q, &smem_bytes),
       "refine launch");
    if (timing) CK(cudaEventRecord(idx->ev[2], s), "event");
    CK(cudaMemcpyAsync(hout_i + ooff, (char*)idx->d_ws[WS_OUT_IDS32] + ooff, olen, cudaMemcpyDeviceToHost, s), "D2H ids");
    if (out_dist)
      CK(cudaMemcpyAsync(hout_d + ooff, (char*)idx->d_ws[WS_OUT_SCORES] + ooff, olen, cudaMemcpyDeviceToHost, s), "D2H dist");
  }
  CK(cudaEventRecord(idx->ev[3], s), "event");
  if (dbg_dev) CK(cudaMemcpyAsync(idx->h_pin[2], dbg_dev, (size_t)dbg_q * 24, cudaMemcpyDeviceToHost, s), "D2H dbg");
  CK(cudaStreamSynchronize(s), "refine sync");
  if (!out_pinned) {
    std::memcpy(out_ids, hout_i, ob);
    if (out_dist) std::memcpy(out_dist, hout_d, ob);
  }
  if (timing) {
    cudaEventElapsedTime(&timing->h2d_ms, idx->ev[0], idx->ev[1]);
    cudaEventElapsedTime(&timing->kernel_ms, idx->ev[1], idx->ev[2]);
    cudaEventElapsedTime(&timing->d2h_ms, idx->ev[2], idx->ev[3]);
    cudaEventElapsedTime(&timing->total_ms, idx->ev[0], idx->ev[3]);
    timing->kernel_launches = 1;
    timing->threads = (uint32_t)kRefineThreadsPerCta;
    timing->shmem_bytes = smem_bytes;
    if (dbg_q) {
      const unsigned long long* h = (const unsigned long long*)idx->h_pin[2];
      double w = 0, d = 0, m = 0;
      for (uint32_t i = 0; i < dbg_q; ++i) {
        w += (double)h[(size_t)i * 3 + 0];
        d += (double)h[(size_t)i * 3 + 1];
        m += (double)h[(size_t)i * 3 + 2];
      }
      timing->dbg_q = dbg_q;
      timing->dbg_write_cycles_avg = (float)(w / dbg_q);
      timing->dbg_dist_cycles_avg = (float)(d / dbg_q);
      timing->dbg_merge_cycles_avg = (float)(m / dbg_q);
    }
  }
  return NVDB_B200_OK;
}

}  // extern "C"

// =====================================================================================================
// per-kernel event timing for the harness
// =====================================================================================================
extern "C" {

int nvdb_b200_index_profile(nvdb_b200_index* idx, int enable) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  idx->profiling = enable != 0;
  idx->prof_used = 0;
  return NVDB_B200_OK;
}

int nvdb_b200_index_profile_read(nvdb_b200_index* idx, double* scan_ms_sum, double* merge_ms_sum, uint32_t* calls) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  double a = 0, b = 0;
  for (size_t i = 0; i + 2 < idx->prof_used + 0 && i + 2 < idx->prof_ev.size() + 0; i += 3) {
    CK(cudaEventSynchronize(idx->prof_ev[i + 2]), "profile sync");
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, idx->prof_ev[i], idx->prof_ev[i + 1]), "profile elapsed");
    a += ms;
    CK(cudaEventElapsedTime(&ms, idx->prof_ev[i + 1], idx->prof_ev[i + 2]), "profile elapsed");
    b += ms;
  }
  if (scan_ms_sum) *scan_ms_sum = a;
  if (merge_ms_sum) *merge_ms_sum = b;
  if (calls) *calls = (uint32_t)(idx->prof_used / 3);
  idx->prof_used = 0;
  return NVDB_B200_OK;
}

// =====================================================================================================
// allocation-free streaming scan (no counterpart in the reference; see include/nvdb_b200.h)
// =====================================================================================================
// Scratch of nvdb_b200_search_device_static: the per-CTA partial top-k lists [nq][grid][k] of the streaming scan, which
// the ordinary entry points keep in a grow-only cudaMalloc workspace of the index. Lives in every GPU that loads the
// library until the process exits. In host code `g_static_scratch` is only a stand-in for the symbol: the device address
// comes from cudaGetSymbolAddress (static_scratch_ptr), once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[NVDB_B200_STATIC_SCRATCH_BYTES];

static std::mutex g_static_mu;
static void* g_static_ptr[64] = {};

// device address of g_static_scratch on `device` (the caller has made it current); no allocation
static int static_scratch_ptr(const nvdb_b200_index* idx, int device, void** out) {
  if (device < 0 || device >= 64) return fail(idx, NVDB_B200_ERR_UNSUPPORTED, "search_device_static: device ordinal >= 64");
  std::lock_guard<std::mutex> g(g_static_mu);
  if (!g_static_ptr[device]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_static_scratch), "cudaGetSymbolAddress (static scratch)");
    if (!p) return fail(idx, NVDB_B200_ERR_CUDA, "cudaGetSymbolAddress returned NULL for the static scratch");
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return NVDB_B200_OK;
}

int nvdb_b200_search_device_static(nvdb_b200_index* idx, const float* queries_dev, uint32_t nq, uint32_t k,
                                   uint64_t* out_ids_dev, float* out_scores_dev, void* stream) {
  if (!idx) return fail(nullptr, NVDB_B200_ERR_ARG, "index is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!queries_dev) return fail(idx, NVDB_B200_ERR_ARG, "Null query");
  if (!out_ids_dev || !out_scores_dev) return fail(idx, NVDB_B200_ERR_ARG, "output buffer is NULL");
  std::lock_guard<std::mutex> g(idx->mu);
  CK(cudaSetDevice(idx->device), "cudaSetDevice");
  ScanPlan plan;
  int rc = plan_scan(idx, nq, k, &plan);
  if (rc) return rc;
  const size_t need = (size_t)nq * (size_t)plan.grid * k * 8u;
  if (need > (size_t)NVDB_B200_STATIC_SCRATCH_BYTES)
    return fail(idx, NVDB_B200_ERR_UNSUPPORTED,
                "search_device_static: nq * " + std::to_string(plan.grid) + " lists * k * 8 = " + std::to_string(need) +
                    " bytes of scratch exceed the static maximum of " + std::to_string((size_t)NVDB_B200_STATIC_SCRATCH_BYTES) +
                    " (use nvdb_b200_search_device, or fewer queries / a smaller k per call)");
  void* scratch = nullptr;
  if ((rc = static_scratch_ptr(idx, idx->device, &scratch))) return rc;
  uint64_t* partial = (uint64_t*)scratch;
  cudaStream_t s = stream ? (cudaStream_t)stream : idx->stream;

  ScanLaunch L;
  L.dtype = (int)idx->dtype;
  L.use_tma = idx->tma_ok;
  L.tmap = &idx->tmap;
  L.grid = plan.grid;
  L.stream = s;
  ScanArgs a;
  a.scales = idx->scales;
  a.base = idx->payload;
  a.n_rows = (uint32_t)idx->n;
  a.dim = idx->dim;
  a.k = k;
  a.cap = plan.cap;
  a.num_tiles = (uint32_t)((idx->n + kScanRowsPerTile - 1) / kScanRowsPerTile);
  a.kchunks = (uint32_t)(((size_t)idx->dim * elem_size(idx->dtype) + 127) / 128);
  a.stages = plan.stages;
  a.done_counter = nullptr;  // unfused: every CTA rewrites its k slots of every query, so nothing has to start at zero
  a.out_ids = nullptr;
  a.out_scores = nullptr;
  a.row_offset = idx->row_offset;
  a.xchg = ScanXchg{};
  for (uint32_t q0 = 0; q0 < nq; q0 += (uint32_t)plan.nq_template) {
    const uint32_t left = nq - q0;
    int t = plan.nq_template;
    while (t > 1 && (uint32_t)(t / 2) >= left) t >>= 1;
    a.nq = left < (uint32_t)t ? left : (uint32_t)t;
    a.queries = queries_dev + (size_t)q0 * idx->dim;
    a.partial = partial + (size_t)q0 * plan.grid * k;
    const cudaError_t le = scan_gemv_launch(L, a, t);
    if (le != cudaSuccess) return fail_cuda(idx, le, "scan_gemv launch (static scratch)");
    count_launch();
  }
  CK(merge_partials_launch(partial, (uint32_t)plan.grid, nq, k, idx->row_offset, out_ids_dev, out_scores_dev, s),
     "merge_partials launch (static scratch)");
  return NVDB_B200_OK;
}

}  // extern "C"
