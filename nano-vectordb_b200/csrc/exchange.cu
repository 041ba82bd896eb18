// exchange.cu — the one exchange step of the row-sharded index over NVLink peer memory, fused with the final merge.
//
// The reference merges the per-thread top-k lists of its row blocks serially on the calling thread
// (src/flat_index_pool.cpp:212-214, src/flat_index_omp.cpp:67-75). Here a row block is a GPU: every
// rank owns a WINDOW in its own HBM that its peers map (CUDA IPC between the per-GPU processes, plain pointers inside
// one process); ONE kernel per rank
//   1. stores this rank's packed local top-k ([nq*k u64 ids][nq*k f32 scores]) into slot `rank` of every peer's window
//      with plain 8-byte stores over NVLink (and into its own window),
//   2. the last CTA to finish publishes `epoch` in flag `rank` of every window (release, system scope),
//   3. every CTA waits until the `world` flags of its OWN window reach `epoch` (acquire, system scope; bounded by a
//      timeout so that a dead peer cannot hang the GPU), then
//   4. merges the world*k candidates of its queries under (score desc, global id asc) — the rule of
//      merge_lists_kernel (topk_merge.cu), so the result does not depend on the number of GPUs.
// No NCCL call, no second launch, no host round trip. Slots are double-buffered by epoch parity: a peer can only
// overwrite parity p two epochs later, which needs this rank's flag of the epoch in between, which this rank publishes
// from a kernel that is stream-ordered after the one still reading p. All calls of one exchange must therefore be
// issued on ONE stream per rank, by every rank, in the same order.
#include <math.h>

#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/nvdb_b200.h"
#include "common.cuh"
#include "exchange.cuh"
#include "exchange.h"
#include "scan.h"

namespace nvdb_b200 {
void count_launch(uint32_t n);
namespace {

constexpr uint32_t kMaxWorld = kXchgMaxWorld;
constexpr uint32_t kHeaderBytes = kXchgHeaderBytes;
constexpr int kThreads = 256;

struct XchgParams {
  ScanXchg x;             // windows, status, timeout, parity offset, world / rank / epoch
  const uint64_t* local;  // this rank's packed result, `words` 8-byte words; null = an empty shard (padding)
  uint64_t* out_ids;      // [nq][k]
  float* out_scores;      // [nq][k]
  uint32_t* counter;      // CTA arrival counter (own device, zero between launches)
  size_t words;
  uint32_t nq, k, phases;  // phases: 1 push, 2 wait + merge, 3 both
};

__global__ void __launch_bounds__(kThreads) exchange_merge_kernel(const XchgParams p) {
  extern __shared__ uint64_t sm[];
  __shared__ uint32_t s_last, s_fail;
  const uint32_t tid = threadIdx.x;
  const size_t nk = (size_t)p.nq * p.k;
  auto sync = [] { __syncthreads(); };

  if (p.phases & 1u) {
    // ---- 1. push: every word goes to slot `rank` of every window
    const size_t my_slot = p.x.parity_off + (size_t)p.x.rank * p.x.slot_bytes;
    for (size_t i = (size_t)blockIdx.x * kThreads + tid; i < p.words; i += (size_t)gridDim.x * kThreads) {
      const uint64_t v = p.local ? p.local[i] : (i < nk ? ~0ull : 0xFF800000FF800000ull);  // id = UINT64_MAX, score = -inf
      for (uint32_t r = 0; r < p.x.world; ++r) reinterpret_cast<uint64_t*>(p.x.win[r] + my_slot)[i] = v;
    }
    // ---- 2. publish: the last CTA has observed every other CTA's arrival, hence (fence cumulativity) their stores
    __threadfence_system();
    __syncthreads();
    if (tid == 0) {
      const uint32_t old = atomicAdd(p.counter, 1u);
      s_last = (old == gridDim.x - 1u) ? 1u : 0u;
    }
    __syncthreads();
    if (s_last) {
      __threadfence_system();
      if (tid == 0) *p.counter = 0u;
      if (tid < p.x.world) st_release_sys(reinterpret_cast<uint32_t*>(p.x.win[tid]) + p.x.rank, p.x.epoch);
    }
  }
  if (!(p.phases & 2u)) return;

  // ---- 3. wait for the `world` flags of the own window
  const bool ok = xchg_wait(reinterpret_cast<const uint32_t*>(p.x.win[p.x.rank]), p.x.world, p.x.epoch, p.x.status,
                            p.x.timeout_ns, tid, &s_fail, sync);
  // ---- 4. merge world*k candidates per query
  const uint32_t L = p.x.world * p.k;
  const unsigned char* slots = p.x.win[p.x.rank] + p.x.parity_off;
  for (uint32_t q = blockIdx.x; q < p.nq; q += gridDim.x)
    xchg_merge_query(slots, p.x.slot_bytes, nk, p.x.world, p.k, q, sm, reinterpret_cast<uint32_t*>(sm + L),
                     p.out_ids + (size_t)q * p.k, p.out_scores + (size_t)q * p.k, ok, tid, (uint32_t)kThreads, sync);
}

thread_local std::string t_xerr;

}  // namespace
}  // namespace nvdb_b200

using namespace nvdb_b200;

void nvdb_b200::exchange_begin_epoch(nvdb_b200_exchange* x, ScanXchg* out) {
  x->epoch += 1u;
  for (uint32_t r = 0; r < kXchgMaxWorld; ++r) out->win[r] = r < x->world ? x->peers[r] : nullptr;
  out->status = x->counter + 1;
  out->timeout_ns = x->timeout_ns;
  out->parity_off = kXchgHeaderBytes + (size_t)(x->epoch & 1u) * x->world * x->slot_bytes;
  out->slot_bytes = x->slot_bytes;
  out->world = x->world;
  out->rank = x->rank;
  out->epoch = x->epoch;
}

static int xfail(nvdb_b200_exchange* x, int code, const std::string& msg) {
  t_xerr = msg;
  if (x) x->err = msg;
  return code;
}
#define XCK(call, what)                                                                                      \
  do {                                                                                                       \
    cudaError_t _e = (call);                                                                                 \
    if (_e != cudaSuccess)                                                                                   \
      return xfail(x, _e == cudaErrorMemoryAllocation ? NVDB_B200_ERR_OOM : NVDB_B200_ERR_CUDA,             \
                   std::string(what) + ": " + cudaGetErrorString(_e));                                       \
  } while (0)

extern "C" {

int nvdb_b200_exchange_create(int device, uint32_t world, uint32_t rank, uint64_t slot_bytes, nvdb_b200_exchange** out) {
  nvdb_b200_exchange* x = nullptr;
  if (!out) return xfail(nullptr, NVDB_B200_ERR_ARG, "exchange_create: out is NULL");
  *out = nullptr;
  if (world == 0 || world > kMaxWorld || rank >= world || slot_bytes == 0)
    return xfail(nullptr, NVDB_B200_ERR_ARG, "exchange_create: need 1 <= world <= 16, rank < world, slot_bytes > 0");
  XCK(cudaSetDevice(device), "cudaSetDevice");
  x = new nvdb_b200_exchange();
  x->device = device;
  x->world = world;
  x->rank = rank;
  x->slot_bytes = (size_t)((slot_bytes + 255u) & ~(uint64_t)255u);
  x->window_bytes = kHeaderBytes + 2u * (size_t)world * x->slot_bytes;
  if (const char* t = getenv("NVDB_B200_XCHG_TIMEOUT_MS")) {
    const long ms = atol(t);
    if (ms > 0) x->timeout_ns = (unsigned long long)ms * 1000000ull;
  }
  cudaError_t e = cudaMalloc(&x->window, x->window_bytes);
  if (e == cudaSuccess) e = cudaMemset(x->window, 0, kHeaderBytes);
  if (e == cudaSuccess) e = cudaMalloc(&x->staging, x->slot_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&x->counter, 2 * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(x->counter, 0, 2 * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&x->num_sms, cudaDevAttrMultiProcessorCount, device);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(x->window);
    cudaFree(x->staging);
    cudaFree(x->counter);
    delete x;
    return xfail(nullptr, e == cudaErrorMemoryAllocation ? NVDB_B200_ERR_OOM : NVDB_B200_ERR_CUDA,
                 std::string("exchange_create: ") + cudaGetErrorString(e));
  }
  x->peers[rank] = x->window;
  *out = x;
  return NVDB_B200_OK;
}

int nvdb_b200_exchange_ipc_handle(nvdb_b200_exchange* x, unsigned char* handle64) {
  if (!x || !handle64) return xfail(x, NVDB_B200_ERR_ARG, "exchange_ipc_handle: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == NVDB_B200_IPC_HANDLE_BYTES, "IPC handle size");
  XCK(cudaSetDevice(x->device), "cudaSetDevice");
  cudaIpcMemHandle_t h;
  XCK(cudaIpcGetMemHandle(&h, x->window), "cudaIpcGetMemHandle");
  memcpy(handle64, &h, sizeof(h));
  return NVDB_B200_OK;
}

int nvdb_b200_exchange_window(nvdb_b200_exchange* x, void** window_dev, uint64_t* window_bytes) {
  if (!x) return xfail(x, NVDB_B200_ERR_ARG, "exchange_window: exchange is NULL");
  if (window_dev) *window_dev = x->window;
  if (window_bytes) *window_bytes = x->window_bytes;
  return NVDB_B200_OK;
}

int nvdb_b200_exchange_connect(nvdb_b200_exchange* x, const unsigned char* handles, void* const* windows) {
  if (!x) return xfail(x, NVDB_B200_ERR_ARG, "exchange_connect: exchange is NULL");
  if (x->connected) return xfail(x, NVDB_B200_ERR_ARG, "exchange_connect: already connected");
  if (x->world > 1 && !handles && !windows)
    return xfail(x, NVDB_B200_ERR_ARG, "exchange_connect: need IPC handles or window pointers of the peers");
  XCK(cudaSetDevice(x->device), "cudaSetDevice");
  for (uint32_t r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    if (windows) {
      if (!windows[r]) return xfail(x, NVDB_B200_ERR_ARG, "exchange_connect: NULL peer window");
      cudaPointerAttributes at;
      XCK(cudaPointerGetAttributes(&at, windows[r]), "cudaPointerGetAttributes(peer window)");
      if (at.type != cudaMemoryTypeDevice) return xfail(x, NVDB_B200_ERR_ARG, "exchange_connect: peer window is not device memory");
      if (at.device != x->device) {
        cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        XCK(e, "cudaDeviceEnablePeerAccess");
      }
      x->peers[r] = static_cast<unsigned char*>(windows[r]);
    } else {
      cudaIpcMemHandle_t h;
      memcpy(&h, handles + (size_t)r * sizeof(h), sizeof(h));
      void* ptr = nullptr;
      XCK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle(peer window)");
      x->peers[r] = static_cast<unsigned char*>(ptr);
      x->opened[r] = true;
    }
  }
  x->connected = true;
  return NVDB_B200_OK;
}

int nvdb_b200_exchange_merge(nvdb_b200_exchange* x, const void* local_packed_dev, uint32_t nq, uint32_t k,
                             uint64_t* out_ids_dev, float* out_scores_dev, int phases, void* stream) {
  if (!x) return xfail(x, NVDB_B200_ERR_ARG, "exchange_merge: exchange is NULL");
  if (nq == 0 || k == 0) return NVDB_B200_OK;
  if (!x->connected && x->world > 1) return xfail(x, NVDB_B200_ERR_ARG, "exchange_merge: not connected");
  if (phases < 1 || phases > 3) return xfail(x, NVDB_B200_ERR_ARG, "exchange_merge: phases must be 1 (push), 2 (merge) or 3");
  if ((phases & 2) && (!out_ids_dev || !out_scores_dev))   // local_packed_dev == NULL: an empty shard, padding is pushed
    return xfail(x, NVDB_B200_ERR_ARG, "exchange_merge: NULL output buffer");
  const size_t nk = (size_t)nq * k, words = nk + (nk + 1) / 2;
  if (words * 8u > x->slot_bytes) return xfail(x, NVDB_B200_ERR_ARG, "exchange_merge: packed result larger than the slot");
  const size_t smem = (size_t)x->world * k * 12u + 8u;
  if (smem > 200u * 1024u) return xfail(x, NVDB_B200_ERR_UNSUPPORTED, "exchange_merge: world * k too large for one CTA");
  XCK(cudaSetDevice(x->device), "cudaSetDevice");
  if (smem > 48u * 1024u)
    XCK(cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
        "cudaFuncSetAttribute(exchange_merge)");
  // A push (phase 1) starts an epoch; a merge-only call (phase 2) completes the epoch of the preceding push.
  XchgParams p;
  if (phases & 1) {
    exchange_begin_epoch(x, &p.x);
  } else {
    x->epoch -= 1u;
    exchange_begin_epoch(x, &p.x);
  }
  p.local = static_cast<const uint64_t*>(local_packed_dev);
  p.out_ids = out_ids_dev;
  p.out_scores = out_scores_dev;
  p.counter = x->counter;
  p.words = words;
  p.nq = nq;
  p.k = k;
  p.phases = (uint32_t)phases;
  // Every CTA spins in step 3, so all of them must be resident at once: at most one CTA per SM.
  const size_t push_ctas = (words + (size_t)kThreads * 4 - 1) / ((size_t)kThreads * 4);
  size_t grid = (phases & 2) ? nq : 1;
  if ((phases & 1) && push_ctas > grid) grid = push_ctas;
  if (grid > (size_t)x->num_sms) grid = (size_t)x->num_sms;
  exchange_merge_kernel<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  const cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess) {
    if (phases & 1) x->epoch -= 1u;  // nothing was launched for this epoch: stay in step with the peers
    return xfail(x, NVDB_B200_ERR_CUDA, std::string("exchange_merge launch: ") + cudaGetErrorString(le));
  }
  count_launch(1);
  return NVDB_B200_OK;
}

// Collective: after a timed-out exchange (status 1) every rank calls reset before the next search. Clears the sticky
// status word, the CTA counter and this rank's flags, and restarts the epochs at 0 — callers must make sure (barrier)
// that no rank is still inside an exchange call and that every rank resets before any rank searches again.
int nvdb_b200_exchange_reset(nvdb_b200_exchange* x) {
  if (!x) return xfail(x, NVDB_B200_ERR_ARG, "exchange_reset: exchange is NULL");
  XCK(cudaSetDevice(x->device), "cudaSetDevice");
  XCK(cudaDeviceSynchronize(), "exchange_reset sync");
  XCK(cudaMemset(x->counter, 0, 2 * sizeof(uint32_t)), "exchange_reset status");
  XCK(cudaMemset(x->window, 0, kHeaderBytes), "exchange_reset flags");
  XCK(cudaDeviceSynchronize(), "exchange_reset sync");
  x->epoch = 0;
  x->err.clear();
  return NVDB_B200_OK;
}

int nvdb_b200_exchange_status(nvdb_b200_exchange* x, uint32_t* status) {
  if (!x || !status) return xfail(x, NVDB_B200_ERR_ARG, "exchange_status: NULL argument");
  XCK(cudaSetDevice(x->device), "cudaSetDevice");
  XCK(cudaMemcpy(status, x->counter + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost), "read exchange status");
  if (*status != 0u) x->err = "exchange: timed out waiting for a peer's result (a rank did not take part in the call)";
  return NVDB_B200_OK;
}

const char* nvdb_b200_exchange_last_error(const nvdb_b200_exchange* x) { return x ? x->err.c_str() : t_xerr.c_str(); }

int nvdb_b200_exchange_destroy(nvdb_b200_exchange* x) {
  if (!x) return NVDB_B200_OK;
  cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  for (uint32_t r = 0; r < x->world; ++r)
    if (x->opened[r]) cudaIpcCloseMemHandle(x->peers[r]);
  cudaFree(x->window);
  cudaFree(x->staging);
  cudaFree(x->counter);
  delete x;
  return NVDB_B200_OK;
}

}  // extern "C"
