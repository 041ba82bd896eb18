// ivfpq.cu — IVF-PQ candidate generation on the device: the step in front of the exact refine.
//
// Behaviour follows the reference's use of FAISS (apps/nvdb_ivfpq_build.cpp:78-135 build; apps/nvdb_ivf_eval.cpp:
// 398-408,482 search: IndexIVFPQ over an IndexFlatL2 coarse quantiser, METRIC_L2, 8-bit codes, residual encoding,
// `nprobe` lists per query, top-k_search candidates by ADC distance handed to the refine stage). FAISS itself is an
// un-vendored third-party dependency (absent here: PARITY UNPINNED), so the algorithm is restated
// from its publication (Jegou, Douze, Schmid: "Product quantization for nearest neighbor search", IVFADC):
//
//   build : coarse k-means (nlist centroids) on the first ntrain rows; PQ codebooks (M sub-spaces x 256 codewords)
//           by k-means on the training residuals x - c(x); every row -> (list, M code bytes); rows grouped by list
//           (stable: ascending row id inside a list).
//   search: per query the nprobe nearest centroids; a CTA per (query, group of probed lists) builds the M x 256 table
//           of inner products q_m . codeword in shared memory and streams its lists' codes (M byte look-ups + adds
//           per vector, fixed m order; distance = |q|^2 - 2 q.c + |y|^2 - 2 sum) through the CTA top-R selector
//           (topk_select.cuh); the groups are merged (topk_merge.cu) into R candidates ordered by
//           (distance asc, id asc). Candidates stay on the device for nvdb_b200_refine_l2_device.
//
// Checked against oracle/ivfpq_ref.py (numpy restatement fed with the trained parameters exported from here).
#include <cub/device/device_radix_sort.cuh>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/nvdb_b200.h"
#include "common.cuh"
#include "ivf.h"
#include "scan.h"
#include "topk_select.cuh"

using namespace nvdb_b200;

struct nvdb_b200_ivfpq {
  int device = 0;
  uint32_t d = 0, nlist = 0, M = 0, dsub = 0;
  uint64_t n = 0;
  int num_sms = 148, max_smem = 227 * 1024;
  float* coarse = nullptr;     // [nlist x d]
  float* coarse_hn = nullptr;  // [nlist] 0.5 |c|^2
  float* codebooks = nullptr;  // [M x 256 x dsub]
  float* cb_hn = nullptr;      // [M x 256]
  uint8_t* codes = nullptr;    // [n x M], grouped by list
  uint32_t* ids = nullptr;     // [n] row id of every grouped position
  float* rnorm = nullptr;      // [n] |c + reconstruction|^2 of every grouped position
  uint32_t* list_off = nullptr;  // [nlist + 1]
  float* rot_t = nullptr;      // OPQ: [d x d] R^T (row j = column j of the rotation R; x' = x R), NULL without OPQ
  float* zeros = nullptr;      // [max(d, 256)] zeros (GEMM bias of plain products)
  cudaStream_t stream = nullptr;
  // grow-only search workspaces
  void* ws[10] = {};
  size_t ws_bytes[10] = {};
  void* h_pin = nullptr;
  size_t h_pin_bytes = 0;
  // Small host-API searches (<= kGraphMaxQ queries) replay a captured graph of the call's copies and kernels: one
  // launch instead of 7-9 API calls. A shape is captured the second time it is seen in a row (its workspaces exist by
  // then); a workspace regrow (ws_gen) drops the graphs that baked the old pointers in.
  struct Graph {
    uint64_t key[8] = {};
    uint64_t ws_gen = 0;
    cudaGraphExec_t exec = nullptr;
    uint32_t launches = 0;
  } graphs[4];
  uint32_t graph_next = 0;
  uint64_t ws_gen = 1, last_key[8] = {};
  int graph_mode = -1;  // -1 = read NVDB_B200_IVF_GRAPH on first use (default on)
  uint64_t graph_replays = 0;
  std::mutex mu;
  mutable std::string err;
};

namespace {

thread_local std::string t_ivf_err;
enum { IWS_Q = 0, IWS_SCORES = 1, IWS_PROBES = 2, IWS_PARTIAL = 3, IWS_IDS64 = 4, IWS_SC = 5, IWS_CAND = 6, IWS_OUT = 7, IWS_LUT = 8,
       IWS_QROT = 9 };

int ifail(const nvdb_b200_ivfpq* h, int code, const std::string& msg) {
  t_ivf_err = msg;
  if (h) h->err = msg;
  return code;
}
#define ICK(call, what)                                                                                         \
  do {                                                                                                          \
    cudaError_t _e = (call);                                                                                    \
    if (_e != cudaSuccess)                                                                                      \
      return ifail(h, _e == cudaErrorMemoryAllocation ? NVDB_B200_ERR_OOM : NVDB_B200_ERR_CUDA,                 \
                   std::string(what) + ": " + cudaGetErrorString(_e));                                          \
  } while (0)

int iws(nvdb_b200_ivfpq* h, int which, size_t bytes) {
  if (h->ws_bytes[which] >= bytes) return NVDB_B200_OK;
  if (h->ws[which]) cudaFree(h->ws[which]);
  h->ws[which] = nullptr;
  h->ws_bytes[which] = 0;
  h->ws_gen++;
  ICK(cudaMalloc(&h->ws[which], bytes), "cudaMalloc (ivfpq workspace)");
  h->ws_bytes[which] = bytes;
  return NVDB_B200_OK;
}

// ---- build-time kernels ------------------------------------------------------------------------------------------

// rows [row0, row0 + nrows) of the base as FP32 (f16 widened exactly, i8 de-quantised with its row scale)
__global__ void rows_to_f32_kernel(uint32_t dtype, const void* __restrict__ payload, const float* __restrict__ scales,
                                   uint64_t row0, uint32_t nrows, uint32_t d, float* __restrict__ out) {
  const size_t total = (size_t)nrows * d;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / d, src = (size_t)row0 * d + e;
    float v;
    if (dtype == kDtF32) v = reinterpret_cast<const float*>(payload)[src];
    else if (dtype == kDtF16) v = __half2float(reinterpret_cast<const __half*>(payload)[src]);
    else v = __fmul_rn((float)reinterpret_cast<const int8_t*>(payload)[src], scales[row0 + r]);
    out[e] = v;
  }
}

__global__ void subtract_centroid_kernel(float* __restrict__ x, uint32_t nrows, uint32_t d, const uint32_t* __restrict__ assign,
                                         const float* __restrict__ coarse) {
  const size_t total = (size_t)nrows * d;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / d, t = e - r * d;
    x[e] = __fsub_rn(x[e], coarse[(size_t)assign[r] * d + t]);
  }
}

__global__ void write_code_kernel(const uint32_t* __restrict__ assign, uint32_t nrows, uint32_t M, uint32_t m,
                                  uint8_t* __restrict__ codes /* already offset to the chunk's first row */) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nrows) codes[(size_t)i * M + m] = (uint8_t)assign[i];
}

// y[i][m*dsub .. +dsub) = codeword_m[assign[i]]   (OPQ training: reconstruction of the rotated training rows)
__global__ void gather_codewords_kernel(const uint32_t* __restrict__ assign, uint32_t nrows, const float* __restrict__ cb_m,
                                        uint32_t dsub, float* __restrict__ y_m, uint32_t ldy) {
  const size_t total = (size_t)nrows * dsub;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t r = e / dsub, t = e - r * dsub;
    y_m[r * ldy + t] = cb_m[(size_t)assign[r] * dsub + t];
  }
}
__global__ void identity_kernel(float* __restrict__ a, uint32_t d) {
  const size_t total = (size_t)d * d;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x)
    a[e] = (e / d == e % d) ? 1.f : 0.f;
}

__global__ void iota_kernel(uint32_t* __restrict__ v, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = i;
}

// list_off[l] = first grouped position whose list id is >= l (sorted_lists ascending)
__global__ void list_offsets_kernel(const uint32_t* __restrict__ sorted_lists, uint32_t n, uint32_t nlist,
                                    uint32_t* __restrict__ list_off) {
  const uint32_t l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l > nlist) return;
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (sorted_lists[mid] < l) lo = mid + 1;
    else hi = mid;
  }
  list_off[l] = lo;
}

__global__ void gather_codes_kernel(const uint8_t* __restrict__ src, const uint32_t* __restrict__ ids, uint32_t n, uint32_t M,
                                    uint8_t* __restrict__ dst) {
  const size_t total = (size_t)n * M;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t pos = e / M, m = e - pos * M;
    dst[e] = src[(size_t)ids[pos] * M + m];
  }
}

// ---- search kernels ----------------------------------------------------------------------------------------------

// probes[q][0..nprobe) = the nprobe best-scoring lists of query q (score desc == L2 asc, id asc among equals)
constexpr int kProbeThreads = 1024;
constexpr uint32_t kProbeBuf = 1024;  // survivors of the radix select that are sorted exactly
// Top-nprobe of nlist scores without sorting all of them: a 4 x 8-bit radix select on the ordered score finds the
// nprobe-th largest score value, the (few) keys at or above it are compacted and bitonic-sorted on the full
// (score, id) key. A full sort of 4096 keys cost 38 us per query (78 barrier-separated stages); this is ~4 histogram
// passes + a sort of next_pow2(nprobe + ties) keys. Falls back to the full sort when more than kProbeBuf keys survive
// (nprobe >= 1024, or a pathological number of equal scores).
__global__ void __launch_bounds__(kProbeThreads)
select_probes_kernel(const float* __restrict__ scores, uint32_t nlist, uint32_t bufn, uint32_t nprobe,
                     uint32_t* __restrict__ probes) {
  extern __shared__ uint64_t pk[];              // [bufn] all keys, then [kProbeBuf] survivors
  uint64_t* top = pk + bufn;
  __shared__ uint32_t hist[256];
  __shared__ uint32_t s_prefix, s_remaining, s_count, s_ties;
  const uint32_t q = blockIdx.x, tid = threadIdx.x, lane = tid & 31u;
  auto sync = [] { __syncthreads(); };
  for (uint32_t j = tid; j < bufn; j += kProbeThreads)
    pk[j] = j < nlist ? make_key(scores[(size_t)q * nlist + j], j) : 0ull;
  if (tid == 0) {
    s_prefix = 0;
    s_remaining = nprobe;
    s_count = 0;
  }
  uint32_t mask = 0;
  for (int shift = 24; shift >= 0; shift -= 8) {
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    for (uint32_t j = tid; j < nlist; j += kProbeThreads) {
      const uint32_t sc = (uint32_t)(pk[j] >> 32);
      if ((sc & mask) == prefix) atomicAdd(&hist[(sc >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {  // bins 8*lane .. 8*lane+7; find the bin where the count from the top reaches `remaining`
      uint32_t local[8], sum = 0;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        local[u] = hist[lane * 8 + u];
        sum += local[u];
      }
      uint32_t above = 0;  // keys in the bins of higher lanes (exclusive suffix sum over the warp)
#pragma unroll
      for (int l2 = 31; l2 > 0; --l2) {
        const uint32_t v = __shfl_sync(0xffffffffu, sum, l2);
        if ((int)lane < l2) above += v;
      }
      const uint32_t remaining = s_remaining;
      if (above < remaining && remaining <= above + sum) {
        uint32_t acc = above;
        for (int u = 7; u >= 0; --u) {
          if (acc + local[u] >= remaining) {
            s_prefix = prefix | ((uint32_t)(lane * 8 + u) << shift);
            s_remaining = remaining - acc;  // still to take among the keys of this bin
            s_ties = local[u];
            break;
          }
          acc += local[u];
        }
      }
    }
    mask |= 0xFFu << shift;
    __syncthreads();
  }
  // s_prefix = the nprobe-th largest score; keys above it: nprobe - s_remaining, keys equal to it: s_ties
  const uint32_t thr = s_prefix, total = nprobe - s_remaining + s_ties;
  if (nprobe <= nlist && total <= kProbeBuf) {
    for (uint32_t j = tid; j < nlist; j += kProbeThreads) {
      const uint64_t key = pk[j];
      if ((uint32_t)(key >> 32) >= thr) top[atomicAdd(&s_count, 1u)] = key;
    }
    __syncthreads();
    if (tid < total) {  // rank counting: total <= kProbeBuf == kProbeThreads keys, all distinct
      const uint64_t key = top[tid];
      uint32_t rank = 0;
      for (uint32_t i = 0; i < total; ++i) rank += (top[i] > key) ? 1u : 0u;
      if (rank < nprobe) probes[(size_t)q * nprobe + rank] = key_row(key);
    }
    return;
  }
  bitonic_sort_desc(pk, bufn, tid, kProbeThreads, sync);
  for (uint32_t p = tid; p < nprobe; p += kProbeThreads)
    probes[(size_t)q * nprobe + p] = pk[p] ? key_row(pk[p]) : 0xFFFFFFFFu;
}

constexpr int kScanThreads = 256;
constexpr uint32_t kScanVecs = 4;  // stored vectors per thread between two barriers of the list scan

struct IvfScanArgs {
  const float* queries;   // [nq x d]
  const float* coarse;    // [nlist x d]
  const float* codebooks; // [M x 256 x dsub]
  const uint8_t* codes;
  const uint32_t* ids;
  const float* rnorm;     // [n] |c + reconstruction|^2 of every grouped position
  const uint32_t* list_off;
  const uint32_t* probes; // [nq x nprobe]
  const float* lut;       // [nq x M x 256] q_m . codeword (ivfpq_lut_kernel)
  uint64_t* partial;      // [nq x G x R] ordered keys (score = -distance)
  uint32_t d, M, dsub, nprobe, R, cap, G;
};

// sum over the CTA in a fixed order (shuffle tree per warp, then the 8 warp sums left to right): every thread returns
// the same bits, independent of the batch the query came in
__device__ __forceinline__ float block_sum(float v, float* red, uint32_t tid) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, off));
  __syncthreads();
  if ((tid & 31u) == 0) red[tid >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < kScanThreads / 32; ++w) s = __fadd_rn(s, red[w]);
  return s;
}

// T[q][m][j] = q_m . codeword_m[j]: one CTA per (sub-space, query), thread j owns codeword j. The table depends on the
// query only (see ivfpq_scan_kernel), so it is built ONCE per query here — M x nq small CTAs that each read 256 x dsub
// codebook words — instead of once per scan CTA (the whole M x 256 x dsub codebook through one SM: 33 us of L2 latency).
__global__ void __launch_bounds__(256)
ivfpq_lut_kernel(const float* __restrict__ queries, const float* __restrict__ codebooks, uint32_t d, uint32_t M, uint32_t dsub,
                 float* __restrict__ lut) {
  const uint32_t m = blockIdx.x, q = blockIdx.y, j = threadIdx.x;
  const float* qm = queries + (size_t)q * d + (size_t)m * dsub;  // warp-uniform loads
  const float* cw = codebooks + ((size_t)m * 256 + j) * dsub;
  float s = 0.f;
  for (uint32_t t = 0; t < dsub; ++t) s = __fmaf_rn(__ldg(qm + t), __ldg(cw + t), s);
  lut[((size_t)q * M + m) * 256 + j] = s;
}

// CTA-level running top-R for the list scan. Same admission idea as topk_select.cuh (compare against the current R-th
// best, append survivors with a shared-memory atomic), but compaction is by RANK COUNTING instead of a bitonic sort:
// every held key counts the keys that beat it (broadcast shared-memory reads, no barriers inside) and, if fewer than R
// do, drops itself at its rank in the other half of a double buffer. Keys are unique ((distance, row id), every row is
// visited once), so ranks are a permutation. For the few hundred keys a CTA holds this is one barrier and ~c compares
// per thread; the bitonic network needs log^2(c)/2 barrier-separated stages at ~0.2 us each — it was 40 % of the scan
// kernel and most of the merge (profiles/r1_ivfpq_latency_breakdown.txt).
struct RankTopR {
  uint64_t* keys;  // [2][cap]
  uint32_t* cnt;
  float* thr;
  uint32_t cap, R, cur, bound;  // bound: uniform upper bound on *cnt
  bool have_thr;

  __device__ __forceinline__ void init(uint64_t* k, uint32_t* c, float* t, uint32_t cap_, uint32_t R_, uint32_t tid) {
    keys = k;
    cnt = c;
    thr = t;
    cap = cap_;
    R = R_;
    cur = 0;
    bound = 0;
    have_thr = false;
    if (tid == 0) {
      *cnt = 0;
      *thr = -INFINITY;
    }
  }
  __device__ __forceinline__ bool offer(float score, uint32_t id) {
    if (score >= *thr) {
      const uint32_t pos = atomicAdd(cnt, 1u);
      keys[cur * cap + pos] = make_key(score, id);
      return true;
    }
    return false;
  }
  // all threads of the CTA; ends with the survivors (best first) in the current half
  __device__ void compact(uint32_t tid, uint32_t nthreads) {
    __syncthreads();
    const uint32_t c = *cnt;
    const uint64_t* src = keys + cur * cap;
    uint64_t* dst = keys + (cur ^ 1u) * cap;
    if (tid == 0 && (c & 1u)) keys[cur * cap + c] = 0ull;  // pad to a whole 16-byte pair (cap is even)
    __syncthreads();
    const ulonglong2* src2 = reinterpret_cast<const ulonglong2*>(src);
    const uint32_t pairs = (c + 1u) >> 1;
    for (uint32_t j = tid; j < c; j += nthreads) {
      const uint64_t key = src[j];
      uint32_t rank = 0;
#pragma unroll 4
      for (uint32_t i = 0; i < pairs; ++i) {  // broadcast 16-byte loads: two held keys per shared-memory wavefront
        const ulonglong2 o = src2[i];
        rank += (o.x > key ? 1u : 0u) + (o.y > key ? 1u : 0u);
      }
      if (rank < R) {
        dst[rank] = key;
        if (rank == R - 1) *thr = key_score(key);
      }
    }
    __syncthreads();
    if (tid == 0) *cnt = c < R ? c : R;
    cur ^= 1u;
    bound = c < R ? c : R;
    have_thr = c >= R;
    __syncthreads();
  }
  // after a round of offers (all threads): refresh the exact count; compact when the next round of up to `next` offers
  // could overflow, or — the first time R keys exist — to install the threshold that makes later pushes rare
  __device__ __forceinline__ void after_offers(uint32_t next, uint32_t tid, uint32_t nthreads) {
    __syncthreads();
    bound = *cnt;
    if (bound + next > cap || (!have_thr && bound >= R)) compact(tid, nthreads);
  }
};

// One CTA per (probe group g, query q). ADC distance of a stored vector y = c + reconstruction:
//   |q - y|^2 = |q|^2 - 2 q.c + |y|^2 - 2 sum_m q_m . codeword_m[code_m]
// The table T (ivfpq_lut_kernel) is copied into shared memory once and serves all the lists of the group
// (p = g, g + G, ...); |y|^2 is stored per vector at build time, q.c is one block reduction per list.
// (A table of squared residual sub-distances would have to be rebuilt for every probed list: measured 14 ms for
// 1000 queries x 64 probes, profiles/r1_ivfpq_1M_v3.jsonl.)
__global__ void __launch_bounds__(kScanThreads) ivfpq_scan_kernel(const IvfScanArgs a) {
  extern __shared__ __align__(16) uint8_t sm[];
  float* lut = reinterpret_cast<float*>(sm);                    // [M x 256]
  float* qv = lut + (size_t)a.M * 256;                          // [d]
  uint64_t* keys = reinterpret_cast<uint64_t*>(qv + ((a.d + 3u) & ~3u));  // [2][cap]
  uint32_t* cnt = reinterpret_cast<uint32_t*>(keys + 2u * a.cap);
  float* thr = reinterpret_cast<float*>(cnt + 1);
  float* red = thr + 1;                                         // [kScanThreads / 32]
  const uint32_t tid = threadIdx.x, g = blockIdx.x, q = blockIdx.y;
  RankTopR sel;
  sel.init(keys, cnt, thr, a.cap, a.R, tid);
  {
    const float4* src = reinterpret_cast<const float4*>(a.lut + (size_t)q * a.M * 256);
    float4* dst = reinterpret_cast<float4*>(lut);
    for (uint32_t e = tid; e < a.M * 64u; e += kScanThreads) dst[e] = __ldg(src + e);
  }
  float qpart = 0.f;
  for (uint32_t t = tid; t < a.d; t += kScanThreads) {
    const float v = a.queries[(size_t)q * a.d + t];
    qv[t] = v;
    qpart = __fmaf_rn(v, v, qpart);
  }
  const float qq = block_sum(qpart, red, tid);  // (its barriers also publish lut, qv and the selector state)
  const bool vec16 = (a.M & 15u) == 0;
  constexpr uint32_t U = kScanVecs, TILE = kScanThreads * kScanVecs;
  for (uint32_t p = g; p < a.nprobe; p += a.G) {
    const uint32_t list = a.probes[(size_t)q * a.nprobe + p];
    if (list == 0xFFFFFFFFu) continue;  // nprobe > nlist (uniform over the CTA)
    const uint32_t begin = a.list_off[list], end = a.list_off[list + 1];
    float cpart = 0.f;
    for (uint32_t t = tid; t < a.d; t += kScanThreads) cpart = __fmaf_rn(qv[t], a.coarse[(size_t)list * a.d + t], cpart);
    const float qc = block_sum(cpart, red, tid);
    const float base_term = __fmaf_rn(-2.f, qc, qq);
    // U vectors per thread and barrier interval: their code words are requested together and the next 16 bytes of
    // every vector are in flight while the current 16 look-ups run, so a tile costs ~one memory round trip instead
    // of one per vector chunk (a 256-vector tile with dependent loads took 5.8 us: profiles/r1_ivfpq_latency_*.txt)
    for (uint32_t base = begin; base < end; base += TILE) {
      uint32_t idx[U], id[U];
      float rn[U], acc[U];
      uint4 w[U];
#pragma unroll
      for (uint32_t u = 0; u < U; ++u) {
        idx[u] = base + u * kScanThreads + tid;
        const bool ok = idx[u] < end;
        const uint32_t i = ok ? idx[u] : begin;  // clamp: harmless duplicate read, never offered
        rn[u] = __ldg(a.rnorm + i);
        id[u] = __ldg(a.ids + i);
        acc[u] = 0.f;
        if (vec16) w[u] = __ldg(reinterpret_cast<const uint4*>(a.codes + (size_t)i * a.M));
        idx[u] = ok ? i : 0xFFFFFFFFu;
      }
      if (vec16) {
        for (uint32_t m0 = 0; m0 < a.M; m0 += 16) {
          uint32_t ww[U][4];
#pragma unroll
          for (uint32_t u = 0; u < U; ++u) {
            ww[u][0] = w[u].x;
            ww[u][1] = w[u].y;
            ww[u][2] = w[u].z;
            ww[u][3] = w[u].w;
            if (m0 + 16 < a.M) {
              const uint32_t i = idx[u] != 0xFFFFFFFFu ? idx[u] : begin;
              w[u] = __ldg(reinterpret_cast<const uint4*>(a.codes + (size_t)i * a.M + m0 + 16));
            }
          }
#pragma unroll
          for (int e = 0; e < 16; ++e)
#pragma unroll
            for (uint32_t u = 0; u < U; ++u)
              acc[u] = __fadd_rn(acc[u], lut[(m0 + e) * 256 + ((ww[u][e >> 2] >> ((e & 3) * 8)) & 0xFFu)]);
        }
      } else {
#pragma unroll
        for (uint32_t u = 0; u < U; ++u) {
          const uint8_t* code = a.codes + (size_t)(idx[u] != 0xFFFFFFFFu ? idx[u] : begin) * a.M;
          for (uint32_t m = 0; m < a.M; ++m) acc[u] = __fadd_rn(acc[u], lut[m * 256 + code[m]]);
        }
      }
      float score[U];
#pragma unroll
      for (uint32_t u = 0; u < U; ++u) score[u] = -__fmaf_rn(-2.f, acc[u], __fadd_rn(base_term, rn[u]));
      if (!sel.have_thr) {  // (uniform) no threshold yet: offer one vector per thread at a time so that the first
                            // compaction — which installs it — ranks a few hundred keys, not a whole tile
#pragma unroll
        for (uint32_t u = 0; u < U; ++u) {
          if (idx[u] != 0xFFFFFFFFu) sel.offer(score[u], id[u]);
          sel.after_offers(u + 1 < U ? kScanThreads : TILE, tid, kScanThreads);
        }
      } else {
#pragma unroll
        for (uint32_t u = 0; u < U; ++u)
          if (idx[u] != 0xFFFFFFFFu) sel.offer(score[u], id[u]);
        sel.after_offers(TILE, tid, kScanThreads);
      }
    }
  }
  sel.compact(tid, kScanThreads);  // survivors sorted best-first in the current half
  const uint32_t held = *cnt;
  const uint64_t* fin = keys + sel.cur * a.cap;
  uint64_t* out = a.partial + ((size_t)q * a.G + g) * a.R;
  for (uint32_t j = tid; j < a.R; j += kScanThreads) out[j] = (j < held) ? fin[j] : 0ull;
}

// G x R candidate keys per query -> the R best as (u32 id, distance), by rank counting (total <= kMergeMax keys)
constexpr uint32_t kMergeMax = 1024;
__global__ void __launch_bounds__(kMergeMax / 2)
ivfpq_merge_kernel(const uint64_t* __restrict__ partial, uint32_t total, uint32_t R, uint32_t* __restrict__ out_ids,
                   float* __restrict__ out_dist) {
  __shared__ __align__(16) uint64_t mk[kMergeMax];
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  uint64_t key[2];
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const uint32_t j = tid + (uint32_t)u * (kMergeMax / 2);
    key[u] = j < total ? partial[(size_t)q * total + j] : 0ull;
    mk[j] = key[u];
  }
  const uint32_t valid = (uint32_t)__syncthreads_count(key[0] != 0ull) + (uint32_t)__syncthreads_count(key[1] != 0ull);
  // two own keys against every held key, two held keys per (broadcast) 16-byte load: rank counting is bound by
  // shared-memory wavefronts = warps x loads, so few warps with several keys each
  uint32_t rank[2] = {0, 0};
  const uint32_t pairs = (total + 1u) >> 1;
  const ulonglong2* mk2 = reinterpret_cast<const ulonglong2*>(mk);
#pragma unroll 4
  for (uint32_t i = 0; i < pairs; ++i) {
    const ulonglong2 o = mk2[i];
#pragma unroll
    for (int u = 0; u < 2; ++u) rank[u] += (o.x > key[u] ? 1u : 0u) + (o.y > key[u] ? 1u : 0u);
    if ((key[0] | key[1]) == 0ull) break;  // nothing to rank (a group's list is zero-padded at ITS end, so key[1] may be
                                           // real while key[0] is padding)
  }
#pragma unroll
  for (int u = 0; u < 2; ++u)
    if (key[u] != 0ull && rank[u] < R) {
      out_ids[(size_t)q * R + rank[u]] = key_row(key[u]);
      if (out_dist) out_dist[(size_t)q * R + rank[u]] = -key_score(key[u]);
    }
  for (uint32_t j = valid + tid; j < R; j += kMergeMax / 2) {  // fewer than R candidates exist
    out_ids[(size_t)q * R + j] = 0xFFFFFFFFu;
    if (out_dist) out_dist[(size_t)q * R + j] = 1e30f;
  }
}

// sorted_lists[pos] = list that owns grouped position pos (inverse of list_offsets_kernel)
__global__ void expand_lists_kernel(const uint32_t* __restrict__ list_off, uint32_t nlist, uint32_t n,
                                    uint32_t* __restrict__ sorted_lists) {
  const uint32_t pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n) return;
  uint32_t lo = 0, hi = nlist;  // last l with list_off[l] <= pos
  while (lo + 1 < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (list_off[mid] <= pos) lo = mid;
    else hi = mid;
  }
  sorted_lists[pos] = lo;
}

// |c_list + reconstruction|^2 of every grouped position (thread per vector, fixed order)
__global__ void recon_norms_kernel(const uint8_t* __restrict__ codes, const uint32_t* __restrict__ sorted_lists,
                                   const float* __restrict__ coarse, const float* __restrict__ codebooks, uint32_t n, uint32_t d,
                                   uint32_t M, uint32_t dsub, float* __restrict__ rnorm) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* c = coarse + (size_t)sorted_lists[i] * d;
  float s = 0.f;
  for (uint32_t m = 0; m < M; ++m) {
    const float* cw = codebooks + ((size_t)m * 256 + codes[(size_t)i * M + m]) * dsub;
    for (uint32_t t = 0; t < dsub; ++t) {
      const float v = __fadd_rn(c[m * dsub + t], cw[t]);
      s = __fmaf_rn(v, v, s);
    }
  }
  rnorm[i] = s;
}

// merged (u64 id, score = -dist) -> (u32 id, dist); empty slots as the refine kernel expects them
__global__ void finalize_candidates_kernel(const uint64_t* __restrict__ ids64, const float* __restrict__ scores, size_t total,
                                           uint32_t* __restrict__ out_ids, float* __restrict__ out_dist) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= total) return;
  const bool ok = ids64[e] != ~0ull;
  out_ids[e] = ok ? (uint32_t)ids64[e] : 0xFFFFFFFFu;
  if (out_dist) out_dist[e] = ok ? -scores[e] : 1e30f;
}

}  // namespace

// =================================================================================================================
// C ABI
// =================================================================================================================
namespace {

void ivf_free(nvdb_b200_ivfpq* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaFree(h->coarse);
  cudaFree(h->coarse_hn);
  cudaFree(h->codebooks);
  cudaFree(h->cb_hn);
  cudaFree(h->codes);
  cudaFree(h->ids);
  cudaFree(h->rnorm);
  cudaFree(h->list_off);
  cudaFree(h->rot_t);
  cudaFree(h->zeros);
  for (auto& g : h->graphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  for (void* w : h->ws) cudaFree(w);
  if (h->h_pin) cudaFreeHost(h->h_pin);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

struct DevBuf {  // build-time scratch, freed on scope exit
  void* p = nullptr;
  ~DevBuf() { cudaFree(p); }
  cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
  template <typename T>
  T* as() { return reinterpret_cast<T*>(p); }
};

// out[n x d] = x[n x d] R  (rot_t = R^T, so this is the "row . row" product of the tile GEMM)
cudaError_t rotate_rows(const nvdb_b200_ivfpq* h, const float* x, uint32_t n, float* out, cudaStream_t st) {
  return dot_bias_scores_launch(x, h->d, n, h->rot_t, h->d, h->zeros, h->d, h->d, out, st);
}

// OPQ (Ge, He, Ke, Sun: "Optimized Product Quantization", non-parametric solution), what the reference enables with
// USE_OPQ=1 (faiss::OPQMatrix, apps/nvdb_ivfpq_build.cpp:78-91): alternate (a) PQ of the rotated training rows X R,
// (b) R = argmin |X R - Y|_F over orthogonal R with Y the PQ reconstruction = polar factor of X^T Y. All on the device:
// rotation and X^T Y by the tile GEMM, the polar factor by Newton-Schulz (kmeans.cu), PQ by the k-means above, warm
// started after the first round. x: [nt x d] training rows (unchanged). Leaves h->rot_t = R^T.
int opq_train(nvdb_b200_ivfpq* h, const float* x, uint32_t nt, int opq_niter, uint64_t seed, void* assign_ws, uint32_t* assign) {
  const uint32_t d = h->d, M = h->M, dsub = h->dsub;
  cudaStream_t st = h->stream;
  const size_t nd = (size_t)nt * d, dd = (size_t)d * d;
  DevBuf xr, y, xt, yt, R, cb, cbhn, pws;
  ICK(xr.alloc(nd * 4), "cudaMalloc OPQ rotated rows");
  ICK(y.alloc(nd * 4), "cudaMalloc OPQ reconstruction");
  ICK(xt.alloc(nd * 4), "cudaMalloc OPQ X^T");
  ICK(yt.alloc(nd * 4), "cudaMalloc OPQ Y^T");
  ICK(R.alloc(dd * 4), "cudaMalloc OPQ rotation");
  ICK(cb.alloc((size_t)M * 256 * dsub * 4), "cudaMalloc OPQ codebooks");
  ICK(cbhn.alloc((size_t)M * 256 * 4), "cudaMalloc OPQ codebook norms");
  ICK(pws.alloc((3 * dd + 1) * 4), "cudaMalloc polar workspace");
  ICK(cudaMalloc(&h->rot_t, dd * 4), "cudaMalloc rotation");
  identity_kernel<<<256, 256, 0, st>>>(h->rot_t, d);
  count_launch();
  ICK(transpose_launch(x, nt, d, xt.as<float>(), st), "transpose X");
  const int grid = h->num_sms * 8;
  for (int it = 0; it < opq_niter; ++it) {
    ICK(rotate_rows(h, x, nt, xr.as<float>(), st), "OPQ rotate");
    for (uint32_t m = 0; m < M; ++m) {
      float* cbm = cb.as<float>() + (size_t)m * 256 * dsub;
      ICK(kmeans_train(xr.as<float>() + (size_t)m * dsub, d, nt, dsub, 256, it == 0 ? 6 : 1, seed + 1000 + m, cbm, h->num_sms, st,
                       it > 0),
          "OPQ PQ k-means");
    }
    ICK(half_norms_launch(cb.as<float>(), dsub, M * 256, dsub, cbhn.as<float>(), st), "OPQ codebook norms");
    for (uint32_t m = 0; m < M; ++m) {
      const float* cbm = cb.as<float>() + (size_t)m * 256 * dsub;
      ICK(nearest_centroid_launch(xr.as<float>() + (size_t)m * dsub, d, nt, cbm, dsub, cbhn.as<float>() + (size_t)m * 256, 256,
                                  dsub, assign, assign_ws, h->num_sms, st),
          "OPQ PQ assignment");
      gather_codewords_kernel<<<grid, 256, 0, st>>>(assign, nt, cbm, dsub, y.as<float>() + (size_t)m * dsub, d);
      count_launch();
    }
    ICK(transpose_launch(y.as<float>(), nt, d, yt.as<float>(), st), "transpose Y");
    // X^T Y  (d x d, inner dimension nt)
    ICK(dot_bias_scores_launch(xt.as<float>(), nt, d, yt.as<float>(), nt, h->zeros, d, nt, R.as<float>(), st), "X^T Y");
    ICK(polar_factor_launch(R.as<float>(), d, 40, pws.as<float>(), h->zeros, st), "polar factor");
    ICK(transpose_launch(R.as<float>(), d, d, h->rot_t, st), "transpose R");
  }
  ICK(cudaStreamSynchronize(st), "OPQ training");
  return NVDB_B200_OK;
}

int build_impl(nvdb_b200_ivfpq* h, const void* payload, const float* scales, uint32_t dtype, uint64_t ntrain, int niter,
               int opq_niter, uint64_t seed) {
  const uint32_t d = h->d, M = h->M, dsub = h->dsub, nlist = h->nlist;
  const uint32_t n = (uint32_t)h->n;
  cudaStream_t st = h->stream;
  const uint32_t nt = (uint32_t)(ntrain < h->n ? ntrain : h->n);
  ICK(cudaMalloc(&h->coarse, (size_t)nlist * d * 4), "cudaMalloc coarse centroids");
  ICK(cudaMalloc(&h->coarse_hn, (size_t)nlist * 4), "cudaMalloc coarse norms");
  ICK(cudaMalloc(&h->codebooks, (size_t)M * 256 * dsub * 4), "cudaMalloc codebooks");
  ICK(cudaMalloc(&h->cb_hn, (size_t)M * 256 * 4), "cudaMalloc codebook norms");
  ICK(cudaMalloc(&h->codes, (size_t)n * M), "cudaMalloc codes");
  ICK(cudaMalloc(&h->ids, (size_t)n * 4), "cudaMalloc ids");
  ICK(cudaMalloc(&h->rnorm, (size_t)n * 4), "cudaMalloc reconstruction norms");
  ICK(cudaMalloc(&h->list_off, (size_t)(nlist + 1) * 4), "cudaMalloc list offsets");
  ICK(cudaMalloc(&h->zeros, (size_t)(d > 256 ? d : 256) * 4), "cudaMalloc zeros");
  ICK(cudaMemsetAsync(h->zeros, 0, (size_t)(d > 256 ? d : 256) * 4, st), "memset zeros");

  const uint32_t chunk = 131072;
  const uint32_t xrows = nt > chunk ? nt : chunk;
  DevBuf x, assign, ws, lists, codes_tmp;
  ICK(x.alloc((size_t)xrows * d * 4), "cudaMalloc f32 rows");
  ICK(assign.alloc((size_t)xrows * 4), "cudaMalloc assignments");
  size_t wsb = nearest_centroid_ws_bytes(xrows, nlist, h->num_sms);
  const size_t wsb2 = nearest_centroid_ws_bytes(xrows, 256, h->num_sms);
  if (wsb2 > wsb) wsb = wsb2;
  ICK(ws.alloc(wsb), "cudaMalloc assignment workspace");
  ICK(lists.alloc((size_t)n * 4), "cudaMalloc list ids");
  ICK(codes_tmp.alloc((size_t)n * M), "cudaMalloc codes (row order)");
  const int fill_grid = h->num_sms * 8;

  // ---- train: (OPQ rotation,) coarse centroids on the first nt rows, PQ codebooks on their residuals
  rows_to_f32_kernel<<<fill_grid, 256, 0, st>>>(dtype, payload, scales, 0, nt, d, x.as<float>());
  count_launch();
  DevBuf xrot;  // rotated rows of the current chunk (OPQ only)
  if (opq_niter > 0) {
    const int rc = opq_train(h, x.as<float>(), nt, opq_niter, seed, ws.p, assign.as<uint32_t>());
    if (rc) return rc;
    ICK(xrot.alloc((size_t)xrows * d * 4), "cudaMalloc rotated rows");
    ICK(rotate_rows(h, x.as<float>(), nt, xrot.as<float>(), st), "rotate training rows");
    ICK(cudaMemcpyAsync(x.p, xrot.p, (size_t)nt * d * 4, cudaMemcpyDeviceToDevice, st), "copy rotated rows");
  }
  ICK(kmeans_train(x.as<float>(), d, nt, d, nlist, niter, seed, h->coarse, h->num_sms, st), "coarse k-means");
  ICK(half_norms_launch(h->coarse, d, nlist, d, h->coarse_hn, st), "coarse norms");
  ICK(nearest_centroid_launch(x.as<float>(), d, nt, h->coarse, d, h->coarse_hn, nlist, d, assign.as<uint32_t>(), ws.p,
                              h->num_sms, st),
      "coarse assignment (train)");
  subtract_centroid_kernel<<<fill_grid, 256, 0, st>>>(x.as<float>(), nt, d, assign.as<uint32_t>(), h->coarse);
  count_launch();
  for (uint32_t m = 0; m < M; ++m)
    ICK(kmeans_train(x.as<float>() + (size_t)m * dsub, d, nt, dsub, 256, niter, seed + 1 + m,
                     h->codebooks + (size_t)m * 256 * dsub, h->num_sms, st),
        "PQ k-means");
  ICK(half_norms_launch(h->codebooks, dsub, M * 256, dsub, h->cb_hn, st), "codebook norms");

  // ---- encode every row: list id + M code bytes
  for (uint64_t r0 = 0; r0 < n; r0 += chunk) {
    const uint32_t nr = (uint32_t)((n - r0) < chunk ? (n - r0) : chunk);
    rows_to_f32_kernel<<<fill_grid, 256, 0, st>>>(dtype, payload, scales, r0, nr, d, x.as<float>());
    count_launch();
    if (h->rot_t) {
      ICK(rotate_rows(h, x.as<float>(), nr, xrot.as<float>(), st), "rotate rows");
      ICK(cudaMemcpyAsync(x.p, xrot.p, (size_t)nr * d * 4, cudaMemcpyDeviceToDevice, st), "copy rotated rows");
    }
    ICK(nearest_centroid_launch(x.as<float>(), d, nr, h->coarse, d, h->coarse_hn, nlist, d, lists.as<uint32_t>() + r0, ws.p,
                                h->num_sms, st),
        "coarse assignment");
    subtract_centroid_kernel<<<fill_grid, 256, 0, st>>>(x.as<float>(), nr, d, lists.as<uint32_t>() + r0, h->coarse);
    count_launch();
    for (uint32_t m = 0; m < M; ++m) {
      ICK(nearest_centroid_launch(x.as<float>() + (size_t)m * dsub, d, nr, h->codebooks + (size_t)m * 256 * dsub, dsub,
                                  h->cb_hn + (size_t)m * 256, 256, dsub, assign.as<uint32_t>(), ws.p, h->num_sms, st),
          "PQ assignment");
      write_code_kernel<<<(nr + 255) / 256, 256, 0, st>>>(assign.as<uint32_t>(), nr, M, m, codes_tmp.as<uint8_t>() + r0 * M);
      count_launch();
    }
  }
  // ---- group by list (stable radix sort on the list id keeps ascending row ids inside a list)
  DevBuf iota, sorted_lists, tmp;
  ICK(iota.alloc((size_t)n * 4), "cudaMalloc iota");
  ICK(sorted_lists.alloc((size_t)n * 4), "cudaMalloc sorted list ids");
  iota_kernel<<<(n + 255) / 256, 256, 0, st>>>(iota.as<uint32_t>(), n);
  count_launch();
  size_t tmp_bytes = 0;
  int end_bit = 1;
  while ((1u << end_bit) < nlist && end_bit < 32) ++end_bit;
  ICK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, lists.as<uint32_t>(), sorted_lists.as<uint32_t>(),
                                      iota.as<uint32_t>(), h->ids, (int)n, 0, end_bit, st),
      "radix sort (size query)");
  ICK(tmp.alloc(tmp_bytes), "cudaMalloc sort scratch");
  ICK(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, lists.as<uint32_t>(), sorted_lists.as<uint32_t>(),
                                      iota.as<uint32_t>(), h->ids, (int)n, 0, end_bit, st),
      "radix sort");
  list_offsets_kernel<<<(nlist + 1 + 255) / 256, 256, 0, st>>>(sorted_lists.as<uint32_t>(), n, nlist, h->list_off);
  gather_codes_kernel<<<fill_grid, 256, 0, st>>>(codes_tmp.as<uint8_t>(), h->ids, n, M, h->codes);
  recon_norms_kernel<<<(n + 255) / 256, 256, 0, st>>>(h->codes, sorted_lists.as<uint32_t>(), h->coarse, h->codebooks, n, d, M,
                                                      dsub, h->rnorm);
  count_launch(3);
  ICK(cudaGetLastError(), "ivfpq build kernels");
  ICK(cudaStreamSynchronize(st), "ivfpq build");
  return NVDB_B200_OK;
}

// candidates of nq device-resident queries -> cand_ids [nq x R] (u32, device), cand_dist optional (device)
int search_device_impl(nvdb_b200_ivfpq* h, const float* dq, uint32_t nq, uint32_t nprobe, uint32_t R, uint32_t* cand_ids,
                       float* cand_dist, cudaStream_t st) {
  if (nq == 0) return NVDB_B200_OK;
  if (nprobe == 0 || R == 0 || R > 2048) return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search: need nprobe >= 1 and 1 <= R <= 2048");
  if (nprobe > h->nlist) nprobe = h->nlist;
  const uint32_t bufn = next_pow2(h->nlist);
  if (((size_t)bufn + kProbeBuf) * 8 > (size_t)h->max_smem)
    return ifail(h, NVDB_B200_ERR_UNSUPPORTED, "nlist too large for the probe selector");
  int rc;
  if ((rc = iws(h, IWS_SCORES, (size_t)nq * h->nlist * 4))) return rc;
  if ((rc = iws(h, IWS_PROBES, (size_t)nq * nprobe * 4))) return rc;
  // probe groups per query: enough CTAs to fill the machine for a few queries, one CTA per query for big batches
  uint32_t G = ((uint32_t)h->num_sms * 2u + nq - 1u) / nq;
  if (G > nprobe) G = nprobe;
  // the merge ranks G x R keys by counting (cost ~ (G R)^2): a few hundred keys (else >= 8 groups; > kMergeMax keys take
  // the bitonic merge kernel). Fewer groups = more lists per scan CTA, which costs less than a quadratic merge.
  const uint32_t g_merge = 640u / R > 8u ? 640u / R : 8u;
  if (G > g_merge) G = g_merge;
  if (G < 1u) G = 1u;
  if ((rc = iws(h, IWS_PARTIAL, (size_t)nq * G * R * 8))) return rc;
  if ((rc = iws(h, IWS_LUT, (size_t)nq * h->M * 1024))) return rc;
  if ((rc = iws(h, IWS_IDS64, (size_t)nq * R * 8))) return rc;
  if ((rc = iws(h, IWS_SC, (size_t)nq * R * 4))) return rc;
  float* scores = (float*)h->ws[IWS_SCORES];
  uint32_t* probes = (uint32_t*)h->ws[IWS_PROBES];
  if (h->rot_t) {  // OPQ: the index lives in the rotated space (distances are rotation invariant)
    if ((rc = iws(h, IWS_QROT, (size_t)nq * h->d * 4))) return rc;
    ICK(rotate_rows(h, dq, nq, (float*)h->ws[IWS_QROT], st), "rotate queries");
    dq = (const float*)h->ws[IWS_QROT];
  }
  ICK(dot_bias_scores_launch(dq, h->d, nq, h->coarse, h->d, h->coarse_hn, h->nlist, h->d, scores, st), "coarse scores");
  const size_t probe_smem = ((size_t)bufn + kProbeBuf) * 8;
  if (probe_smem > 48u * 1024u)
    ICK(cudaFuncSetAttribute(select_probes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)probe_smem), "smem attr");
  select_probes_kernel<<<nq, kProbeThreads, probe_smem, st>>>(scores, h->nlist, bufn, nprobe, probes);
  count_launch();
  for (uint32_t q0 = 0; q0 < nq; q0 += 65535u) {
    const uint32_t nqp = nq - q0 < 65535u ? nq - q0 : 65535u;
    ivfpq_lut_kernel<<<dim3(h->M, nqp), 256, 0, st>>>(dq + (size_t)q0 * h->d, h->codebooks, h->d, h->M, h->dsub,
                                                      (float*)h->ws[IWS_LUT] + (size_t)q0 * h->M * 256);
    count_launch();
  }
  IvfScanArgs a;
  a.queries = dq;
  a.lut = (const float*)h->ws[IWS_LUT];
  a.coarse = h->coarse;
  a.codebooks = h->codebooks;
  a.codes = h->codes;
  a.ids = h->ids;
  a.rnorm = h->rnorm;
  a.G = G;
  a.list_off = h->list_off;
  a.probes = probes;
  a.partial = (uint64_t*)h->ws[IWS_PARTIAL];
  a.d = h->d;
  a.M = h->M;
  a.dsub = h->dsub;
  a.nprobe = nprobe;
  a.R = R;
  a.cap = (R + 2 * kScanThreads * kScanVecs + 1u) & ~1u;  // even: both halves of the double buffer 16-byte aligned
  const size_t smem = (size_t)h->M * 1024 + (size_t)((h->d + 3u) & ~3u) * 4 + (size_t)a.cap * 16 + 8 + 4 * (kScanThreads / 32) + 8;
  if (smem > (size_t)h->max_smem) return ifail(h, NVDB_B200_ERR_UNSUPPORTED, "PQ look-up table does not fit shared memory");
  ICK(cudaFuncSetAttribute(ivfpq_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "smem attr");
  for (uint32_t q0 = 0; q0 < nq; q0 += 65535u) {  // grid.y limit
    const uint32_t nqp = nq - q0 < 65535u ? nq - q0 : 65535u;
    IvfScanArgs b = a;
    b.queries = dq + (size_t)q0 * h->d;
    b.probes = probes + (size_t)q0 * nprobe;
    b.partial = a.partial + (size_t)q0 * G * R;
    b.lut = a.lut + (size_t)q0 * h->M * 256;
    ivfpq_scan_kernel<<<dim3(G, nqp), kScanThreads, smem, st>>>(b);
    count_launch();
  }
  if (G * R <= kMergeMax) {
    ivfpq_merge_kernel<<<nq, kMergeMax / 2, 0, st>>>(a.partial, G * R, R, cand_ids, cand_dist);
    count_launch();
    ICK(cudaGetLastError(), "ivfpq search kernels");
    return NVDB_B200_OK;
  }
  ICK(merge_partials_launch(a.partial, G, nq, R, 0, (uint64_t*)h->ws[IWS_IDS64], (float*)h->ws[IWS_SC], st),
      "candidate merge");
  const size_t total = (size_t)nq * R;
  finalize_candidates_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>((const uint64_t*)h->ws[IWS_IDS64],
                                                                              (const float*)h->ws[IWS_SC], total, cand_ids,
                                                                              cand_dist);
  count_launch();
  ICK(cudaGetLastError(), "ivfpq search kernels");
  return NVDB_B200_OK;
}

int pin(nvdb_b200_ivfpq* h, size_t bytes) {
  if (h->h_pin_bytes >= bytes) return NVDB_B200_OK;
  if (h->h_pin) cudaFreeHost(h->h_pin);
  h->h_pin = nullptr;
  h->h_pin_bytes = 0;
  h->ws_gen++;
  ICK(cudaMallocHost(&h->h_pin, bytes), "cudaMallocHost");
  h->h_pin_bytes = bytes;
  return NVDB_B200_OK;
}

constexpr uint32_t kGraphMaxQ = 8;

// Runs `enqueue` (the asynchronous copies and kernels of one host call on h->stream, no synchronisation inside) either
// directly or, for a small shape seen twice in a row, as a captured graph that later calls replay with one launch.
template <class F>
int run_or_replay(nvdb_b200_ivfpq* h, const uint64_t (&key)[8], uint32_t nq, F&& enqueue) {
  if (h->graph_mode < 0) {
    const char* e = getenv("NVDB_B200_IVF_GRAPH");
    h->graph_mode = (e && e[0] == '0') ? 0 : 1;
  }
  if (!h->graph_mode || nq > kGraphMaxQ) return enqueue();
  for (auto& g : h->graphs)
    if (g.exec && g.ws_gen == h->ws_gen && !memcmp(g.key, key, sizeof key)) {
      ICK(cudaGraphLaunch(g.exec, h->stream), "ivfpq graph launch");
      count_launch(g.launches);
      h->graph_replays++;
      return NVDB_B200_OK;
    }
  if (memcmp(h->last_key, key, sizeof key)) {  // first sighting: plain run (grows and warms every workspace)
    memcpy(h->last_key, key, sizeof key);
    return enqueue();
  }
  const uint64_t gen0 = h->ws_gen;
  const uint64_t l0 = nvdb_b200_launch_count();
  if (cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) {
    cudaGetLastError();
    h->graph_mode = 0;
    return enqueue();
  }
  const int rc = enqueue();
  cudaGraph_t graph = nullptr;
  const cudaError_t ee = cudaStreamEndCapture(h->stream, &graph);
  const uint32_t nl = (uint32_t)(nvdb_b200_launch_count() - l0);
  cudaGraphExec_t exec = nullptr;
  if (rc == NVDB_B200_OK && ee == cudaSuccess && graph && h->ws_gen == gen0 &&
      cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess) {
    cudaGraphDestroy(graph);
    auto& g = h->graphs[h->graph_next++ % 4u];
    if (g.exec) cudaGraphExecDestroy(g.exec);
    memcpy(g.key, key, sizeof key);
    g.ws_gen = h->ws_gen;
    g.exec = exec;
    g.launches = nl;
    ICK(cudaGraphLaunch(g.exec, h->stream), "ivfpq graph launch");  // the captured work has not run yet
    return NVDB_B200_OK;
  }
  if (graph) cudaGraphDestroy(graph);
  cudaGetLastError();
  h->graph_mode = 0;  // not capturable here: stay on the plain path from now on
  if (rc != NVDB_B200_OK) return rc;
  return enqueue();
}

}  // namespace

extern "C" {

const char* nvdb_b200_ivfpq_last_error(const nvdb_b200_ivfpq* h) { return h ? h->err.c_str() : t_ivf_err.c_str(); }

int nvdb_b200_ivfpq_build(nvdb_b200_index* base, uint32_t nlist, uint32_t pq_m, uint32_t pq_bits, uint64_t ntrain,
                          int niter, uint64_t seed, nvdb_b200_ivfpq** out) {
  return nvdb_b200_ivfpq_build_opq(base, nlist, pq_m, pq_bits, ntrain, niter, 0, seed, out);
}

int nvdb_b200_ivfpq_rotation(const nvdb_b200_ivfpq* h, float* rotation, int* has_rotation) {
  if (!h) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq handle is NULL");
  if (has_rotation) *has_rotation = h->rot_t ? 1 : 0;
  if (rotation && h->rot_t) {
    ICK(cudaSetDevice(h->device), "cudaSetDevice");
    std::vector<float> rt((size_t)h->d * h->d);
    ICK(cudaMemcpy(rt.data(), h->rot_t, rt.size() * 4, cudaMemcpyDeviceToHost), "export rotation");
    for (uint32_t i = 0; i < h->d; ++i)
      for (uint32_t j = 0; j < h->d; ++j) rotation[(size_t)i * h->d + j] = rt[(size_t)j * h->d + i];
  }
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_set_rotation(nvdb_b200_ivfpq* h, const float* rotation) {
  if (!h) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq handle is NULL");
  std::lock_guard<std::mutex> lock(h->mu);
  ICK(cudaSetDevice(h->device), "cudaSetDevice");
  if (!rotation) {
    cudaFree(h->rot_t);
    h->rot_t = nullptr;
    return NVDB_B200_OK;
  }
  std::vector<float> rt((size_t)h->d * h->d);
  for (uint32_t i = 0; i < h->d; ++i)
    for (uint32_t j = 0; j < h->d; ++j) rt[(size_t)j * h->d + i] = rotation[(size_t)i * h->d + j];
  if (!h->rot_t) ICK(cudaMalloc(&h->rot_t, rt.size() * 4), "cudaMalloc rotation");
  ICK(cudaMemcpy(h->rot_t, rt.data(), rt.size() * 4, cudaMemcpyHostToDevice), "H2D rotation");
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_build_opq(nvdb_b200_index* base, uint32_t nlist, uint32_t pq_m, uint32_t pq_bits, uint64_t ntrain,
                              int niter, int opq_niter, uint64_t seed, nvdb_b200_ivfpq** out) {
  nvdb_b200_ivfpq* h = nullptr;
  if (!out) return ifail(nullptr, NVDB_B200_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (!base) return ifail(nullptr, NVDB_B200_ERR_ARG, "base index is NULL");
  uint64_t n = 0, row_offset = 0;
  uint32_t dim = 0, dtype = 0;
  int device = 0;
  const void* payload = nullptr;
  const float* scales = nullptr;
  if (nvdb_b200_index_info(base, &n, &dim, &dtype, &device, &row_offset, &payload, &scales) != NVDB_B200_OK)
    return ifail(nullptr, NVDB_B200_ERR_ARG, "cannot read the base index");
  if (pq_bits != 8) return ifail(nullptr, NVDB_B200_ERR_UNSUPPORTED, "PQ_BITS must be 8");
  if (pq_m == 0 || dim % pq_m != 0) return ifail(nullptr, NVDB_B200_ERR_ARG, "PQ_M must divide dim");  // nvdb_ivfpq_build.cpp:54
  if (nlist == 0 || n == 0) return ifail(nullptr, NVDB_B200_ERR_EMPTY, "empty base or nlist = 0");
  if (n > 0xFFFFFFF0ull) return ifail(nullptr, NVDB_B200_ERR_UNSUPPORTED, "more than 2^32-16 rows per IVF-PQ shard");
  const uint64_t nt = ntrain == 0 || ntrain > n ? n : ntrain;
  if (nt < nlist || nt < 256) return ifail(nullptr, NVDB_B200_ERR_ARG, "training set smaller than nlist / 256 codewords");
  if (niter <= 0) niter = 25;  // the clustering default of the library the reference calls
  h = new nvdb_b200_ivfpq();
  h->device = device;
  h->d = dim;
  h->nlist = nlist;
  h->M = pq_m;
  h->dsub = dim / pq_m;
  h->n = n;
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  cudaDeviceProp prop;
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) {
    const int rc = ifail(nullptr, NVDB_B200_ERR_CUDA, std::string("ivfpq init: ") + cudaGetErrorString(e));
    ivf_free(h);
    return rc;
  }
  h->num_sms = prop.multiProcessorCount;
  h->max_smem = (int)prop.sharedMemPerBlockOptin;
  const int rc = build_impl(h, payload, scales, dtype, nt, niter, opq_niter < 0 ? 0 : opq_niter, seed);
  if (rc != NVDB_B200_OK) {
    t_ivf_err = h->err;
    ivf_free(h);
    return rc;
  }
  *out = h;
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_destroy(nvdb_b200_ivfpq* h) {
  ivf_free(h);
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_info(const nvdb_b200_ivfpq* h, uint64_t* n, uint32_t* dim, uint32_t* nlist, uint32_t* pq_m,
                         uint64_t* index_bytes) {
  if (!h) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq handle is NULL");
  if (n) *n = h->n;
  if (dim) *dim = h->d;
  if (nlist) *nlist = h->nlist;
  if (pq_m) *pq_m = h->M;
  if (index_bytes)
    *index_bytes = (uint64_t)h->nlist * h->d * 4 + (uint64_t)h->M * 256 * h->dsub * 4 + h->n * h->M + h->n * 8 +
                   (uint64_t)(h->nlist + 1) * 4;
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_import(int device, uint32_t dim, uint32_t nlist, uint32_t pq_m, uint64_t n, const float* coarse,
                           const float* codebooks, const uint32_t* list_off, const uint32_t* ids, const uint8_t* codes,
                           nvdb_b200_ivfpq** out) {
  nvdb_b200_ivfpq* h = nullptr;
  if (!out) return ifail(nullptr, NVDB_B200_ERR_ARG, "out is NULL");
  *out = nullptr;
  if (!coarse || !codebooks || !list_off || !ids || !codes) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: NULL array");
  if (pq_m == 0 || dim == 0 || dim % pq_m != 0 || nlist == 0 || n == 0 || n > 0xFFFFFFF0ull)
    return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: bad shape");
  if (list_off[0] != 0 || list_off[nlist] != n) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: list offsets do not cover n");
  for (uint32_t l = 0; l < nlist; ++l)
    if (list_off[l] > list_off[l + 1]) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: list offsets not ascending");
  {  // the selectors rank candidates by (distance, id) keys and assume them unique: every row id once, all below n
    std::vector<uint64_t> seen((n + 63) / 64, 0ull);
    for (uint64_t i = 0; i < n; ++i) {
      const uint32_t id = ids[i];
      if (id >= n) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: row id out of range");
      if (seen[id >> 6] & (1ull << (id & 63))) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq import: duplicate row id");
      seen[id >> 6] |= 1ull << (id & 63);
    }
  }
  h = new nvdb_b200_ivfpq();
  h->device = device;
  h->d = dim;
  h->nlist = nlist;
  h->M = pq_m;
  h->dsub = dim / pq_m;
  h->n = n;
  auto bail = [&](int rc) {
    t_ivf_err = h->err;
    ivf_free(h);
    return rc;
  };
  auto body = [&]() -> int {
    ICK(cudaSetDevice(device), "cudaSetDevice");
    ICK(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    cudaDeviceProp prop;
    ICK(cudaGetDeviceProperties(&prop, device), "cudaGetDeviceProperties");
    h->num_sms = prop.multiProcessorCount;
    h->max_smem = (int)prop.sharedMemPerBlockOptin;
    const size_t cb = (size_t)nlist * dim * 4, pb = (size_t)pq_m * 256 * h->dsub * 4;
    ICK(cudaMalloc(&h->coarse, cb), "cudaMalloc coarse centroids");
    ICK(cudaMalloc(&h->coarse_hn, (size_t)nlist * 4), "cudaMalloc coarse norms");
    ICK(cudaMalloc(&h->codebooks, pb), "cudaMalloc codebooks");
    ICK(cudaMalloc(&h->cb_hn, (size_t)pq_m * 256 * 4), "cudaMalloc codebook norms");
    ICK(cudaMalloc(&h->codes, (size_t)n * pq_m), "cudaMalloc codes");
    ICK(cudaMalloc(&h->ids, (size_t)n * 4), "cudaMalloc ids");
    ICK(cudaMalloc(&h->rnorm, (size_t)n * 4), "cudaMalloc reconstruction norms");
    ICK(cudaMalloc(&h->list_off, (size_t)(nlist + 1) * 4), "cudaMalloc list offsets");
    ICK(cudaMalloc(&h->zeros, (size_t)(dim > 256 ? dim : 256) * 4), "cudaMalloc zeros");
    ICK(cudaMemset(h->zeros, 0, (size_t)(dim > 256 ? dim : 256) * 4), "memset zeros");
    ICK(cudaMemcpy(h->coarse, coarse, cb, cudaMemcpyHostToDevice), "H2D coarse");
    ICK(cudaMemcpy(h->codebooks, codebooks, pb, cudaMemcpyHostToDevice), "H2D codebooks");
    ICK(cudaMemcpy(h->codes, codes, (size_t)n * pq_m, cudaMemcpyHostToDevice), "H2D codes");
    ICK(cudaMemcpy(h->ids, ids, (size_t)n * 4, cudaMemcpyHostToDevice), "H2D ids");
    ICK(cudaMemcpy(h->list_off, list_off, (size_t)(nlist + 1) * 4, cudaMemcpyHostToDevice), "H2D list offsets");
    DevBuf sl;
    ICK(sl.alloc((size_t)n * 4), "cudaMalloc list ids");
    ICK(half_norms_launch(h->coarse, dim, nlist, dim, h->coarse_hn, h->stream), "coarse norms");
    ICK(half_norms_launch(h->codebooks, h->dsub, pq_m * 256, h->dsub, h->cb_hn, h->stream), "codebook norms");
    const uint32_t n32 = (uint32_t)n;
    expand_lists_kernel<<<(n32 + 255) / 256, 256, 0, h->stream>>>(h->list_off, nlist, n32, sl.as<uint32_t>());
    recon_norms_kernel<<<(n32 + 255) / 256, 256, 0, h->stream>>>(h->codes, sl.as<uint32_t>(), h->coarse, h->codebooks, n32, dim,
                                                                 pq_m, h->dsub, h->rnorm);
    count_launch(2);
    ICK(cudaGetLastError(), "ivfpq import kernels");
    ICK(cudaStreamSynchronize(h->stream), "ivfpq import");
    return NVDB_B200_OK;
  };
  const int rc = body();
  if (rc != NVDB_B200_OK) return bail(rc);
  *out = h;
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_export(const nvdb_b200_ivfpq* h, float* coarse, float* codebooks, uint32_t* list_off, uint32_t* ids,
                           uint8_t* codes) {
  if (!h) return ifail(nullptr, NVDB_B200_ERR_ARG, "ivfpq handle is NULL");
  ICK(cudaSetDevice(h->device), "cudaSetDevice");
  if (coarse) ICK(cudaMemcpy(coarse, h->coarse, (size_t)h->nlist * h->d * 4, cudaMemcpyDeviceToHost), "export coarse");
  if (codebooks)
    ICK(cudaMemcpy(codebooks, h->codebooks, (size_t)h->M * 256 * h->dsub * 4, cudaMemcpyDeviceToHost), "export codebooks");
  if (list_off) ICK(cudaMemcpy(list_off, h->list_off, (size_t)(h->nlist + 1) * 4, cudaMemcpyDeviceToHost), "export offsets");
  if (ids) ICK(cudaMemcpy(ids, h->ids, (size_t)h->n * 4, cudaMemcpyDeviceToHost), "export ids");
  if (codes) ICK(cudaMemcpy(codes, h->codes, (size_t)h->n * h->M, cudaMemcpyDeviceToHost), "export codes");
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_search_device(nvdb_b200_ivfpq* h, const float* queries_dev, uint32_t nq, uint32_t nprobe, uint32_t R,
                                  uint32_t* cand_ids_dev, float* cand_dist_dev, void* stream) {
  if (!h || !queries_dev || !cand_ids_dev) return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search: NULL argument");
  std::lock_guard<std::mutex> lock(h->mu);
  ICK(cudaSetDevice(h->device), "cudaSetDevice");
  return search_device_impl(h, queries_dev, nq, nprobe, R, cand_ids_dev, cand_dist_dev,
                            stream ? (cudaStream_t)stream : h->stream);  // NULL = the handle's own stream, as in the index API
}

int nvdb_b200_ivfpq_search(nvdb_b200_ivfpq* h, const float* queries_host, uint32_t nq, uint32_t nprobe, uint32_t R,
                           uint32_t* out_ids, float* out_dist) {
  if (!h || !queries_host || !out_ids) return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search: NULL argument");
  if (nq == 0) return NVDB_B200_OK;
  std::lock_guard<std::mutex> lock(h->mu);
  ICK(cudaSetDevice(h->device), "cudaSetDevice");
  int rc;
  const size_t qb = (size_t)nq * h->d * 4, ob = (size_t)nq * R * 4;
  if ((rc = iws(h, IWS_Q, qb))) return rc;
  if ((rc = iws(h, IWS_CAND, ob))) return rc;
  if ((rc = iws(h, IWS_OUT, ob))) return rc;
  if ((rc = pin(h, qb + 2 * ob))) return rc;
  memcpy(h->h_pin, queries_host, qb);  // pinned staging: asynchronous copies, one sync per call
  uint8_t* hp = (uint8_t*)h->h_pin + qb;
  // small results are written by the last kernel straight into the pinned buffer (device-addressable): a copy-engine
  // round trip behind the kernels costs ~6 us of a 50-90 us call
  const bool direct = 2 * ob <= 4096u;
  auto enqueue = [&]() -> int {
    ICK(cudaMemcpyAsync(h->ws[IWS_Q], h->h_pin, qb, cudaMemcpyHostToDevice, h->stream), "H2D queries");
    if (int rc2 = search_device_impl(h, (const float*)h->ws[IWS_Q], nq, nprobe, R,
                                     direct ? (uint32_t*)hp : (uint32_t*)h->ws[IWS_CAND],
                                     out_dist ? (direct ? (float*)(hp + ob) : (float*)h->ws[IWS_OUT]) : nullptr, h->stream))
      return rc2;
    if (!direct) {
      ICK(cudaMemcpyAsync(hp, h->ws[IWS_CAND], ob, cudaMemcpyDeviceToHost, h->stream), "D2H candidate ids");
      if (out_dist) ICK(cudaMemcpyAsync(hp + ob, h->ws[IWS_OUT], ob, cudaMemcpyDeviceToHost, h->stream), "D2H candidate distances");
    }
    return NVDB_B200_OK;
  };
  const uint64_t key[8] = {out_dist ? 3u : 1u, nq, nprobe, R, 0, 0, 0, 0};
  if ((rc = run_or_replay(h, key, nq, enqueue))) return rc;
  ICK(cudaStreamSynchronize(h->stream), "ivfpq search");
  memcpy(out_ids, hp, ob);
  if (out_dist) memcpy(out_dist, hp + ob, ob);
  return NVDB_B200_OK;
}

int nvdb_b200_ivfpq_search_refine(nvdb_b200_ivfpq* h, nvdb_b200_index* base, const float* queries_host, uint32_t nq,
                                  uint32_t nprobe, uint32_t R, uint32_t k, uint32_t* out_ids, float* out_dist) {
  if (!h || !base || !queries_host || !out_ids) return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search_refine: NULL argument");
  if (nq == 0) return NVDB_B200_OK;
  if (k == 0 || k > R) return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search_refine: need 1 <= k <= R");
  const void* b_payload = nullptr;
  const float* b_scales = nullptr;
  uint32_t b_dtype = 0;
  {  // `base` must be the rows this index was built over, on this index's device (its stream and workspaces are used)
    uint64_t bn = 0;
    uint32_t bd = 0;
    int bdev = -1;
    if (nvdb_b200_index_info(base, &bn, &bd, &b_dtype, &bdev, nullptr, &b_payload, &b_scales) != NVDB_B200_OK || bdev != h->device ||
        bd != h->d || bn != h->n)
      return ifail(h, NVDB_B200_ERR_ARG, "ivfpq search_refine: base does not match the index (device, dim or row count)");
  }
  std::lock_guard<std::mutex> lock(h->mu);
  ICK(cudaSetDevice(h->device), "cudaSetDevice");
  int rc;
  const size_t qb = (size_t)nq * h->d * 4, cb = (size_t)nq * R * 4, ob = (size_t)nq * k * 4;
  if ((rc = iws(h, IWS_Q, qb))) return rc;
  if ((rc = iws(h, IWS_CAND, cb))) return rc;
  if ((rc = iws(h, IWS_OUT, 2 * ob))) return rc;
  if ((rc = pin(h, qb + 2 * ob))) return rc;
  // staged through pinned memory so that the copies are truly asynchronous and the call costs one sync
  memcpy(h->h_pin, queries_host, qb);
  uint8_t* hp = (uint8_t*)h->h_pin + qb;
  auto enqueue = [&]() -> int {
    ICK(cudaMemcpyAsync(h->ws[IWS_Q], h->h_pin, qb, cudaMemcpyHostToDevice, h->stream), "H2D queries");
    if (int rc2 = search_device_impl(h, (const float*)h->ws[IWS_Q], nq, nprobe, R, (uint32_t*)h->ws[IWS_CAND], nullptr, h->stream))
      return rc2;
    // small results: the refine kernel writes straight into the pinned buffer (no copy-engine round trip behind it)
    const bool direct = 2 * ob <= 4096u;
    uint32_t* o_ids = direct ? (uint32_t*)hp : (uint32_t*)h->ws[IWS_OUT];
    float* o_dist = direct ? (float*)(hp + ob) : (float*)((uint8_t*)h->ws[IWS_OUT] + ob);
    // candidates never leave the device (the reference copies them to the host and back: src/cuda_refine.cu:926-929)
    if (nvdb_b200_refine_l2_device(base, (const float*)h->ws[IWS_Q], (const uint32_t*)h->ws[IWS_CAND], nq, R, k, o_ids, o_dist,
                                   h->stream) != NVDB_B200_OK)
      return ifail(h, NVDB_B200_ERR_CUDA, std::string("refine: ") + nvdb_b200_last_error(base));
    if (!direct) ICK(cudaMemcpyAsync(hp, h->ws[IWS_OUT], 2 * ob, cudaMemcpyDeviceToHost, h->stream), "D2H results");
    return NVDB_B200_OK;
  };
  // the graph bakes the base's device pointers in: they are part of the key
  const uint64_t key[8] = {2, nq, nprobe, R, k, (uint64_t)(uintptr_t)b_payload, (uint64_t)(uintptr_t)b_scales, b_dtype};
  if ((rc = run_or_replay(h, key, nq, enqueue))) return rc;
  ICK(cudaStreamSynchronize(h->stream), "ivfpq search_refine");
  memcpy(out_ids, hp, ob);
  if (out_dist) memcpy(out_dist, hp + ob, ob);
  return NVDB_B200_OK;
}

}  // extern "C"
