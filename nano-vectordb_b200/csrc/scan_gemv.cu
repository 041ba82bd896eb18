// scan_gemv.cu — bandwidth-bound flat scan (batch 1 .. 8 queries per pass) with top-k fused into the stream.
//
// Replaces the hot loop of FlatIndex::search_topk_dot (reference src/flat_index.cpp:31-38) and the bench-side
// batched scan (apps/nvdb_bench.cpp:99-131) together with the AVX2 kernels they call
// (src/simd_dot.cpp:26-49, 102-124, 160-199).
//
// Shape of the kernel (one persistent CTA per SM):
//   * warp 8 (one elected lane) is the TMA producer: it walks the CTA's row tiles (256 rows) and, per tile,
//     the row's 128-byte column chunks, issuing one 2-D tensor-map load (box = 256 rows x 128 B, SWIZZLE_128B)
//     per ring slot, mbarrier-tracked (full/empty), STAGES deep — so ~STAGES*32 KB per SM are always in flight;
//   * warps 0..7 are consumers, ONE THREAD PER BASE ROW: a thread reads its row's eight 16-byte pieces of the
//     slot (the swizzle makes the 8-thread LDS.128 phases conflict-free) and FMAs them into 8 accumulators
//     per query in exactly the reference's order: chain j takes elements j, j+8, j+16, ... ascending, then
//     ((a0+a4)+(a1+a5))+((a2+a6)+(a3+a7)) — so FP32 scores are BIT-IDENTICAL to the AVX2 kernels
//     (SURVEY.md Appendix A.2). Queries sit in shared memory and are read as warp-broadcast LDS.128;
//   * after a tile the thread owns the finished score of its row for each query and offers it to the CTA's
//     top-k selector (topk_select.cuh); scores never touch HBM;
//   * at the end each CTA writes its k best keys per query; merge_partials_kernel (topk_merge.cu) reduces
//     gridDim.x * k keys to the final k.
//
// HBM traffic = the base payload once per pass (plus 4 B/row of scales for int8): the roofline of this kernel.
#include <atomic>
#include "common.cuh"
#include "exchange.cuh"
#include "scan.h"
#include "topk_select.cuh"
#include "dot_chains.cuh"

namespace nvdb_b200 {

namespace {

constexpr int kRows = kScanRowsPerTile;        // 256 rows per tile == consumer threads
constexpr int kConsumers = kRows;
constexpr int kThreads = kConsumers + 32;      // + producer warp
constexpr int kStageBytes = kRows * 128;       // 32 KB

struct SmemLayout {
  uint32_t stages_off, queries_off, keys_off, cnt_off, thr_off, bars_off, total;
};

__host__ __device__ inline SmemLayout make_layout(int stages, int nq, uint32_t dim, uint32_t cap) {
  SmemLayout L;
  uint32_t off = 0;
  L.stages_off = off;
  off += (uint32_t)stages * kStageBytes;
  L.queries_off = off;
  off += (uint32_t)nq * dim * 4u;
  off = (off + 15u) & ~15u;
  L.keys_off = off;
  off += (uint32_t)nq * cap * 8u;
  L.cnt_off = off;
  off += (uint32_t)nq * 4u;
  L.thr_off = off;
  off += (uint32_t)nq * 4u;
  off = (off + 7u) & ~7u;
  L.bars_off = off;
  off += (uint32_t)stages * 16u;
  L.total = off + 1024u;  // slack for the 1024-byte alignment of the swizzled stages
  return L;
}

// Single-query passes can be compiled for two resident CTAs per SM (<= 113 registers; -DNVDB_GEMV_TWO_CTAS=1) so that
// the TUNE_GEMV_CTAS knob has something to select; measured slower than one CTA with the deep ring for every dtype
// (profiles/r1_gemv_ffma2_ctas.jsonl), so the default build does not constrain the register allocation.
#ifndef NVDB_GEMV_TWO_CTAS
#define NVDB_GEMV_TWO_CTAS 0
#endif
template <int DT, int NQ>
__global__ void __launch_bounds__(kThreads, (NQ == 1 && NVDB_GEMV_TWO_CTAS) ? 2 : 1)
scan_gemv_tma_kernel(const __grid_constant__ CUtensorMap tmap, const ScanArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const SmemLayout L = make_layout(a.stages, NQ, a.dim, a.cap);
  float* sq = reinterpret_cast<float*>(smem + L.queries_off);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* empty = full + a.stages;

  const uint32_t tid = threadIdx.x;
  const uint32_t warp = tid >> 5, lane = tid & 31;
  constexpr int ES = Elem<DT>::size, P16 = Elem<DT>::per16, EPC = 128 / ES;

  // queries of this pass -> shared memory (zero for the unused slots of a partial pass)
  for (uint32_t i = tid; i < (uint32_t)NQ * a.dim; i += kThreads) {
    const uint32_t qi = i / a.dim;
    sq[i] = (qi < a.nq) ? a.queries[i] : 0.f;
  }
  if (tid == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], kConsumers / 32);
    }
    mbar_fence_init();
  }
  TopKSelect<NQ> sel;
  sel.init(reinterpret_cast<uint64_t*>(smem + L.keys_off), reinterpret_cast<uint32_t*>(smem + L.cnt_off),
           reinterpret_cast<float*>(smem + L.thr_off), a.cap, a.k, tid);
  __syncthreads();

  if (warp == kConsumers / 32) {
    // ===== TMA producer =====
    if (lane == 0) {
      tma_prefetch_desc(&tmap);
      const uint64_t pol = l2_policy_evict_first();
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
        for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], kStageBytes);
          tma_load_2d_hint(smem + L.stages_off + stage * kStageBytes, &tmap, &full[stage], (int)(kc * EPC),
                           (int)(tile * kRows), pol);
          if (++stage == (uint32_t)a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    return;
  }

  // ===== consumers: one thread per row of the tile =====
  const uint32_t row_in_tile = tid;
  const uint32_t swz = (row_in_tile & 7u) << 4;
  const uint32_t stages_u32 = smem_u32(smem + L.stages_off) + row_in_tile * 128u;
  const uint32_t row_bytes = a.dim * ES;
  const uint32_t sq_u32 = smem_u32(sq), qstride_b = a.dim * 4u;
  uint32_t stage = 0, phase = 0;
  auto csync = [] { named_bar_sync(1, kConsumers); };

  for (uint32_t tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
    const uint32_t row = tile * kRows + row_in_tile;
    float scale = 1.f;
    if constexpr (DT == kDtI8) scale = (row < a.n_rows) ? __ldg(a.scales + row) : 0.f;

    uint64_t acc[NQ][4];  // packed chain pairs, +0.0f in both halves
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
      for (int p = 0; p < 4; ++p) acc[qi][p] = 0ull;

    for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
      const uint32_t rem = row_bytes - kc * 128u;
      const int npieces = rem >= 128u ? 8 : (int)(rem >> 4);
      const uint32_t qk = sq_u32 + kc * (EPC * 4u);
      mbar_wait(&full[stage], phase);
      const uint32_t base = stages_u32 + stage * kStageBytes;
      if (npieces == 8) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 v = lds128(base + (((uint32_t)c << 4) ^ swz));
          consume16<DT, NQ>(acc, v, qk + c * (P16 * 4u), qstride_b, c);
        }
      } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          if (c < npieces) {
            const uint4 v = lds128(base + (((uint32_t)c << 4) ^ swz));
            consume16<DT, NQ>(acc, v, qk + c * (P16 * 4u), qstride_b, c);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
      if (++stage == (uint32_t)a.stages) {
        stage = 0;
        phase ^= 1;
      }
    }

    bool pushed = false;
    if (row < a.n_rows) {
#pragma unroll
      for (int qi = 0; qi < NQ; ++qi) {
        float s = hsum8_ref_order(acc[qi]);
        if constexpr (DT == kDtI8) s = __fmul_rn(s, scale);  // src/simd_dot.cpp:198
        // unused slots of a partial pass (3 queries in a 4-query pass) score 0 for every row: without this test each
        // row ties with their threshold, is appended and forces a compaction per tile (2.6 ms instead of 0.9)
        if (NQ == 1 || (uint32_t)qi < a.nq) pushed |= sel.offer(qi, s, row);
      }
    }
    // barrier + uniform count of threads that appended (bar.red.popc on the consumer-only barrier)
    uint32_t npushed;
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t"
        "bar.red.popc.u32 %0, 1, %2, p;\n\t}"
        : "=r"(npushed)
        : "r"((uint32_t)pushed), "r"(kConsumers)
        : "memory");
    sel.end_tile(npushed, kRows, tid, kConsumers, csync);
  }

  sel.compact(tid, kConsumers, csync);
  // partial layout: [query][cta][k]
  sel.store(a.partial + (size_t)blockIdx.x * a.k, (size_t)gridDim.x * a.k, a.nq, tid, kConsumers);

  // ---- fused final merge: the CTA that finishes last reduces the gridDim.x lists of every query (replaces the serial
  // heap merges of src/flat_index_omp.cpp:67-75 and a separate merge launch: ~25 us of a 680 us batch-1 step).
  if (a.done_counter == nullptr) return;
  __shared__ uint32_t s_ticket, s_cnt;
  __shared__ uint64_t s_floor;
  __threadfence();  // this thread's partial keys are visible device-wide before the ticket is taken
  csync();
  if (tid == 0) s_ticket = atomicAdd(a.done_counter, 1u);
  csync();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  // all TMA loads of this CTA were consumed, so the ring is free: it becomes the survivors' buffer
  uint64_t* buf = reinterpret_cast<uint64_t*>(smem + L.stages_off);
  const uint32_t nlists = gridDim.x, total = nlists * a.k;
  for (uint32_t qi = 0; qi < a.nq; ++qi) {
    const uint64_t* in = a.partial + (size_t)qi * total;
    // Every list is sorted descending, so the final k-th best is (a) >= the best of the lists' own k-th entries and
    // (b) >= the m-th largest of the lists' j-th entries whenever m * j >= k (m lists hold j rows each that reach it):
    // with j = ceil(k / nlists) (1 when there are at least k lists) and m = ceil(k / j) this is far tighter than (a):
    // ~k survivors instead of hundreds (631 of 1480 keys at 148 lists and k = 10, whose 1024-key bitonic sort took 10 us
    // of a 635 us step; k = 100: ~170 of 14,800). The j-th entries are ranked by counting in the idle ring. When the
    // keys fit the ring twice they are staged there with ONE round of L2 loads and everything else reads shared memory.
    if (tid == 0) {
      s_floor = 0ull;
      s_cnt = 0u;
    }
    const size_t ring_bytes = (size_t)a.stages * kStageBytes;
    const bool staged = (size_t)total * 16u <= ring_bytes;
    const bool heads_fit = (size_t)(total + nlists) * 8u <= ring_bytes;
    uint64_t* all = buf + total;  // staged copy of the lists (survivors go to buf[0 .. total))
    uint64_t* heads = staged ? buf : buf + total;  // the lists' j-th entries (read before the survivors are written)
    const uint32_t jth = (a.k + nlists - 1u) / nlists, mth = (a.k + jth - 1u) / jth;
    if (staged)
      for (uint32_t i = tid; i < total; i += kConsumers) all[i] = __ldcg(in + i);
    csync();
    uint64_t fl = 0ull;
    for (uint32_t l = tid; l < nlists; l += kConsumers) {
      const uint64_t v = staged ? all[(size_t)l * a.k + (a.k - 1)] : __ldcg(in + (size_t)l * a.k + (a.k - 1));
      fl = v > fl ? v : fl;
      if (heads_fit) heads[l] = staged ? all[(size_t)l * a.k + (jth - 1u)] : __ldcg(in + (size_t)l * a.k + (jth - 1u));
    }
    csync();
    if (heads_fit) {  // (b)
      for (uint32_t l = tid; l < nlists; l += kConsumers) {
        const uint64_t mine = heads[l];
        uint32_t r = 0;
        for (uint32_t m = 0; m < nlists; ++m) {
          const uint64_t o = heads[m];
          r += (o > mine || (o == mine && m < l)) ? 1u : 0u;  // empty lists: equal zero entries, ranked by position
        }
        if (r == mth - 1u && mine > fl) fl = mine;
      }
    }
    csync();  // the j-th entries have been read: their place may take survivors
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const uint64_t o = __shfl_xor_sync(0xffffffffu, fl, off);
      fl = o > fl ? o : fl;
    }
    if ((tid & 31u) == 0 && fl) atomicMax(reinterpret_cast<unsigned long long*>(&s_floor), (unsigned long long)fl);
    csync();
    const uint64_t floor_key = s_floor;
    for (uint32_t i = tid; i < total; i += kConsumers) {
      const uint64_t v = staged ? all[i] : __ldcg(in + i);
      if (v != 0ull && v >= floor_key) buf[atomicAdd(&s_cnt, 1u)] = v;
    }
    csync();
    const uint32_t cnt = s_cnt;
    if (cnt <= 2u * (uint32_t)kConsumers) {
      // few survivors: rank them by counting (distinct keys: one per row) and store each at its rank
      const uint64_t m0 = tid < cnt ? buf[tid] : 0ull, m1 = tid + kConsumers < cnt ? buf[tid + kConsumers] : 0ull;
      uint32_t r0 = 0, r1 = 0;
      if (tid < cnt) {
        for (uint32_t m = 0; m < cnt; ++m) {
          const uint64_t o = buf[m];
          r0 += o > m0 ? 1u : 0u;
          r1 += o > m1 ? 1u : 0u;
        }
      }
      csync();
      if (tid < cnt) buf[r0] = m0;
      if (tid + kConsumers < cnt) buf[r1] = m1;
      csync();
    } else {
      const uint32_t bufn = next_pow2(cnt);
      for (uint32_t i = cnt + tid; i < bufn; i += kConsumers) buf[i] = 0ull;
      csync();
      bitonic_sort_desc(buf, bufn, tid, kConsumers, csync);
    }
    if (a.xchg.world == 0) {
      for (uint32_t j = tid; j < a.k; j += kConsumers) {
        const uint64_t key = j < cnt ? buf[j] : 0ull;
        const bool ok = key != 0ull;
        a.out_ids[(size_t)qi * a.k + j] = ok ? a.row_offset + (uint64_t)key_row(key) : ~0ull;
        a.out_scores[(size_t)qi * a.k + j] = ok ? key_score(key) : -INFINITY;
      }
    } else {
      // row-sharded index: this shard's list goes straight into slot `rank` of every GPU's window (NVLink stores),
      // packed as [nq*k u64 ids][nq*k f32 scores] — no local copy, no exchange launch
      const size_t slot = a.xchg.parity_off + (size_t)a.xchg.rank * a.xchg.slot_bytes;
      const size_t nk = (size_t)a.nq * a.k;
      for (uint32_t j = tid; j < a.k; j += kConsumers) {
        const uint64_t key = j < cnt ? buf[j] : 0ull;
        const bool ok = key != 0ull;
        const uint64_t id = ok ? a.row_offset + (uint64_t)key_row(key) : ~0ull;
        const float sc = ok ? key_score(key) : -INFINITY;
        for (uint32_t r = 0; r < a.xchg.world; ++r) {
          uint64_t* ids = reinterpret_cast<uint64_t*>(a.xchg.win[r] + slot);
          ids[(size_t)qi * a.k + j] = id;
          reinterpret_cast<float*>(ids + nk)[(size_t)qi * a.k + j] = sc;
        }
      }
    }
    csync();
  }
  if (tid == 0) *a.done_counter = 0u;  // ready for the next launch on this stream
  if (a.xchg.world == 0) return;

  // ---- exchange tail (exchange.cuh): publish this shard's flag in every window, wait for the peers' flags in the own
  // window, merge world*k candidates per query under (score desc, global id asc). One CTA, the rest of the GPU is idle
  // or already running the next kernel's prologue.
  __shared__ uint32_t s_fail;
  __threadfence_system();
  csync();
  if (tid < a.xchg.world) st_release_sys(reinterpret_cast<uint32_t*>(a.xchg.win[tid]) + a.xchg.rank, a.xchg.epoch);
  const bool ok = xchg_wait(reinterpret_cast<const uint32_t*>(a.xchg.win[a.xchg.rank]), a.xchg.world, a.xchg.epoch,
                            a.xchg.status, a.xchg.timeout_ns, tid, &s_fail, csync);
  const uint32_t Lx = a.xchg.world * a.k;
  const unsigned char* slots = a.xchg.win[a.xchg.rank] + a.xchg.parity_off;
  for (uint32_t qi = 0; qi < a.nq; ++qi)
    xchg_merge_query(slots, a.xchg.slot_bytes, (size_t)a.nq * a.k, a.xchg.world, a.k, qi, buf,
                     reinterpret_cast<uint32_t*>(buf + Lx), a.out_ids + (size_t)qi * a.k, a.out_scores + (size_t)qi * a.k,
                     ok, tid, (uint32_t)kConsumers, csync);
}

// ---- generic fallback for shapes the tensor map cannot describe (row bytes not a multiple of 16, or dims with a
// scalar tail). Thread per row straight from global memory, arithmetic = oracle/nvdb_oracle.c incl. the tails.
template <int DT, int NQ>
__global__ void __launch_bounds__(kConsumers, 1) scan_gemv_generic_kernel(const ScanArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const SmemLayout L = make_layout(0, NQ, a.dim, a.cap);
  uint8_t* smem = smem_raw;
  float* sq = reinterpret_cast<float*>(smem + L.queries_off);
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < (uint32_t)NQ * a.dim; i += kConsumers) sq[i] = (i / a.dim < a.nq) ? a.queries[i] : 0.f;
  TopKSelect<NQ> sel;
  sel.init(reinterpret_cast<uint64_t*>(smem + L.keys_off), reinterpret_cast<uint32_t*>(smem + L.cnt_off),
           reinterpret_cast<float*>(smem + L.thr_off), a.cap, a.k, tid);
  __syncthreads();
  auto csync = [] { __syncthreads(); };
  constexpr int ES = Elem<DT>::size;
  const uint32_t step = (DT == kDtI8) ? 16u : 8u;
  const uint32_t body = a.dim - a.dim % step;

  auto elem = [&](const uint8_t* rowp, uint32_t i) -> float {
    if constexpr (DT == kDtF32) return reinterpret_cast<const float*>(rowp)[i];
    else if constexpr (DT == kDtF16) return __half2float(reinterpret_cast<const __half*>(rowp)[i]);
    else return (float)reinterpret_cast<const int8_t*>(rowp)[i];
  };

  for (uint32_t tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
    const uint32_t row = tile * kRows + tid;
    bool pushed = false;
    if (row < a.n_rows) {
      const uint8_t* rowp = reinterpret_cast<const uint8_t*>(a.base) + (size_t)row * a.dim * ES;
#pragma unroll 1
      for (int qi = 0; qi < NQ; ++qi) {
        const float* q = sq + qi * a.dim;
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (uint32_t i = 0; i < body; i += 8)
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = __fmaf_rn(q[i + j], elem(rowp, i + j), acc[j]);
        float s = hsum8_ref_order(acc);
        uint32_t i = body;
        if constexpr (DT == kDtF32) {
          // tail as compiled in the reference build (oracle/nvdb_oracle.c, orc_dot_f32)
          if (a.dim - i >= 4)
            for (int j = 0; j < 4; ++j, ++i) s = __fadd_rn(s, __fmul_rn(q[i], elem(rowp, i)));
          for (; i < a.dim; ++i) s = __fmaf_rn(q[i], elem(rowp, i), s);
        } else {
          for (; i < a.dim; ++i) s = __fadd_rn(s, __fmul_rn(q[i], elem(rowp, i)));
        }
        if constexpr (DT == kDtI8) s = __fmul_rn(s, a.scales[row]);
        if (NQ == 1 || (uint32_t)qi < a.nq) pushed |= sel.offer(qi, s, row);
      }
    }
    const uint32_t npushed = (uint32_t)__syncthreads_count(pushed);
    sel.end_tile(npushed, kRows, tid, kConsumers, csync);
  }
  sel.compact(tid, kConsumers, csync);
  sel.store(a.partial + (size_t)blockIdx.x * a.k, (size_t)gridDim.x * a.k, a.nq, tid, kConsumers);
}

template <int DT, int NQ>
cudaError_t launch_one(const ScanLaunch& l, const ScanArgs& a) {
  if (l.use_tma) {
    const SmemLayout L = make_layout(a.stages, NQ, a.dim, a.cap);
    auto kern = scan_gemv_tma_kernel<DT, NQ>;
    // the attribute is per device and sticky: set it only when this launch needs more than any before (1 us per call)
    static std::atomic<uint32_t> smem_set[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64 || smem_set[dev].load(std::memory_order_relaxed) < L.total) {
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
      if (e != cudaSuccess) return e;
      if (dev >= 0 && dev < 64) smem_set[dev].store(L.total, std::memory_order_relaxed);
    }
    kern<<<l.grid, kThreads, L.total, l.stream>>>(*l.tmap, a);
  } else {
    const SmemLayout L = make_layout(0, NQ, a.dim, a.cap);
    auto kern = scan_gemv_generic_kernel<DT, NQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) return e;
    kern<<<l.grid, kConsumers, L.total, l.stream>>>(a);
  }
  return cudaGetLastError();
}

template <int DT>
cudaError_t launch_dt(const ScanLaunch& l, const ScanArgs& a, int nq_template) {
  switch (nq_template) {
    case 1: return launch_one<DT, 1>(l, a);
    case 2: return launch_one<DT, 2>(l, a);
    case 4: return launch_one<DT, 4>(l, a);
    case 8: return launch_one<DT, 8>(l, a);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace

uint32_t scan_gemv_smem_bytes(bool use_tma, int stages, int nq_template, uint32_t dim, uint32_t cap) {
  return make_layout(use_tma ? stages : 0, nq_template, dim, cap).total;
}

cudaError_t scan_gemv_launch(const ScanLaunch& l, const ScanArgs& a, int nq_template) {
  switch (l.dtype) {
    case kDtF32: return launch_dt<kDtF32>(l, a, nq_template);
    case kDtF16: return launch_dt<kDtF16>(l, a, nq_template);
    case kDtI8: return launch_dt<kDtI8>(l, a, nq_template);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace nvdb_b200
