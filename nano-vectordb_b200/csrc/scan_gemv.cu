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
#include "common.cuh"
#include "scan.h"
#include "topk_select.cuh"

namespace nvdb_b200 {

namespace {

constexpr int kRows = kScanRowsPerTile;        // 256 rows per tile == consumer threads
constexpr int kConsumers = kRows;
constexpr int kThreads = kConsumers + 32;      // + producer warp
constexpr int kStageBytes = kRows * 128;       // 32 KB

template <int DT>
struct Elem;
template <>
struct Elem<kDtF32> { static constexpr int size = 4, per16 = 4; };
template <>
struct Elem<kDtF16> { static constexpr int size = 2, per16 = 8; };
template <>
struct Elem<kDtI8> { static constexpr int size = 1, per16 = 16; };

// ---- 16 bytes of one row -> FMAs into the 8 chains of every query -------------------------------------------
// `c` is the 16-byte piece index inside the 128-byte chunk (compile-time after unrolling).

template <int NQ>
__device__ __forceinline__ void fma8(float (&acc)[NQ][8], const float (&x)[8], const float* __restrict__ q,
                                     uint32_t qstride) {
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    const float4 qa = *reinterpret_cast<const float4*>(q + qi * qstride);
    const float4 qb = *reinterpret_cast<const float4*>(q + qi * qstride + 4);
    acc[qi][0] = __fmaf_rn(qa.x, x[0], acc[qi][0]);
    acc[qi][1] = __fmaf_rn(qa.y, x[1], acc[qi][1]);
    acc[qi][2] = __fmaf_rn(qa.z, x[2], acc[qi][2]);
    acc[qi][3] = __fmaf_rn(qa.w, x[3], acc[qi][3]);
    acc[qi][4] = __fmaf_rn(qb.x, x[4], acc[qi][4]);
    acc[qi][5] = __fmaf_rn(qb.y, x[5], acc[qi][5]);
    acc[qi][6] = __fmaf_rn(qb.z, x[6], acc[qi][6]);
    acc[qi][7] = __fmaf_rn(qb.w, x[7], acc[qi][7]);
  }
}

template <int DT, int NQ>
__device__ __forceinline__ void consume16(float (&acc)[NQ][8], const uint4 v, const float* __restrict__ q,
                                          uint32_t qstride, int c) {
  if constexpr (DT == kDtF16) {
    float x[8];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));  // exact widening
      x[2 * i] = f.x;
      x[2 * i + 1] = f.y;
    }
    fma8<NQ>(acc, x, q, qstride);
  } else if constexpr (DT == kDtF32) {
    // 4 elements: chains 0..3 for even pieces, 4..7 for odd pieces
    const float x[4] = {__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)};
    const int j0 = (c & 1) * 4;
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi) {
      const float4 qa = *reinterpret_cast<const float4*>(q + qi * qstride);
      acc[qi][j0 + 0] = __fmaf_rn(qa.x, x[0], acc[qi][j0 + 0]);
      acc[qi][j0 + 1] = __fmaf_rn(qa.y, x[1], acc[qi][j0 + 1]);
      acc[qi][j0 + 2] = __fmaf_rn(qa.z, x[2], acc[qi][j0 + 2]);
      acc[qi][j0 + 3] = __fmaf_rn(qa.w, x[3], acc[qi][j0 + 3]);
    }
  } else {
    // 16 int8: elements 0..7 then 8..15 into the same 8 chains (src/simd_dot.cpp:175-187).
    // int8 -> f32 exactly: byte ^ 0x80 placed in the mantissa of 2^23 gives 2^23 + (x+128); subtracting
    // (2^23 + 128) is exact. One LOP3 per 4 bytes + one PRMT and one FADD per element, no I2F.
    const uint32_t w[4] = {v.x ^ 0x80808080u, v.y ^ 0x80808080u, v.z ^ 0x80808080u, v.w ^ 0x80808080u};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float x[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t word = w[h * 2 + (i >> 2)];
        const uint32_t bits = __byte_perm(word, 0x4B000000u, 0x7650 + (i & 3));  // {0x4B,0x00,0x00,byte}
        x[i] = __fsub_rn(__uint_as_float(bits), 8388736.0f);                     // 2^23 + 128
      }
      fma8<NQ>(acc, x, q + h * 8, qstride);
    }
  }
}

__device__ __forceinline__ float hsum8_ref_order(const float (&a)[8]) {
  const float s0 = __fadd_rn(a[0], a[4]), s1 = __fadd_rn(a[1], a[5]);
  const float s2 = __fadd_rn(a[2], a[6]), s3 = __fadd_rn(a[3], a[7]);
  return __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
}

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

template <int DT, int NQ>
__global__ void __launch_bounds__(kThreads, 1)
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
  uint32_t stage = 0, phase = 0;
  auto csync = [] { named_bar_sync(1, kConsumers); };

  for (uint32_t tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x) {
    const uint32_t row = tile * kRows + row_in_tile;
    float scale = 1.f;
    if constexpr (DT == kDtI8) scale = (row < a.n_rows) ? __ldg(a.scales + row) : 0.f;

    float acc[NQ][8];
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[qi][j] = 0.f;

    for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
      const uint32_t rem = row_bytes - kc * 128u;
      const int npieces = rem >= 128u ? 8 : (int)(rem >> 4);
      const float* qk = sq + kc * EPC;
      mbar_wait(&full[stage], phase);
      const uint32_t base = stages_u32 + stage * kStageBytes;
      if (npieces == 8) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 v = lds128(base + (((uint32_t)c << 4) ^ swz));
          consume16<DT, NQ>(acc, v, qk + c * P16, a.dim, c);
        }
      } else {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          if (c < npieces) {
            const uint4 v = lds128(base + (((uint32_t)c << 4) ^ swz));
            consume16<DT, NQ>(acc, v, qk + c * P16, a.dim, c);
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
        pushed |= sel.offer(qi, s, row);
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
        pushed |= sel.offer(qi, s, row);
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
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
    if (e != cudaSuccess) return e;
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
