// scan_mma.cu — tensor-core flat scan for query batches (tcgen05 / TMEM), top-k fused into the epilogue.
//
// Replaces the bench-side batched scan of the reference (apps/nvdb_bench.cpp:47-159: for each row tile, for each
// row, for each query of the batch: score + TopKBuffer::consider) for batches where re-reading the base once
// per 8 queries (scan_gemv.cu) stops paying.
//
// GEMM orientation: QUERIES ON M (TMEM lanes), BASE ROWS ON N (TMEM columns). One CTA owns 128 queries
// (an "M-tile") and sweeps base tiles of BLOCK_N rows; every epilogue thread owns ONE query: it reads its TMEM
// lane (the scores of its query against the BLOCK_N rows of the tile), compares against a threshold it keeps in
// a register and, rarely, inserts into its private sorted top-k list in shared memory — no atomics, no
// cross-thread selection, scores never reach HBM.
//
//   warp 0 (1 lane) : TMA producer — per 128-byte K chunk: query limb tile(s) [128 x 128 B] + base tile
//                     [BLOCK_N x 128 B], SWIZZLE_128B, STAGES-deep mbarrier ring
//   warp 1 (1 lane) : tcgen05.mma issuer — accumulators in TMEM, double buffered (2 x 256 columns)
//   warps 2..5      : epilogue — tcgen05.ld 32x32b.x32, threshold filter, private list insert
//
// Numerics
//   f16 base (kind::f16, FP32 accumulate): the FP32 query is scaled by a power of two (max|q| -> [0.5,1)) and
//     split into LIMBS fp16 limbs (hi [+ lo = fp16(q - hi)]); both limbs accumulate into the same accumulator.
//     Tolerance-mode parity (accumulation order inside the tensor core is unspecified): <= 1e-3 relative
//     required, ~1e-6 observed with 2 limbs.
//   i8 base (kind::i8, S32 accumulate): the FP32 query is decomposed into two int8 limbs,
//     q ~= s1*q1 + s2*q2, s1 = max|q|/127, s2 = s1/254; I1 = q1.x and I2 = q2.x are EXACT
//     int32; score = fmaf(s2, I2, s1*I1) * scale_row in a fixed FP32 order -> bit-exact against the CPU
//     restatement of this scheme (oracle/nvdb_oracle.c orc_dot_i8_limbs), ~3e-6 relative to the AVX2 kernel.
#include "common.cuh"
#include <cstdlib>
#include "scan.h"
#include "tc_common.cuh"
#include "scan_mma_common.cuh"

namespace nvdb_b200 {
namespace {

// BLOCK_N (base rows per tile == accumulator columns): f16 256 (or 128 when shared memory is short: large k with two
// limbs); i8 128 (two limb accumulators of 128 columns per buffer).

struct MmaSmem {
  uint32_t stages_off, res_off, lists_off, rscale_off, bars_off, tmem_off, total, stage_bytes, a_limb_bytes, b_bytes;
};

// res_chunks > 0: limb 0 of the query operand stays RESIDENT in shared memory for the whole kernel (res_chunks K chunks of
// 16 KB, loaded once); the ring then carries only the remaining limb(s) and the base tile.
__host__ __device__ inline MmaSmem mma_layout(uint32_t block_n, int limbs, int stages, uint32_t k, uint32_t res_chunks = 0) {
  MmaSmem L;
  L.a_limb_bytes = kBlockM * 128u;
  L.b_bytes = block_n * 128u;
  L.stage_bytes = (uint32_t)(limbs - (res_chunks ? 1 : 0)) * L.a_limb_bytes + L.b_bytes;
  uint32_t off = 0;
  L.stages_off = off;
  off += (uint32_t)stages * L.stage_bytes;
  L.res_off = off;
  off += res_chunks * L.a_limb_bytes;
  L.lists_off = off;
  off += k * kBlockM * 8u;
  L.rscale_off = off;
  off += 2u * block_n * 4u;
  L.bars_off = off;
  off += (2u * (uint32_t)stages + 5u) * 8u;  // full[], empty[], tfull[2], tempty[2], resident-operand barrier
  L.tmem_off = off;
  off += 16u;
  L.total = off + 1024u;  // slack for the 1024-byte alignment of the swizzled stages
  return L;
}

// ---------------------------------------------------------------------------------------------------------------
// query preparation: FP32 queries -> tensor-core A operand (limb-major: [LIMBS][mpad][dim]) + per-query scales
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float block_max_abs(const float* q, uint32_t dim, float* red) {
  float m = 0.f;
  for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) m = fmaxf(m, fabsf(q[i]));
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = 0.f;
  for (uint32_t w = 0; w < (blockDim.x >> 5); ++w) m = fmaxf(m, red[w]);
  return m;
}

__global__ void __launch_bounds__(128) prep_queries_f16_kernel(const float* __restrict__ queries, uint32_t nq,
                                                                uint32_t dim, uint32_t mpad, int limbs,
                                                                __half* __restrict__ A, float* __restrict__ qscale) {
  __shared__ float red[4];
  const uint32_t q = blockIdx.x;
  __half* hi = A + (size_t)q * dim;
  __half* lo = A + ((size_t)mpad + q) * dim;
  if (q >= nq) {
    for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) {
      hi[i] = __float2half_rn(0.f);
      if (limbs > 1) lo[i] = __float2half_rn(0.f);
    }
    if (threadIdx.x == 0) qscale[q] = 0.f;
    return;
  }
  const float* src = queries + (size_t)q * dim;
  const float amax = block_max_abs(src, dim, red);
  int e = 0;
  if (amax > 0.f && amax < INFINITY) {
    e = -(ilogbf(amax) + 1);  // amax * 2^e in [0.5, 1)
    e = e > 100 ? 100 : (e < -100 ? -100 : e);
  }
  for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) {
    const float v = scalbnf(src[i], e);  // exact
    const __half h = __float2half_rn(v);
    hi[i] = h;
    if (limbs > 1) lo[i] = __float2half_rn(__fsub_rn(v, __half2float(h)));
  }
  if (threadIdx.x == 0) qscale[q] = scalbnf(1.0f, -e);
}

// q ~= s1*q1 + s2*q2 (oracle/nvdb_oracle.c: orc_query_limbs_i8 restates this exactly)
__global__ void __launch_bounds__(128) prep_queries_i8_kernel(const float* __restrict__ queries, uint32_t nq,
                                                               uint32_t dim, uint32_t mpad, int8_t* __restrict__ A,
                                                               float* __restrict__ s1_out, float* __restrict__ s2_out) {
  __shared__ float red[4];
  const uint32_t q = blockIdx.x;
  int8_t* l1 = A + (size_t)q * dim;
  int8_t* l2 = A + ((size_t)mpad + q) * dim;
  if (q >= nq) {
    for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) l1[i] = l2[i] = 0;
    if (threadIdx.x == 0) s1_out[q] = s2_out[q] = 0.f;
    return;
  }
  const float* src = queries + (size_t)q * dim;
  const float amax = block_max_abs(src, dim, red);
  const float s1 = (amax > 0.f && amax < INFINITY) ? __fdiv_rn(amax, 127.f) : 1.f;
  const float inv1 = __fdiv_rn(1.f, s1);
  const float s2 = __fdiv_rn(s1, 254.f);
  const float inv2 = __fdiv_rn(1.f, s2);
  for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) {
    const float v = src[i];
    int a = __float2int_rn(__fmul_rn(v, inv1));
    a = a > 127 ? 127 : (a < -127 ? -127 : a);
    const float r = __fmaf_rn(-(float)a, s1, v);
    int b = __float2int_rn(__fmul_rn(r, inv2));
    b = b > 127 ? 127 : (b < -127 ? -127 : b);
    l1[i] = (int8_t)a;
    l2[i] = (int8_t)b;
  }
  if (threadIdx.x == 0) {
    s1_out[q] = s1;
    s2_out[q] = s2;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// the scan kernel
// ---------------------------------------------------------------------------------------------------------------

template <int DT, int LIMBS, int BLOCK_N, int KREG, int RES = 0>
__global__ void __launch_bounds__(kMmaThreads, 1)
scan_mma_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const MmaArgs a) {
  constexpr int ES = (DT == kDtF16) ? 2 : 1, EPC = 128 / ES;  // elements per 128-byte chunk
  constexpr bool kI8 = (DT == kDtI8);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  static_assert(RES == 0 || LIMBS == 2, "resident limb 0 + streamed limb 1");
  const MmaSmem L = mma_layout(BLOCK_N, LIMBS, a.stages, KREG > 0 ? 0u : a.k, RES ? a.kchunks : 0u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* empty = full + a.stages;
  uint64_t* tfull = empty + a.stages;   // [2] accumulator buffer ready for the epilogue
  uint64_t* tempty = tfull + 2;         // [2] accumulator buffer drained by the epilogue
  uint64_t* resbar = tempty + 2;        // resident query limb landed (RES)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);

  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t m_tile = blockIdx.x % a.G, j = blockIdx.x / a.G;

  if (tid == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1);
      mbar_init(&tempty[b], 4);  // one arrival per epilogue warp
    }
    mbar_init(resbar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tc::tmem_alloc(tmem_slot, kTmemCols);
  tc::fence_before_thread_sync();
  __syncthreads();
  tc::fence_after_thread_sync();
  // broadcast through a shuffle so that the compiler keeps the TMEM base in a uniform register: a value it cannot
  // prove warp-uniform costs an ELECT / R2UR.BROADCAST / BRA.U.ANY waterfall around EVERY tcgen05.mma issue
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ===== TMA producer =====
    {
      if (lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
      }
      const uint64_t pol = l2_policy_evict_first();
      if constexpr (RES) {  // limb 0 of this M-tile: all K chunks, once
        if (lane == 0) {
          mbar_arrive_expect_tx(resbar, a.kchunks * L.a_limb_bytes);
          for (uint32_t kc = 0; kc < a.kchunks; ++kc)
            tma_load_2d(smem + L.res_off + kc * L.a_limb_bytes, &tmap_a, resbar, (int)(kc * EPC), (int)(m_tile * kBlockM));
        }
        __syncwarp();
      }
      uint32_t stage = 0, phase = 0;
      for (uint32_t tile = a.tile0 + j; tile < a.num_tiles; tile += a.P) {
        for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
          mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* st = smem + L.stages_off + stage * L.stage_bytes;
          if (tc::elect_one_sync()) {
          mbar_arrive_expect_tx(&full[stage], L.stage_bytes);
#pragma unroll
          for (int l = RES; l < LIMBS; ++l)
            tma_load_2d(st + (l - RES) * L.a_limb_bytes, &tmap_a, &full[stage], (int)(kc * EPC),
                        (int)(l * a.mpad + m_tile * kBlockM));
          uint8_t* sb = st + (LIMBS - RES) * L.a_limb_bytes;
          if (a.evict_first_base)
            tma_load_2d_hint(sb, &tmap_b, &full[stage], (int)(kc * EPC), (int)(tile * BLOCK_N), pol);
          else
            tma_load_2d(sb, &tmap_b, &full[stage], (int)(kc * EPC), (int)(tile * BLOCK_N));
          }
          __syncwarp();
          if (++stage == (uint32_t)a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the pipeline, one elected lane issues =====
    {
      constexpr uint32_t idesc = kI8 ? tc::idesc_s8_s32(kBlockM, BLOCK_N) : tc::idesc_f16_f32(kBlockM, BLOCK_N);
      uint32_t stage = 0, phase = 0, it = 0;
      if constexpr (RES) {
        mbar_wait(resbar, 0);
        tc::fence_after_thread_sync();
      }
      const uint32_t res_u32 = smem_u32(smem + L.res_off);
      for (uint32_t tile = a.tile0 + j; tile < a.num_tiles; tile += a.P, ++it) {
        const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
        mbar_wait(&tempty[buf], bphase ^ 1);
        tc::fence_after_thread_sync();
        for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
          mbar_wait(&full[stage], phase);
          tc::fence_after_thread_sync();
          const uint32_t st = smem_u32(smem + L.stages_off + stage * L.stage_bytes);
          const uint64_t bdesc = tc::smem_desc_k128(st + (LIMBS - RES) * L.a_limb_bytes);
          if (tc::elect_one_sync()) {
#pragma unroll
          for (int l = 0; l < LIMBS; ++l) {
            const uint64_t adesc = (RES && l == 0) ? tc::smem_desc_k128(res_u32 + kc * L.a_limb_bytes)
                                                   : tc::smem_desc_k128(st + (l - RES) * L.a_limb_bytes);
            const uint32_t acc = tmem_base + buf * kAccColsPerBuf + (kI8 ? (uint32_t)l * BLOCK_N : 0u);
#pragma unroll
            for (uint32_t k4 = 0; k4 < 4; ++k4) {  // 4 x 32 bytes of K per 128-byte chunk (UMMA_K = 16 f16 / 32 i8)
              const uint32_t accumulate = kI8 ? (uint32_t)(kc > 0 || k4 > 0) : (uint32_t)(kc > 0 || l > 0 || k4 > 0);
              if constexpr (kI8)
                tc::mma_i8_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, accumulate);
              else
                tc::mma_f16_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, accumulate);
            }
          }
          tc::mma_commit(&empty[stage]);  // frees the smem slot when these MMAs have read it
          if (kc + 1 == a.kchunks) tc::mma_commit(&tfull[buf]);  // accumulator complete -> epilogue
          }
          __syncwarp();
          if (++stage == (uint32_t)a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else {
    // ===== epilogue: one thread per query (== TMEM lane) =====
    const uint32_t qd = warp & 3;                  // the TMEM lane quadrant this warp may read
    const uint32_t ql = qd * 32 + lane;            // query within the M-tile
    const uint32_t et = (warp - 2) * 32 + lane;    // 0..127 among the epilogue threads
    const uint32_t qg = m_tile * kBlockM + ql;     // < mpad
    const bool valid = qg < a.nq;
    ThreadTopK<KREG> top;
    top.init(reinterpret_cast<uint64_t*>(smem + L.lists_off) + ql, a.k);  // smem variant: entry jj at list[jj * 128]
    const uint32_t rscale_u32 = smem_u32(smem + L.rscale_off);              // [2][BLOCK_N] row scales (i8)
    // admission threshold: score of the current k-th best, optionally seeded with the k-th best of an earlier pass over
    // other rows (a valid lower bound of the final k-th best); padding lanes never insert
    float thr = valid ? (a.thr_init ? a.thr_init[qg] : -INFINITY) : INFINITY;
    const float qs1 = a.qscale[qg];
    const float qs2 = kI8 ? a.qscale2[qg] : 0.f;
    const uint32_t lane_addr = tmem_base + ((qd * 32u) << 16);
    uint32_t it = 0;
    for (uint32_t tile = a.tile0 + j; tile < a.num_tiles; tile += a.P, ++it) {
      const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
      const uint32_t row0 = tile * BLOCK_N;
      if constexpr (kI8) {
        static_assert(!kI8 || BLOCK_N == 128, "one row scale per epilogue thread");
        const float rs_mine = (row0 + et < a.n_rows) ? __ldg(a.scales + row0 + et) : 0.f;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(rscale_u32 + (buf * BLOCK_N + et) * 4u), "f"(rs_mine) : "memory");
      }
      mbar_wait(&tfull[buf], bphase);
      tc::fence_after_thread_sync();
      if constexpr (kI8) named_bar_sync(2, 128);
#pragma unroll 1
      for (uint32_t c0 = 0; c0 < (uint32_t)BLOCK_N; c0 += 32) {
        uint32_t v[32];
        float s[32];
        tc::tmem_ld32(lane_addr + buf * kAccColsPerBuf + c0, v);
        if constexpr (kI8) {
          uint32_t w[32];
          tc::tmem_ld32(lane_addr + buf * kAccColsPerBuf + BLOCK_N + c0, w);
          tc::tmem_ld_wait();
          const uint32_t rs4 = rscale_u32 + (buf * BLOCK_N + c0) * 4u;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            float4 r;  // warp-broadcast LDS.128 (explicit: the aligned smem base has lost its address space)
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                         : "r"(rs4 + (uint32_t)i4 * 16u));
            const float rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = i4 * 4 + u;
              const float t = __fmul_rn(qs1, (float)(int)v[i]);           // |I| < 2^24: exact conversions
              s[i] = __fmul_rn(__fmaf_rn(qs2, (float)(int)w[i], t), rr[u]);
            }
          }
          if (a.dbg) {
            int* d = reinterpret_cast<int*>(a.dbg);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (row0 + c0 + i < a.n_rows) {
                d[(size_t)qg * a.n_rows + row0 + c0 + i] = (int)v[i];
                d[((size_t)a.mpad + qg) * a.n_rows + row0 + c0 + i] = (int)w[i];
              }
          }
        } else {
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) s[i] = __fmul_rn(__uint_as_float(v[i]), qs1);  // power-of-two rescale: exact
          if (a.dbg) {
            float* d = reinterpret_cast<float*>(a.dbg);
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (row0 + c0 + i < a.n_rows) d[(size_t)qg * a.n_rows + row0 + c0 + i] = __uint_as_float(v[i]);
          }
        }
        uint32_t mask = 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) mask |= (s[i] >= thr) ? (1u << i) : 0u;
        while (mask) {  // rare after the first tiles: P(admit) ~ k / rows seen
          const int i = __ffs(mask) - 1;
          mask &= mask - 1;
          float sv = s[0];
#pragma unroll
          for (int u = 1; u < 32; ++u) sv = (i == u) ? s[u] : sv;
          const uint32_t row = row0 + c0 + (uint32_t)i;
          if (sv >= thr && row < a.n_rows) thr = fmaxf(thr, top.insert(make_key(sv, row), thr));
        }
      }
      tc::fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
    }
    if (valid) {
      top.store(a.partial + ((size_t)qg * a.lists + j) * a.k);
    }
  }

  tc::fence_before_thread_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, kTmemCols);
}

template <int DT, int LIMBS, int BLOCK_N, int KREG, int RES = 0>
cudaError_t launch_mma_k(const CUtensorMap* ta, const CUtensorMap* tb, const MmaArgs& a, cudaStream_t stream) {
  const MmaSmem L = mma_layout(BLOCK_N, LIMBS, a.stages, KREG > 0 ? 0u : a.k, RES ? a.kchunks : 0u);
  auto kern = scan_mma_kernel<DT, LIMBS, BLOCK_N, KREG, RES>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
  if (e != cudaSuccess) return e;
  kern<<<a.G * a.P, kMmaThreads, L.total, stream>>>(*ta, *tb, a);
  count_launch();
  return cudaGetLastError();
}

template <int DT, int LIMBS, int BLOCK_N>
cudaError_t launch_mma(const CUtensorMap* ta, const CUtensorMap* tb, const MmaArgs& a, cudaStream_t stream) {
  if (a.k <= (uint32_t)kMmaRegK) return launch_mma_k<DT, LIMBS, BLOCK_N, kMmaRegK>(ta, tb, a, stream);
  return launch_mma_k<DT, LIMBS, BLOCK_N, 0>(ta, tb, a, stream);
}

}  // namespace

// Picks the base tile height and the ring depth that fit next to the k-entry lists: f16 prefers 256-row tiles.
bool scan_mma_plan(int dtype, int limbs, uint32_t k, int max_smem, uint32_t* block_n, int* stages) {
  const uint32_t cand[2] = {dtype == kDtF16 ? 256u : 128u, 128u};
  for (int c = 0; c < 2; ++c) {
    const int min_stages = (c == 0 && cand[0] != cand[1]) ? 3 : 2;
    int st_max = 6;
    if (const char* e = getenv("NVDB_B200_MMA_MAX_STAGES")) {  // diagnostic: ring depth experiments
      const int v = atoi(e);
      if (v >= 3 && v < st_max) st_max = v;
    }
    for (int st = st_max; st >= min_stages; --st)
      if (mma_layout(cand[c], limbs, st, k <= (uint32_t)kMmaRegK ? 0u : k).total <= (uint32_t)max_smem) {
        *block_n = cand[c];
        *stages = st;
        return true;
      }
  }
  return false;
}

cudaError_t scan_mma_prep_launch(int dtype, int limbs, const float* queries, uint32_t nq, uint32_t dim, uint32_t mpad,
                                 void* A, float* qs1, float* qs2, cudaStream_t stream) {
  if (dtype == kDtF16)
    prep_queries_f16_kernel<<<mpad, 128, 0, stream>>>(queries, nq, dim, mpad, limbs, reinterpret_cast<__half*>(A), qs1);
  else if (dtype == kDtI8)
    prep_queries_i8_kernel<<<mpad, 128, 0, stream>>>(queries, nq, dim, mpad, reinterpret_cast<int8_t*>(A), qs1, qs2);
  else
    return cudaErrorInvalidValue;
  count_launch();
  return cudaGetLastError();
}

cudaError_t scan_mma_launch(int dtype, int limbs, const CUtensorMap* tmap_a, const CUtensorMap* tmap_b, const MmaArgs& a,
                            cudaStream_t stream) {
  if (dtype == kDtF16 && limbs == 1 && a.block_n == 256) return launch_mma<kDtF16, 1, 256>(tmap_a, tmap_b, a, stream);
  if (dtype == kDtF16 && limbs == 2 && a.block_n == 256) return launch_mma<kDtF16, 2, 256>(tmap_a, tmap_b, a, stream);
  if (dtype == kDtF16 && limbs == 1 && a.block_n == 128) return launch_mma<kDtF16, 1, 128>(tmap_a, tmap_b, a, stream);
  if (dtype == kDtF16 && limbs == 2 && a.block_n == 128) return launch_mma<kDtF16, 2, 128>(tmap_a, tmap_b, a, stream);
  if (dtype == kDtI8 && limbs == 2 && a.block_n == 128) {
    // resident first query limb (per-step operand traffic 48 -> 32 KB; the i8 scan is L2->SM feed-bound, ncu: tensor pipe
    // 52 % busy at 13.9 TB/s of operand traffic) whenever it fits next to a >= 4-deep ring and register lists
    if (a.resident_a && a.k <= (uint32_t)kMmaRegK) {
      for (int st = 6; st >= 4; --st)
        if (mma_layout(128, 2, st, 0u, a.kchunks).total <= (uint32_t)a.max_smem) {
          MmaArgs b = a;
          b.stages = st;
          return launch_mma_k<kDtI8, 2, 128, kMmaRegK, 1>(tmap_a, tmap_b, b, stream);
        }
    }
    return launch_mma<kDtI8, 2, 128>(tmap_a, tmap_b, a, stream);
  }
  return cudaErrorInvalidValue;
}

}  // namespace nvdb_b200
