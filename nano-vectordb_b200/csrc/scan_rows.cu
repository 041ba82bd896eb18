// scan_rows.cu — tensor-core flat scan for SMALL query batches (3 .. ~80 queries): BASE ROWS ON M, QUERIES ON N.
//
// Replaces the bench-side batched scan of the reference in the regime the reference itself measured (batch <= 8:
// apps/nvdb_bench.cpp:99-131, Performance.md:289-290,395-402) and up to the point where a batch stops fitting in one
// SM's shared memory. scan_mma.cu puts the queries on M: a batch of 8 still pays for a 128-row query tile, the query
// tile(s) are re-streamed from L2 with every base tile (INT8: two limbs = twice the base bytes) and every admitted row
// costs a 16-entry register-list insert in a divergent warp. Here the orientation is swapped:
//
//   A operand (M = 128, TMEM lanes)   = a tile of 128 BASE ROWS, streamed once from HBM by TMA (16 KB per 128-byte K chunk)
//   B operand (N = 16 .. 256 columns) = ALL query limbs of the batch, converted once per CTA and RESIDENT in shared memory
//                                       (f16: one fp16 limb per query; i8: two int8 limbs per query, limb 2 at column bp + q)
//   D (TMEM, FP32 / S32)              = 128 rows x N scores, double buffered
//
// so the tensor work is 128 x N x D per tile instead of 128 x 256 x D, L2 -> SM traffic is the base bytes only, and the
// kernel is bound by the HBM stream like the streaming scan (scan_gemv.cu) — one pass over the base for the whole batch.
//
//   warp 0      : TMA producer (base tiles only), STAGES-deep mbarrier ring, L2 evict_first
//   warp 1      : tcgen05.mma issuer (one elected lane), 4 MMAs per K chunk
//   warps 2..5  : epilogue, ONE THREAD PER BASE ROW: tcgen05.ld of its lane (its row's scores against every query), compare
//                 with the query's threshold (a broadcast shared-memory word) and — rarely — append an ordered key to the
//                 query's candidate list in shared memory (atomic slot). An append that finds its list full is kept as a
//                 pending bit and starts a maintenance round: every list that is at least half full is compacted by one
//                 warp (bitonic sort, keep the best k', raise the threshold), then the pending appends are retried from
//                 the accumulators, which stay in TMEM until the tile is done: nothing is ever dropped, whatever the
//                 data (first tile: every row qualifies for every query).
//
// Numerics are those of scan_mma.cu (same limb schemes, same FP32 epilogue order), so the candidates carry the same
// approximate scores and the exact path's certificate (rescore.cu) holds unchanged.
#include "common.cuh"
#include <cstdlib>
#include "scan.h"
#include "scan_mma_common.cuh"
#include "tc_common.cuh"

namespace nvdb_b200 {
namespace {

constexpr int kRowsEpiWarps = 8;             // two per TMEM lane quadrant: each takes every other group of 16 queries
constexpr int kRowsEpiThreads = kRowsEpiWarps * 32;
constexpr int kRowsThreads = 64 + kRowsEpiThreads;
constexpr uint32_t kRowsTile = 128;          // base rows per tile == TMEM lanes
constexpr uint32_t kRowsChunkBytes = kRowsTile * 128u;  // one 128-byte K chunk of a tile: one TMA box
constexpr uint32_t kRowsCps = 2;             // K chunks per ring stage (one barrier round trip per 32 KB)
constexpr uint32_t kRowsStageBytes = kRowsCps * kRowsChunkBytes;
constexpr uint32_t kRowsMaxStages = 6;

struct RowsSmem {
  uint32_t stages_off, b_off, keys_off, par_off, cnt_off, wl_off, bars_off, tmem_off, total, b_chunk_bytes;
};

__host__ __device__ inline RowsSmem rows_layout(uint32_t nb, uint32_t bp, uint32_t cap, uint32_t kchunks, int stages) {
  RowsSmem L;
  uint32_t off = 0;
  L.stages_off = off;
  off += (uint32_t)stages * kRowsStageBytes;
  L.b_off = off;  // kchunks blocks of nb rows x 128 B, SWIZZLE_128B atoms of 8 rows (nb % 16 == 0 -> 1024-byte aligned blocks)
  L.b_chunk_bytes = nb * 128u;
  off += kchunks * L.b_chunk_bytes;
  L.keys_off = off;  // [bp][cap] u64 candidate keys
  off += bp * cap * 8u;
  L.par_off = off;   // [bp] float4 {qs1, qs2, thr, -}
  off += bp * 16u;
  L.cnt_off = off;   // [bp] u32
  off += bp * 4u;
  L.wl_off = off;    // compaction work lists: [2][4] {count, head, -, -} then [2][bp] query numbers
  off += 32u + 2u * bp * 4u;
  off = (off + 7u) & ~7u;
  L.bars_off = off;  // full[stages], empty[stages], tfull[8], tempty[8]
  off += (2u * (uint32_t)stages + 16u) * 8u;
  L.tmem_off = off;
  off += 16u;
  L.total = off + 1024u;  // slack for the 1024-byte alignment of the swizzled regions
  return L;
}

// byte offset of (row r, byte c of the 128-byte K chunk) inside a SWIZZLE_128B block: 16-byte piece index XOR (r & 7)
__device__ __forceinline__ uint32_t swz128(uint32_t r, uint32_t c) {
  return (r >> 3) * 1024u + (r & 7u) * 128u + ((((c >> 4) ^ (r & 7u)) << 4) | (c & 15u));
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ uint32_t atom_add_shared(uint32_t addr, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(addr), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void sts_f32(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// barrier over the epilogue threads that also counts a predicate
__device__ __forceinline__ uint32_t epi_bar_count(bool p) {
  uint32_t n;
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\t"
      "bar.red.popc.u32 %0, 1, %2, p;\n\t}"
      : "=r"(n)
      : "r"((uint32_t)p), "n"(kRowsEpiThreads)
      : "memory");
  return n;
}

// One warp sorts list q (cap keys, entries >= c cleared) descending in shared memory, keeps the best kp and publishes the
// new count and threshold. All 32 lanes call with the same arguments.
__device__ __noinline__ void warp_compact(uint32_t keys_q, uint32_t cap, uint32_t c, uint32_t kp, uint32_t cnt_addr,
                                          uint32_t thr_addr, uint32_t lane) {
  for (uint32_t i = c + lane; i < cap; i += 32) sts_u64(keys_q + i * 8u, 0ull);
  __syncwarp();
  for (uint32_t size = 2; size <= cap; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      for (uint32_t t = lane; t < (cap >> 1); t += 32) {
        const uint32_t lo = 2 * t - (t & (stride - 1));
        const uint32_t hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const uint64_t a = lds_u64(keys_q + lo * 8u), b = lds_u64(keys_q + hi * 8u);
        if ((a < b) == desc) {
          sts_u64(keys_q + lo * 8u, b);
          sts_u64(keys_q + hi * 8u, a);
        }
      }
      __syncwarp();
    }
  }
  if (lane == 0) {
    const uint32_t nc = c < kp ? c : kp;
    sts_u32(cnt_addr, nc);
    // admission is `>=` so that rows tying with the k'-th best reach the exact (score, id) ordering of the next sort
    // (never lowered: a bound adopted from the other CTAs - threshold seeding - stays in force)
    if (nc == kp) {
      const float t = key_score(lds_u64(keys_q + (kp - 1u) * 8u));
      float cur;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(cur) : "r"(thr_addr));
      if (t > cur) sts_f32(thr_addr, t);
    }
  }
  __syncwarp();
}

// cap == 64 (k' <= 16): the same compaction in registers — two keys per lane, bitonic sort over shuffles (no shared-memory
// round trip per stage: ~4x faster than the loop above, and compactions are what the epilogue waits for at its barrier).
__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, uint32_t m) {
  const uint32_t lo = __shfl_xor_sync(0xffffffffu, (uint32_t)v, m), hi = __shfl_xor_sync(0xffffffffu, (uint32_t)(v >> 32), m);
  return ((uint64_t)hi << 32) | lo;
}
__device__ __noinline__ void warp_compact64(uint32_t keys_q, uint32_t c, uint32_t kp, uint32_t cnt_addr, uint32_t thr_addr,
                                            uint32_t lane) {
  uint64_t k0 = lane < c ? lds_u64(keys_q + lane * 8u) : 0ull;              // element lane
  uint64_t k1 = lane + 32u < c ? lds_u64(keys_q + (lane + 32u) * 8u) : 0ull;  // element lane + 32
#pragma unroll
  for (uint32_t size = 2; size <= 64; size <<= 1) {
#pragma unroll
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride == 32) {  // size == 64: partner lives in the same lane; descending -> the larger key at the lower element
        if (k0 < k1) {
          const uint64_t t = k0;
          k0 = k1;
          k1 = t;
        }
      } else {
        const bool lower = (lane & stride) == 0;
        const uint64_t o0 = shfl_xor_u64(k0, stride), o1 = shfl_xor_u64(k1, stride);
        const bool desc0 = (lane & size) == 0 || size == 64;           // element lane:      (e & size) == 0
        const bool desc1 = ((lane + 32u) & size) == 0 || size == 64;   // element lane + 32
        const bool max0 = lower == desc0, max1 = lower == desc1;
        k0 = max0 ? (k0 > o0 ? k0 : o0) : (k0 < o0 ? k0 : o0);
        k1 = max1 ? (k1 > o1 ? k1 : o1) : (k1 < o1 ? k1 : o1);
      }
    }
  }
  const uint32_t nc = c < kp ? c : kp;
  if (lane < kp) sts_u64(keys_q + lane * 8u, k0);                       // kp <= 16: the best kp sit in lanes 0 .. kp-1
  const uint64_t kth = __shfl_sync(0xffffffffu, (uint32_t)(k0 >> 32), kp - 1u);
  if (lane == 0) {
    sts_u32(cnt_addr, nc);
    if (nc == kp) {  // never lowered (see warp_compact)
      const float t = ordered_to_float((uint32_t)kth);
      float cur;
      asm volatile("ld.shared.f32 %0, [%1];" : "=f"(cur) : "r"(thr_addr));
      if (t > cur) sts_f32(thr_addr, t);
    }
  }
  __syncwarp();
}

template <int DT>
__global__ void __launch_bounds__(kRowsThreads, 1) scan_rows_kernel(const __grid_constant__ CUtensorMap tmap_base, const RowsArgs a) {
  constexpr bool kI8 = (DT == kDtI8);
  constexpr uint32_t ES = kI8 ? 1u : 2u, EPC = 128u / ES;  // elements per 128-byte chunk
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const RowsSmem L = rows_layout(a.nb, a.bp, a.cap, a.kchunks, a.stages);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* empty = full + a.stages;
  uint64_t* tfull = empty + a.stages;
  uint64_t* tempty = tfull + 8;
  // Accumulator buffers in TMEM: as many as 512 columns hold (2 for 256 accumulator columns ... 8 from 64 down), a power
  // of two. The MMA warp may run that many tiles ahead of the epilogue, which hides a maintenance round of the candidate
  // lists (a few microseconds during which no epilogue warp releases a buffer) from the HBM stream.
  uint32_t nbuf_log2 = 1;
  while (nbuf_log2 < 3u && (a.nb << (nbuf_log2 + 1u)) <= kTmemCols) ++nbuf_log2;
  const uint32_t nbuf = 1u << nbuf_log2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);
  const uint32_t b_u32 = smem_u32(smem + L.b_off), keys_u32 = smem_u32(smem + L.keys_off);
  const uint32_t par_u32 = smem_u32(smem + L.par_off), cnt_u32 = smem_u32(smem + L.cnt_off);
  const uint32_t wl_u32 = smem_u32(smem + L.wl_off);   // [2] {count, head, -, -}, then [2][bp] queued query numbers

  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t j = blockIdx.x, P = gridDim.x;
  const uint32_t nsg = (a.kchunks + kRowsCps - 1u) / kRowsCps;  // ring stages per tile

  if (tid == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (uint32_t b = 0; b < nbuf; ++b) {
      mbar_init(&tfull[b], 1);
      mbar_init(&tempty[b], kRowsEpiWarps);  // one arrival per epilogue warp
    }
    mbar_fence_init();
  }
  if (tid < 8) sts_u32(wl_u32 + tid * 4u, 0u);
  if (warp == 2) tc::tmem_alloc(tmem_slot, kTmemCols);
  // the resident query operand starts as zeros: padding rows (columns past the batch), and the tail of a ragged last chunk
  {
    const uint32_t pieces = a.kchunks * L.b_chunk_bytes / 16u;
    for (uint32_t i = tid; i < pieces; i += kRowsThreads) sts_v4(b_u32 + i * 16u, make_uint4(0u, 0u, 0u, 0u));
    // the candidate lists too: a slot that was never written must read as "no key" (threshold seeding looks at slots that
    // another warp may have reserved but not yet written)
    const uint32_t kpieces = a.bp * a.cap / 2u;
    for (uint32_t i = tid; i < kpieces; i += kRowsThreads) sts_v4(keys_u32 + i * 16u, make_uint4(0u, 0u, 0u, 0u));
  }
  tc::fence_before_thread_sync();
  __syncthreads();
  tc::fence_after_thread_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  // One ring stage = up to kRowsCps K chunks of one tile (one box each). The producer position is kept by all lanes.
  const uint64_t pol = l2_policy_evict_first();
  uint32_t p_stage = 0, p_phase = 0, p_tile = j, p_sg = 0;
  auto produce = [&]() {  // issue the loads of stage (p_tile, p_sg) into ring slot p_stage and advance; whole warp
    const uint32_t c0 = p_sg * kRowsCps;
    const uint32_t nch = a.kchunks - c0 < kRowsCps ? a.kchunks - c0 : kRowsCps;
    if (tc::elect_one_sync()) {
      mbar_arrive_expect_tx(&full[p_stage], nch * kRowsChunkBytes);
      for (uint32_t i = 0; i < nch; ++i)
        tma_load_2d_hint(smem + L.stages_off + p_stage * kRowsStageBytes + i * kRowsChunkBytes, &tmap_base, &full[p_stage],
                         (int)((c0 + i) * EPC), (int)(p_tile * kRowsTile), pol);
    }
    __syncwarp();
    if (++p_stage == (uint32_t)a.stages) {
      p_stage = 0;
      p_phase ^= 1;
    }
    if (++p_sg == nsg) {
      p_sg = 0;
      p_tile += P;
    }
  };
  // The HBM stream starts before the queries are converted: every ring slot is free, so the first `stages` loads need
  // no wait.
  if (warp == 0) {
    if (lane == 0) tma_prefetch_desc(&tmap_base);
    for (int s = 0; s < a.stages && p_tile < a.num_tiles; ++s) produce();
  }

  // ---- query preparation: FP32 queries -> limb rows of the B operand (swizzled, K-major) + per-query scales. Same
  // arithmetic as scan_mma.cu's prep kernels (oracle/nvdb_oracle.c: orc_query_limbs_f16 / _i8 restate it).
  // The MMA cannot start before every query is converted, so this loop is pure start-up latency (20 us for 64 queries
  // when every query cost two dependent global round trips). dim <= 1024: a lane keeps its share of the query in registers
  // (8 float4: the pieces it converts) and the NEXT query of the warp is loaded while the current one is converted.
  constexpr uint32_t kEPP = kI8 ? 16u : 8u;  // FP32 elements per 16-byte operand piece
  constexpr uint32_t kF4P = kEPP / 4u;       // float4 loads per piece
  auto load_query = [&](uint32_t q, float4 (&r)[8]) {
    const float* src = a.queries + (size_t)q * a.dim;
#pragma unroll
    for (uint32_t c = 0; c < 8u; ++c) {
      const uint32_t e = ((c / kF4P) * 32u + lane) * kEPP + (c % kF4P) * 4u;
      r[c] = (q < a.nq && e < a.dim) ? *reinterpret_cast<const float4*>(src + e) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  const bool in_regs = a.dim <= 1024u;
  float4 nxt[8];
  if (in_regs) load_query(warp, nxt);
  for (uint32_t q = warp; q < a.bp; q += kRowsThreads / 32) {
    float qs1 = 0.f, qs2 = 0.f, thr0 = INFINITY;  // padding columns: zero rows, nothing is ever admitted
    float4 cur[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) cur[c] = nxt[c];
    if (in_regs && q + kRowsThreads / 32 < a.nq) load_query(q + kRowsThreads / 32, nxt);
    if (q < a.nq) {
      const float* src = a.queries + (size_t)q * a.dim;
      float amax = 0.f;
      if (in_regs) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 v = cur[c];
          amax = fmaxf(fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
        }
      } else {
        for (uint32_t i = lane * 4u; i < a.dim; i += 128u) {
          const float4 v = *reinterpret_cast<const float4*>(src + i);
          amax = fmaxf(fmaxf(amax, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
        }
      }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
      thr0 = a.thr_init ? a.thr_init[q] : -INFINITY;
      if constexpr (!kI8) {
        int e = 0;
        if (amax > 0.f && amax < INFINITY) {
          e = -(ilogbf(amax) + 1);  // amax * 2^e in [0.5, 1)
          e = e > 100 ? 100 : (e < -100 ? -100 : e);
        }
        qs1 = scalbnf(1.0f, -e);
        auto piece = [&](uint32_t p8, const float4& v0, const float4& v1) {  // 8 elements = one 16-byte piece of the fp16 row
          const __half2 h0 = __floats2half2_rn(scalbnf(v0.x, e), scalbnf(v0.y, e));
          const __half2 h1 = __floats2half2_rn(scalbnf(v0.z, e), scalbnf(v0.w, e));
          const __half2 h2 = __floats2half2_rn(scalbnf(v1.x, e), scalbnf(v1.y, e));
          const __half2 h3 = __floats2half2_rn(scalbnf(v1.z, e), scalbnf(v1.w, e));
          uint4 w;
          w.x = *reinterpret_cast<const uint32_t*>(&h0);
          w.y = *reinterpret_cast<const uint32_t*>(&h1);
          w.z = *reinterpret_cast<const uint32_t*>(&h2);
          w.w = *reinterpret_cast<const uint32_t*>(&h3);
          const uint32_t byte = p8 * 16u, kc = byte >> 7, c = byte & 127u;
          sts_v4(b_u32 + kc * L.b_chunk_bytes + swz128(q, c), w);
        };
        if (in_regs) {
#pragma unroll
          for (uint32_t pc = 0; pc < 4u; ++pc) {
            const uint32_t p8 = pc * 32u + lane;
            if (p8 * 8u < a.dim) piece(p8, cur[(pc * 2u) & 7u], cur[(pc * 2u + 1u) & 7u]);
          }
        } else {
          for (uint32_t p8 = lane; p8 * 8u < a.dim; p8 += 32u)
            piece(p8, *reinterpret_cast<const float4*>(src + p8 * 8u), *reinterpret_cast<const float4*>(src + p8 * 8u + 4u));
        }
      } else {
        const float s1 = (amax > 0.f && amax < INFINITY) ? __fdiv_rn(amax, 127.f) : 1.f;
        const float inv1 = __fdiv_rn(1.f, s1);
        const float s2 = __fdiv_rn(s1, 254.f);
        const float inv2 = __fdiv_rn(1.f, s2);
        qs1 = s1;
        qs2 = s2;
        auto piece = [&](uint32_t p16, const float4 (&vv4)[4]) {  // 16 elements = one 16-byte piece of each int8 limb row
          uint32_t w1[4], w2[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const float vv[4] = {vv4[g].x, vv4[g].y, vv4[g].z, vv4[g].w};
            uint32_t b1 = 0, b2 = 0;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              int x = __float2int_rn(__fmul_rn(vv[u], inv1));
              x = x > 127 ? 127 : (x < -127 ? -127 : x);
              const float r = __fmaf_rn(-(float)x, s1, vv[u]);
              int y = __float2int_rn(__fmul_rn(r, inv2));
              y = y > 127 ? 127 : (y < -127 ? -127 : y);
              b1 |= ((uint32_t)x & 0xFFu) << (8 * u);
              b2 |= ((uint32_t)y & 0xFFu) << (8 * u);
            }
            w1[g] = b1;
            w2[g] = b2;
          }
          const uint32_t byte = p16 * 16u, kc = byte >> 7, c = byte & 127u;
          sts_v4(b_u32 + kc * L.b_chunk_bytes + swz128(q, c), make_uint4(w1[0], w1[1], w1[2], w1[3]));
          sts_v4(b_u32 + kc * L.b_chunk_bytes + swz128(a.bp + q, c), make_uint4(w2[0], w2[1], w2[2], w2[3]));
        };
        if (in_regs) {
#pragma unroll
          for (uint32_t pc = 0; pc < 2u; ++pc) {
            const uint32_t p16 = pc * 32u + lane;
            const float4 vv4[4] = {cur[(pc * 4u) & 7u], cur[(pc * 4u + 1u) & 7u], cur[(pc * 4u + 2u) & 7u], cur[(pc * 4u + 3u) & 7u]};
            if (p16 * 16u < a.dim) piece(p16, vv4);
          }
        } else {
          for (uint32_t p16 = lane; p16 * 16u < a.dim; p16 += 32u) {
            const float4 vv4[4] = {*reinterpret_cast<const float4*>(src + p16 * 16u), *reinterpret_cast<const float4*>(src + p16 * 16u + 4u),
                                   *reinterpret_cast<const float4*>(src + p16 * 16u + 8u), *reinterpret_cast<const float4*>(src + p16 * 16u + 12u)};
            piece(p16, vv4);
          }
        }
      }
    }
    if (lane == 0) {
      asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(par_u32 + q * 16u), "f"(qs1), "f"(qs2), "f"(thr0), "f"(0.f) : "memory");
      sts_u32(cnt_u32 + q * 4u, 0u);
    }
  }
  // generic-proxy writes of the B operand -> visible to the tensor core's (async proxy) shared-memory reads
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();

  if (warp == 0) {
    // ===== TMA producer (continues after the prologue's first loads) =====
    while (p_tile < a.num_tiles) {
      mbar_wait(&empty[p_stage], p_phase ^ 1);
      produce();
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the whole warp walks the pipeline, one elected lane issues =====
    const uint32_t idesc = kI8 ? tc::idesc_s8_s32(kRowsTile, a.nb) : tc::idesc_f16_f32(kRowsTile, a.nb);
    uint32_t stage = 0, phase = 0, it = 0;
    for (uint32_t tile = j; tile < a.num_tiles; tile += P, ++it) {
      const uint32_t buf = it & (nbuf - 1u), bphase = (it >> nbuf_log2) & 1;
      mbar_wait(&tempty[buf], bphase ^ 1);
      tc::fence_after_thread_sync();
      const uint32_t acc = tmem_base + buf * a.nb;
      for (uint32_t sg = 0; sg < nsg; ++sg) {
        const uint32_t c0 = sg * kRowsCps;
        const uint32_t nch = a.kchunks - c0 < kRowsCps ? a.kchunks - c0 : kRowsCps;
        mbar_wait(&full[stage], phase);
        tc::fence_after_thread_sync();
        const uint32_t st_u32 = smem_u32(smem + L.stages_off + stage * kRowsStageBytes);
        if (tc::elect_one_sync()) {
          for (uint32_t i = 0; i < nch; ++i) {
            const uint64_t adesc = tc::smem_desc_k128(st_u32 + i * kRowsChunkBytes);
            const uint64_t bdesc = tc::smem_desc_k128(b_u32 + (c0 + i) * L.b_chunk_bytes);
#pragma unroll
            for (uint32_t k4 = 0; k4 < 4; ++k4) {  // 4 x 32 bytes of K per 128-byte chunk (UMMA_K = 16 f16 / 32 i8)
              const uint32_t accumulate = (uint32_t)(c0 + i > 0 || k4 > 0);
              if constexpr (kI8)
                tc::mma_i8_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, accumulate);
              else
                tc::mma_f16_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, accumulate);
            }
          }
          tc::mma_commit(&empty[stage]);
          if (sg + 1 == nsg) tc::mma_commit(&tfull[buf]);
        }
        __syncwarp();
        if (++stage == (uint32_t)a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // ===== epilogue: one thread per base row (== TMEM lane); the two warps of a lane quadrant split the query groups =====
    const uint32_t ew = warp - 2;                // 0..7
    const uint32_t qd = warp & 3;                // the TMEM lane quadrant this warp may read
    const uint32_t hf = ew >> 2;                 // groups hf, hf + 2, ... of 16 queries are this warp's
    const uint32_t rl = qd * 32 + lane;          // row within the tile
    const uint32_t lane_addr = tmem_base + ((qd * 32u) << 16);
    const uint32_t ngroups = a.bp >> 4;          // groups of 16 queries
    const uint32_t half_cap = a.cap >> 1;
    // cross-CTA threshold seeding (see below): off when there is no workspace or fewer CTAs than list entries; round r is
    // published after tile 0 / 8 / 64 when every CTA has at least 2 more tiles after it
    uint32_t nrounds = 0;
    if (a.seed_ws != nullptr && P >= a.kp && P <= 256u) {
      const uint32_t tiles_min = a.num_tiles / P;
      for (uint32_t pit = 0; nrounds < a.seed_rounds && pit + 2u < tiles_min; pit = pit == 0u ? 8u : pit * 8u) ++nrounds;
      if (nrounds == 0) nrounds = 1;  // the single early round of small bases
    }
    uint32_t my_round = 0, pub_round = 0, pub_it = 0;  // rounds adopted by this warp / published by it; tile of the next one
    uint32_t arrived_round = 0, pcur = 0;             // thread 64 only: rounds whose P publications it has seen
    bool polled = false;

    // score of (this row, query q) from the raw accumulators, in scan_mma.cu's fixed FP32 order
    auto score_of = [&](const float4& p, uint32_t v, uint32_t w, float rs) -> float {
      if constexpr (kI8) {
        const float t = __fmul_rn(p.x, (float)(int)v);  // |I| < 2^24: exact conversions
        return __fmul_rn(__fmaf_rn(p.y, (float)(int)w, t), rs);
      } else {
        return __fmul_rn(__uint_as_float(v), p.x);      // power-of-two rescale: exact
      }
    };
    auto compact = [&](uint32_t q) {  // whole warp
      const uint32_t c0 = lds_u32(cnt_u32 + q * 4u);
      const uint32_t c = c0 < a.cap ? c0 : a.cap;
      if (a.cap == 64u)
        warp_compact64(keys_u32 + q * 512u, c, a.kp, cnt_u32 + q * 4u, par_u32 + q * 16u + 8u, lane);
      else
        warp_compact(keys_u32 + q * a.cap * 8u, a.cap, c, a.kp, cnt_u32 + q * 4u, par_u32 + q * 16u + 8u, lane);
    };

    uint32_t it = 0;
    float rs_next = 0.f;
    if constexpr (kI8) {
      const uint32_t r0 = j * kRowsTile + rl;
      rs_next = (j < a.num_tiles && r0 < a.n_rows) ? __ldg(a.scales + r0) : 0.f;
    }
    for (uint32_t tile = j; tile < a.num_tiles; tile += P, ++it) {
      const uint32_t buf = it & (nbuf - 1u), bphase = (it >> nbuf_log2) & 1;
      const uint32_t row = tile * kRowsTile + rl;
      const bool row_ok = row < a.n_rows;
      if (tid == 64u && arrived_round < pub_round) {  // poll issued here, looked at after the tile: its latency is hidden
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(pcur) : "l"(a.seed_ws + arrived_round) : "memory");
        polled = true;
      }
      // int8 row scale: loaded one tile ahead (under a saturated HBM stream a global load takes ~1000 cycles, which the
      // first multiply of the tile's offers would otherwise wait for)
      float rs = 1.f;
      if constexpr (kI8) {
        rs = rs_next;
        const uint32_t nrow = row + P * kRowsTile;
        rs_next = (tile + P < a.num_tiles && nrow < a.n_rows) ? __ldg(a.scales + nrow) : 0.f;
      }
      mbar_wait(&tfull[buf], bphase);
      tc::fence_after_thread_sync();
      const uint32_t acc = lane_addr + buf * a.nb;
      uint64_t pend_lo = 0ull, pend_hi = 0ull;  // queries whose append found the list full (bit q)

      // An append takes a slot of the query's list (one shared-memory atomic per admitted row; a warp-aggregated variant -
      // one atomic per warp and query, slots handed out by ballot - measured 20 % slower at 64 queries, interleaved on
      // one box); one that finds the list full stays pending and triggers a maintenance round (below).
      auto append = [&](uint32_t q, float s) {
        const uint32_t slot = atom_add_shared(cnt_u32 + q * 4u, 1u);
        if (slot < a.cap) sts_u64(keys_u32 + (q * a.cap + slot) * 8u, make_key(s, row));
        else if (q < 64u) pend_lo |= 1ull << q;
        else pend_hi |= 1ull << (q - 64u);
      };

      // ---- offers. Fast path, branch free: thresholds / scales of 8 queries at a time come from shared memory with
      // back-to-back broadcast loads, the compares build a bit mask; the append path runs only for set bits (rare
      // after the first tiles: P(admit) ~ k' / rows seen).
#pragma unroll 1
      for (uint32_t g = hf; g < ngroups; g += 2) {
        uint32_t v[16], w[16];
        tmem_ld16(acc + g * 16u, v);
        if constexpr (kI8) tmem_ld16(acc + a.bp + g * 16u, w);
        float s[16];
        uint32_t mask = 0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float4 p[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) p[i] = lds_f4(par_u32 + (g * 16u + (uint32_t)(h * 8 + i)) * 16u);
          if (h == 0) tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            s[h * 8 + i] = score_of(p[i], v[h * 8 + i], kI8 ? w[h * 8 + i] : 0u, rs);
            mask |= (s[h * 8 + i] >= p[i].z) ? (1u << (h * 8 + i)) : 0u;
          }
        }
        if (!row_ok) mask = 0u;
        if (mask) {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (mask & (1u << i)) append(g * 16u + (uint32_t)i, s[i]);
        }
      }

      // ---- list maintenance, only when an append is pending (its list was full). A round costs every epilogue warp two
      // barriers, so it must be rare and do parallel work: ALL lists that are at least half full are compacted in it (warp
      // w takes lists w, w + 8, ...), which raises their thresholds; then the pending appends are retried from TMEM.
      // (Queueing every list for a round of its own as soon as it was half full made a round nearly every tile from 48
      // queries on - one warp compacting one list, seven waiting: INT8 64 queries 0.60 ms against 0.35 ms with the appends
      // switched off.)
      for (;;) {
        if (epi_bar_count((pend_lo | pend_hi) != 0ull) == 0u) break;
        for (uint32_t q = ew; q < a.nq; q += kRowsEpiWarps)
          if (lds_u32(cnt_u32 + q * 4u) >= half_cap) compact(q);
        named_bar_sync(1, kRowsEpiThreads);  // compactions visible
#pragma unroll 1
        for (uint32_t g = hf; g < ngroups; g += 2) {
          const uint32_t bits = (uint32_t)(((g < 4u) ? (pend_lo >> (g * 16u)) : (pend_hi >> ((g - 4u) * 16u))) & 0xFFFFull);
          if (!__any_sync(0xffffffffu, bits != 0u)) continue;  // tcgen05.ld is warp-collective: a warp-uniform decision
          uint32_t v[16], w[16];
          tmem_ld16(acc + g * 16u, v);
          if constexpr (kI8) tmem_ld16(acc + a.bp + g * 16u, w);
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            if (bits & (1u << i)) {
              const uint32_t q = g * 16u + (uint32_t)i;
              if (q < 64u) pend_lo &= ~(1ull << q);
              else pend_hi &= ~(1ull << (q - 64u));
              const float4 p = lds_f4(par_u32 + q * 16u);
              const float s = score_of(p, v[i], kI8 ? w[i] : 0u, rs);
              if (s >= p.z) append(q, s);  // sets the pending bit again when the list is still full
            }
          }
        }
      }
      tc::fence_before_thread_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);

      // ---- threshold seeding across the CTAs. A CTA sees 1/P of the rows, so its own k'-th best stays far below the
      // final one and (k' ln) rows per query keep passing it. In round r (after tile 0, 8, 64 of the CTA) every CTA
      // publishes, per query, the best score it has seen so far; a bound computed from the P published values (below) is a
      // lower bound of the final k'-th best, about as tight as the k'-th best of ALL rows the grid has seen by then. Each
      // round cuts the admission rate ~8x: after the first tile a query admits ~1 row per CTA instead of ~k' ln(rows).
      // Nobody waits and there is no CTA barrier in it: a warp publishes its own queries (the lists are stable below their
      // counts between maintenance rounds; a slot another warp has reserved but not yet written holds an older key of the
      // same query or zero, both harmless for a maximum that only has to be reached by one real row), the eighth warp
      // to finish bumps the round's arrival counter, thread 64 polls it with a load issued at the top of the next tiles,
      // and every warp adopts the bound for its own queries when it sees the flag. Epochs (seed_target) make the counters
      // reusable without a memset.
      if (nrounds != 0u) {
        if (pub_round < nrounds && it == pub_it) {
          uint32_t* ws_val = a.seed_ws + 16u + pub_round * (P * a.bp);
          for (uint32_t q = ew; q < a.nq; q += kRowsEpiWarps) {
            const uint32_t c0 = lds_u32(cnt_u32 + q * 4u), c = c0 < a.cap ? c0 : a.cap;
            uint32_t best = 0u;
            for (uint32_t i = lane; i < c; i += 32) {
              const uint32_t h = (uint32_t)(lds_u64(keys_u32 + (q * a.cap + i) * 8u) >> 32);
              best = h > best ? h : best;
            }
            best = __reduce_max_sync(0xffffffffu, best);
            if (lane == 0) ws_val[j * a.bp + q] = best;
          }
          if (lane == 0) {
            __threadfence();
            if (atom_add_shared(wl_u32 + 12u, 1u) + 1u == kRowsEpiWarps * (pub_round + 1u)) atomicAdd(a.seed_ws + pub_round, 1u);
          }
          ++pub_round;
          pub_it = pub_it == 0u ? 8u : pub_it * 8u;
        }
        if (tid == 64u && polled) {
          polled = false;
          if ((int32_t)(pcur - a.seed_target) >= 0) sts_u32(wl_u32 + 8u, ++arrived_round);
        }
        const uint32_t avail = __shfl_sync(0xffffffffu, lds_u32(wl_u32 + 8u), 0);
        if (my_round < avail) {
          const uint32_t* ws_val = a.seed_ws + 16u + my_round * (P * a.bp);
          ++my_round;
          // bound of query q from the P published maxima. k' <= 32 (the usual case): the CTAs form 16 or 32 groups (CTA
          // c -> group c mod 16 / 32 = its lane mod 16 / 32), the smallest group maximum is reached by one row of every
          // group, i.e. by >= k' distinct rows: one shuffle and one warp reduction per query. Larger k': the exact k'-th
          // largest of the P values by a 32-step bit search. Four queries per pass so that their L2 loads overlap.
          for (uint32_t q0 = ew * 4u; q0 < a.nq; q0 += kRowsEpiWarps * 4u) {
            uint32_t h[4][8];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t q = q0 + (uint32_t)u < a.nq ? q0 + (uint32_t)u : a.nq - 1u;
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const uint32_t l = lane + 32u * (uint32_t)i;
                h[u][i] = l < P ? __ldcg(ws_val + l * a.bp + q) : 0u;
              }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t q = q0 + (uint32_t)u;
              if (q >= a.nq) break;  // warp-uniform
              uint32_t ans = 0u;
              if (a.kp <= 32u && P >= 32u) {
                uint32_t m = 0u;
#pragma unroll
                for (int i = 0; i < 8; ++i) m = h[u][i] > m ? h[u][i] : m;
                if (a.kp <= 16u) {
                  const uint32_t o = __shfl_xor_sync(0xffffffffu, m, 16);
                  m = o > m ? o : m;
                }
                ans = __reduce_min_sync(0xffffffffu, m);
              } else {  // the largest v with at least kp published values >= v: the kp-th largest of them
#pragma unroll 1
                for (int bit = 31; bit >= 0; --bit) {
                  const uint32_t cand = ans | (1u << bit);
                  uint32_t n = 0;
#pragma unroll
                  for (int i = 0; i < 8; ++i) n += h[u][i] >= cand ? 1u : 0u;
                  if (__reduce_add_sync(0xffffffffu, n) >= a.kp) ans = cand;
                }
              }
              if (lane == 0 && ans != 0u) {  // this warp is the only one that adopts for q; a compaction only raises too
                const float t = ordered_to_float(ans);
                float cur;
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(cur) : "r"(par_u32 + q * 16u + 8u));
                if (t > cur) sts_f32(par_u32 + q * 16u + 8u, t);
              }
            }
          }
        }
      }
    }

    // ---- this CTA's best k' per query -> partial[q][cta][k'] (sorted descending, 0 = empty)
    named_bar_sync(1, kRowsEpiThreads);
    for (uint32_t q = ew; q < a.nq; q += kRowsEpiWarps) {
      compact(q);
      const uint32_t nc = lds_u32(cnt_u32 + q * 4u);
      uint64_t* out = a.partial + ((size_t)q * a.lists + j) * a.kp;
      for (uint32_t jj = lane; jj < a.kp; jj += 32) out[jj] = jj < nc ? lds_u64(keys_u32 + (q * a.cap + jj) * 8u) : 0ull;
    }
  }

  tc::fence_before_thread_sync();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace

// Shapes this kernel takes: nq queries whose limb rows (+ candidate lists) fit next to a >= 3-deep ring of 32 KB stages.
bool scan_rows_plan(int dtype, uint32_t nq, uint32_t kp, uint32_t dim, int max_smem, RowsPlan* out) {
  if (dtype != kDtF16 && dtype != kDtI8) return false;
  if (nq == 0 || nq > 128u) return false;
  const uint32_t es = dtype == kDtF16 ? 2u : 1u;
  const uint32_t bp = (nq + 15u) & ~15u;
  const uint32_t nb = dtype == kDtF16 ? bp : 2u * bp;
  if (nb > kAccColsPerBuf) return false;
  uint32_t cap = 64;
  while (cap < 4u * kp) cap <<= 1;  // a list is compacted when half full: at least k' free slots afterwards
  if (cap > 4096u) return false;
  const uint32_t kchunks = (dim * es + 127u) / 128u;
  // Ring depth: 4 x 32 KB in flight per SM already covers the HBM latency, and a deeper ring is SLOWER (FP16 8 queries:
  // 6 stages 0.721 ms, 5 0.697, 4 0.665, 3 0.674; INT8 0.376 / 0.373 / 0.365 / 0.375: profiles/r2_ring_depth_sweep.jsonl) -
  // the requests of 148 SMs then spread over more DRAM pages at a time. NVDB_B200_ROWS_MAX_STAGES=3..6 overrides (A/B).
  int st_max = 4;
  if (const char* e = getenv("NVDB_B200_ROWS_MAX_STAGES")) {
    const int v = atoi(e);
    if (v >= 3 && v <= (int)kRowsMaxStages) st_max = v;
  }
  for (int st = st_max; st >= 3; --st) {
    if (rows_layout(nb, bp, cap, kchunks, st).total <= (uint32_t)max_smem) {
      out->bp = bp;
      out->nb = nb;
      out->cap = cap;
      out->stages = st;
      out->smem_bytes = rows_layout(nb, bp, cap, kchunks, st).total;
      return true;
    }
  }
  return false;
}

cudaError_t scan_rows_launch(int dtype, const CUtensorMap* tmap_base128, const RowsArgs& a, int grid, cudaStream_t stream) {
  const RowsSmem L = rows_layout(a.nb, a.bp, a.cap, a.kchunks, a.stages);
  cudaError_t e;
  if (dtype == kDtF16) {
    auto kern = scan_rows_kernel<kDtF16>;
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total)) != cudaSuccess) return e;
    kern<<<grid, kRowsThreads, L.total, stream>>>(*tmap_base128, a);
  } else if (dtype == kDtI8) {
    auto kern = scan_rows_kernel<kDtI8>;
    if ((e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total)) != cudaSuccess) return e;
    kern<<<grid, kRowsThreads, L.total, stream>>>(*tmap_base128, a);
  } else {
    return cudaErrorInvalidValue;
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace nvdb_b200
