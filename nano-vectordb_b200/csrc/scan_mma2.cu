// scan_mma2.cu — CTA-pair (tcgen05 cta_group::2) variant of the tensor-core flat scan for f16 bases.
//
// Same job as scan_mma.cu (the reference's bench-side batched scan, apps/nvdb_bench.cpp:47-159, with top-k fused into
// the epilogue), same orientation (queries on M / TMEM lanes, base rows on N / TMEM columns), but two CTAs on the two
// SMs of a TPC work on one 256-query x 256-row tile:
//
//   * CTA r of the pair owns queries [ (2*mp + r) * 128, +128 ) — its A tile, its TMEM lanes, its epilogue threads;
//   * each CTA loads only HALF of the base tile (rows tile*256 + r*128 .. +128): the tensor cores of both SMs read both
//     halves, so per 128x256x64 MMA step a CTA moves 16 KB (A) + 16 KB (B/2) through L2 -> shared memory instead of
//     16 + 32 KB. The single-CTA kernel with one query limb is bound by exactly that feed (profiles/: 4.1 ms at batch
//     1024 against 3.0 ms of tensor time); this is the remedy the 2-SM MMA exists for;
//   * the leader (rank 0) issues tcgen05.mma.cta_group::2 (M = 256, N = 256, K = 16) for both; completion is
//     multicast to both CTAs' barriers; each CTA's epilogue reads its own 128 lanes x 256 columns.
//
// Barriers (per CTA unless noted):
//   full[s]   leader only, count 1 (+ tx bytes of BOTH CTAs' TMA loads, which signal the leader's barrier)
//   empty[s]  count 1, arrived by the leader's tcgen05.commit multicast (slot consumed by the pair's MMA)
//   tfull[b]  count 1, arrived by the commit multicast (accumulator buffer b complete)
//   tempty[b] leader only, count 8: the four epilogue warps of both CTAs (the peer's arrive remotely)
#include <cstdlib>

#include "common.cuh"
#include "scan.h"
#include "scan_mma_common.cuh"
#include "tc_common.cuh"

namespace nvdb_b200 {
namespace {

// tile geometry per dtype: f16 — 256 base rows per pair tile (one 256-column accumulator per buffer);
// i8 — two query limbs with separate accumulators (2 x 128 columns per buffer), so 128 base rows per pair tile.
template <int DT>
struct PairGeom {
  static constexpr uint32_t kPairN = (DT == kDtI8) ? 128u : 256u;  // base rows per tile (accumulator columns per limb)
  static constexpr uint32_t kHalfN = kPairN / 2;                   // base rows each CTA loads per tile
  static constexpr int kEpc = (DT == kDtI8) ? 128 : 64;            // elements per 128-byte K chunk
};

struct Mma2Smem {
  uint32_t stages_off, res_off, lists_off, rscale_off, bars_off, tmem_off, total, stage_bytes, a_limb_bytes;
};
// res_limbs > 0: that many leading limbs of the CTA's query tile stay resident (all `kchunks` K chunks each), the ring
// carries the remaining limbs + B/2
__host__ __device__ inline Mma2Smem mma2_layout(uint32_t half_n, int limbs, int stages, uint32_t k_smem_lists,
                                                int res_limbs, uint32_t kchunks) {
  Mma2Smem L;
  L.a_limb_bytes = kBlockM * 128u;
  L.stage_bytes = (uint32_t)(limbs - res_limbs) * L.a_limb_bytes + half_n * 128u;
  uint32_t off = 0;
  L.res_off = off;
  off += (uint32_t)res_limbs * kchunks * L.a_limb_bytes;
  L.stages_off = off;
  off += (uint32_t)stages * L.stage_bytes;
  L.lists_off = off;
  off += k_smem_lists * kBlockM * 8u;
  L.rscale_off = off;
  off += 2u * 128u * 4u;
  L.bars_off = off;
  off += (2u * (uint32_t)stages + 5u) * 8u;
  L.tmem_off = off;
  off += 16u;
  L.total = off + 1024u;
  return L;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load of this CTA's tile into ITS shared memory; the transaction bytes complete on `bar_cluster_addr`
// (the leader's barrier, a shared::cluster address).
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const void* tmap, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void mma2_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma2_i8_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all MMAs issued so far -> one arrival on the barrier at this shared-memory offset in BOTH CTAs of the pair
__device__ __forceinline__ void mma2_commit_both(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}

template <int DT, int LIMBS, int KREG, int RES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kMmaThreads, 1)
scan_mma2_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const MmaArgs a) {
  using Geo = PairGeom<DT>;
  constexpr bool kI8 = (DT == kDtI8);
  constexpr int EPC = Geo::kEpc;
  constexpr uint32_t kPairN = Geo::kPairN, kHalfN = Geo::kHalfN;
  static_assert(!kI8 || LIMBS == 2, "int8 queries are always two limbs");
  static_assert(RES <= LIMBS, "resident limbs are the leading limbs");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const Mma2Smem L = mma2_layout(kHalfN, LIMBS, a.stages, KREG > 0 ? 0u : a.k, RES, a.kchunks);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L.bars_off);
  uint64_t* empty = full + a.stages;
  uint64_t* tfull = empty + a.stages;
  uint64_t* tempty = tfull + 2;
  uint64_t* resbar = tempty + 2;  // leader only: both CTAs' resident limbs have landed (RES)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + L.tmem_off);

  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const uint32_t cid = blockIdx.x >> 1, Gp = a.G >> 1;
  const uint32_t mpair = cid % Gp, j = cid / Gp;
  const uint32_t m_tile = 2 * mpair + rank;

  if (tid == 0) {
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tfull[b], 1);
      mbar_init(&tempty[b], 8);  // four epilogue warps of each CTA of the pair
    }
    mbar_init(resbar, 1);
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc_pair(tmem_slot, kTmemCols);
  tc::fence_before_thread_sync();
  cluster_sync_all();  // barrier inits + TMEM allocation visible to both CTAs before any remote arrive / multicast
  tc::fence_after_thread_sync();
  // broadcast through a shuffle so that the compiler keeps the TMEM base in a uniform register: a value it cannot
  // prove warp-uniform costs an ELECT / R2UR.BROADCAST / BRA.U.ANY waterfall around EVERY tcgen05.mma issue
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own A tile(s) + own half of the base tile; bytes land on the leader's barrier =====
    if (lane == 0) {
      tma_prefetch_desc(&tmap_a);
      tma_prefetch_desc(&tmap_b);
      if constexpr (RES > 0) {  // the leading limb(s) of this CTA's query tile: all K chunks, once
        if (leader) mbar_arrive_expect_tx(resbar, 2u * RES * a.kchunks * L.a_limb_bytes);
        const uint32_t rb = mapa_shared(smem_u32(resbar), 0);
        for (uint32_t l = 0; l < (uint32_t)RES; ++l)
          for (uint32_t kc = 0; kc < a.kchunks; ++kc)
            tma_load_2d_pair(smem + L.res_off + (l * a.kchunks + kc) * L.a_limb_bytes, &tmap_a, rb, (int)(kc * EPC),
                             (int)(l * a.mpad + m_tile * kBlockM));
      }
    }
    __syncwarp();
    uint32_t stage = 0, phase = 0;
    for (uint32_t tile = a.tile0 + j; tile < a.num_tiles; tile += a.P) {
      for (uint32_t kc = 0; kc < a.kchunks; ++kc) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* st = smem + L.stages_off + stage * L.stage_bytes;
        const uint32_t bar = mapa_shared(smem_u32(&full[stage]), 0);
        if (tc::elect_one_sync()) {
          if (leader) mbar_arrive_expect_tx(&full[stage], 2u * L.stage_bytes);
#pragma unroll
          for (int l = RES; l < LIMBS; ++l)
            tma_load_2d_pair(st + (l - RES) * L.a_limb_bytes, &tmap_a, bar, (int)(kc * EPC),
                             (int)(l * a.mpad + m_tile * kBlockM));
          tma_load_2d_pair(st + (LIMBS - RES) * L.a_limb_bytes, &tmap_b, bar, (int)(kc * EPC),
                           (int)(tile * kPairN + rank * kHalfN));
        }
        __syncwarp();
        if (++stage == (uint32_t)a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one elected lane of the LEADER CTA's warp drives the tensor cores of both SMs =====
    if (leader) {
      constexpr uint32_t idesc = kI8 ? tc::idesc_s8_s32(2 * kBlockM, kPairN) : tc::idesc_f16_f32(2 * kBlockM, kPairN);
      uint32_t stage = 0, phase = 0, it = 0;
      if constexpr (RES > 0) {
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
              const uint64_t adesc = (l < RES) ? tc::smem_desc_k128(res_u32 + ((uint32_t)l * a.kchunks + kc) * L.a_limb_bytes)
                                               : tc::smem_desc_k128(st + (l - RES) * L.a_limb_bytes);
              const uint32_t acc = tmem_base + buf * kAccColsPerBuf + (kI8 ? (uint32_t)l * kPairN : 0u);
#pragma unroll
              for (uint32_t k4 = 0; k4 < 4; ++k4) {
                if constexpr (kI8)
                  mma2_i8_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)(kc > 0 || k4 > 0));
                else
                  mma2_f16_ss(acc, adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)(kc > 0 || l > 0 || k4 > 0));
              }
            }
            mma2_commit_both(&empty[stage]);
            if (kc + 1 == a.kchunks) mma2_commit_both(&tfull[buf]);
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
    // ===== epilogue (both CTAs): one thread per query (== TMEM lane of this CTA) =====
    const uint32_t qd = warp & 3;
    const uint32_t ql = qd * 32 + lane;
    const uint32_t et = (warp - 2) * 32 + lane;  // 0..127 among the epilogue threads
    const uint32_t qg = m_tile * kBlockM + ql;
    const bool valid = qg < a.nq;
    ThreadTopK<KREG> top;
    top.init(reinterpret_cast<uint64_t*>(smem + L.lists_off) + ql, a.k);
    const uint32_t rscale_u32 = smem_u32(smem + L.rscale_off);  // [2][128] row scales (i8)
    float thr = valid ? (a.thr_init ? a.thr_init[qg] : -INFINITY) : INFINITY;
    const float qs1 = a.qscale[qg];
    const float qs2 = kI8 ? a.qscale2[qg] : 0.f;
    const uint32_t lane_addr = tmem_base + ((qd * 32u) << 16);
    uint32_t tempty_remote[2];
    tempty_remote[0] = mapa_shared(smem_u32(&tempty[0]), 0);
    tempty_remote[1] = mapa_shared(smem_u32(&tempty[1]), 0);
    uint32_t it = 0;
    for (uint32_t tile = a.tile0 + j; tile < a.num_tiles; tile += a.P, ++it) {
      const uint32_t buf = it & 1, bphase = (it >> 1) & 1;
      const uint32_t row0 = tile * kPairN;
      if constexpr (kI8) {
        const float rs_mine = (row0 + et < a.n_rows) ? __ldg(a.scales + row0 + et) : 0.f;
        asm volatile("st.shared.f32 [%0], %1;" ::"r"(rscale_u32 + (buf * 128u + et) * 4u), "f"(rs_mine) : "memory");
      }
      mbar_wait(&tfull[buf], bphase);
      tc::fence_after_thread_sync();
      if constexpr (kI8) named_bar_sync(2, 128);
#pragma unroll 1
      for (uint32_t c0 = 0; c0 < kPairN; c0 += 32) {
        uint32_t v[32];
        float s[32];
        tc::tmem_ld32(lane_addr + buf * kAccColsPerBuf + c0, v);
        if constexpr (kI8) {
          uint32_t w[32];
          tc::tmem_ld32(lane_addr + buf * kAccColsPerBuf + kPairN + c0, w);
          tc::tmem_ld_wait();
          const uint32_t rs4 = rscale_u32 + (buf * 128u + c0) * 4u;
#pragma unroll
          for (int i4 = 0; i4 < 8; ++i4) {
            float4 r;  // warp-broadcast LDS.128
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                         : "r"(rs4 + (uint32_t)i4 * 16u));
            const float rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i = i4 * 4 + u;
              const float t = __fmul_rn(qs1, (float)(int)v[i]);  // |I| < 2^24: exact conversions
              s[i] = __fmul_rn(__fmaf_rn(qs2, (float)(int)w[i], t), rr[u]);
            }
          }
        } else {
          tc::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) s[i] = __fmul_rn(__uint_as_float(v[i]), qs1);  // power-of-two rescale: exact
        }
        uint32_t mask = 0;
#pragma unroll
        for (int i = 0; i < 32; ++i) mask |= (s[i] >= thr) ? (1u << i) : 0u;
        while (mask) {
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
      if (lane == 0) mbar_arrive_remote(tempty_remote[buf]);
    }
    if (valid) top.store(a.partial + ((size_t)qg * a.lists + j) * a.k);
  }

  tc::fence_before_thread_sync();
  cluster_sync_all();  // nobody leaves (or frees TMEM) while the peer may still signal / read
  if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

template <int DT, int LIMBS, int KREG, int RES>
cudaError_t launch_mma2_k(const CUtensorMap* ta, const CUtensorMap* tb, const MmaArgs& a, cudaStream_t stream) {
  const Mma2Smem L = mma2_layout(PairGeom<DT>::kHalfN, LIMBS, a.stages, KREG > 0 ? 0u : a.k, RES, a.kchunks);
  auto kern = scan_mma2_kernel<DT, LIMBS, KREG, RES>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total);
  if (e != cudaSuccess) return e;
  kern<<<a.G * a.P, kMmaThreads, L.total, stream>>>(*ta, *tb, a);  // (G/2 pair columns) x P clusters x 2 CTAs
  count_launch();
  return cudaGetLastError();
}

// deepest ring that fits; 0 when not even `min_stages` do
int mma2_fit_stages(uint32_t half_n, int limbs, uint32_t k, int res_limbs, uint32_t kchunks, int max_smem, int min_stages) {
  int st_max = 8;
  if (const char* e = getenv("NVDB_B200_PAIR_MAX_STAGES")) {  // ring depth experiments (A/B)
    const int v = atoi(e);
    if (v >= 2 && v < st_max) st_max = v;
  }
  if (st_max < min_stages) st_max = min_stages;
  for (int st = st_max; st >= min_stages; --st)
    if (mma2_layout(half_n, limbs, st, k <= (uint32_t)kMmaRegK ? 0u : k, res_limbs, kchunks).total <= (uint32_t)max_smem)
      return st;
  return 0;
}

}  // namespace

// rows of the base tile each CTA of the pair loads (the base tensor map's box rows) and rows per pair tile
uint32_t scan_mma2_half_n(int dtype) { return dtype == kDtI8 ? PairGeom<kDtI8>::kHalfN : PairGeom<kDtF16>::kHalfN; }

// ring depth next to the k-entry lists (k <= 16: lists in registers); false when not even 3 stages fit
bool scan_mma2_plan(int dtype, int limbs, uint32_t k, int max_smem, int* stages) {
  // smallest ring accepted: 2 stages — only reached when 3 do not fit, i.e. by the wide shared-memory lists of the second
  // certificate tier (150+ entries): the pair kernel with a 2-stage ring still beats the single-CTA kernel there
  // (clustered base, k = 100, batch 1024: 11.32 -> 10.07 ms). NVDB_B200_PAIR_MIN_STAGES=3 restores the old rule (A/B).
  static const int min_stages = [] {
    const char* e = getenv("NVDB_B200_PAIR_MIN_STAGES");
    const int v = e ? atoi(e) : 2;
    return v < 2 ? 2 : v;
  }();
  *stages = mma2_fit_stages(scan_mma2_half_n(dtype), limbs, k, 0, 0u, max_smem, min_stages);
  return *stages != 0;
}

// a.G even (pairs of 128-query tiles), a.P = clusters per pair column, tmap_b with scan_mma2_half_n(dtype)-row boxes.
// a.stages comes from scan_mma2_plan; with a.resident_a (i8, register lists) the ring is re-planned around the resident limb.
cudaError_t scan_mma2_launch(int dtype, int limbs, const CUtensorMap* tmap_a, const CUtensorMap* tmap_b, const MmaArgs& a,
                             cudaStream_t stream) {
  if ((a.G & 1u) || a.G == 0 || a.P == 0) return cudaErrorInvalidValue;
  const bool reg = a.k <= (uint32_t)kMmaRegK;
  if (dtype == kDtI8) {
    if (limbs != 2) return cudaErrorInvalidValue;
    // resident query limbs (register lists only); knob MMA_RESIDENT: 0 none, 1 (default) the first limb, 2 both when
    // the 8 KB base ring still gets >= 4 slots (dim <= 768: 192 KB + 32 KB). Both resident measured no faster than one
    // (profiles/r1_ab_i8_pair_resident.jsonl): with one limb resident the kernel is already at the tensor rate the
    // power cap allows.
    if (a.resident_a && reg) {
      MmaArgs r = a;
      if (a.resident_a == 2 && (r.stages = mma2_fit_stages(PairGeom<kDtI8>::kHalfN, 2, a.k, 2, a.kchunks, a.max_smem, 4)))
        return launch_mma2_k<kDtI8, 2, kMmaRegK, 2>(tmap_a, tmap_b, r, stream);
      if ((r.stages = mma2_fit_stages(PairGeom<kDtI8>::kHalfN, 2, a.k, 1, a.kchunks, a.max_smem, 4)))
        return launch_mma2_k<kDtI8, 2, kMmaRegK, 1>(tmap_a, tmap_b, r, stream);
    }
    return reg ? launch_mma2_k<kDtI8, 2, kMmaRegK, 0>(tmap_a, tmap_b, a, stream)
               : launch_mma2_k<kDtI8, 2, 0, 0>(tmap_a, tmap_b, a, stream);
  }
  if (limbs == 1)
    return reg ? launch_mma2_k<kDtF16, 1, kMmaRegK, 0>(tmap_a, tmap_b, a, stream)
               : launch_mma2_k<kDtF16, 1, 0, 0>(tmap_a, tmap_b, a, stream);
  if (limbs == 2)
    return reg ? launch_mma2_k<kDtF16, 2, kMmaRegK, 0>(tmap_a, tmap_b, a, stream)
               : launch_mma2_k<kDtF16, 2, 0, 0>(tmap_a, tmap_b, a, stream);
  return cudaErrorInvalidValue;
}

}  // namespace nvdb_b200
