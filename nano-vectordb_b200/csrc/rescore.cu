// rescore.cu — exact FP32 re-score + certificate for the candidates of the tensor-core scan.
//
// The tensor-core scan (scan_mma.cu) ranks rows by an APPROXIMATE score (fp16 / int8 query limbs). This file turns
// its output into the reference's exact answer (FlatIndex::search_topk_dot, src/flat_index.cpp:16-48):
//
//   1. the scan keeps k' > k candidates per query (ranked by approximate score);
//   2. rescore_exact_kernel recomputes each candidate's score with the reference arithmetic (dot_chains.cuh: the AVX2
//      kernels' FP32 FMA chain order, bit for bit), sorts by (exact score desc, id asc) and emits the best k;
//   3. certificate: every row that is NOT a candidate has approximate score <= a' (the smallest candidate's), hence
//      exact score <= a' + margin, where margin bounds |approx - exact| for this query. If a' + margin < (k-th exact
//      score) no outsider can enter — or tie into — the top k, so the answer is exactly the reference's. Otherwise the
//      query is flagged and the caller re-runs it on the bit-exact streaming scan (capi.cu).
//
// Error bounds (rows b, query q, D = dim; all norms over one row / the query):
//   f16, one fp16 limb: hi = RN16(q * 2^e); delta = q - hi * 2^-e is EXACT in FP32 and recomputed here per query, so
//     |approx - real| <= |delta|_2 |b|_2 (Cauchy-Schwarz; ~3x tighter than the worst case 2^-11 |q|_2 |b|_2);
//     tensor-core FP32 accumulation <= D 2^-23 |q|_2 |b|_2 (one ulp of the running sum per addition, any order);
//     reference chain rounding <= (D/8 + 4) 2^-24 |q|_2 |b|_2.
//   i8, two int8 limbs: |q_i - s1 q1_i - s2 q2_i| <= s2/2, so |approx - real| <= (s2/2) * scale_b * |b|_1; FP32
//     evaluation of fmaf(s2, I2, s1 I1) * scale and the reference chain together <= (D/8 + 8) 2^-24 |q|_2 scale_b |b|_2.
//   max over rows of |b|_2 and |b|_1 (times the row scale for int8) is computed once per index (row_norm_max_kernel).
#include <math.h>

#include "common.cuh"
#include "dot_chains.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

constexpr int kRescoreThreads = 128;

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float m = red[0];
  for (uint32_t w = 1; w < (blockDim.x >> 5); ++w) m = fmaxf(m, red[w]);
  __syncthreads();
  return m;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
  for (uint32_t w = 0; w < (blockDim.x >> 5); ++w) s += red[w];
  __syncthreads();
  return s;
}

// exact score of one row, reference arithmetic. Rows are whole 16-byte pieces (the tensor-core scan's eligibility rule).
template <int DT>
__device__ __forceinline__ float exact_dot_row(const uint8_t* __restrict__ rowp, uint32_t row_bytes, uint32_t sq_u32) {
  constexpr int P16 = Elem<DT>::per16;
  uint64_t acc[1][4] = {{0ull, 0ull, 0ull, 0ull}};
  const uint32_t pieces = row_bytes >> 4;
  uint32_t p = 0;
  for (; p + 4 <= pieces; p += 4) {  // four 16-byte loads in flight
    uint4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = ldg128_stream(rowp + (size_t)(p + u) * 16u);
#pragma unroll
    for (int u = 0; u < 4; ++u) consume16<DT, 1>(acc, v[u], sq_u32 + (p + u) * (P16 * 4u), 0u, (int)((p + u) & 1u));
  }
  for (; p < pieces; ++p) {
    const uint4 v = ldg128_stream(rowp + (size_t)p * 16u);
    consume16<DT, 1>(acc, v, sq_u32 + p * (P16 * 4u), 0u, (int)(p & 1u));
  }
  return hsum8_ref_order(acc[0]);
}

template <int DT>
__global__ void __launch_bounds__(kRescoreThreads)
rescore_exact_kernel(const RescoreArgs a) {
  extern __shared__ __align__(16) uint8_t rs_smem[];
  float* sq = reinterpret_cast<float*>(rs_smem);                                   // [dim]
  uint64_t* keys = reinterpret_cast<uint64_t*>(rs_smem + ((a.dim * 4u + 15u) & ~15u));  // [bufn]
  __shared__ float red[kRescoreThreads / 32];
  __shared__ float s_amin;
  const uint32_t q = blockIdx.x, tid = threadIdx.x;
  const float* qsrc = a.queries + (size_t)q * a.dim;
  float sumsq = 0.f, amax = 0.f;
  for (uint32_t i = tid; i < a.dim; i += kRescoreThreads) {
    const float v = qsrc[i];
    sq[i] = v;
    sumsq = fmaf(v, v, sumsq);
    amax = fmaxf(amax, fabsf(v));
  }
  sumsq = block_reduce_sum(sumsq, red);  // also orders the sq[] writes before the reads below
  amax = block_reduce_max(amax, red);
  float dsq = 0.f;  // f16: |delta|_2^2 of the one-limb query split (same arithmetic as prep_queries_f16_kernel)
  if constexpr (DT == kDtF16) {
    int e = 0;
    if (amax > 0.f && amax < INFINITY) {
      e = -(ilogbf(amax) + 1);
      e = e > 100 ? 100 : (e < -100 ? -100 : e);
    }
    for (uint32_t i = tid; i < a.dim; i += kRescoreThreads) {
      const float v = scalbnf(sq[i], e);
      const float dl = scalbnf(__fsub_rn(v, __half2float(__float2half_rn(v))), -e);  // exact
      dsq = fmaf(dl, dl, dsq);
    }
    dsq = block_reduce_sum(dsq, red);
  }

  const uint32_t bufn = next_pow2(a.kp);
  const uint32_t sq_u32 = smem_u32(sq);
  const uint32_t row_bytes = a.dim * (uint32_t)Elem<DT>::size;
  const uint64_t* cid = a.cand_ids ? a.cand_ids + (size_t)q * a.kp : nullptr;
  const float* capx = a.cand_approx ? a.cand_approx + (size_t)q * a.kp : nullptr;
  // Fused merge (a.partial): the candidates are the kp best keys of this query's per-CTA lists [nlists][kp] (each sorted
  // descending, 0 = empty) — selected here instead of by a separate merge launch. Every list's last entry is a lower
  // bound of the final kp-th best key, so only keys >= the largest of them can matter: those few survivors are sorted.
  uint64_t* sel = nullptr;
  if (a.partial) {
    __shared__ unsigned long long s_floor;
    __shared__ uint32_t s_nsel;
    sel = reinterpret_cast<uint64_t*>(rs_smem + a.sel_off);
    const uint64_t* part = a.partial + (size_t)q * a.nlists * a.kp;
    const uint32_t total = a.nlists * a.kp;
    if (tid == 0) {
      s_floor = 0ull;
      s_nsel = 0u;
    }
    __syncthreads();
    // two lower bounds of the final kp-th best key: the largest of the lists' LAST entries (that list alone holds kp keys
    // at least as good), and — much tighter when there are many lists — the kp-th largest of the lists' FIRST entries
    // (kp distinct keys at least as good); with 148 lists of 16 the first leaves ~1200 survivors, the second ~2 kp.
    uint64_t fl = 0ull;
    for (uint32_t l = tid; l < a.nlists; l += kRescoreThreads) {
      const uint64_t v = part[(size_t)l * a.kp + (a.kp - 1u)];
      fl = v > fl ? v : fl;
      sel[l] = part[(size_t)l * a.kp];
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const uint64_t o = __shfl_xor_sync(0xffffffffu, fl, off);
      fl = o > fl ? o : fl;
    }
    if ((tid & 31u) == 0u && fl) atomicMax(&s_floor, (unsigned long long)fl);
    __syncthreads();
    if (a.nlists >= a.kp) {
      for (uint32_t l = tid; l < a.nlists; l += kRescoreThreads) {
        const uint64_t h = sel[l];
        if (h == 0ull) continue;
        uint32_t rank = 0;
        for (uint32_t m = 0; m < a.nlists; ++m) rank += sel[m] > h ? 1u : 0u;
        if (rank == a.kp - 1u) atomicMax(&s_floor, (unsigned long long)h);   // keys are unique: at most one thread
      }
    }
    __syncthreads();
    const uint64_t floor_key = s_floor;
    for (uint32_t i = tid; i < total; i += kRescoreThreads) {
      const uint64_t v = part[i];
      if (v != 0ull && v >= floor_key) sel[atomicAdd(&s_nsel, 1u)] = v;   // <= total entries: the buffer holds them all
    }
    __syncthreads();
    const uint32_t nsel = s_nsel, seln = next_pow2(nsel < 2u ? 2u : nsel);
    for (uint32_t i = nsel + tid; i < seln; i += kRescoreThreads) sel[i] = 0ull;
    __syncthreads();
    auto ssync = [] { __syncthreads(); };
    bitonic_sort_desc(sel, seln, tid, kRescoreThreads, ssync);
    for (uint32_t i = seln + tid; i < bufn; i += kRescoreThreads) sel[i] = 0ull;   // sel[c] is read for c < bufn
    __syncthreads();
  }
  // candidate c of this query: local row (0xFFFFFFFF = none) and approximate score
  auto candidate = [&](uint32_t c, float* apx) -> uint32_t {
    if (c >= a.kp) return 0xFFFFFFFFu;
    if (sel) {
      const uint64_t key = sel[c];
      if (key == 0ull) return 0xFFFFFFFFu;
      *apx = key_score(key);
      return key_row(key);
    }
    const uint64_t raw = cid[c];
    const uint64_t row64 = raw - a.cand_offset;
    if (raw == ~0ull || raw < a.cand_offset || row64 >= (uint64_t)a.n_rows) return 0xFFFFFFFFu;
    *apx = capx[c];
    return (uint32_t)row64;
  };
  float amin = INFINITY;  // smallest approximate score among the valid candidates
  if (a.stage_rows) {
    // Candidate rows are scattered over the base: a thread walking one row alone pays a DRAM round trip per 128-byte
    // line, twelve in a row for 1536 bytes. Instead the CTA first stages 16 candidate rows in shared memory with every
    // thread's loads in flight at once, then 8 lanes share a candidate, lane j running chain j of the reference (elements
    // j, j + 8, ... ascending: src/simd_dot.cpp:26-49,102-124,160-199) and the fixed horizontal sum runs over shuffles.
    uint8_t* rows = rs_smem + ((a.dim * 4u + 15u) & ~15u) + ((bufn * 8u + 15u) & ~15u);  // [16][row_bytes + 16]
    __shared__ uint32_t s_row[16];
    __shared__ float s_apx[16];
    const uint32_t rstride = row_bytes + 16u, ppr = row_bytes >> 4;
    const uint32_t grp = tid >> 3, jl = tid & 7u;
    for (uint32_t c0 = 0; c0 < bufn; c0 += 16u) {
      if (tid < 16u) {
        float apx = 0.f;
        s_row[tid] = candidate(c0 + tid, &apx);
        s_apx[tid] = apx;
      }
      __syncthreads();
      for (uint32_t pc = tid; pc < 16u * ppr; pc += kRescoreThreads) {
        const uint32_t c = pc / ppr, p16 = pc - c * ppr;
        const uint32_t row = s_row[c];
        if (row != 0xFFFFFFFFu) {
          const uint4 v = ldg128_stream(reinterpret_cast<const uint8_t*>(a.base) + (size_t)row * row_bytes + (size_t)p16 * 16u);
          *reinterpret_cast<uint4*>(rows + c * rstride + p16 * 16u) = v;
        }
      }
      __syncthreads();
      {
        const uint32_t row = s_row[grp];
        float acc = 0.f;
        if (row != 0xFFFFFFFFu) {
          const uint8_t* rp = rows + grp * rstride;
          for (uint32_t i = jl; i < a.dim; i += 8u) {
            float x;
            if constexpr (DT == kDtF16) x = __half2float(reinterpret_cast<const __half*>(rp)[i]);
            else if constexpr (DT == kDtI8) x = (float)reinterpret_cast<const int8_t*>(rp)[i];
            else x = reinterpret_cast<const float*>(rp)[i];
            acc = __fmaf_rn(sq[i], x, acc);
          }
        }
        // ((a0+a4)+(a1+a5))+((a2+a6)+(a3+a7)): lanes j and j^4, then j^1, then j^2 (FP32 addition commutes exactly)
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 4));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 1));
        acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, 2));
        if (jl == 0) {
          uint64_t key = 0ull;
          if (row != 0xFFFFFFFFu) {
            float sc = acc;
            if constexpr (DT == kDtI8) sc = __fmul_rn(sc, __ldg(a.scales + row));  // src/simd_dot.cpp:198
            key = make_key(sc, row);
            amin = fminf(amin, s_apx[grp]);
          }
          if (c0 + grp < bufn) keys[c0 + grp] = key;
        }
      }
      __syncthreads();
    }
  } else {
  for (uint32_t c = tid; c < bufn; c += kRescoreThreads) {
    uint64_t key = 0ull;
    float apx = 0.f;
    const uint32_t row = candidate(c, &apx);
    if (row != 0xFFFFFFFFu) {
      float s = exact_dot_row<DT>(reinterpret_cast<const uint8_t*>(a.base) + (size_t)row * row_bytes, row_bytes, sq_u32);
      if constexpr (DT == kDtI8) s = __fmul_rn(s, __ldg(a.scales + row));  // src/simd_dot.cpp:198
      key = make_key(s, row);
      amin = fminf(amin, apx);
    }
    keys[c] = key;
  }
  }
  __syncthreads();
  auto sync = [] { __syncthreads(); };
  bitonic_sort_desc(keys, bufn, tid, kRescoreThreads, sync);
  amin = -block_reduce_max(-amin, red);
  // count valid candidates (keys != 0 after the sort are contiguous at the front)
  uint32_t cnt = 0;
  for (uint32_t c = tid; c < bufn; c += kRescoreThreads) cnt += keys[c] != 0ull ? 1u : 0u;
  const float cntf = block_reduce_sum((float)cnt, red);  // kp <= 8192: exact in float
  if (tid == 0) s_amin = amin;
  __syncthreads();

  for (uint32_t j = tid; j < a.k; j += kRescoreThreads) {
    const uint64_t key = (j < bufn) ? keys[j] : 0ull;
    const bool ok = key != 0ull;
    a.out_ids[(size_t)q * a.k + j] = ok ? a.row_offset + (uint64_t)key_row(key) : ~0ull;
    a.out_scores[(size_t)q * a.k + j] = ok ? key_score(key) : -INFINITY;
  }

  if (tid == 0) {
    bool pass;
    const uint32_t valid = (uint32_t)cntf;
    if (valid < a.kp) {
      pass = true;  // the scan returned every row it saw: nothing was excluded
    } else {
      const float qn = sqrtf(sumsq) * 1.000001f;
      float margin;
      if constexpr (DT == kDtI8) {
        const float s1 = (amax > 0.f && amax < INFINITY) ? __fdiv_rn(amax, 127.f) : 1.f;  // scan_mma.cu prep
        const float s2 = __fdiv_rn(s1, 254.f);
        margin = 0.5f * s2 * a.max_l1 * 1.001f + a.rel_err * qn * a.max_l2;
      } else {
        margin = (sqrtf(dsq) * 1.00001f + a.rel_err * qn) * a.max_l2;
      }
      margin = margin * 1.0001f + 1e-37f;
      const uint32_t kk = a.k < valid ? a.k : valid;
      const float ek = key_score(keys[kk - 1]);
      const float bound = __fadd_ru(s_amin, margin);
      pass = (bound < ek) && isfinite(ek) && isfinite(bound);
    }
    a.flags[q] = pass ? 0u : 1u;
    if (!pass) atomicAdd(a.nflag, 1u);
    if (q == 0 && a.nflag_next) *a.nflag_next = 0u;  // the counter of the NEXT call on this stream (no memset launch)
    if (a.host_nflag) {  // the last CTA hands the verdict to the host itself: no copy-engine round trip behind the kernel
      __threadfence();
      if (atomicAdd(a.done_ctr, 1u) == gridDim.x - 1u) {
        __threadfence();
        const uint32_t c = atomicAdd(a.nflag, 0u);
        *a.done_ctr = 0u;
        *reinterpret_cast<volatile uint32_t*>(a.host_nflag) = c;
      }
    }
  }
}

// max over rows of |row|_2 and |row|_1 (times the row scale for int8), rounded up. One warp per row.
template <int DT>
__global__ void __launch_bounds__(256) row_norm_max_kernel(const void* __restrict__ base, const float* __restrict__ scales,
                                                           uint64_t n, uint32_t dim, uint32_t* __restrict__ out2) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  float m2 = 0.f, m1 = 0.f;
  for (uint64_t row = warp0; row < n; row += nwarps) {
    float s2 = 0.f, s1 = 0.f;
    if constexpr (DT == kDtF16) {
      const __half* r = reinterpret_cast<const __half*>(base) + row * dim;
      for (uint32_t i = lane; i < dim; i += 32) {
        const float v = __half2float(r[i]);
        s2 = fmaf(v, v, s2);
        s1 += fabsf(v);
      }
    } else if constexpr (DT == kDtI8) {
      const int8_t* r = reinterpret_cast<const int8_t*>(base) + row * dim;
      for (uint32_t i = lane; i < dim; i += 32) {
        const float v = (float)r[i];
        s2 = fmaf(v, v, s2);
        s1 += fabsf(v);
      }
    } else {
      const float* r = reinterpret_cast<const float*>(base) + row * dim;
      for (uint32_t i = lane; i < dim; i += 32) {
        const float v = r[i];
        s2 = fmaf(v, v, s2);
        s1 += fabsf(v);
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      s2 += __shfl_xor_sync(0xffffffffu, s2, off);
      s1 += __shfl_xor_sync(0xffffffffu, s1, off);
    }
    float l2 = sqrtf(s2), l1 = s1;
    if constexpr (DT == kDtI8) {
      const float sc = fabsf(scales[row]);
      l2 *= sc;
      l1 *= sc;
    }
    m2 = fmaxf(m2, l2);
    m1 = fmaxf(m1, l1);
  }
  if (lane == 0) {  // non-negative floats order like their bit patterns; NaN/inf propagate as "huge" -> always flagged
    atomicMax(out2 + 0, __float_as_uint(m2 * 1.0001f));
    atomicMax(out2 + 1, __float_as_uint(m1 * 1.0001f));
  }
}

// flagged queries of the exact path: gather their FP32 rows into a dense batch / scatter the re-run results back.
// One CTA per listed query; `list` holds indices into the source batch (ascending).
__global__ void __launch_bounds__(128)
gather_queries_kernel(const float* __restrict__ src, const uint32_t* __restrict__ list, uint32_t dim, float* __restrict__ dst) {
  const float* s = src + (size_t)list[blockIdx.x] * dim;
  float* d = dst + (size_t)blockIdx.x * dim;
  for (uint32_t i = threadIdx.x; i < dim; i += 128u) d[i] = s[i];
}
__global__ void __launch_bounds__(128)
scatter_topk_kernel(const uint64_t* __restrict__ src_ids, const float* __restrict__ src_scores,
                    const uint32_t* __restrict__ list, uint32_t k, uint64_t* __restrict__ dst_ids,
                    float* __restrict__ dst_scores) {
  const size_t so = (size_t)blockIdx.x * k, dof = (size_t)list[blockIdx.x] * k;
  for (uint32_t i = threadIdx.x; i < k; i += 128u) {
    dst_ids[dof + i] = src_ids[so + i];
    dst_scores[dof + i] = src_scores[so + i];
  }
}

}  // namespace

cudaError_t gather_queries_launch(const float* src, const uint32_t* list, uint32_t n, uint32_t dim, float* dst,
                                  cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  gather_queries_kernel<<<n, 128, 0, stream>>>(src, list, dim, dst);
  count_launch();
  return cudaGetLastError();
}

cudaError_t scatter_topk_launch(const uint64_t* src_ids, const float* src_scores, const uint32_t* list, uint32_t n,
                                uint32_t k, uint64_t* dst_ids, float* dst_scores, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  scatter_topk_kernel<<<n, 128, 0, stream>>>(src_ids, src_scores, list, k, dst_ids, dst_scores);
  count_launch();
  return cudaGetLastError();
}

cudaError_t rescore_exact_launch(int dtype, const RescoreArgs& a, uint32_t nq, cudaStream_t stream) {
  if (nq == 0 || a.k == 0) return cudaSuccess;
  size_t smem = ((a.dim * 4u + 15u) & ~15u) + (((size_t)next_pow2(a.kp) * 8u + 15u) & ~(size_t)15u);
  if (smem > 200u * 1024u) return cudaErrorInvalidValue;
  // staged form: whole 8-chains only (the shapes the tensor-core scans take), 16 rows of the base next to the query
  RescoreArgs b = a;
  const uint32_t es = dtype == kDtF32 ? 4u : dtype == kDtF16 ? 2u : 1u;
  const size_t stage = 16u * ((size_t)a.dim * es + 16u);
  b.stage_rows = (a.dim % 8u == 0u) && ((a.dim * es) % 16u == 0u) && (smem + stage <= 160u * 1024u) ? 1 : 0;
  if (b.stage_rows) smem += stage;
  if (a.partial) {  // selection buffer of the fused merge: every key of the lists in the worst case
    const size_t seln = next_pow2(a.nlists * a.kp < 2u ? 2u : a.nlists * a.kp);
    const size_t need = seln > next_pow2(a.kp) ? seln : next_pow2(a.kp);
    b.sel_off = (uint32_t)smem;
    smem += need * 8u;
    if (smem > 200u * 1024u) return cudaErrorInvalidValue;
  }
  auto launch = [&](auto kern) -> cudaError_t {
    if (smem > 48u * 1024u) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    kern<<<nq, kRescoreThreads, smem, stream>>>(b);
    count_launch();
    return cudaGetLastError();
  };
  switch (dtype) {
    case kDtF16: return launch(rescore_exact_kernel<kDtF16>);
    case kDtI8: return launch(rescore_exact_kernel<kDtI8>);
    case kDtF32: return launch(rescore_exact_kernel<kDtF32>);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t row_norm_max_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim, uint32_t* out2,
                                int num_sms, cudaStream_t stream) {
  cudaError_t e = cudaMemsetAsync(out2, 0, 8, stream);
  if (e != cudaSuccess) return e;
  const int grid = num_sms * 8;
  switch (dtype) {
    case kDtF16: row_norm_max_kernel<kDtF16><<<grid, 256, 0, stream>>>(base, scales, n, dim, out2); break;
    case kDtI8: row_norm_max_kernel<kDtI8><<<grid, 256, 0, stream>>>(base, scales, n, dim, out2); break;
    case kDtF32: row_norm_max_kernel<kDtF32><<<grid, 256, 0, stream>>>(base, scales, n, dim, out2); break;
    default: return cudaErrorInvalidValue;
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace nvdb_b200
