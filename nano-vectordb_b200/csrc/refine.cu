// refine.cu — exact L2 re-rank of candidate rows (gather-bound).
//
// Replaces nvdb::cuda_l2_topk_batch's kernels (reference src/cuda_refine.cu:405-838). The reference gives one
// THREAD a whole candidate row (32 lanes -> 32 different rows, uncoalesced) and merges K-lists on thread 0
// (58 % of its kernel time, Performance_CUDA.md:267). Here:
//   * one CTA per query, one WARP per candidate row: a row is read as consecutive 16-byte pieces by
//     consecutive lanes (every request is a fully used 512-byte span), two rows in flight per warp;
//   * the query lives in shared memory; per-lane partial sums use 4 independent chains, then a fixed
//     shuffle tree — deterministic, and closer to the double-precision CPU refine
//     (apps/nvdb_ivf_eval.cpp:232-240) than the reference kernel's 4 x 192-long chains;
//   * selection = block bitonic sort of R (distance, id) keys in shared memory; K smallest written in
//     ascending order (ties: id ascending). No serial merge.
// Semantics kept: ids >= N (incl. 0xFFFFFFFF) are skipped; unfilled output slots are 0xFFFFFFFF / 1e30f
// (src/cuda_refine.cu:435-437, 891-894).
#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

constexpr int kRefineThreads = 256;
constexpr int kRefineWarps = kRefineThreads / 32;

template <int DT>
__device__ __forceinline__ void l2_piece(const uint4 v, const float* __restrict__ q, float scale, float (&acc)[4]) {
  if constexpr (DT == kDtF16) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    const float4 qa = *reinterpret_cast<const float4*>(q);
    const float4 qb = *reinterpret_cast<const float4*>(q + 4);
    const float qv[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      const float d0 = __fsub_rn(qv[2 * i], f.x), d1 = __fsub_rn(qv[2 * i + 1], f.y);
      acc[i] = __fmaf_rn(d0, d0, acc[i]);
      acc[i] = __fmaf_rn(d1, d1, acc[i]);
    }
  } else if constexpr (DT == kDtF32) {
    const float4 qa = *reinterpret_cast<const float4*>(q);
    const float x[4] = {__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)};
    const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float d = __fsub_rn(qv[i], x[i]);
      acc[i] = __fmaf_rn(d, d, acc[i]);
    }
  } else {
    // int8 rows are dequantised as float(x) * scale (one FP32 multiply), include/nvdb/to_f32_row.h:26-31
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 qa = *reinterpret_cast<const float4*>(q + 4 * i);
      const float qv[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const float x = __fmul_rn((float)(int8_t)(w[i] >> (8 * b)), scale);
        const float d = __fsub_rn(qv[b], x);
        acc[b] = __fmaf_rn(d, d, acc[b]);
      }
    }
  }
}

template <int DT>
__global__ void __launch_bounds__(kRefineThreads)
refine_l2_kernel(const uint8_t* __restrict__ base, const float* __restrict__ scales, uint64_t n, uint32_t dim,
                 const float* __restrict__ queries, const uint32_t* __restrict__ cand, uint32_t R, uint32_t K,
                 uint32_t sortn, uint32_t* __restrict__ out_ids, float* __restrict__ out_dist) {
  extern __shared__ uint64_t smem64[];
  uint64_t* skeys = smem64;                                  // [sortn]
  float* sq = reinterpret_cast<float*>(smem64 + sortn);      // [dim]
  constexpr int ES = (DT == kDtF32) ? 4 : (DT == kDtF16 ? 2 : 1);
  constexpr int P16 = 16 / ES;
  const uint32_t q = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const size_t row_bytes = (size_t)dim * ES;
  const uint32_t npieces = (uint32_t)(row_bytes >> 4);       // dim * ES is a multiple of 16 (checked by the host)

  for (uint32_t i = tid; i < dim; i += kRefineThreads) sq[i] = queries[(size_t)q * dim + i];
  for (uint32_t i = R + tid; i < sortn; i += kRefineThreads) skeys[i] = 0ull;
  __syncthreads();

  const uint32_t* cq = cand + (size_t)q * R;
  for (uint32_t r0 = warp * 2; r0 < R; r0 += kRefineWarps * 2) {
    const uint32_t idA = cq[r0];
    const uint32_t idB = (r0 + 1 < R) ? cq[r0 + 1] : 0xFFFFFFFFu;
    const bool okA = (uint64_t)idA < n, okB = (uint64_t)idB < n;
    const uint8_t* rowA = base + (size_t)(okA ? idA : 0) * row_bytes;
    const uint8_t* rowB = base + (size_t)(okB ? idB : 0) * row_bytes;
    float scA = 1.f, scB = 1.f;
    if constexpr (DT == kDtI8) {
      scA = okA ? __ldg(scales + idA) : 0.f;
      scB = okB ? __ldg(scales + idB) : 0.f;
    }
    float accA[4] = {0.f, 0.f, 0.f, 0.f}, accB[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 3
    for (uint32_t c = lane; c < npieces; c += 32) {
      uint4 va = make_uint4(0, 0, 0, 0), vb = make_uint4(0, 0, 0, 0);
      if (okA) va = ldg128_stream(rowA + (size_t)c * 16);
      if (okB) vb = ldg128_stream(rowB + (size_t)c * 16);
      l2_piece<DT>(va, sq + c * P16, scA, accA);
      l2_piece<DT>(vb, sq + c * P16, scB, accB);
    }
    float dA = __fadd_rn(__fadd_rn(accA[0], accA[1]), __fadd_rn(accA[2], accA[3]));
    float dB = __fadd_rn(__fadd_rn(accB[0], accB[1]), __fadd_rn(accB[2], accB[3]));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      dA = __fadd_rn(dA, __shfl_xor_sync(0xffffffffu, dA, off));
      dB = __fadd_rn(dB, __shfl_xor_sync(0xffffffffu, dB, off));
    }
    if (lane == 0) {
      // distance >= +0, so its bit pattern orders like the value; smaller (dist, id) -> larger stored key
      skeys[r0] = okA ? ~(((uint64_t)__float_as_uint(dA) << 32) | idA) : 0ull;
      if (r0 + 1 < R) skeys[r0 + 1] = okB ? ~(((uint64_t)__float_as_uint(dB) << 32) | idB) : 0ull;
    }
  }
  __syncthreads();
  auto sync = [] { __syncthreads(); };
  bitonic_sort_desc(skeys, sortn, tid, kRefineThreads, sync);
  for (uint32_t j = tid; j < K; j += kRefineThreads) {
    const uint64_t sk = (j < sortn) ? skeys[j] : 0ull;
    const bool ok = sk != 0ull;
    const uint64_t key = ~sk;
    out_ids[(size_t)q * K + j] = ok ? (uint32_t)key : 0xFFFFFFFFu;
    if (out_dist) out_dist[(size_t)q * K + j] = ok ? __uint_as_float((uint32_t)(key >> 32)) : 1e30f;
  }
}

}  // namespace

cudaError_t refine_l2_launch(int dtype, const void* base, const float* scales, uint64_t n, uint32_t dim,
                             const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                             uint32_t* out_ids, float* out_dist, cudaStream_t stream) {
  if (Q == 0 || K == 0) return cudaSuccess;
  const uint32_t sortn = next_pow2(R < 2 ? 2 : R);
  const size_t smem = (size_t)sortn * 8u + (size_t)dim * 4u;
  if (smem > 220u * 1024u) return cudaErrorInvalidValue;
  auto go = [&](auto kern) -> cudaError_t {
    if (smem > 48u * 1024u) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
    }
    kern<<<Q, kRefineThreads, smem, stream>>>(reinterpret_cast<const uint8_t*>(base), scales, n, dim, queries, cand, R,
                                              K, sortn, out_ids, out_dist);
    count_launch();
    return cudaGetLastError();
  };
  switch (dtype) {
    case kDtF32: return go(refine_l2_kernel<kDtF32>);
    case kDtF16: return go(refine_l2_kernel<kDtF16>);
    case kDtI8: return go(refine_l2_kernel<kDtI8>);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace nvdb_b200
