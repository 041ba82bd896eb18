// dot_chains.cuh — the reference's dot-product arithmetic, reproduced bit for bit.
//
// src/simd_dot.cpp:26-49 (f32), 102-124 (f16 base), 160-199 (int8 base): 8 interleaved FP32 FMA chains (chain j takes
// elements j, j+8, j+16, ... in ascending order), horizontal sum ((a0+a4)+(a1+a5))+((a2+a6)+(a3+a7)), and for int8 one
// FP32 multiply by the row scale at the end. Shared by the streaming scan (scan_gemv.cu) and the exact re-score of
// tensor-core candidates (rescore.cu).
#pragma once
#include "common.cuh"

namespace nvdb_b200 {

template <int DT>
struct Elem;
template <>
struct Elem<kDtF32> { static constexpr int size = 4, per16 = 4; };
template <>
struct Elem<kDtF16> { static constexpr int size = 2, per16 = 8; };
template <>
struct Elem<kDtI8> { static constexpr int size = 1, per16 = 16; };

// ---- 16 bytes of one row -> FMAs into the 8 chains of every query -------------------------------------------
// The 8 chains of a query live in four 64-bit register pairs (chains 2p, 2p+1) and are advanced with the packed
// FP32 instructions of sm_100 (fma.rn.f32x2 / add.rn.f32x2 -> FFMA2 / FADD2): each half is an ordinary IEEE
// round-to-nearest FMA, so the chain values stay bit-identical to the scalar form while the FMA-pipe instruction
// count halves. `c` is the 16-byte piece index inside the 128-byte chunk (compile-time after unrolling).

#ifndef NVDB_DOT_PACKED_FMA
#define NVDB_DOT_PACKED_FMA 1
#endif
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// Same two FMAs as scalar instructions (A/B knob: -DNVDB_DOT_PACKED_FMA=0 builds the f16 / f32 chains with FFMA).
__device__ __forceinline__ uint64_t ffma_pair(uint64_t a, uint64_t b, uint64_t c) {
#if NVDB_DOT_PACKED_FMA
  return ffma2(a, b, c);
#else
  float a0, a1, b0, b1, c0, c1;
  unpack2(a, a0, a1);
  unpack2(b, b0, b1);
  unpack2(c, c0, c1);
  return pack2(__fmaf_rn(a0, b0, c0), __fmaf_rn(a1, b1, c1));
#endif
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// 4 query elements from shared memory (warp-broadcast LDS.128). Not volatile: the queries are written once before
// the first barrier, so the compiler may schedule these loads freely.
__device__ __forceinline__ float4 ldsq(uint32_t addr) {
  float4 v;
  asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// x[p] = row elements of chains (2p, 2p+1); q = shared-memory byte address of the 8 query elements of the same
// columns (query 0), qstride = bytes between consecutive queries.
template <int NQ, int PK = 1>
__device__ __forceinline__ void fma8(uint64_t (&acc)[NQ][4], const uint64_t (&x)[4], uint32_t q, uint32_t qstride) {
#pragma unroll
  for (int qi = 0; qi < NQ; ++qi) {
    const float4 qa = ldsq(q + qi * qstride);
    const float4 qb = ldsq(q + qi * qstride + 16u);
    if constexpr (PK) {
      acc[qi][0] = ffma2(pack2(qa.x, qa.y), x[0], acc[qi][0]);
      acc[qi][1] = ffma2(pack2(qa.z, qa.w), x[1], acc[qi][1]);
      acc[qi][2] = ffma2(pack2(qb.x, qb.y), x[2], acc[qi][2]);
      acc[qi][3] = ffma2(pack2(qb.z, qb.w), x[3], acc[qi][3]);
    } else {
      acc[qi][0] = ffma_pair(pack2(qa.x, qa.y), x[0], acc[qi][0]);
      acc[qi][1] = ffma_pair(pack2(qa.z, qa.w), x[1], acc[qi][1]);
      acc[qi][2] = ffma_pair(pack2(qb.x, qb.y), x[2], acc[qi][2]);
      acc[qi][3] = ffma_pair(pack2(qb.z, qb.w), x[3], acc[qi][3]);
    }
  }
}

template <int DT, int NQ>
__device__ __forceinline__ void consume16(uint64_t (&acc)[NQ][4], const uint4 v, uint32_t q, uint32_t qstride,
                                          int c) {
  if constexpr (DT == kDtF16) {
    uint64_t x[4];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));  // exact widening
      x[i] = pack2(f.x, f.y);
    }
    fma8<NQ, NVDB_DOT_PACKED_FMA>(acc, x, q, qstride);
  } else if constexpr (DT == kDtF32) {
    // 4 elements: chains 0..3 for even pieces, 4..7 for odd pieces
    const uint64_t x0 = pack2(__uint_as_float(v.x), __uint_as_float(v.y));
    const uint64_t x1 = pack2(__uint_as_float(v.z), __uint_as_float(v.w));
    const int p0 = (c & 1) * 2;
#pragma unroll
    for (int qi = 0; qi < NQ; ++qi) {
      const float4 qa = ldsq(q + qi * qstride);
      acc[qi][p0 + 0] = ffma_pair(pack2(qa.x, qa.y), x0, acc[qi][p0 + 0]);
      acc[qi][p0 + 1] = ffma_pair(pack2(qa.z, qa.w), x1, acc[qi][p0 + 1]);
    }
  } else {
    // 16 int8: elements 0..7 then 8..15 into the same 8 chains (src/simd_dot.cpp:175-187).
    // int8 -> f32 exactly: byte ^ 0x80 placed in the mantissa of 2^23 gives 2^23 + (x+128); adding
    // -(2^23 + 128) is exact. One LOP3 per 4 bytes, one PRMT per element, one FADD2 per two elements, no I2F.
    const uint32_t w[4] = {v.x ^ 0x80808080u, v.y ^ 0x80808080u, v.z ^ 0x80808080u, v.w ^ 0x80808080u};
    const uint64_t magic = pack2(-8388736.0f, -8388736.0f);  // -(2^23 + 128)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint64_t x[4];
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const uint32_t word = w[h * 2 + (p >> 1)];
        const uint32_t b0 = __byte_perm(word, 0x4B000000u, 0x7650 + ((2 * p) & 3));      // {0x4B,0x00,0x00,byte}
        const uint32_t b1 = __byte_perm(word, 0x4B000000u, 0x7650 + ((2 * p + 1) & 3));
        x[p] = fadd2(pack2(__uint_as_float(b0), __uint_as_float(b1)), magic);
      }
      fma8<NQ>(acc, x, q + h * 32u, qstride);
    }
  }
}

__device__ __forceinline__ float hsum8_ref_order(const uint64_t (&acc)[4]) {
  float a[8];
#pragma unroll
  for (int p = 0; p < 4; ++p) unpack2(acc[p], a[2 * p], a[2 * p + 1]);
  const float s0 = __fadd_rn(a[0], a[4]), s1 = __fadd_rn(a[1], a[5]);
  const float s2 = __fadd_rn(a[2], a[6]), s3 = __fadd_rn(a[3], a[7]);
  return __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
}
__device__ __forceinline__ float hsum8_ref_order(const float (&a)[8]) {
  const float s0 = __fadd_rn(a[0], a[4]), s1 = __fadd_rn(a[1], a[5]);
  const float s2 = __fadd_rn(a[2], a[6]), s3 = __fadd_rn(a[3], a[7]);
  return __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
}

}  // namespace nvdb_b200
