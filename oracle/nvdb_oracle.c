/*
 * oracle/nvdb_oracle.c — CPU restatement of the reference's flat-scan / refine arithmetic.
 *
 * THIS IS TEST INFRASTRUCTURE. It is the checker for the CUDA path (tests/, __graft_entry__.smoke(),
 * bench.py's cpu_baseline leg). It is never imported, linked or executed by the product path
 * (nano-vectordb_b200/), which fails loudly when its CUDA library is missing.
 *
 * Parity status: PINNED. Every function below is checked (tests/test_oracle.py) against
 *   (a) the unmodified reference compiled from /root/reference by oracle/Makefile `ref`
 *       (oracle/_ref/libnvdb_ref.so and the reference's own tools), when present, and
 *   (b) golden vectors those binaries produced in the build container, committed under tests/golden/
 *       together with the generating script tests/golden/make_golden.py.
 * The reference itself ships no tests / golden vectors for this path.
 *
 * File:line citations are into /root/reference.
 */
#include "nvdb_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------- element conversions */

/* include/nvdb/f16_scalar.h:8-36, src/simd_dot.cpp:68-99 — exact half -> float widening. */
float orc_f16_to_f32(uint16_t h) {
  const uint32_t sign = ((uint32_t)(h & 0x8000u)) << 16;
  uint32_t exp = (h >> 10) & 0x1Fu;
  uint32_t mant = h & 0x3FFu;
  uint32_t out;
  if (exp == 0) {
    if (mant == 0) {
      out = sign;
    } else {
      int e = 1;
      while ((mant & 0x400u) == 0) {
        mant <<= 1;
        e--;
      }
      mant &= 0x3FFu;
      out = sign | ((uint32_t)(e - 15 + 127) << 23) | (mant << 13);
    }
  } else if (exp == 0x1F) {
    out = sign | 0x7F800000u | (mant << 13);
  } else {
    out = sign | ((exp - 15 + 127) << 23) | (mant << 13);
  }
  float f;
  memcpy(&f, &out, 4);
  return f;
}

/* tools/nvdb_convert_f16.cpp:95-119 — on an F16C host the tool uses _mm256_cvtps_ph(v, 0), i.e. IEEE
 * round-to-nearest-even for every input (including f32 subnormals, which the scalar fallback at :40-43
 * flushes). This restates the F16C behaviour, which is what runs on AVX2 hosts. */
uint16_t orc_f32_to_f16_rne(float f) {
  uint32_t x;
  memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  const uint32_t ex = (x >> 23) & 0xFFu;
  uint32_t mant = x & 0x7FFFFFu;
  if (ex == 0xFF) {
    if (mant == 0) return (uint16_t)(sign | 0x7C00u);
    return (uint16_t)(sign | 0x7C00u | 0x200u | (mant >> 13)); /* quiet NaN, payload kept */
  }
  int32_t e = (int32_t)ex - 127;
  if (e > 15) return (uint16_t)(sign | 0x7C00u);
  if (e >= -14) {
    uint32_t m = mant >> 13;
    uint32_t rem = mant & 0x1FFFu;
    uint32_t out = ((uint32_t)(e + 15) << 10) | m;
    if (rem > 0x1000u || (rem == 0x1000u && (m & 1u))) out++; /* carries into exponent / inf correctly */
    return (uint16_t)(sign | out);
  }
  if (ex == 0) return (uint16_t)sign; /* f32 subnormal: far below half subnormal range */
  /* half subnormal: value = mant24 * 2^(e-23); unit = 2^-24 */
  int shift = (-14 - e) + 13; /* >= 14 */
  if (shift > 24) return (uint16_t)sign;
  uint32_t m24 = mant | 0x800000u;
  uint32_t m = m24 >> shift;
  uint32_t rem = m24 & ((1u << shift) - 1u);
  uint32_t half = 1u << (shift - 1);
  if (rem > half || (rem == half && (m & 1u))) m++;
  return (uint16_t)(sign | m);
}

void orc_convert_f16(const float* in, uint64_t count, uint16_t* out) {
  for (uint64_t i = 0; i < count; ++i) out[i] = orc_f32_to_f16_rne(in[i]);
}

/* ---------------------------------------------------------------- dot kernels */

/* src/simd_dot.cpp:39-44: lo+hi, hadd, hadd  ==  ((a0+a4)+(a1+a5)) + ((a2+a6)+(a3+a7)) */
static inline float hsum8(const float a[8]) {
  const float s0 = a[0] + a[4], s1 = a[1] + a[5], s2 = a[2] + a[6], s3 = a[3] + a[7];
  const float t0 = s0 + s1, t1 = s2 + s3;
  return t0 + t1;
}

/* src/simd_dot.cpp:26-49 (dot_avx2_fma): one 8-lane accumulator, FMA per 8 elements in ascending order,
 * hsum, then a sequential tail `out += a[i]*b[i]` (see the note at the tail). */
float orc_dot_f32(const float* a, const float* b, uint32_t dim) {
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t i = 0;
  for (; i + 8 <= dim; i += 8)
    for (int j = 0; j < 8; ++j) acc[j] = fmaf(a[i + j], b[i + j], acc[j]);
  float out = hsum8(acc);
  /* tail (:47): as compiled by GCC 13 -O3 in target("avx2,fma") — a group of 4 is vectorised as an unfused
   * multiply followed by in-order adds, the remaining <=3 elements are contracted to FMAs. Compiler artefact,
   * pinned empirically against oracle/_ref; irrelevant for the reference's own dims (384, 768). */
  if (dim - i >= 4) {
    for (int j = 0; j < 4; ++j, ++i) out = out + a[i] * b[i];
  }
  for (; i < dim; ++i) out = fmaf(a[i], b[i], out);
  return out;
}

/* src/simd_dot.cpp:102-124 (dot_f32_f16base_avx2): widen 8 halves exactly, same FMA / hsum structure. */
float orc_dot_f16(const float* q, const uint16_t* x, uint32_t dim) {
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t i = 0;
  for (; i + 8 <= dim; i += 8)
    for (int j = 0; j < 8; ++j) acc[j] = fmaf(q[i + j], orc_f16_to_f32(x[i + j]), acc[j]);
  float out = hsum8(acc);
  for (; i < dim; ++i) out = out + q[i] * orc_f16_to_f32(x[i]); /* :122 — compiled unfused (scalar decode call) */
  return out;
}

/* src/simd_dot.cpp:160-199 (dot_f32_i8_avx2): 16 int8 per iteration, two FMAs into the SAME accumulator
 * (elements i..i+7 then i+8..i+15), hsum, sequential tail, then ONE multiply by the row scale (:198). */
float orc_dot_i8(const float* q, const int8_t* x, uint32_t dim, float scale) {
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t i = 0;
  for (; i + 16 <= dim; i += 16) {
    for (int j = 0; j < 8; ++j) acc[j] = fmaf(q[i + j], (float)x[i + j], acc[j]);
    for (int j = 0; j < 8; ++j) acc[j] = fmaf(q[i + 8 + j], (float)x[i + 8 + j], acc[j]);
  }
  float out = hsum8(acc);
  for (; i < dim; ++i) out = out + q[i] * (float)x[i]; /* :196 — compiled unfused */
  return out * scale;
}

/* include/nvdb/score_dispatch.h:25-48 */
float orc_score_row(const void* payload, const float* scales, uint32_t dtype, uint64_t row, uint32_t dim,
                    const float* q) {
  const size_t off = (size_t)row * dim;
  if (dtype == ORC_F32) return orc_dot_f32(q, (const float*)payload + off, dim);
  if (dtype == ORC_F16) return orc_dot_f16(q, (const uint16_t*)payload + off, dim);
  return orc_dot_i8(q, (const int8_t*)payload + off, dim, scales[row]);
}

void orc_score_all(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                   const float* q, float* out_scores) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) out_scores[i] = orc_score_row(payload, scales, dtype, (uint64_t)i, dim, q);
}

/* ---------------------------------------------------------------- top-k */

typedef struct {
  float score;
  uint64_t id;
} node_t;

/* "a is worse than b" under the canonical order (score desc, id asc). The reference admits a row only on a
 * STRICTLY greater score (src/flat_index.cpp:36-37) while scanning ids in ascending order, which is this
 * order at the admission boundary; inside exact-tie groups the reference's own variants disagree with each
 * other, so the canonical order is what the tie-aware comparator in tests/ checks. */
static inline int worse(const node_t* a, const node_t* b) {
  if (a->score < b->score) return 1;
  if (a->score > b->score) return 0;
  return a->id > b->id;
}

static void heap_sift_down(node_t* h, uint32_t n, uint32_t i) {
  for (;;) {
    uint32_t l = 2 * i + 1, r = l + 1, m = i;
    if (l < n && worse(&h[l], &h[m])) m = l;
    if (r < n && worse(&h[r], &h[m])) m = r;
    if (m == i) return;
    node_t t = h[i];
    h[i] = h[m];
    h[m] = t;
    i = m;
  }
}
static void heap_sift_up(node_t* h, uint32_t i) {
  while (i > 0) {
    uint32_t p = (i - 1) / 2;
    if (!worse(&h[i], &h[p])) return;
    node_t t = h[i];
    h[i] = h[p];
    h[p] = t;
    i = p;
  }
}

/* min-heap of the k best; strict-greater admission (src/flat_index.cpp:28-38) */
static inline void topk_offer(node_t* heap, uint32_t* size, uint32_t k, float s, uint64_t id) {
  if (*size < k) {
    heap[*size].score = s;
    heap[*size].id = id;
    heap_sift_up(heap, (*size)++);
  } else {
    node_t c = {s, id};
    if (worse(&heap[0], &c)) { /* for ascending ids this is exactly `s > heap.top().score` */
      heap[0] = c;
      heap_sift_down(heap, k, 0);
    }
  }
}

static uint32_t topk_drain_sorted(node_t* heap, uint32_t size, uint64_t* out_ids, float* out_scores) {
  /* pop worst -> best, fill from the back (src/flat_index.cpp:40-47: pop, push_back, reverse) */
  uint32_t n = size;
  while (size > 0) {
    out_ids[size - 1] = heap[0].id;
    out_scores[size - 1] = heap[0].score;
    heap[0] = heap[size - 1];
    size--;
    heap_sift_down(heap, size, 0);
  }
  return n;
}

/* src/flat_index.cpp:16-48 (FlatIndex::search_topk_dot): full ascending scan, k clamped to n. */
uint32_t orc_flat_search(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                         const float* q, uint32_t k, uint64_t* out_ids, float* out_scores) {
  if (k == 0 || n == 0) return 0;
  if (k > n) k = (uint32_t)n;
  node_t* heap = (node_t*)malloc(sizeof(node_t) * k);
  uint32_t size = 0;
  for (uint64_t i = 0; i < n; ++i) topk_offer(heap, &size, k, orc_score_row(payload, scales, dtype, i, dim, q), i);
  uint32_t r = topk_drain_sorted(heap, size, out_ids, out_scores);
  free(heap);
  return r;
}

/* src/flat_index_omp.cpp:16-85 (FlatIndexOMP): static row chunks, one heap per thread, serial merge in
 * thread order. With the canonical order the merge result does not depend on the thread count. */
uint32_t orc_flat_search_mt(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                            const float* q, uint32_t k, uint64_t* out_ids, float* out_scores) {
  if (k == 0 || n == 0) return 0;
  if (k > n) k = (uint32_t)n;
  int T = 1;
#ifdef _OPENMP
  T = omp_get_max_threads();
#endif
  node_t* heaps = (node_t*)malloc(sizeof(node_t) * (size_t)k * T);
  uint32_t* sizes = (uint32_t*)calloc(T, sizeof(uint32_t));
#pragma omp parallel num_threads(T)
  {
    int tid = 0;
#ifdef _OPENMP
    tid = omp_get_thread_num();
#endif
    node_t* h = heaps + (size_t)tid * k;
    uint32_t sz = 0;
#pragma omp for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; ++i)
      topk_offer(h, &sz, k, orc_score_row(payload, scales, dtype, (uint64_t)i, dim, q), (uint64_t)i);
    sizes[tid] = sz;
  }
  node_t* g = (node_t*)malloc(sizeof(node_t) * k);
  uint32_t gs = 0;
  for (int t = 0; t < T; ++t)
    for (uint32_t j = 0; j < sizes[t]; ++j) topk_offer(g, &gs, k, heaps[(size_t)t * k + j].score, heaps[(size_t)t * k + j].id);
  uint32_t r = topk_drain_sorted(g, gs, out_ids, out_scores);
  free(g);
  free(sizes);
  free(heaps);
  return r;
}

/* apps/nvdb_bench.cpp:47-159 (batched_scan_omp_or_st): per row, every query of the batch is scored with the
 * same score_query_base_at and offered to its own top-k; by construction == nq independent searches.
 * Output layout [nq][k], padded with id=UINT64_MAX / score=-inf when k > n. */
void orc_flat_search_batch(const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                           const float* queries, uint32_t nq, uint32_t k, uint64_t* out_ids, float* out_scores) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t qi = 0; qi < (int64_t)nq; ++qi) {
    uint64_t* oi = out_ids + (size_t)qi * k;
    float* os = out_scores + (size_t)qi * k;
    uint32_t r = orc_flat_search(payload, scales, dtype, n, dim, queries + (size_t)qi * dim, k, oi, os);
    for (uint32_t j = r; j < k; ++j) {
      oi[j] = UINT64_MAX;
      os[j] = -INFINITY;
    }
  }
}

/* ---------------------------------------------------------------- converters (fixtures) */

/* apps/nvdb_quantize_i8.cpp:12-16,72-83: scale = max|x|/127 (1 if the row is all zero), inv = 1.0f/scale,
 * q = clamp(lrint(x*inv), -127, 127), all in FP32; lrint = current rounding mode = nearest-even. */
void orc_quantize_i8_row(const float* row, uint32_t dim, int8_t* out, float* out_scale) {
  float max_abs = 0.f;
  for (uint32_t j = 0; j < dim; ++j) {
    float a = fabsf(row[j]);
    if (a > max_abs) max_abs = a;
  }
  const float scale = (max_abs > 0.f) ? (max_abs / 127.f) : 1.f;
  const float inv = 1.0f / scale;
  for (uint32_t j = 0; j < dim; ++j) {
    const float p = row[j] * inv;
    long v = lrintf(p);
    if (v > 127) v = 127;
    if (v < -127) v = -127;
    out[j] = (int8_t)v;
  }
  *out_scale = scale;
}

void orc_quantize_i8(const float* rows, uint64_t n, uint32_t dim, int8_t* out, float* out_scales) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i)
    orc_quantize_i8_row(rows + (size_t)i * dim, dim, out + (size_t)i * dim, out_scales + i);
}

/* ---------------------------------------------------------------- refine */

/* include/nvdb/to_f32_row.h:10-34 + apps/nvdb_ivf_eval.cpp:232-240: row -> f32 (i8: float(x)*scale, one FP32
 * multiply), then L2^2 accumulated in double, cast to float. */
float orc_l2_cpu(const float* q, const void* payload, const float* scales, uint32_t dtype, uint64_t row,
                 uint32_t dim) {
  const size_t off = (size_t)row * dim;
  double s = 0.0;
  for (uint32_t j = 0; j < dim; ++j) {
    float v;
    if (dtype == ORC_F32)
      v = ((const float*)payload)[off + j];
    else if (dtype == ORC_F16)
      v = orc_f16_to_f32(((const uint16_t*)payload)[off + j]);
    else
      v = (float)((const int8_t*)payload)[off + j] * scales[row];
    const double d = (double)q[j] - (double)v;
    s += d * d;
  }
  return (float)s;
}

/* src/cuda_refine.cu:326-382 (l2_fp16_base_half2): half2 pair p = j/2 goes to accumulator p mod 4 for the
 * unrolled body (4 half2 per iteration), tail pairs to acc0; diff rounded to FP32 first; two dependent FMAs
 * per pair; dist = (acc0+acc1)+(acc2+acc3). D even. */
float orc_l2_cudaorder_f16(const float* q, const uint16_t* row, uint32_t dim) {
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  const uint32_t D2 = dim / 2;
  uint32_t p = 0;
  for (; p + 3 < D2; p += 4) {
    for (int m = 0; m < 4; ++m) {
      const float dx = q[2 * (p + m)] - orc_f16_to_f32(row[2 * (p + m)]);
      const float dy = q[2 * (p + m) + 1] - orc_f16_to_f32(row[2 * (p + m) + 1]);
      acc[m] = fmaf(dx, dx, acc[m]);
      acc[m] = fmaf(dy, dy, acc[m]);
    }
  }
  for (p = (D2 & ~3u); p < D2; ++p) {
    const float dx = q[2 * p] - orc_f16_to_f32(row[2 * p]);
    const float dy = q[2 * p + 1] - orc_f16_to_f32(row[2 * p + 1]);
    acc[0] = fmaf(dx, dx, acc[0]);
    acc[0] = fmaf(dy, dy, acc[0]);
  }
  return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

typedef struct {
  float dist;
  uint32_t id;
  uint32_t pos;
} cand_t;

static int cand_cmp(const void* a, const void* b) {
  const cand_t* x = (const cand_t*)a;
  const cand_t* y = (const cand_t*)b;
  if (x->dist < y->dist) return -1;
  if (x->dist > y->dist) return 1;
  if (x->id < y->id) return -1;
  if (x->id > y->id) return 1;
  return (x->pos > y->pos) - (x->pos < y->pos);
}

/* apps/nvdb_ivf_eval.cpp:278-307 (refine_topk_l2_ids, mode 0) and src/cuda_refine.cu:405-502 (mode 1):
 * score every valid candidate (id < n; 0xFFFFFFFF / id<0 skipped), keep the K smallest (strict `<`
 * replacement), output ascending distance. Unfilled slots keep id=0xFFFFFFFF, dist=1e30f
 * (src/cuda_refine.cu:891-894). Canonical tie order: (dist asc, id asc). */
void orc_refine_l2(int mode, const void* payload, const float* scales, uint32_t dtype, uint64_t n, uint32_t dim,
                   const float* queries, const uint32_t* cand, uint32_t Q, uint32_t R, uint32_t K,
                   uint32_t* out_ids, float* out_dist) {
#pragma omp parallel for schedule(dynamic, 4)
  for (int64_t qi = 0; qi < (int64_t)Q; ++qi) {
    cand_t* c = (cand_t*)malloc(sizeof(cand_t) * (R ? R : 1));
    uint32_t m = 0;
    const float* q = queries + (size_t)qi * dim;
    for (uint32_t r = 0; r < R; ++r) {
      const uint32_t id = cand[(size_t)qi * R + r];
      if ((uint64_t)id >= n) continue;
      float d;
      if (mode == 1 && dtype == ORC_F16)
        d = orc_l2_cudaorder_f16(q, (const uint16_t*)payload + (size_t)id * dim, dim);
      else
        d = orc_l2_cpu(q, payload, scales, dtype, id, dim);
      c[m].dist = d;
      c[m].id = id;
      c[m].pos = r;
      m++;
    }
    qsort(c, m, sizeof(cand_t), cand_cmp);
    for (uint32_t j = 0; j < K; ++j) {
      out_ids[(size_t)qi * K + j] = (j < m) ? c[j].id : 0xFFFFFFFFu;
      out_dist[(size_t)qi * K + j] = (j < m) ? c[j].dist : 1e30f;
    }
    free(c);
  }
}

/* ---------------------------------------------------------------- synthetic generator (ours) */

/* splitmix64 finaliser over (seed, element index): pure 64-bit integer arithmetic, so the CUDA generator
 * (nano-vectordb_b200/csrc/synth.cu) and this function agree bit-for-bit on every platform. The value is a
 * signed 16-bit integer; the f32 element is (float)v * c with c = (float)(sqrt(3/D)/32768) so that rows have
 * norm ~= 1 (the reference's data are L2-normalised embeddings, Performance.md:15). */
static inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
int32_t orc_synth_i16(uint64_t seed, uint64_t global_index) {
  const uint64_t h = mix64(seed * 0x9E3779B97F4A7C15ULL + global_index * 0xD1B54A32D192ED03ULL + 0x2545F4914F6CDD1DULL);
  return (int32_t)(h >> 48) - 32768;
}
void orc_synth_rows_f32(uint64_t seed, uint64_t row0, uint64_t nrows, uint32_t dim, float* out) {
  const float c = (float)(sqrt(3.0 / (double)dim) / 32768.0);
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < (int64_t)nrows; ++r)
    for (uint32_t j = 0; j < dim; ++j)
      out[(size_t)r * dim + j] = (float)orc_synth_i16(seed, (row0 + (uint64_t)r) * dim + j) * c;
}

/* ---------------------------------------------------------------- tensor-core scan: query limb schemes (ours)
 * Restatement of nano-vectordb_b200/csrc/scan_mma.cu's query preparation and epilogue arithmetic, so that the
 * kind::i8 path can be checked bit-for-bit and the kind::f16 path against its exact real-number value. These are
 * not reference functions (the reference has no quantised-query path); the parity of the final
 * scores against the reference's AVX2 kernels is checked with a tolerance in tests/test_gpu_scan_mma.py. */

/* q ~= s1*q1 + s2*q2 with int8 limbs (two limbs) */
void orc_query_limbs_i8(const float* q, uint32_t dim, int8_t* q1, int8_t* q2, float* s1_out, float* s2_out) {
  float amax = 0.f;
  for (uint32_t i = 0; i < dim; ++i) amax = fmaxf(amax, fabsf(q[i]));
  const float s1 = (amax > 0.f && amax < INFINITY) ? amax / 127.f : 1.f;
  const float inv1 = 1.f / s1;
  const float s2 = s1 / 254.f;
  const float inv2 = 1.f / s2;
  for (uint32_t i = 0; i < dim; ++i) {
    long a = lrintf(q[i] * inv1);
    a = a > 127 ? 127 : (a < -127 ? -127 : a);
    const float r = fmaf(-(float)a, s1, q[i]);
    long b = lrintf(r * inv2);
    b = b > 127 ? 127 : (b < -127 ? -127 : b);
    q1[i] = (int8_t)a;
    q2[i] = (int8_t)b;
  }
  *s1_out = s1;
  *s2_out = s2;
}

/* exact int32 limb dot products + the epilogue's fixed FP32 combination */
float orc_dot_i8_limbs(const int8_t* q1, const int8_t* q2, float s1, float s2, const int8_t* x, uint32_t dim,
                       float scale, int32_t* i1_out, int32_t* i2_out) {
  int32_t i1 = 0, i2 = 0;
  for (uint32_t i = 0; i < dim; ++i) {
    i1 += (int32_t)q1[i] * (int32_t)x[i];
    i2 += (int32_t)q2[i] * (int32_t)x[i];
  }
  if (i1_out) *i1_out = i1;
  if (i2_out) *i2_out = i2;
  const float t = s1 * (float)i1;
  return fmaf(s2, (float)i2, t) * scale;
}

/* all rows of an i8 base against one query under the limb scheme */
void orc_score_all_i8_limbs(const int8_t* payload, const float* scales, uint64_t n, uint32_t dim, const float* q,
                            float* out_scores, int32_t* out_i1, int32_t* out_i2) {
  int8_t* q1 = (int8_t*)malloc(dim);
  int8_t* q2 = (int8_t*)malloc(dim);
  float s1, s2;
  orc_query_limbs_i8(q, dim, q1, q2, &s1, &s2);
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < (int64_t)n; ++r) {
    int32_t a, b;
    out_scores[r] = orc_dot_i8_limbs(q1, q2, s1, s2, payload + (size_t)r * dim, dim, scales[r], &a, &b);
    if (out_i1) out_i1[r] = a;
    if (out_i2) out_i2[r] = b;
  }
  free(q1);
  free(q2);
}

/* f16 limbs: v = q * 2^e with max|v| in [0.5,1); hi = fp16(v), lo = fp16(v - hi); qscale = 2^-e */
void orc_query_limbs_f16(const float* q, uint32_t dim, uint16_t* hi, uint16_t* lo, float* qscale_out) {
  float amax = 0.f;
  for (uint32_t i = 0; i < dim; ++i) amax = fmaxf(amax, fabsf(q[i]));
  int e = 0;
  if (amax > 0.f && amax < INFINITY) {
    e = -(ilogbf(amax) + 1);
    e = e > 100 ? 100 : (e < -100 ? -100 : e);
  }
  for (uint32_t i = 0; i < dim; ++i) {
    const float v = scalbnf(q[i], e);
    const uint16_t h = orc_f32_to_f16_rne(v);
    hi[i] = h;
    lo[i] = orc_f32_to_f16_rne(v - orc_f16_to_f32(h));
  }
  *qscale_out = scalbnf(1.0f, -e);
}
