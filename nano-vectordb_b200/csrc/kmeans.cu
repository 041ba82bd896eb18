// kmeans.cu — nearest-centroid assignment and Lloyd k-means for the IVF-PQ candidate generator.
//
// The reference delegates this to FAISS (apps/nvdb_ivfpq_build.cpp:78-135: IndexFlatL2 coarse quantiser + IndexIVFPQ,
// train() = k-means for the nlist centroids and for the M x 256 PQ codewords). FAISS is a third-party dependency that is
// not in /root/reference and not in this image (parity unpinned), so this file restates the published
// algorithm (Lloyd iterations; L2 assignment via x.c - |c|^2/2) and is checked against oracle/ivfpq_ref.py.
//
// dot_bias_kernel is a plain FP32 SIMT tile GEMM (128 x 128 x 8, 8 x 8 per thread) with the bias subtraction and, in
// assignment mode, the arg-max fused into the epilogue (scores never reach memory). FP32 on purpose: this is the
// build-time path (north star: no TF32), and a fixed fmaf order keeps training deterministic.
#include <math.h>

#include "common.cuh"
#include "ivf.h"
#include "scan.h"

namespace nvdb_b200 {
namespace {

constexpr int BM = 128, BN = 128, BK = 8, NT = 256;

__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ uint32_t frag_index(uint32_t t4, int u) {  // 8 rows/cols of a thread: {t*4..+3} U {64+t*4..+3}
  return (u < 4) ? t4 * 4u + (uint32_t)u : 64u + t4 * 4u + (uint32_t)(u - 4);
}

// MODE 0: scores[i][j] = x_i.c_j - hn[j];  MODE 1: per (point, split) running arg-max over the CTA's centroid range
template <int MODE>
__global__ void __launch_bounds__(NT)
dot_bias_kernel(const float* __restrict__ x, size_t ldx, uint32_t n, const float* __restrict__ c, size_t ldc,
                const float* __restrict__ hn, uint32_t k, uint32_t d, uint32_t cols_per_cta, uint32_t splits,
                float* __restrict__ scores, float* __restrict__ best_s, uint32_t* __restrict__ best_j) {
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];
  const uint32_t tid = threadIdx.x, tx = tid & 15u, ty = tid >> 4;
  const uint32_t m0 = blockIdx.x * BM;
  const uint32_t col_begin = blockIdx.y * cols_per_cta;
  const uint32_t col_end = min(k, col_begin + cols_per_cta);
  const uint32_t lr = tid >> 1, lk = (tid & 1u) * 4u;  // tile loader: row lr, elements lk..lk+3 of the K slice

  float run_s[8];
  uint32_t run_j[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    run_s[u] = -INFINITY;
    run_j[u] = 0xFFFFFFFFu;
  }

  for (uint32_t n0 = col_begin; n0 < col_end; n0 += BN) {
    float acc[8][8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int v = 0; v < 8; ++v) acc[u][v] = 0.f;
    for (uint32_t k0 = 0; k0 < d; k0 += BK) {
      const uint32_t gr = m0 + lr, gc = n0 + lr;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint32_t kk = k0 + lk + e;
        As[lk + e][lr] = (gr < n && kk < d) ? x[(size_t)gr * ldx + kk] : 0.f;
        Bs[lk + e][lr] = (gc < col_end && kk < d) ? c[(size_t)gc * ldc + kk] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
        const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
          for (int v = 0; v < 8; ++v) acc[u][v] = __fmaf_rn(a[u], b[v], acc[u][v]);
      }
      __syncthreads();
    }
    // epilogue of this centroid tile
    float bias[8];
    uint32_t col[8];
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      col[v] = n0 + frag_index(tx, v);
      bias[v] = col[v] < col_end ? hn[col[v]] : 0.f;
    }
    if constexpr (MODE == 0) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t row = m0 + frag_index(ty, u);
        if (row >= n) continue;
#pragma unroll
        for (int v = 0; v < 8; ++v)
          if (col[v] < col_end) scores[(size_t)row * k + col[v]] = __fsub_rn(acc[u][v], bias[v]);
      }
    } else {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float bs = -INFINITY;
        uint32_t bj = 0xFFFFFFFFu;
#pragma unroll
        for (int v = 0; v < 8; ++v) {  // this thread's columns are not ascending across the two groups: compare ids
          const float s = __fsub_rn(acc[u][v], bias[v]);
          if (col[v] < col_end && (s > bs || (s == bs && col[v] < bj))) {
            bs = s;
            bj = col[v];
          }
        }
#pragma unroll
        for (int off = 8; off > 0; off >>= 1) {  // the 16 lanes that share this point row
          const float os = __shfl_xor_sync(0xffffffffu, bs, off);
          const uint32_t oj = __shfl_xor_sync(0xffffffffu, bj, off);
          if (os > bs || (os == bs && oj < bj)) {
            bs = os;
            bj = oj;
          }
        }
        if (bs > run_s[u] || (bs == run_s[u] && bj < run_j[u])) {
          run_s[u] = bs;
          run_j[u] = bj;
        }
      }
    }
  }
  if constexpr (MODE == 1) {
    if (tx == 0) {
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const uint32_t row = m0 + frag_index(ty, u);
        if (row < n) {
          best_s[(size_t)row * splits + blockIdx.y] = run_s[u];
          best_j[(size_t)row * splits + blockIdx.y] = run_j[u];
        }
      }
    }
  }
}

// few points (queries): a CTA takes 32 centroids x up to 8 points; thread (centroid, point) runs the SAME sequential
// fmaf chain over the dimensions as the tile kernel above, so a query scores bit-identically in a batch of 1 or 1000.
constexpr int kSmallN = 8;
__global__ void __launch_bounds__(256)
dot_bias_small_kernel(const float* __restrict__ x, size_t ldx, uint32_t n, const float* __restrict__ c, size_t ldc,
                      const float* __restrict__ hn, uint32_t k, uint32_t d, float* __restrict__ scores) {
  extern __shared__ float xs[];  // [n][d]
  __shared__ float cs[32][33];
  const uint32_t tid = threadIdx.x, cj = tid & 31u, qi = tid >> 5, j0 = blockIdx.x * 32u;
  for (uint32_t e = tid; e < n * d; e += 256) xs[e] = x[(size_t)(e / d) * ldx + (e % d)];
  float acc = 0.f;
  float nxt[4];  // the next 32-column slab of the 32 centroid rows is fetched while the current one is consumed
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const uint32_t r = qi + 8u * u;
    nxt[u] = (j0 + r < k && cj < d) ? c[(size_t)(j0 + r) * ldc + cj] : 0.f;
  }
  for (uint32_t k0 = 0; k0 < d; k0 += 32) {
    __syncthreads();
#pragma unroll
    for (int u = 0; u < 4; ++u) cs[qi + 8u * u][cj] = nxt[u];
    if (k0 + 32 < d) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t r = qi + 8u * u, col = k0 + 32 + cj;
        nxt[u] = (j0 + r < k && col < d) ? c[(size_t)(j0 + r) * ldc + col] : 0.f;
      }
    }
    __syncthreads();
    if (qi < n) {
      const uint32_t lim = min(32u, d - k0);
      for (uint32_t t = 0; t < lim; ++t) acc = __fmaf_rn(xs[qi * d + k0 + t], cs[cj][t], acc);
    }
  }
  if (qi < n && j0 + cj < k) scores[(size_t)qi * k + j0 + cj] = __fsub_rn(acc, hn[j0 + cj]);
}

__global__ void argmax_splits_kernel(const float* __restrict__ best_s, const uint32_t* __restrict__ best_j, uint32_t n,
                                     uint32_t splits, uint32_t* __restrict__ assign) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float bs = -INFINITY;
  uint32_t bj = 0xFFFFFFFFu;
  for (uint32_t s = 0; s < splits; ++s) {
    const float v = best_s[(size_t)i * splits + s];
    const uint32_t j = best_j[(size_t)i * splits + s];
    if (v > bs || (v == bs && j < bj)) {
      bs = v;
      bj = j;
    }
  }
  assign[i] = bj;
}

__global__ void half_norms_kernel(const float* __restrict__ c, size_t ldc, uint32_t k, uint32_t d, float* __restrict__ hn) {
  const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
  if (j >= k) return;
  float s = 0.f;
  for (uint32_t t = lane; t < d; t += 32) {
    const float v = c[(size_t)j * ldc + t];
    s = __fmaf_rn(v, v, s);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
  if (lane == 0) hn[j] = 0.5f * s;
}

__global__ void init_centroids_kernel(const float* __restrict__ x, size_t ldx, uint32_t n, uint32_t d, uint32_t k,
                                      uint64_t seed, float* __restrict__ cent) {
  const uint32_t j = blockIdx.x;
  const uint64_t lo = (uint64_t)j * n / k, hi = (uint64_t)(j + 1) * n / k;  // disjoint buckets -> distinct rows
  const uint64_t span = hi > lo ? hi - lo : 1;
  const uint64_t row = lo + splitmix64(seed ^ (0x9E3779B97F4A7C15ull * (j + 1))) % span;
  for (uint32_t t = threadIdx.x; t < d; t += blockDim.x) cent[(size_t)j * d + t] = x[(size_t)row * ldx + t];
}

constexpr int kUpdThreads = 256, kUpdRegs = 16;  // d <= 4096

// One CTA per centroid: walk the assignment array in point order (ballot compaction keeps the order), sum member
// rows in that order -> deterministic means without atomics.
__global__ void __launch_bounds__(kUpdThreads)
centroid_update_kernel(const float* __restrict__ x, size_t ldx, uint32_t n, uint32_t d, const uint32_t* __restrict__ assign,
                       float* __restrict__ cent, uint32_t* __restrict__ counts) {
  __shared__ uint32_t members[kUpdThreads];
  __shared__ uint32_t warp_cnt[kUpdThreads / 32];
  const uint32_t j = blockIdx.x, tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  float acc[kUpdRegs];
#pragma unroll
  for (int u = 0; u < kUpdRegs; ++u) acc[u] = 0.f;
  uint32_t total = 0;
  for (uint32_t base = 0; base < n; base += kUpdThreads) {
    const uint32_t i = base + tid;
    const bool mine = i < n && assign[i] == j;
    const uint32_t cnt = (uint32_t)__syncthreads_count(mine);
    if (cnt == 0) continue;
    const uint32_t bal = __ballot_sync(0xffffffffu, mine);
    if (lane == 0) warp_cnt[warp] = __popc(bal);
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w = 0; w < warp; ++w) off += warp_cnt[w];
    if (mine) members[off + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    for (uint32_t e = 0; e < cnt; ++e) {
      const float* row = x + (size_t)members[e] * ldx;
#pragma unroll
      for (int u = 0; u < kUpdRegs; ++u) {
        const uint32_t t = tid + (uint32_t)u * kUpdThreads;
        if (t < d) acc[u] = __fadd_rn(acc[u], row[t]);
      }
    }
    total += cnt;
    __syncthreads();
  }
  if (total > 0) {
    const float inv_n = (float)total;
#pragma unroll
    for (int u = 0; u < kUpdRegs; ++u) {
      const uint32_t t = tid + (uint32_t)u * kUpdThreads;
      if (t < d) cent[(size_t)j * d + t] = __fdiv_rn(acc[u], inv_n);
    }
  }
  if (tid == 0) counts[j] = total;
}

// Empty clusters take a perturbed copy of the currently largest cluster's centroid, which is perturbed the other way
// (the usual split rule of k-means implementations, FAISS' included); counts are split so that one big cluster can
// feed several empty ones.
__global__ void __launch_bounds__(256)
fix_empty_kernel(float* __restrict__ cent, uint32_t* __restrict__ counts, uint32_t k, uint32_t d) {
  __shared__ uint32_t big;
  constexpr float eps = 1.f / 1024.f;
  for (uint32_t e = 0; e < k; ++e) {
    if (counts[e] != 0) continue;  // uniform: every thread reads the same word
    if (threadIdx.x == 0) {
      uint32_t b = 0, bc = 0;
      for (uint32_t j = 0; j < k; ++j)
        if (counts[j] > bc) {
          bc = counts[j];
          b = j;
        }
      big = bc >= 2 ? b : 0xFFFFFFFFu;
    }
    __syncthreads();
    const uint32_t b = big;
    if (b == 0xFFFFFFFFu) return;
    for (uint32_t t = threadIdx.x; t < d; t += blockDim.x) {
      const float v = cent[(size_t)b * d + t];
      const float up = v * (1.f + eps), dn = v * (1.f - eps);
      cent[(size_t)e * d + t] = (t & 1u) ? dn : up;
      cent[(size_t)b * d + t] = (t & 1u) ? up : dn;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const uint32_t half = counts[b] / 2;
      counts[e] = half;
      counts[b] -= half;
    }
    __syncthreads();
  }
}

uint32_t pick_splits(uint32_t n, uint32_t k, int num_sms) {
  const uint32_t row_tiles = (n + BM - 1) / BM, col_tiles = (k + BN - 1) / BN;
  if (row_tiles >= (uint32_t)num_sms) return 1;
  uint32_t s = ((uint32_t)num_sms + row_tiles - 1) / row_tiles;
  return s < col_tiles ? s : col_tiles;
}

}  // namespace

cudaError_t half_norms_launch(const float* c, size_t ldc, uint32_t k, uint32_t d, float* hn, cudaStream_t stream) {
  if (k == 0) return cudaSuccess;
  half_norms_kernel<<<(k * 32u + 255u) / 256u, 256, 0, stream>>>(c, ldc, k, d, hn);
  count_launch();
  return cudaGetLastError();
}

cudaError_t dot_bias_scores_launch(const float* x, size_t ldx, uint32_t n, const float* c, size_t ldc, const float* hn,
                                   uint32_t k, uint32_t d, float* scores, cudaStream_t stream) {
  if (n == 0 || k == 0) return cudaSuccess;
  if (n <= (uint32_t)kSmallN && (size_t)n * d * 4 <= 48u * 1024u) {
    dot_bias_small_kernel<<<(k + 31u) / 32u, 256, (size_t)n * d * 4, stream>>>(x, ldx, n, c, ldc, hn, k, d, scores);
    count_launch();
    return cudaGetLastError();
  }
  // a few CTAs per point tile so that a handful of queries still spreads over the SMs
  const uint32_t col_tiles = (k + BN - 1) / BN;
  const uint32_t row_tiles = (n + BM - 1) / BM;
  uint32_t splits = row_tiles >= 148u ? 1u : (148u + row_tiles - 1) / row_tiles;
  if (splits > col_tiles) splits = col_tiles;
  const uint32_t cols_per_cta = ((col_tiles + splits - 1) / splits) * BN;
  splits = (k + cols_per_cta - 1) / cols_per_cta;
  dot_bias_kernel<0><<<dim3(row_tiles, splits), NT, 0, stream>>>(x, ldx, n, c, ldc, hn, k, d, cols_per_cta, splits, scores,
                                                                 nullptr, nullptr);
  count_launch();
  return cudaGetLastError();
}

size_t nearest_centroid_ws_bytes(uint32_t n, uint32_t k, int num_sms) {
  return (size_t)n * pick_splits(n, k, num_sms) * 8u;
}

cudaError_t nearest_centroid_launch(const float* x, size_t ldx, uint32_t n, const float* c, size_t ldc, const float* hn,
                                    uint32_t k, uint32_t d, uint32_t* assign, void* ws, int num_sms, cudaStream_t stream) {
  if (n == 0 || k == 0) return cudaSuccess;
  const uint32_t col_tiles = (k + BN - 1) / BN, row_tiles = (n + BM - 1) / BM;
  uint32_t splits = pick_splits(n, k, num_sms);
  const uint32_t cols_per_cta = ((col_tiles + splits - 1) / splits) * BN;
  const uint32_t used = (k + cols_per_cta - 1) / cols_per_cta;  // <= splits
  float* bs = reinterpret_cast<float*>(ws);
  uint32_t* bj = reinterpret_cast<uint32_t*>(bs + (size_t)n * splits);
  dot_bias_kernel<1><<<dim3(row_tiles, used), NT, 0, stream>>>(x, ldx, n, c, ldc, hn, k, d, cols_per_cta, used, nullptr, bs,
                                                               bj);
  argmax_splits_kernel<<<(n + 255u) / 256u, 256, 0, stream>>>(bs, bj, n, used, assign);
  count_launch(2);
  return cudaGetLastError();
}

cudaError_t kmeans_train(const float* x, size_t ldx, uint32_t n, uint32_t d, uint32_t k, int niter, uint64_t seed,
                         float* centroids, int num_sms, cudaStream_t stream, bool warm_start) {
  if (n == 0 || k == 0 || d == 0 || d > (uint32_t)(kUpdThreads * kUpdRegs) || n < k) return cudaErrorInvalidValue;
  uint32_t *assign = nullptr, *counts = nullptr;
  float* hn = nullptr;
  void* ws = nullptr;
  cudaError_t e = cudaSuccess;
  auto done = [&](cudaError_t r) {
    cudaStreamSynchronize(stream);
    cudaFree(assign);
    cudaFree(counts);
    cudaFree(hn);
    cudaFree(ws);
    return r;
  };
  if ((e = cudaMalloc(&assign, (size_t)n * 4)) != cudaSuccess) return done(e);
  if ((e = cudaMalloc(&counts, (size_t)k * 4)) != cudaSuccess) return done(e);
  if ((e = cudaMalloc(&hn, (size_t)k * 4)) != cudaSuccess) return done(e);
  if ((e = cudaMalloc(&ws, nearest_centroid_ws_bytes(n, k, num_sms))) != cudaSuccess) return done(e);
  if (!warm_start) {
    init_centroids_kernel<<<k, 128, 0, stream>>>(x, ldx, n, d, k, seed, centroids);
    count_launch();
  }
  for (int it = 0; it < niter; ++it) {
    if ((e = half_norms_launch(centroids, d, k, d, hn, stream)) != cudaSuccess) return done(e);
    if ((e = nearest_centroid_launch(x, ldx, n, centroids, d, hn, k, d, assign, ws, num_sms, stream)) != cudaSuccess)
      return done(e);
    centroid_update_kernel<<<k, kUpdThreads, 0, stream>>>(x, ldx, n, d, assign, centroids, counts);
    fix_empty_kernel<<<1, 256, 0, stream>>>(centroids, counts, k, d);
    count_launch(2);
    if ((e = cudaGetLastError()) != cudaSuccess) return done(e);
  }
  return done(cudaStreamSynchronize(stream));
}

// ---- small dense helpers for the OPQ rotation (ivfpq.cu): transposes, A <- a*A + b*B, Frobenius norm ---------------
namespace {
__global__ void transpose_kernel(const float* __restrict__ in, uint32_t rows, uint32_t cols, float* __restrict__ out) {
  __shared__ float tile[32][33];
  const uint32_t bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (uint32_t r = threadIdx.y; r < 32; r += 8) {
    const uint32_t gr = by + r, gc = bx + threadIdx.x;
    tile[r][threadIdx.x] = (gr < rows && gc < cols) ? in[(size_t)gr * cols + gc] : 0.f;
  }
  __syncthreads();
  for (uint32_t r = threadIdx.y; r < 32; r += 8) {
    const uint32_t gr = bx + r, gc = by + threadIdx.x;  // out is [cols][rows]
    if (gr < cols && gc < rows) out[(size_t)gr * rows + gc] = tile[threadIdx.x][r];
  }
}
__global__ void axpby_kernel(float a, const float* __restrict__ A, float b, const float* __restrict__ B, size_t n,
                             float* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    out[i] = __fmaf_rn(a, A[i], __fmul_rn(b, B[i]));
}
// deterministic: one CTA, fixed strides and a fixed tree
__global__ void __launch_bounds__(1024) sumsq_kernel(const float* __restrict__ A, size_t n, float* __restrict__ out) {
  __shared__ float red[32];
  float s = 0.f;
  for (size_t i = threadIdx.x; i < n; i += 1024) s = __fmaf_rn(A[i], A[i], s);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 32; ++w) t = __fadd_rn(t, red[w]);
    *out = t;
  }
}
__global__ void scale_by_rsqrt_kernel(float* __restrict__ A, size_t n, const float* __restrict__ sumsq) {
  const float inv = rsqrtf(fmaxf(*sumsq, 1e-30f));
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) A[i] *= inv;
}
}  // namespace

cudaError_t transpose_launch(const float* in, uint32_t rows, uint32_t cols, float* out, cudaStream_t stream) {
  transpose_kernel<<<dim3((cols + 31) / 32, (rows + 31) / 32), dim3(32, 8), 0, stream>>>(in, rows, cols, out);
  count_launch();
  return cudaGetLastError();
}

// Orthogonal polar factor of the square matrix M (d x d, row-major, overwritten): the solution U V^T of the orthogonal
// Procrustes problem for M = U S V^T, by Newton-Schulz iterations A <- 1.5 A - 0.5 (A A^T) A on A = M / |M|_F — matrix
// products only (the tile GEMM above), no SVD library. ws: 3 * d * d floats + 1.
cudaError_t polar_factor_launch(float* M, uint32_t d, int iters, float* ws, const float* zeros_d, cudaStream_t stream) {
  const size_t nn = (size_t)d * d;
  float *G = ws, *AT = ws + nn, *T = ws + 2 * nn, *ss = ws + 3 * nn;
  sumsq_kernel<<<1, 1024, 0, stream>>>(M, nn, ss);
  scale_by_rsqrt_kernel<<<256, 256, 0, stream>>>(M, nn, ss);
  count_launch(2);
  for (int it = 0; it < iters; ++it) {
    cudaError_t e = dot_bias_scores_launch(M, d, d, M, d, zeros_d, d, d, G, stream);        // G = A A^T
    if (e != cudaSuccess) return e;
    if ((e = transpose_launch(M, d, d, AT, stream)) != cudaSuccess) return e;               // AT = A^T
    if ((e = dot_bias_scores_launch(G, d, d, AT, d, zeros_d, d, d, T, stream)) != cudaSuccess) return e;  // T = G A
    axpby_kernel<<<256, 256, 0, stream>>>(1.5f, M, -0.5f, T, nn, M);
    count_launch();
  }
  return cudaGetLastError();
}

}  // namespace nvdb_b200
