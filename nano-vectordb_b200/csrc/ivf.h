// ivf.h — launchers shared by the IVF-PQ candidate generator (kmeans.cu, ivfpq.cu). SURVEY.md §8(f) f-1.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nvdb_b200 {

// hn[j] = 0.5 * ||c_j||^2   (c: [k x d] f32, row pitch ldc elements)
cudaError_t half_norms_launch(const float* c, size_t ldc, uint32_t k, uint32_t d, float* hn, cudaStream_t stream);

// S[i][j] = x_i . c_j - hn[j] for i < n, j < k  (== -0.5*||x_i - c_j||^2 + 0.5*||x_i||^2: the same ranking over j as L2,
// descending). x: [n x d] f32 with row pitch ldx; c: [k x d] f32 with row pitch ldc. scores: [n x k].
cudaError_t dot_bias_scores_launch(const float* x, size_t ldx, uint32_t n, const float* c, size_t ldc, const float* hn,
                                   uint32_t k, uint32_t d, float* scores, cudaStream_t stream);

// assign[i] = argmax_j S[i][j] (lowest j among equal scores) = nearest centroid in L2. `ws` holds 2 * n * splits 32-bit
// words when the launcher decides to split the centroids over several CTAs (small n); query the size first.
size_t nearest_centroid_ws_bytes(uint32_t n, uint32_t k, int num_sms);
cudaError_t nearest_centroid_launch(const float* x, size_t ldx, uint32_t n, const float* c, size_t ldc, const float* hn,
                                    uint32_t k, uint32_t d, uint32_t* assign, void* ws, int num_sms, cudaStream_t stream);

// Lloyd k-means, deterministic (fixed seed -> fixed centroids): evenly spaced seeded initial rows, assignment by
// nearest_centroid, centroid update in ascending point order, empty clusters re-seeded from the largest cluster.
// centroids: [k x d] dense. Allocates and frees its own scratch (build-time path).
cudaError_t kmeans_train(const float* x, size_t ldx, uint32_t n, uint32_t d, uint32_t k, int niter, uint64_t seed,
                         float* centroids, int num_sms, cudaStream_t stream, bool warm_start = false);

// out[c][r] = in[r][c]
cudaError_t transpose_launch(const float* in, uint32_t rows, uint32_t cols, float* out, cudaStream_t stream);
// M <- orthogonal polar factor of M (d x d) by `iters` Newton-Schulz steps; ws: 3*d*d + 1 floats; zeros_d: d zeros
cudaError_t polar_factor_launch(float* M, uint32_t d, int iters, float* ws, const float* zeros_d, cudaStream_t stream);

}  // namespace nvdb_b200
