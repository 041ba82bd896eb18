"""TEST INFRASTRUCTURE — CPU restatement of IVF-PQ (IVFADC) candidate generation. Never imported by the product.

PARITY UNPINNED. The reference obtains this stage from FAISS (apps/nvdb_ivfpq_build.cpp:78-135: IndexIVFPQ over an
IndexFlatL2 quantiser, METRIC_L2, PQ_BITS=8, residual encoding; apps/nvdb_ivf_eval.cpp:398-408,482: nprobe +
idx->search(1, q, k_search)). FAISS is a third-party dependency that the reference does not vendor or pin
(CMake `find_package(faiss)`), it is not in /root/reference and not in this image, and the reference holds no golden
vectors for this stage. This file therefore restates the published algorithm — H. Jegou, M. Douze, C. Schmid,
"Product Quantization for Nearest Neighbor Search", TPAMI 2011, section IV (IVFADC) — in float64 numpy:

  * coarse quantiser: nearest of `nlist` centroids in L2 (lowest index among exact ties);
  * residual r = x - c(x) (computed in float32, as the kernels do), split into M sub-vectors of dsub dims, each coded by
    the index of its nearest of 256 codewords;
  * search: the nprobe nearest lists of q; for a probed list with centroid c the ADC distance of a stored vector is
    sum_m |(q - c)_m - codeword_m[code_m]|^2; the R smallest over all probed lists, ordered (distance, id).

The CUDA path accumulates in float32 with a fixed order, this file in float64, so comparisons are by tolerance
(tests/test_gpu_ivfpq.py states it). Lloyd k-means is restated too (`kmeans`) with the kernels' initialisation rule so that
training quality can be compared, not bits.
"""
from __future__ import annotations

import numpy as np

_MASK = (1 << 64) - 1


def splitmix64(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & _MASK
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
    return z ^ (z >> 31)


def init_rows(n: int, k: int, seed: int) -> np.ndarray:
    """rows the device k-means starts from (kmeans.cu: init_centroids_kernel): one seeded row per bucket j*n/k..(j+1)*n/k"""
    rows = np.empty(k, np.int64)
    for j in range(k):
        lo, hi = j * n // k, (j + 1) * n // k
        span = max(hi - lo, 1)
        rows[j] = lo + splitmix64(seed ^ ((0x9E3779B97F4A7C15 * (j + 1)) & _MASK)) % span
    return rows


def sq_dists(x: np.ndarray, c: np.ndarray) -> np.ndarray:
    x = np.asarray(x, np.float64)
    c = np.asarray(c, np.float64)
    return (x * x).sum(1)[:, None] - 2.0 * (x @ c.T) + (c * c).sum(1)[None, :]


def nearest(x: np.ndarray, c: np.ndarray) -> np.ndarray:
    return np.argmin(sq_dists(x, c), axis=1)  # argmin returns the lowest index among equal minima


def objective(x: np.ndarray, c: np.ndarray) -> float:
    return float(np.maximum(sq_dists(x, c).min(1), 0.0).mean())


def kmeans(x: np.ndarray, k: int, niter: int, seed: int) -> np.ndarray:
    x = np.asarray(x, np.float64)
    cent = x[init_rows(len(x), k, seed)].copy()
    for _ in range(niter):
        a = nearest(x, cent)
        counts = np.bincount(a, minlength=k)
        sums = np.zeros_like(cent)
        np.add.at(sums, a, x)
        nz = counts > 0
        cent[nz] = sums[nz] / counts[nz, None]
        for e in np.flatnonzero(counts == 0):  # split the largest cluster (kmeans.cu: fix_empty_kernel)
            b = int(np.argmax(counts))
            if counts[b] < 2:
                break
            eps = 1.0 / 1024.0
            sign = np.where(np.arange(cent.shape[1]) % 2 == 1, -1.0, 1.0)
            v = cent[b].copy()
            cent[e] = v * (1.0 + sign * eps)
            cent[b] = v * (1.0 - sign * eps)
            counts[e] = counts[b] // 2
            counts[b] -= counts[e]
    return cent


def train(x: np.ndarray, nlist: int, pq_m: int, niter: int, seed: int):
    """(coarse [nlist,d], codebooks [M,256,dsub]) with the build's seeding scheme (seed, seed+1+m)"""
    x = np.asarray(x, np.float32)
    d = x.shape[1]
    dsub = d // pq_m
    coarse = kmeans(x, nlist, niter, seed).astype(np.float32)
    res = x - coarse[nearest(x, coarse)]
    cb = np.stack([kmeans(res[:, m * dsub:(m + 1) * dsub], 256, niter, seed + 1 + m) for m in range(pq_m)]).astype(np.float32)
    return coarse, cb


def encode(x: np.ndarray, coarse: np.ndarray, codebooks: np.ndarray):
    """(list id [n], codes [n,M]) of float32 rows x"""
    x = np.asarray(x, np.float32)
    coarse = np.asarray(coarse, np.float32)
    lists = nearest(x, coarse)
    res = x - coarse[lists]  # float32 subtraction, as the kernel
    M, _, dsub = codebooks.shape
    codes = np.empty((len(x), M), np.uint8)
    for m in range(M):
        codes[:, m] = nearest(res[:, m * dsub:(m + 1) * dsub], codebooks[m])
    return lists.astype(np.uint32), codes


def group(lists: np.ndarray, codes: np.ndarray, nlist: int):
    """inverted-list layout of the device index: (list_off [nlist+1], ids [n], grouped codes [n,M])"""
    order = np.argsort(lists, kind="stable")
    off = np.zeros(nlist + 1, np.uint32)
    off[1:] = np.cumsum(np.bincount(lists, minlength=nlist))
    return off, order.astype(np.uint32), codes[order]


def adc_distances(q: np.ndarray, coarse, codebooks, list_off, ids, codes, nprobe: int):
    """ADC distance (float64) of every vector stored in the nprobe nearest lists of ONE query: (row ids, dists, probed lists)"""
    q = np.asarray(q, np.float32)
    coarse = np.asarray(coarse, np.float32)
    d2 = sq_dists(q[None, :], coarse)[0]
    probes = np.lexsort((np.arange(len(coarse)), d2))[:nprobe]
    M, _, dsub = codebooks.shape
    out_ids, out_d = [], []
    for l in probes:
        b, e = int(list_off[l]), int(list_off[l + 1])
        if e == b:
            continue
        r = (q - coarse[l]).reshape(M, 1, dsub)                       # float32
        diff = (r - np.asarray(codebooks, np.float32)).astype(np.float64)  # float32 subtraction, then widened
        lut = (diff * diff).sum(2)                                    # [M,256]
        cc = codes[b:e]
        out_d.append(lut[np.arange(M)[None, :], cc].sum(1))
        out_ids.append(ids[b:e])
    if not out_ids:
        return np.empty(0, np.uint32), np.empty(0, np.float64), probes
    return np.concatenate(out_ids), np.concatenate(out_d), probes


def search(queries, coarse, codebooks, list_off, ids, codes, nprobe: int, R: int):
    """(ids u32 [nq,R], dist f64 [nq,R]) ordered (distance asc, id asc); empty slots 0xFFFFFFFF / 1e30"""
    queries = np.asarray(queries, np.float32)
    out_i = np.full((len(queries), R), 0xFFFFFFFF, np.uint32)
    out_d = np.full((len(queries), R), 1e30, np.float64)
    for qi, q in enumerate(queries):
        vid, vd, _ = adc_distances(q, coarse, codebooks, list_off, ids, codes, nprobe)
        order = np.lexsort((vid, vd))[:R]
        out_i[qi, :len(order)] = vid[order]
        out_d[qi, :len(order)] = vd[order]
    return out_i, out_d


def exact_l2_topk(base: np.ndarray, queries: np.ndarray, k: int) -> np.ndarray:
    d2 = sq_dists(queries, base)
    return np.argsort(d2, axis=1, kind="stable")[:, :k]


def recall_at_k(found: np.ndarray, truth: np.ndarray) -> float:
    """fraction of the true top-k present anywhere in `found` (candidate recall when found has R > k columns)"""
    k = truth.shape[1]
    return float(np.mean([len(set(f.tolist()) & set(t.tolist())) / k for f, t in zip(found, truth)]))


def opq_train(x: np.ndarray, pq_m: int, opq_niter: int, seed: int, pq_iters_first=6, pq_iters_next=1):
    """Non-parametric OPQ (Ge, He, Ke, Sun, CVPR 2013; what faiss::OPQMatrix implements and the reference switches on with
    USE_OPQ=1, apps/nvdb_ivfpq_build.cpp:78-91): alternate PQ of X R (warm-started Lloyd steps, the build's seeds
    seed+1000+m) with the orthogonal Procrustes solution R = U V^T of X^T Y = U S V^T. Returns (R, quantisation error
    per round). float64; the device version uses Newton-Schulz instead of an SVD for the polar factor."""
    x = np.asarray(x, np.float64)
    n, d = x.shape
    dsub = d // pq_m
    R = np.eye(d)
    cbs = [None] * pq_m
    errs = []
    for it in range(opq_niter):
        xr = x @ R
        y = np.empty_like(xr)
        for m in range(pq_m):
            sub = xr[:, m * dsub:(m + 1) * dsub]
            if it == 0:
                cbs[m] = kmeans(sub, 256, pq_iters_first, seed + 1000 + m)
            else:
                for _ in range(pq_iters_next):
                    a = nearest(sub, cbs[m])
                    counts = np.bincount(a, minlength=256)
                    sums = np.zeros_like(cbs[m])
                    np.add.at(sums, a, sub)
                    nz = counts > 0
                    cbs[m][nz] = sums[nz] / counts[nz, None]
            y[:, m * dsub:(m + 1) * dsub] = cbs[m][nearest(sub, cbs[m])]
        errs.append(float(((xr - y) ** 2).sum() / n))
        u, _, vt = np.linalg.svd(x.T @ y)
        R = u @ vt
    return R, errs


def pq_error(x: np.ndarray, pq_m: int, niter: int, seed: int) -> float:
    """mean squared PQ reconstruction error of rows x with freshly trained codebooks"""
    x = np.asarray(x, np.float64)
    dsub = x.shape[1] // pq_m
    err = 0.0
    for m in range(pq_m):
        sub = x[:, m * dsub:(m + 1) * dsub]
        cb = kmeans(sub, 256, niter, seed + m)
        err += float(((sub - cb[nearest(sub, cb)]) ** 2).sum())
    return err / len(x)
