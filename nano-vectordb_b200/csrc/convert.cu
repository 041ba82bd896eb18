// convert.cu — the reference's fixture converters on the device.
//
//   f32 -> f16 : IEEE round-to-nearest-even, element-wise         (tools/nvdb_convert_f16.cpp:20-119: the scalar and the
//                                                                  F16C path both implement RNE; == cvt.rn.f16.f32)
//   f32|f16 -> i8 (+ one f32 scale per row)                       (apps/nvdb_quantize_i8.cpp:57-85)
//       rowf = float(row)  (f16 widened exactly);  max_abs = max |rowf_j|;  scale = max_abs > 0 ? max_abs / 127.f : 1.f
//       inv = 1.0f / scale;  q_j = clamp(lrint(rowf_j * inv), -127, 127)     — all in FP32, lrint = round-half-even
//
// Both are pure streaming kernels (HBM-bound: read the source once, write the destination once).
#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

__global__ void __launch_bounds__(256) convert_f32_f16_kernel(const float* __restrict__ in, __half* __restrict__ out,
                                                              uint64_t total) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // 4 elements per thread where alignment allows (16-byte loads, 8-byte stores)
  const uint64_t quads = ((reinterpret_cast<uintptr_t>(in) & 15u) == 0 && (reinterpret_cast<uintptr_t>(out) & 7u) == 0) ? total / 4 : 0;
  for (uint64_t q = i; q < quads; q += stride) {
    const float4 v = reinterpret_cast<const float4*>(in)[q];
    __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(out)[q] = o;
  }
  for (uint64_t e = quads * 4 + i; e < total; e += stride) out[e] = __float2half_rn(in[e]);
}

// one warp per row
template <int SRC>
__global__ void __launch_bounds__(256) quantize_i8_kernel(const void* __restrict__ in, uint64_t n, uint32_t dim,
                                                          int8_t* __restrict__ out, float* __restrict__ scales) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  auto load = [&](uint64_t r, uint32_t j) -> float {
    if constexpr (SRC == kDtF32) return reinterpret_cast<const float*>(in)[r * dim + j];
    else return __half2float(reinterpret_cast<const __half*>(in)[r * dim + j]);
  };
  for (uint64_t r = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < n; r += warps) {
    float mx = 0.f;
    for (uint32_t j = lane; j < dim; j += 32) mx = fmaxf(mx, fabsf(load(r, j)));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    const float scale = (mx > 0.f) ? __fdiv_rn(mx, 127.f) : 1.f;
    const float inv = __fdiv_rn(1.0f, scale);
    for (uint32_t j = lane; j < dim; j += 32) {
      int v = __float2int_rn(__fmul_rn(load(r, j), inv));
      v = v > 127 ? 127 : (v < -127 ? -127 : v);
      out[r * dim + j] = (int8_t)v;
    }
    if (lane == 0) scales[r] = scale;
  }
}

}  // namespace

cudaError_t convert_launch(int src_dtype, const void* src, int dst_dtype, uint64_t n, uint32_t dim, void* dst, float* scales,
                           int num_sms, cudaStream_t stream) {
  if (n == 0 || dim == 0) return cudaSuccess;
  const int grid = num_sms * 8;
  if (src_dtype == kDtF32 && dst_dtype == kDtF16) {
    convert_f32_f16_kernel<<<grid, 256, 0, stream>>>(reinterpret_cast<const float*>(src), reinterpret_cast<__half*>(dst), n * dim);
  } else if (dst_dtype == kDtI8 && src_dtype == kDtF32) {
    quantize_i8_kernel<kDtF32><<<grid, 256, 0, stream>>>(src, n, dim, reinterpret_cast<int8_t*>(dst), scales);
  } else if (dst_dtype == kDtI8 && src_dtype == kDtF16) {
    quantize_i8_kernel<kDtF16><<<grid, 256, 0, stream>>>(src, n, dim, reinterpret_cast<int8_t*>(dst), scales);
  } else {
    return cudaErrorInvalidValue;
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace nvdb_b200
