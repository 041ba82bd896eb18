// synth.cu — on-device synthetic base generator + the reference converters' arithmetic.
//
// The generator is ours (integer hash, restated for the CPU in oracle/nvdb_oracle.c: orc_synth_i16 /
// orc_synth_rows_f32) so that any row of a 100 M-row index can be regenerated on the host bit-for-bit.
// The conversions follow the reference tools so that a device-generated base equals "generate f32 ->
// nvdb_convert_f16 / nvdb_quantize_i8":
//   f16: IEEE round-to-nearest-even                       (tools/nvdb_convert_f16.cpp:95-119, F16C path)
//   i8 : scale = max|x|/127 (1 if zero row), inv = 1.0f/scale, q = clamp(lrint(x*inv), -127, 127), all FP32
//                                                         (apps/nvdb_quantize_i8.cpp:12-16,72-83)
#include "common.cuh"
#include "scan.h"

namespace nvdb_b200 {
namespace {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
__device__ __forceinline__ float synth_value(uint64_t seed, uint64_t gidx, float c) {
  const uint64_t h = mix64(seed * 0x9E3779B97F4A7C15ULL + gidx * 0xD1B54A32D192ED03ULL + 0x2545F4914F6CDD1DULL);
  const int v = (int)(h >> 48) - 32768;
  return __fmul_rn((float)v, c);
}

template <int DT>
__global__ void synth_fill_kernel(uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row, float c,
                                  void* __restrict__ payload) {
  const uint64_t total = n * dim;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const float f = synth_value(seed, first_row * dim + i, c);
    if constexpr (DT == kDtF32) reinterpret_cast<float*>(payload)[i] = f;
    else reinterpret_cast<__half*>(payload)[i] = __float2half_rn(f);
  }
}

// one warp per row
__global__ void synth_fill_i8_kernel(uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row, float c,
                                     int8_t* __restrict__ payload, float* __restrict__ scales) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t r = (((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < n; r += warps) {
    const uint64_t g0 = (first_row + r) * dim;
    float mx = 0.f;
    for (uint32_t j = lane; j < dim; j += 32) mx = fmaxf(mx, fabsf(synth_value(seed, g0 + j, c)));
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    const float scale = (mx > 0.f) ? __fdiv_rn(mx, 127.f) : 1.f;
    const float inv = __fdiv_rn(1.0f, scale);
    for (uint32_t j = lane; j < dim; j += 32) {
      int v = __float2int_rn(__fmul_rn(synth_value(seed, g0 + j, c), inv));
      v = v > 127 ? 127 : (v < -127 ? -127 : v);
      payload[r * dim + j] = (int8_t)v;
    }
    if (lane == 0) scales[r] = scale;
  }
}

}  // namespace

cudaError_t synth_fill_launch(int dtype, uint64_t n, uint32_t dim, uint64_t seed, uint64_t first_row, void* payload,
                              float* scales, cudaStream_t stream) {
  if (n == 0) return cudaSuccess;
  const float c = (float)(sqrt(3.0 / (double)dim) / 32768.0);
  const int grid = 148 * 8;
  if (dtype == kDtF32) synth_fill_kernel<kDtF32><<<grid, 256, 0, stream>>>(n, dim, seed, first_row, c, payload);
  else if (dtype == kDtF16) synth_fill_kernel<kDtF16><<<grid, 256, 0, stream>>>(n, dim, seed, first_row, c, payload);
  else if (dtype == kDtI8)
    synth_fill_i8_kernel<<<grid, 256, 0, stream>>>(n, dim, seed, first_row, c, reinterpret_cast<int8_t*>(payload), scales);
  else return cudaErrorInvalidValue;
  count_launch();
  return cudaGetLastError();
}

}  // namespace nvdb_b200
