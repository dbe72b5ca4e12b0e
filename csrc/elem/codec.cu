// Gradient wire codecs on the device: global max-abs, quantize, dequantize(+accumulate).
//
// The reference finds max|g| with one GPU->CPU sync per parameter tensor (R:451-463), quantizes
// each of the 100 tensors with five small torch ops (R:484-487 / R:474) and copies them to the
// host one by one (R:491), then reverses all of that on receive (R:513-545) -- SURVEY K15, K16.
// Here: ONE amax launch over the whole flat gradient buffer, ONE quantize launch, ONE
// dequantize(-accumulate) launch; the payload never leaves the device.
// Codec ids: 0 fp32, 1 bf16, 2 fp16, 3 q201 (round(g/max*100) as fp16), 4 q21 (round(g/max*10) as
// int8), 5 int8s (round(g/max*127)), 6 fp8 e4m3 (g/max*448).
#include "../common.cuh"
#include "b200seg.h"
#include <cuda_fp16.h>
#include <cuda_fp8.h>

namespace b200 {

__global__ void __launch_bounds__(256) amax_kernel(const float* __restrict__ x, long long n,
                                                   float* __restrict__ out) {
  float m = 0.f;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 v = reinterpret_cast<const float4*>(x)[i];
    m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(x[i]));
  m = warp_max(m);
  __shared__ float s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x < 8) {
    m = s[threadIdx.x];
    for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffu, m, o));
    // non-negative floats order like their bit patterns
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));
  }
}

template <int CODEC>
__device__ __forceinline__ void enc_one(float g, float inv_scale_levels, void* out, long long i) {
  if (CODEC == 1) reinterpret_cast<__nv_bfloat16*>(out)[i] = __float2bfloat16(g);
  if (CODEC == 2) reinterpret_cast<__half*>(out)[i] = __float2half(g);
  if (CODEC == 3) reinterpret_cast<__half*>(out)[i] = __float2half(rintf(g * inv_scale_levels));
  if (CODEC == 4 || CODEC == 5) reinterpret_cast<int8_t*>(out)[i] = (int8_t)rintf(g * inv_scale_levels);
  if (CODEC == 6) reinterpret_cast<__nv_fp8_e4m3*>(out)[i] = __nv_fp8_e4m3(fminf(fmaxf(g * inv_scale_levels, -448.f), 448.f));
}

template <int CODEC>
__global__ void __launch_bounds__(256) encode_kernel(const float* __restrict__ x, long long n,
                                                     const float* __restrict__ scale, float levels,
                                                     void* __restrict__ out) {
  const float s = *scale;
  // (g / max) * S as in R:487; a zero max encodes zeros (reference Q4 crash)
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float g = x[i];
    float q = (s > 0.f) ? __fmul_rn(__fdiv_rn(g, s), levels) : 0.f;   // IEEE ops: bit parity with torch
    if (CODEC == 1 || CODEC == 2) q = g;
    enc_one<CODEC>(q, 1.f, out, i);
  }
}

template <int CODEC>
__global__ void __launch_bounds__(256) decode_kernel(const void* __restrict__ in, long long n,
                                                     const float* __restrict__ scale, float levels,
                                                     float* __restrict__ out, int accumulate) {
  const float s = *scale;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    float q;
    if (CODEC == 1) q = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(in)[i]);
    else if (CODEC == 2 || CODEC == 3) q = __half2float(reinterpret_cast<const __half*>(in)[i]);
    else if (CODEC == 4 || CODEC == 5) q = (float)reinterpret_cast<const int8_t*>(in)[i];
    else q = float(reinterpret_cast<const __nv_fp8_e4m3*>(in)[i]);
    float val = (CODEC == 1 || CODEC == 2) ? q : __fmul_rn(__fdiv_rn(q, levels), s);      // q / S * max as in R:543
    out[i] = accumulate ? out[i] + val : val;
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200seg_amax(const float* x, long long n, float* out, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float), st);
  if (e != cudaSuccess) return (int)e;
  amax_kernel<<<grid_for(n / 4 + 1, 256, 4), 256, 0, st>>>(x, n, out);
  return (int)cudaGetLastError();
}

static float codec_levels(int codec) {
  switch (codec) { case 3: return 100.f; case 4: return 10.f; case 5: return 127.f; case 6: return 448.f; }
  return 1.f;
}

extern "C" int b200seg_codec_encode(const float* x, long long n, const float* scale, int codec, void* out,
                                    cudaStream_t st) {
  int grid = grid_for(n, 256, 8);
  float lv = codec_levels(codec);
  switch (codec) {
    case 0: return (int)cudaMemcpyAsync(out, x, n * 4, cudaMemcpyDeviceToDevice, st);
    case 1: encode_kernel<1><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    case 2: encode_kernel<2><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    case 3: encode_kernel<3><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    case 4: encode_kernel<4><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    case 5: encode_kernel<5><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    case 6: encode_kernel<6><<<grid, 256, 0, st>>>(x, n, scale, lv, out); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}

extern "C" int b200seg_codec_decode(const void* in, long long n, const float* scale, int codec, float* out,
                                    int accumulate, cudaStream_t st) {
  int grid = grid_for(n, 256, 8);
  float lv = codec_levels(codec);
  switch (codec) {
    case 1: decode_kernel<1><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    case 2: decode_kernel<2><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    case 3: decode_kernel<3><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    case 4: decode_kernel<4><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    case 5: decode_kernel<5><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    case 6: decode_kernel<6><<<grid, 256, 0, st>>>(in, n, scale, lv, out, accumulate); break;
    default: return (int)cudaErrorInvalidValue;
  }
  return (int)cudaGetLastError();
}
