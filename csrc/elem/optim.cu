// Fused Adam / SGD over the flat parameter storage (one launch for the whole model).
//
// Replaces torch.optim.Adam's per-parameter update loop, optimizer.zero_grad() and the weight
// cast that a mixed-precision step needs (reference: R:704 build, R:437-438 / R:552-553 step and
// zero_grad).  Streams per element: read g, p, m, v; write p, m, v, the bf16
// copy of p that the tensor-core kernels read, and (optionally) g = 0 for the next window.
// `grad_scale` carries the 1/world and 1/K factors of the gradient exchange, so averaging costs
// no separate pass.  Semantics = torch.optim.Adam (no amsgrad, L2-style weight decay).
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

__global__ void __launch_bounds__(256)
fused_adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                  float* __restrict__ v, long long n, float lr, float b1, float b2, float eps, float wd,
                  float gscale, float bc1, float bc2_sqrt, __nv_bfloat16* __restrict__ p_lp,
                  int zero_grad) {
  const long long n4 = n >> 2;
  const float step = lr / bc1;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 gg = reinterpret_cast<float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pa[4] = {pp.x, pp.y, pp.z, pp.w}, ga[4] = {gg.x, gg.y, gg.z, gg.w};
    float ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gr = ga[j] * gscale + wd * pa[j];
      ma[j] = b1 * ma[j] + (1.f - b1) * gr;
      va[j] = b2 * va[j] + (1.f - b2) * gr * gr;
      float denom = __fdiv_rn(__fsqrt_rn(va[j]), bc2_sqrt) + eps;
      pa[j] -= step * __fdiv_rn(ma[j], denom);
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (p_lp != nullptr) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pa[0], pa[1]), hi = __floats2bfloat162_rn(pa[2], pa[3]);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(p_lp)[i] = pk;
    }
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256)
fused_sgd_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ buf, long long n,
                 float lr, float momentum, float wd, float gscale, __nv_bfloat16* __restrict__ p_lp,
                 int zero_grad, int first_step) {
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 gg = reinterpret_cast<float4*>(g)[i];
    float pa[4] = {pp.x, pp.y, pp.z, pp.w}, ga[4] = {gg.x, gg.y, gg.z, gg.w};
    float ba[4] = {0.f, 0.f, 0.f, 0.f};
    if (buf != nullptr && !first_step) {
      float4 bb = reinterpret_cast<float4*>(buf)[i];
      ba[0] = bb.x; ba[1] = bb.y; ba[2] = bb.z; ba[3] = bb.w;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gr = ga[j] * gscale + wd * pa[j];
      if (buf != nullptr) { ba[j] = first_step ? gr : momentum * ba[j] + gr; gr = ba[j]; }
      pa[j] -= lr * gr;
    }
    reinterpret_cast<float4*>(p)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    if (buf != nullptr) reinterpret_cast<float4*>(buf)[i] = make_float4(ba[0], ba[1], ba[2], ba[3]);
    if (p_lp != nullptr) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pa[0], pa[1]), hi = __floats2bfloat162_rn(pa[2], pa[3]);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      reinterpret_cast<uint2*>(p_lp)[i] = pk;
    }
    if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

}  // namespace b200

using namespace b200;

// n must be a multiple of 4 (the flat buffers are padded to 64 elements per tensor).
extern "C" int b200seg_fused_adam(float* p, float* g, float* m, float* v, long long n, float lr, float b1,
                                  float b2, float eps, float wd, float gscale, float bc1, float bc2,
                                  void* p_lowp, int zero_grad, cudaStream_t st) {
  if (n & 3) return (int)cudaErrorInvalidValue;
  int grid = grid_for(n / 4, 256, 8);
  fused_adam_kernel<<<grid, 256, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, wd, gscale, bc1, sqrtf(bc2),
                                          (__nv_bfloat16*)p_lowp, zero_grad);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_fused_sgd(float* p, float* g, float* buf, long long n, float lr, float momentum,
                                 float wd, float gscale, void* p_lowp, int zero_grad, int first_step,
                                 cudaStream_t st) {
  if (n & 3) return (int)cudaErrorInvalidValue;
  int grid = grid_for(n / 4, 256, 8);
  fused_sgd_kernel<<<grid, 256, 0, st>>>(p, g, buf, n, lr, momentum, wd, gscale, (__nv_bfloat16*)p_lowp,
                                         zero_grad, first_step);
  return (int)cudaGetLastError();
}
