// BatchNorm(train) + ReLU (+ 2x2 max-pool) forward and backward on NHWC bf16 tensors.
//
// Replaces, per conv stage of the reference's DoubleConv (R:578-585) and DownBlock (R:595):
// cuDNN/ATen batch_norm (stats pass + normalise pass), the in-place ReLU, max_pool2d and their
// backward kernels.  The per-channel statistics of the forward come out of the
// conv epilogue (csrc/conv) as per-CTA partial sums, so the forward here is ONE pass:
//   y (raw conv output) -> relu(y*scale+shift) -> skip tensor (a channel slice of the decoder's
//   concat buffer, which removes torch.cat, R:616) and, for encoder blocks, the pooled tensor.
// Backward is reduce -> finalize -> apply; the pool's arg-max is recomputed from y (first
// maximum in row-major window order, as ATen), so no index tensor is stored.
#define B200SEG_PDL_CLASS 2
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

// Streaming (evict-first) loads / stores for the forward BN+ReLU(+pool) pass: its input is dead
// afterwards as far as the caches are concerned and its output is far larger than L2 on the levels that
// matter (fwd kernels 1.19 -> 1.01 and 0.49 -> 0.41 ms).  The backward passes keep default caching:
// apply re-reads what reduce just read, and with hints they got slower (profiles/variants_in_situ.txt).
__device__ __forceinline__ bf16x8 ld8s(const __nv_bfloat16* p) {
  const uint4 v = __ldcs(reinterpret_cast<const uint4*>(p));
  return *reinterpret_cast<const bf16x8*>(&v);
}
__device__ __forceinline__ void st8s(__nv_bfloat16* p, const bf16x8& x) {
  __stcs(reinterpret_cast<uint4*>(p), *reinterpret_cast<const uint4*>(&x));
}


constexpr int kThreads = 256;

// ---------------------------------------------------------------------------------------------
// Per-channel sum / sum of squares of a bf16 NHWC tensor -> partial[gridDim.x][2][C].
// (Only needed when the conv did not produce the partials itself, i.e. the cuDNN comparison arm.)
__global__ void __launch_bounds__(kThreads)
channel_stats_kernel(const __nv_bfloat16* __restrict__ y, long long P, int C, long long ld,
                     float* __restrict__ partial) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float sm[];
  const int CV = C >> 3;
  const int rows_pb = kThreads / CV;
  const int cv = threadIdx.x % CV, r = threadIdx.x / CV;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  if (r < rows_pb) {
    for (long long p = (long long)blockIdx.x * rows_pb + r; p < P; p += (long long)gridDim.x * rows_pb) {
      float f[8];
      unpack8(ld8(y + p * ld + cv * 8), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) { s[j] += f[j]; q[j] = fmaf(f[j], f[j], q[j]); }
    }
  }
  float* s_sum = sm;
  float* s_sq = sm + rows_pb * C;
  if (r < rows_pb) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_sum[r * C + cv * 8 + j] = s[j]; s_sq[r * C + cv * 8 + j] = q[j]; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += kThreads) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < rows_pb; ++i) { a += s_sum[i * C + c]; b += s_sq[i * C + c]; }
    partial[((size_t)blockIdx.x * 2 + 0) * C + c] = a;
    partial[((size_t)blockIdx.x * 2 + 1) * C + c] = b;
  }
}

// ---------------------------------------------------------------------------------------------
// partial[R][2][C] -> ss[4][C] = scale, shift, mean, invstd; running stats (momentum, unbiased
// variance) and num_batches_tracked as nn.BatchNorm2d (R:580).
// One warp per channel.  `conv_bias`: the bias of the conv that produced y.  The conv kernels do
// NOT add it: in training mode it cancels inside the normalisation (it only shifts the running
// mean), in eval mode it is folded into `shift`.
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void bn_finalize_kernel(const float* __restrict__ partial, int rows, long long count,
                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                   const float* __restrict__ conv_bias,
                                   float* __restrict__ running_mean, float* __restrict__ running_var,
                                   long long* __restrict__ nbt, float* __restrict__ ss, int C,
                                   float momentum, float eps, int training) {
  pdl_wait();
  pdl_trigger();
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  const double cb = conv_bias != nullptr ? (double)conv_bias[c] : 0.0;
  double mean, var;
  if (training) {
    double s = 0.0, q = 0.0;
    for (int r = lane; r < rows; r += 32) {
      s += (double)partial[((size_t)r * 2 + 0) * C + c];
      q += (double)partial[((size_t)r * 2 + 1) * C + c];
    }
    s = warp_sum_d(s);
    q = warp_sum_d(q);
    mean = s / (double)count;
    var = q / (double)count - mean * mean;
    if (var < 0.0) var = 0.0;
    if (lane == 0 && running_mean != nullptr) {
      double unb = var * ((double)count / (double)(count > 1 ? count - 1 : 1));
      running_mean[c] = (float)((1.0 - momentum) * (double)running_mean[c] + momentum * (mean + cb));
      running_var[c] = (float)((1.0 - momentum) * (double)running_var[c] + momentum * unb);
    }
    if (c == 0 && lane == 0 && nbt != nullptr) *nbt += 1;
  } else {
    mean = (double)running_mean[c] - cb;          // statistics of the bias-free conv output
    var = (double)running_var[c];
  }
  if (lane != 0) return;
  const double invstd = rsqrt(var + (double)eps);
  const double scale = (double)gamma[c] * invstd;
  ss[0 * C + c] = (float)scale;
  ss[1 * C + c] = (float)((double)beta[c] - mean * scale);
  ss[2 * C + c] = (float)mean;
  ss[3 * C + c] = (float)invstd;
}

// ---------------------------------------------------------------------------------------------
// Index math of the elementwise kernels.  A thread owns ONE 8-channel group for its whole life (the
// launchers make the grid stride a multiple of the number of groups), so the group index and the
// per-channel constants are computed once and a trip only advances the pixel.  Nothing in the loops
// divides 64-bit numbers: that division is ~100 emulated instructions and made these kernels
// issue-bound at 60 % of the HBM roofline.
struct PixWalk {
  int cv;                 // 8-channel group of this thread
  long long pix, step;    // first pixel (or 2x2 window) and pixels per trip
};
__device__ __forceinline__ PixWalk pix_walk(int CV) {
  const long long gs = (long long)gridDim.x * blockDim.x;      // multiple of CV (see grid_for_groups)
  const long long idx0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  PixWalk w;
  w.cv = (int)(idx0 % CV);
  w.pix = idx0 / CV;
  w.step = gs / CV;
  return w;
}
// 2x2 window index -> top-left pixel of the window (32-bit arithmetic; the launchers check the range)
__device__ __forceinline__ long long window_origin(long long win, int Wo, int Ho, int H, int W) {
  const unsigned u = (unsigned)win;
  const unsigned t = u / (unsigned)Wo, wo = u - t * (unsigned)Wo;
  const unsigned n = t / (unsigned)Ho, ho = t - n * (unsigned)Ho;
  return ((long long)n * H + 2 * ho) * W + 2 * wo;
}

template <bool POOL>
__global__ void __launch_bounds__(kThreads)
bn_relu_pool_fwd_kernel(const __nv_bfloat16* __restrict__ y, long long ld_y,
                        const float* __restrict__ ss, __nv_bfloat16* __restrict__ out, long long ld_o,
                        __nv_bfloat16* __restrict__ pool, long long ld_p, int N, int H, int W, int C) {
  pdl_wait();
  pdl_trigger();
  const int CV = C >> 3;
  const long long P = (long long)N * H * W;
  const long long items = POOL ? P / 4 : P;
  const int Wo = W >> 1, Ho = H >> 1;
  const PixWalk wk = pix_walk(CV);
  const int cv = wk.cv;
  const long long ps = wk.step;
  long long pix = wk.pix;
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = __ldg(ss + cv * 8 + j); sh[j] = __ldg(ss + C + cv * 8 + j); }
  if (!POOL) {
    const __nv_bfloat16* yp = y + cv * 8;
    __nv_bfloat16* op = out + cv * 8;
    for (; pix + 3 * ps < items; pix += 4 * ps) {           // four loads in flight before any math
      bf16x8 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ld8s(yp + (pix + u * ps) * ld_y);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
        st8s(op + (pix + u * ps) * ld_o, pack8(f));
      }
    }
    for (; pix < items; pix += ps) {
      float f[8];
      unpack8(ld8s(yp + pix * ld_y), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
      st8s(op + pix * ld_o, pack8(f));
    }
  } else {
    for (; pix < items; pix += ps) {
      const long long p00 = window_origin(pix, Wo, Ho, H, W);
      bf16x8 in[4];
      in[0] = ld8s(y + p00 * ld_y + cv * 8);
      in[1] = ld8s(y + (p00 + 1) * ld_y + cv * 8);
      in[2] = ld8s(y + (p00 + W) * ld_y + cv * 8);
      in[3] = ld8s(y + (p00 + W + 1) * ld_y + cv * 8);
      float m[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) m[j] = 0.f;          // relu output >= 0
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float f[8];
        unpack8(in[k], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
        bf16x8 o = pack8(f);
        const long long pp = p00 + (k & 1) + (k >> 1) * W;
        st8s(out + pp * ld_o + cv * 8, o);
        unpack8(o, f);                                    // pool the rounded values, as torch does
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], f[j]);
      }
      st8s(pool + pix * ld_p + cv * 8, pack8(m));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Backward pass 1: g = (d_out + route(d_pool)) * [y*scale+shift > 0]; per-channel partial sums of
// g and g*xhat.  One thread owns one 8-channel group and walks pixels (or 2x2 windows).
template <bool POOL>
__device__ __forceinline__ void load_g(const __nv_bfloat16* __restrict__ d_out, long long ld_do,
                                       const __nv_bfloat16* __restrict__ d_pool, long long ld_dp,
                                       const __nv_bfloat16* __restrict__ y, long long ld_y,
                                       const float (&sc)[8], const float (&sh)[8], long long pix,
                                       long long p00, int W, int cv, float (&yv)[4][8],
                                       float (&g)[4][8]) {
  constexpr int NP = POOL ? 4 : 1;
  float a[NP][8];
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const long long pp = POOL ? (p00 + (k & 1) + (k >> 1) * W) : pix;
    unpack8(ld8(y + pp * ld_y + cv * 8), yv[k]);
    if (d_out != nullptr) {
      unpack8(ld8(d_out + pp * ld_do + cv * 8), g[k]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) g[k][j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) a[k][j] = fmaf(yv[k][j], sc[j], sh[j]);
  }
  if (POOL && d_pool != nullptr) {
    float dp[8];
    unpack8(ld8(d_pool + pix * ld_dp + cv * 8), dp);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // the forward pooled bf16-rounded relu outputs; find the first maximum among them
      float best = -1.f;
      int arg = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float v = __bfloat162float(__float2bfloat16(fmaxf(a[k][j], 0.f)));
        if (v > best) { best = v; arg = k; }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) g[k][j] += (k == arg) ? dp[j] : 0.f;
    }
  }
#pragma unroll
  for (int k = 0; k < NP; ++k)
#pragma unroll
    for (int j = 0; j < 8; ++j) g[k][j] = (a[k][j] > 0.f) ? g[k][j] : 0.f;
}

// Register caps chosen by measurement (profiles/variants_in_situ.txt): the pooled variants keep a 2x2
// window of three tensors live and run best at 2 CTAs / SM; the plain reduce is fastest with the compiler's own choice
// (119 registers, 8 loads in flight per thread; any explicit cap measured 25-45 % slower), the plain apply at 3 CTAs / SM (78 registers, no spills).
template <bool POOL>
__device__ __forceinline__ void bn_bwd_reduce_body(const __nv_bfloat16* __restrict__ d_out, long long ld_do,
                     const __nv_bfloat16* __restrict__ d_pool, long long ld_dp,
                     const __nv_bfloat16* __restrict__ y, long long ld_y, const float* __restrict__ ss,
                     int N, int H, int W, int C, float* __restrict__ partial) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float sm[];
  const int CV = C >> 3;
  const int rows_pb = kThreads / CV;
  const int cv = threadIdx.x % CV, r = threadIdx.x / CV;
  const long long P = (long long)N * H * W;
  const long long items = POOL ? P / 4 : P;
  const int Wo = W >> 1, Ho = H >> 1;
  // partial sums per channel: s = sum g, q = sum g*y  (finalize turns q into sum g*xhat)
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  if (r < rows_pb) {
    float sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { sc[j] = ss[cv * 8 + j]; sh[j] = ss[C + cv * 8 + j]; }
    const long long stride = (long long)gridDim.x * rows_pb;
    long long pix = (long long)blockIdx.x * rows_pb + r;
    if (!POOL) {
      // 4 pixels per trip, all eight 16-byte loads issued before any math (memory-level parallelism)
      for (; pix + 3 * stride < items; pix += 4 * stride) {
        bf16x8 yy[4], dd[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          yy[u] = ld8(y + (pix + u * stride) * ld_y + cv * 8);
          if (d_out != nullptr) dd[u] = ld8(d_out + (pix + u * stride) * ld_do + cv * 8);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float yf[8], gf[8];
          unpack8(yy[u], yf);
          if (d_out != nullptr) unpack8(dd[u], gf);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float gm = (d_out != nullptr && fmaf(yf[j], sc[j], sh[j]) > 0.f) ? gf[j] : 0.f;
            s[j] += gm;
            q[j] = fmaf(gm, yf[j], q[j]);
          }
        }
      }
    }
    for (; pix < items; pix += stride) {
      const long long p00 = POOL ? window_origin(pix, Wo, Ho, H, W) : 0;
      float yv[4][8], g[4][8];
      load_g<POOL>(d_out, ld_do, d_pool, ld_dp, y, ld_y, sc, sh, pix, p00, W, cv, yv, g);
#pragma unroll
      for (int k = 0; k < (POOL ? 4 : 1); ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          s[j] += g[k][j];
          q[j] = fmaf(g[k][j], yv[k][j], q[j]);
        }
    }
  }
  float* s_sum = sm;
  float* s_sq = sm + rows_pb * C;
  if (r < rows_pb) {
#pragma unroll
    for (int j = 0; j < 8; ++j) { s_sum[r * C + cv * 8 + j] = s[j]; s_sq[r * C + cv * 8 + j] = q[j]; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += kThreads) {
    float a = 0.f, b = 0.f;
    for (int i = 0; i < rows_pb; ++i) { a += s_sum[i * C + c]; b += s_sq[i * C + c]; }
    partial[((size_t)blockIdx.x * 2 + 0) * C + c] = a;
    partial[((size_t)blockIdx.x * 2 + 1) * C + c] = b;
  }
}

template <bool POOL>
__global__ void bn_bwd_reduce_kernel(const __nv_bfloat16* d_out, long long ld_do, const __nv_bfloat16* d_pool,
                                     long long ld_dp, const __nv_bfloat16* y, long long ld_y, const float* ss, int N,
                                     int H, int W, int C, float* partial);
template <>
__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_kernel<false>(const __nv_bfloat16* d_out, long long ld_do, const __nv_bfloat16* d_pool, long long ld_dp,
                            const __nv_bfloat16* y, long long ld_y, const float* ss, int N, int H, int W, int C,
                            float* partial) {
  bn_bwd_reduce_body<false>(d_out, ld_do, d_pool, ld_dp, y, ld_y, ss, N, H, W, C, partial);
}
template <>
__global__ void __launch_bounds__(kThreads, 2)
bn_bwd_reduce_kernel<true>(const __nv_bfloat16* d_out, long long ld_do, const __nv_bfloat16* d_pool, long long ld_dp,
                           const __nv_bfloat16* y, long long ld_y, const float* ss, int N, int H, int W, int C,
                           float* partial) {
  bn_bwd_reduce_body<true>(d_out, ld_do, d_pool, ld_dp, y, ld_y, ss, N, H, W, C, partial);
}

// Backward pass 2 (tiny): finish the two sums, accumulate dgamma / dbeta, and fold everything the
// apply pass needs into 3 per-channel coefficients: dy = k1*g + e1*y + e0.
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int rows, long long count,
                                       const float* __restrict__ gamma, const float* __restrict__ ss,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       float* __restrict__ coef, int C) {
  pdl_wait();
  pdl_trigger();
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  double sg = 0.0, sgx = 0.0;
  for (int r = lane; r < rows; r += 32) {
    sg += (double)partial[((size_t)r * 2 + 0) * C + c];
    sgx += (double)partial[((size_t)r * 2 + 1) * C + c];
  }
  sg = warp_sum_d(sg);
  sgx = warp_sum_d(sgx);                               // so far: sum g*y
  if (lane != 0) return;
  const double mean = ss[2 * C + c], inv = ss[3 * C + c];
  sgx = inv * (sgx - mean * sg);                       // sum g*xhat
  dbeta[c] += (float)sg;
  dgamma[c] += (float)sgx;
  const double k1 = (double)gamma[c] * inv;
  const double c1 = sg / (double)count, c2 = sgx / (double)count;
  coef[0 * C + c] = (float)k1;
  coef[1 * C + c] = (float)(-k1 * c2 * inv);
  coef[2 * C + c] = (float)(k1 * (c2 * mean * inv - c1));
}

template <bool POOL>
__global__ void __launch_bounds__(kThreads, POOL ? 2 : 3)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ d_out, long long ld_do,
                    const __nv_bfloat16* __restrict__ d_pool, long long ld_dp,
                    const __nv_bfloat16* __restrict__ y, long long ld_y, const float* __restrict__ ss,
                    const float* __restrict__ coef, int N, int H, int W, int C,
                    __nv_bfloat16* __restrict__ dy, long long ld_dy) {
  pdl_wait();
  pdl_trigger();
  const int CV = C >> 3;
  const long long P = (long long)N * H * W;
  const long long items = POOL ? P / 4 : P;
  const int Wo = W >> 1, Ho = H >> 1;
  const PixWalk wk = pix_walk(CV);
  const int cv = wk.cv;
  const long long ps = wk.step;
  long long pix = wk.pix;
  float sc[8], sh[8], k1[8], e1[8], e0[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = __ldg(ss + cv * 8 + j); sh[j] = __ldg(ss + C + cv * 8 + j);
    k1[j] = __ldg(coef + cv * 8 + j); e1[j] = __ldg(coef + C + cv * 8 + j);
    e0[j] = __ldg(coef + 2 * C + cv * 8 + j);
  }
  if (!POOL) {
    const __nv_bfloat16* yp = y + cv * 8;
    const __nv_bfloat16* gp = d_out == nullptr ? nullptr : d_out + cv * 8;
    __nv_bfloat16* op = dy + cv * 8;
    for (; pix + ps < items; pix += 2 * ps) {              // two pixels per trip, loads first
      const long long px1 = pix + ps;
      const bf16x8 y0 = ld8(yp + pix * ld_y), y1 = ld8(yp + px1 * ld_y);
      bf16x8 d0 = y0, d1 = y1;
      if (gp != nullptr) { d0 = ld8(gp + pix * ld_do); d1 = ld8(gp + px1 * ld_do); }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        float yf[8], gf[8], o[8];
        unpack8(u ? y1 : y0, yf);
        unpack8(u ? d1 : d0, gf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gm = (gp != nullptr && fmaf(yf[j], sc[j], sh[j]) > 0.f) ? gf[j] : 0.f;
          o[j] = fmaf(k1[j], gm, fmaf(e1[j], yf[j], e0[j]));
        }
        st8(op + (u ? px1 : pix) * ld_dy, pack8(o));
      }
    }
  }
  for (; pix < items; pix += ps) {
    const long long p00 = POOL ? window_origin(pix, Wo, Ho, H, W) : 0;
    float yv[4][8], g[4][8];
    load_g<POOL>(d_out, ld_do, d_pool, ld_dp, y, ld_y, sc, sh, pix, p00, W, cv, yv, g);
#pragma unroll
    for (int k = 0; k < (POOL ? 4 : 1); ++k) {
      float o[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = fmaf(k1[j], g[k][j], fmaf(e1[j], yv[k][j], e0[j]));
      const long long pp = POOL ? (p00 + (k & 1) + (k >> 1) * W) : pix;
      st8(dy + pp * ld_dy + cv * 8, pack8(o));
    }
  }
}

}  // namespace b200

using namespace b200;

// Grid for the kernels that use pix_walk(): blocks * kThreads must be a multiple of the number of
// 8-channel groups so that a thread keeps its group from trip to trip.
static int grid_for_groups(long long pixels, int CV) {
  int a = CV, b = kThreads;
  while (b) { const int t = a % b; a = b; b = t; }          // a = gcd(CV, kThreads)
  const int m = CV / a;
  int g = grid_for(pixels * CV, kThreads);
  g = (g / m) * m;
  return g < m ? m : g;
}

extern "C" int b200seg_stats_rows() { return 4 * kNumSMs; }   // CTAs of the reduction kernels (4 per SM)

extern "C" int b200seg_channel_stats(const void* y, long long P, int C, long long ld, float* partial,
                                     int rows, cudaStream_t st) {
  if (C % 8 || C > 2048) return (int)cudaErrorInvalidValue;
  const int CV = C / 8, rows_pb = kThreads / CV;
  if (rows_pb < 1) return (int)cudaErrorInvalidValue;
  size_t smem = (size_t)2 * rows_pb * C * sizeof(float);
  launch_k(channel_stats_kernel, rows, kThreads, smem, st, (const __nv_bfloat16*)y, P, C, ld, partial);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_bn_finalize(const float* partial, int rows, long long count, const float* gamma,
                                   const float* beta, const float* conv_bias, float* running_mean,
                                   float* running_var, long long* nbt, float* ss, int C, float momentum,
                                   float eps, int training, cudaStream_t st) {
  launch_k(bn_finalize_kernel, (C + 3) / 4, 128, 0, st, partial, rows, count, gamma, beta, conv_bias, running_mean,
           running_var, nbt, ss, C, momentum, eps, training);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_bn_relu_pool_fwd(const void* y, long long ld_y, const float* ss, void* out,
                                        long long ld_o, void* pool, long long ld_p, int N, int H, int W,
                                        int C, cudaStream_t st) {
  if (C % 8) return (int)cudaErrorInvalidValue;
  const long long P = (long long)N * H * W;
  if (pool != nullptr) {
    if ((H | W) & 1) return (int)cudaErrorInvalidValue;
    if (P / 4 >= (1ll << 32)) return (int)cudaErrorInvalidValue;
    int grid = grid_for_groups(P / 4, C / 8);
    launch_k(bn_relu_pool_fwd_kernel<true>, grid, kThreads, 0, st,
             (const __nv_bfloat16*)y, ld_y, ss, (__nv_bfloat16*)out, ld_o, (__nv_bfloat16*)pool, ld_p, N, H, W, C);
  } else {
    int grid = grid_for_groups(P, C / 8);
    launch_k(bn_relu_pool_fwd_kernel<false>, grid, kThreads, 0, st,
             (const __nv_bfloat16*)y, ld_y, ss, (__nv_bfloat16*)out, ld_o, nullptr, 0, N, H, W, C);
  }
  return (int)cudaGetLastError();
}

// workspace: float[rows*2*C + 3*C]; rows = b200seg_stats_rows().
extern "C" int b200seg_bn_relu_pool_bwd(const void* d_out, long long ld_do, const void* d_pool,
                                        long long ld_dp, const void* y, long long ld_y, const float* ss,
                                        const float* gamma, float* dgamma, float* dbeta, void* dy,
                                        long long ld_dy, int N, int H, int W, int C, int pooled,
                                        float* workspace, cudaStream_t st) {
  if (C % 8 || C > 2048) return (int)cudaErrorInvalidValue;
  const int CV = C / 8, rows_pb = kThreads / CV;
  if (rows_pb < 1) return (int)cudaErrorInvalidValue;
  const int rows = b200seg_stats_rows();
  const long long P = (long long)N * H * W;
  if (pooled && P / 4 >= (1ll << 32)) return (int)cudaErrorInvalidValue;
  float* partial = workspace;
  float* coef = workspace + (size_t)rows * 2 * C;
  size_t smem = (size_t)2 * rows_pb * C * sizeof(float);
  const __nv_bfloat16 *a = (const __nv_bfloat16*)d_out, *b = (const __nv_bfloat16*)d_pool,
                      *yy = (const __nv_bfloat16*)y;
  if (pooled) {
    launch_k(bn_bwd_reduce_kernel<true>, rows, kThreads, smem, st, a, ld_do, b, ld_dp, yy, ld_y, ss, N, H, W, C, partial);
  } else {
    launch_k(bn_bwd_reduce_kernel<false>, rows, kThreads, smem, st, a, ld_do, nullptr, 0, yy, ld_y, ss, N, H, W, C,
             partial);
  }
  launch_k(bn_bwd_finalize_kernel, (C + 3) / 4, 128, 0, st, partial, rows, P, gamma, ss, dgamma, dbeta, coef, C);
  if (pooled) {
    int grid = grid_for_groups(P / 4, CV);
    launch_k(bn_bwd_apply_kernel<true>, grid, kThreads, 0, st, a, ld_do, b, ld_dp, yy, ld_y, ss, coef, N, H, W, C,
             (__nv_bfloat16*)dy, ld_dy);
  } else {
    int grid = grid_for_groups(P, CV);
    launch_k(bn_bwd_apply_kernel<false>, grid, kThreads, 0, st, a, ld_do, nullptr, 0, yy, ld_y, ss, coef, N, H, W, C,
             (__nv_bfloat16*)dy, ld_dy);
  }
  return (int)cudaGetLastError();
}

// The same backward when the two per-channel sums were already produced by the kernel that wrote d_out
// (csrc/elem/head_tail.cu for the last stage): finalize + apply only.  partial: float[rows][2][C];
// coef: float[3*C] scratch.
extern "C" int b200seg_bn_relu_bwd_from_partial(const float* partial, int rows, const void* d_out, long long ld_do,
                                                const void* y, long long ld_y, const float* ss, const float* gamma,
                                                float* dgamma, float* dbeta, void* dy, long long ld_dy, int N, int H,
                                                int W, int C, float* coef, cudaStream_t st) {
  if (C % 8 || C > 2048 || rows < 1 || kThreads / (C / 8) < 1) return (int)cudaErrorInvalidValue;
  const long long P = (long long)N * H * W;
  launch_k(bn_bwd_finalize_kernel, (C + 3) / 4, 128, 0, st, partial, rows, P, gamma, ss, dgamma, dbeta, coef, C);
  const int grid = grid_for_groups(P, C / 8);
  launch_k(bn_bwd_apply_kernel<false>, grid, kThreads, 0, st, (const __nv_bfloat16*)d_out, ld_do,
           (const __nv_bfloat16*)nullptr, 0ll, (const __nv_bfloat16*)y, ld_y, ss, coef, N, H, W, C, (__nv_bfloat16*)dy,
           ld_dy);
  return (int)cudaGetLastError();
}
