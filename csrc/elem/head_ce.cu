// 1x1 classification head + softmax cross-entropy + arg-max / accuracy + the head's whole backward,
// in one pass over the last activation (bf16 NHWC).
//
// Replaces the reference's conv_last (R:641), nn.CrossEntropyLoss forward and backward (R:703,
// R:755-756), the arg-max / pixel-accuracy line that forces a device sync every micro-batch
// (R:775-776) and the 1x1 conv's dgrad / wgrad.  The fp32 logits never
// touch HBM.  Per pixel p with activation a_p (C channels) and label y_p:
//     z = W a_p + b;  loss += logsumexp(z) - z[y_p];  pred_p = argmax z;
//     dl = (softmax(z) - onehot(y_p)) * grad_scale;   da_p = W^T dl;   dW += dl a_p^T;  db += dl
// Reductions (dW, db, loss, correct) are deterministic: per-CTA partials in a workspace, summed
// in CTA order by the last CTA to finish.
#define B200SEG_PDL_CLASS 2
#include "../common.cuh"
#include "b200seg.h"
#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace b200 {

constexpr int kHeadThreads = 256;
constexpr int kHeadWarps = kHeadThreads / 32;

// Per warp: a tile of 32 pixels staged through shared memory so that every global access is a
// coalesced 16-byte vector (lane = chunk), the per-pixel math runs lane = pixel, and the dW
// accumulation runs lane = channel pair.
template <int KMAX>
__global__ void __launch_bounds__(kHeadThreads)
head_ce_kernel(const __nv_bfloat16* __restrict__ a, long long ld_a, const float* __restrict__ w,
               const float* __restrict__ bias, const uint8_t* __restrict__ labels, long long P, int C,
               int K, float grad_scale, __nv_bfloat16* __restrict__ da, long long ld_da,
               float* __restrict__ dw, float* __restrict__ db, uint8_t* __restrict__ pred,
               float* __restrict__ loss_acc, float* __restrict__ workspace,
               unsigned int* __restrict__ counter, int warps) {
  extern __shared__ __align__(16) uint8_t smraw[];
  const int pitch = C + 8;                                    // bf16 elements per tile row
  const int nred = K * C + K + 2;
  float* s_w = reinterpret_cast<float*>(smraw);              // [K][C]
  float* s_b = s_w + K * C;                                   // [KMAX]
  float* s_red = s_b + KMAX;                                  // [nred]
  float* s_dl = s_red + ((nred + 3) & ~3);                    // [warps][32][KMAX]
  __nv_bfloat16* s_tiles = reinterpret_cast<__nv_bfloat16*>(s_dl + warps * 32 * KMAX);   // [warps][2][32][pitch]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < K * C; i += blockDim.x) s_w[i] = w[i];
  for (int i = threadIdx.x; i < K; i += blockDim.x) s_b[i] = bias[i];
  for (int i = threadIdx.x; i < nred; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();

  const bool bwd = (da != nullptr);
  const int CV = C >> 3;
  const int NP = (C + 63) >> 6;                               // bf16 pairs per lane in the dW phase (<= 4)
  __nv_bfloat16* t_a = s_tiles + (size_t)warp * 2 * 32 * pitch;
  __nv_bfloat16* t_da = t_a + 32 * pitch;
  float* my_dl = s_dl + warp * 32 * KMAX;
  float accw[KMAX][8];
  float accb[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    accb[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) accw[k][j] = 0.f;
  }
  float loss_local = 0.f, correct_local = 0.f;

  const long long chunks = (P + 31) / 32;
  for (long long ch = (long long)blockIdx.x * warps + warp; ch < chunks; ch += (long long)gridDim.x * warps) {
    const long long p0 = ch * 32;
    // 1. coalesced tile load
    for (int i = lane; i < 32 * CV; i += 32) {
      const int px = i / CV, cv = i - px * CV;
      bf16x8 v;
      if (p0 + px < P) v = ld8(a + (p0 + px) * ld_a + cv * 8);
      else { v.v[0] = v.v[1] = v.v[2] = v.v[3] = __floats2bfloat162_rn(0.f, 0.f); }
      *reinterpret_cast<bf16x8*>(t_a + px * pitch + cv * 8) = v;
    }
    __syncwarp();
    // 2. logits, softmax, loss (lane = pixel)
    const long long p = p0 + lane;
    const bool valid = p < P;
    float z[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) z[k] = (k < K) ? s_b[k] : -INFINITY;
    for (int cv = 0; cv < CV; ++cv) {
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(t_a + lane * pitch + cv * 8), f);
#pragma unroll
      for (int k = 0; k < KMAX; ++k)
        if (k < K) {
          const float4 w0 = *reinterpret_cast<const float4*>(s_w + k * C + cv * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(s_w + k * C + cv * 8 + 4);
          z[k] = fmaf(f[0], w0.x, fmaf(f[1], w0.y, fmaf(f[2], w0.z, fmaf(f[3], w0.w, z[k]))));
          z[k] = fmaf(f[4], w1.x, fmaf(f[5], w1.y, fmaf(f[6], w1.z, fmaf(f[7], w1.w, z[k]))));
        }
    }
    float mx = z[0];
    int arg = 0;
#pragma unroll
    for (int k = 1; k < KMAX; ++k)
      if (k < K && z[k] > mx) { mx = z[k]; arg = k; }
    float e[KMAX], sum = 0.f;
#pragma unroll
    for (int k = 0; k < KMAX; ++k) { e[k] = (k < K) ? __expf(z[k] - mx) : 0.f; sum += e[k]; }
    const int lab = valid ? (int)labels[p] : 0;
    if (valid) {
      float zl = 0.f;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) zl = (k == lab) ? z[k] : zl;
      loss_local += __logf(sum) + mx - zl;
      correct_local += (arg == lab) ? 1.f : 0.f;
      if (pred != nullptr) pred[p] = (uint8_t)arg;
    }
    if (bwd) {
      const float inv = grad_scale / sum;
      float dl[KMAX];
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        dl[k] = valid ? (e[k] * inv - ((k == lab) ? grad_scale : 0.f)) : 0.f;
        my_dl[lane * KMAX + k] = dl[k];
        accb[k] += dl[k];
      }
      // 3. da row -> smem tile
      for (int cv = 0; cv < CV; ++cv) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < K) {
            const float4 w0 = *reinterpret_cast<const float4*>(s_w + k * C + cv * 8);
            const float4 w1 = *reinterpret_cast<const float4*>(s_w + k * C + cv * 8 + 4);
            o[0] = fmaf(dl[k], w0.x, o[0]); o[1] = fmaf(dl[k], w0.y, o[1]);
            o[2] = fmaf(dl[k], w0.z, o[2]); o[3] = fmaf(dl[k], w0.w, o[3]);
            o[4] = fmaf(dl[k], w1.x, o[4]); o[5] = fmaf(dl[k], w1.y, o[5]);
            o[6] = fmaf(dl[k], w1.z, o[6]); o[7] = fmaf(dl[k], w1.w, o[7]);
          }
        *reinterpret_cast<bf16x8*>(t_da + lane * pitch + cv * 8) = pack8(o);
      }
      __syncwarp();
      // 4. coalesced tile store
      for (int i = lane; i < 32 * CV; i += 32) {
        const int px = i / CV, cv = i - px * CV;
        if (p0 + px < P) st8(da + (p0 + px) * ld_da + cv * 8, *reinterpret_cast<const bf16x8*>(t_da + px * pitch + cv * 8));
      }
      // 5. dW: lane owns channel pairs (2*lane + 64*j); walks the 32 pixels of the tile
#pragma unroll 4
      for (int i = 0; i < 32; ++i) {
        float av[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int c = j * 64 + lane * 2;
          float2 t = make_float2(0.f, 0.f);
          if (j < NP && c < C) t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(t_a + i * pitch + c));
          av[2 * j] = t.x;
          av[2 * j + 1] = t.y;
        }
        float dv[KMAX];
#pragma unroll
        for (int k = 0; k < KMAX; k += 4) {
          const float4 t4 = *reinterpret_cast<const float4*>(my_dl + i * KMAX + k);
          dv[k] = t4.x; dv[k + 1] = t4.y; dv[k + 2] = t4.z; dv[k + 3] = t4.w;
        }
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
#pragma unroll
          for (int j = 0; j < 8; ++j) accw[k][j] = fmaf(dv[k], av[j], accw[k][j]);
        }
      }
    }
    __syncwarp();
  }
  // ---- block reduction -> per-CTA partial ----------------------------------------------------
  loss_local = warp_sum(loss_local);
  correct_local = warp_sum(correct_local);
#pragma unroll
  for (int k = 0; k < KMAX; ++k) accb[k] = warp_sum(accb[k]);
  for (int wv = 0; wv < warps; ++wv) {               // fixed warp order: deterministic sums
    if (warp == wv) {
      if (bwd) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < K) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int c = j * 64 + lane * 2;
              if (j < NP && c < C) { s_red[k * C + c] += accw[k][2 * j]; s_red[k * C + c + 1] += accw[k][2 * j + 1]; }
            }
            if (lane == 0) s_red[K * C + k] += accb[k];
          }
      }
      if (lane == 0) {
        s_red[K * C + K] += loss_local;
        s_red[K * C + K + 1] += correct_local;
      }
    }
    __syncthreads();
  }
  float* mine = workspace + (size_t)blockIdx.x * nred;
  for (int i = threadIdx.x; i < nred; i += blockDim.x) mine[i] = s_red[i];
  __threadfence();
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < nred; i += blockDim.x) {
    float t = 0.f;
    for (unsigned int b = 0; b < gridDim.x; ++b) t += workspace[(size_t)b * nred + i];
    if (i < K * C) { if (bwd) dw[i] += t; }
    else if (i < K * C + K) { if (bwd) db[i - K * C] += t; }
    else loss_acc[i - K * C - K] = t;
  }
  if (threadIdx.x == 0) *counter = 0u;               // ready for the next launch
}


// -------------------------------------------------------------------------------------------------
// Fast path for the shapes the U-Net actually has (C = 32 / 64 / 128 channels, K <= 8 classes).
// Same arithmetic as head_ce_kernel (bit-identical logits, predictions and da), reorganised around
// what bounds the generic kernel: shared-memory read bandwidth.  A warp-wide 16-byte load costs
// four cycles of the SM's 128 B/clk path even when every lane reads the same address, and the
// generic kernel issues two such broadcast loads of weights per eight FMAs.  Here
//   * a warp owns a tile of 128 pixels and every lane four of them (64 and two for C = 128), so each
//     weight vector fetched from shared memory feeds 32 FMAs instead of 8;
//   * the dW phase maps a lane to (8 channels, a subset of the pixels): it reads distinct 16-byte
//     pieces of the activation tile (no broadcast) and does K*8 FMAs per load; the pixel subsets are
//     folded with shuffles once, at the end of the kernel;
//   * loops are unrolled over compile-time K and C, and the da tile overwrites the activation tile.
constexpr int kHfWarps = 8;
template <int CT> struct HfCfg {
  static constexpr int kPix = CT > 64 ? 2 : 4;                  // pixels per lane (shared memory bounds it)
  static constexpr int kTile = 32 * kPix;                       // pixels per warp tile
};

template <int KT, int CT>
__global__ void __launch_bounds__(kHfWarps * 32, 1)
head_ce_fast_kernel(const __nv_bfloat16* __restrict__ a, long long ld_a, const float* __restrict__ w,
                    const float* __restrict__ bias, const uint8_t* __restrict__ labels, long long P, int K,
                    float grad_scale, __nv_bfloat16* __restrict__ da, long long ld_da, float* __restrict__ dw,
                    float* __restrict__ db, uint8_t* __restrict__ pred, float* __restrict__ loss_acc,
                    float* __restrict__ workspace, unsigned int* __restrict__ counter) {
  extern __shared__ __align__(16) uint8_t smraw[];
  pdl_wait();
  pdl_trigger();
  constexpr int kHfPix = HfCfg<CT>::kPix, kHfTile = HfCfg<CT>::kTile;
  constexpr int pitch = CT + 8;                               // bf16 elements per tile row
  constexpr int CV = CT / 8;                                  // 16-byte pieces per pixel
  constexpr int NSUB = 32 / CV;                               // pixel subsets in the dW phase
  const int nred = K * CT + K + 2;
  float* s_w = reinterpret_cast<float*>(smraw);               // [KT][CT], rows >= K are zero
  float* s_b = s_w + KT * CT;                                 // [8]
  float* s_red = s_b + 8;                                     // [nred]
  float* s_dl = s_red + ((nred + 3) & ~3);                    // [warps][tile][8]
  __nv_bfloat16* s_tiles = reinterpret_cast<__nv_bfloat16*>(s_dl + kHfWarps * kHfTile * 8);   // [warps][tile][pitch]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < KT * CT; i += blockDim.x) s_w[i] = i < K * CT ? w[i] : 0.f;
  for (int i = threadIdx.x; i < 8; i += blockDim.x) s_b[i] = i < K ? bias[i] : 0.f;
  for (int i = threadIdx.x; i < nred; i += blockDim.x) s_red[i] = 0.f;
  __syncthreads();

  const bool bwd = (da != nullptr);
  __nv_bfloat16* t_a = s_tiles + (size_t)warp * kHfTile * pitch;
  float* my_dl = s_dl + warp * kHfTile * 8;
  const int cvl = lane % CV, sub = lane / CV;                 // dW phase: channel piece, pixel subset
  float accw[KT][8];
  float accb[KT];
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    accb[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) accw[k][j] = 0.f;
  }
  float loss_local = 0.f, correct_local = 0.f;

  const long long chunks = (P + kHfTile - 1) / kHfTile;
  for (long long ch = (long long)blockIdx.x * kHfWarps + warp; ch < chunks; ch += (long long)gridDim.x * kHfWarps) {
    const long long p0 = ch * kHfTile;
    // 1. coalesced tile load
#pragma unroll 8
    for (int i0 = 0; i0 < kHfTile * CV; i0 += 32) {
      const int i = i0 + lane;
      const int px = i / CV, cv = i - px * CV;
      bf16x8 v;
      if (p0 + px < P) v = ld8(a + (p0 + px) * ld_a + cv * 8);
      else { v.v[0] = v.v[1] = v.v[2] = v.v[3] = __floats2bfloat162_rn(0.f, 0.f); }
      *reinterpret_cast<bf16x8*>(t_a + px * pitch + cv * 8) = v;
    }
    __syncwarp();
    // 2. logits (lane = four pixels: lane, lane + 32, ...)
    float z[kHfPix][KT];
#pragma unroll
    for (int q = 0; q < kHfPix; ++q)
#pragma unroll
      for (int k = 0; k < KT; ++k) z[q][k] = s_b[k];
#pragma unroll
    for (int cv = 0; cv < CV; ++cv) {
      float f[kHfPix][8];
#pragma unroll
      for (int q = 0; q < kHfPix; ++q)
        unpack8(*reinterpret_cast<const bf16x8*>(t_a + (q * 32 + lane) * pitch + cv * 8), f[q]);
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        const float4 w0 = *reinterpret_cast<const float4*>(s_w + k * CT + cv * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(s_w + k * CT + cv * 8 + 4);
#pragma unroll
        for (int q = 0; q < kHfPix; ++q) {
          float t = z[q][k];
          t = fmaf(f[q][0], w0.x, fmaf(f[q][1], w0.y, fmaf(f[q][2], w0.z, fmaf(f[q][3], w0.w, t))));
          t = fmaf(f[q][4], w1.x, fmaf(f[q][5], w1.y, fmaf(f[q][6], w1.z, fmaf(f[q][7], w1.w, t))));
          z[q][k] = t;
        }
      }
    }
    // 3. softmax, loss, arg-max; dl -> registers and the staging tile
    float dl[kHfPix][KT];
#pragma unroll
    for (int q = 0; q < kHfPix; ++q) {
      const long long p = p0 + q * 32 + lane;
      const bool valid = p < P;
#pragma unroll
      for (int k = 0; k < KT; ++k) z[q][k] = (k < K) ? z[q][k] : -INFINITY;
      float mx = z[q][0];
      int arg = 0;
#pragma unroll
      for (int k = 1; k < KT; ++k)
        if (z[q][k] > mx) { mx = z[q][k]; arg = k; }
      float e[KT], sum = 0.f;
#pragma unroll
      for (int k = 0; k < KT; ++k) { e[k] = (k < K) ? __expf(z[q][k] - mx) : 0.f; sum += e[k]; }
      const int lab = valid ? (int)labels[p] : 0;
      if (valid) {
        float zl = 0.f;
#pragma unroll
        for (int k = 0; k < KT; ++k) zl = (k == lab) ? z[q][k] : zl;
        loss_local += __logf(sum) + mx - zl;
        correct_local += (arg == lab) ? 1.f : 0.f;
        if (pred != nullptr) pred[p] = (uint8_t)arg;
      }
      const float inv = grad_scale / sum;
      float d8[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) d8[k] = 0.f;
#pragma unroll
      for (int k = 0; k < KT; ++k) {
        dl[q][k] = (valid && k < K) ? (e[k] * inv - ((k == lab) ? grad_scale : 0.f)) : 0.f;
        d8[k] = dl[q][k];
        accb[k] += dl[q][k];
      }
      if (bwd) {
        float4* dst = reinterpret_cast<float4*>(my_dl + (q * 32 + lane) * 8);
        dst[0] = make_float4(d8[0], d8[1], d8[2], d8[3]);
        dst[1] = make_float4(d8[4], d8[5], d8[6], d8[7]);
      }
    }
    if (bwd) {
      __syncwarp();
      // 4. dW: lane = (channel piece cvl, pixel subset sub); distinct 16-byte reads of the tile
#pragma unroll 4
      for (int i = 0; i < kHfTile / NSUB; ++i) {
        const int px = i * NSUB + sub;
        float av[8];
        unpack8(*reinterpret_cast<const bf16x8*>(t_a + px * pitch + cvl * 8), av);
        float dv[8];
        *reinterpret_cast<float4*>(dv) = *reinterpret_cast<const float4*>(my_dl + px * 8);
        if (KT > 4) *reinterpret_cast<float4*>(dv + 4) = *reinterpret_cast<const float4*>(my_dl + px * 8 + 4);
#pragma unroll
        for (int k = 0; k < KT; ++k) {
#pragma unroll
          for (int j = 0; j < 8; ++j) accw[k][j] = fmaf(dv[k], av[j], accw[k][j]);
        }
      }
      __syncwarp();
      // 5. da rows overwrite the activation tile (lane = pixel again)
#pragma unroll
      for (int cv = 0; cv < CV; ++cv) {
        float o[kHfPix][8];
#pragma unroll
        for (int q = 0; q < kHfPix; ++q)
#pragma unroll
          for (int j = 0; j < 8; ++j) o[q][j] = 0.f;
#pragma unroll
        for (int k = 0; k < KT; ++k) {
          const float4 w0 = *reinterpret_cast<const float4*>(s_w + k * CT + cv * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(s_w + k * CT + cv * 8 + 4);
#pragma unroll
          for (int q = 0; q < kHfPix; ++q) {
            const float d = dl[q][k];
            o[q][0] = fmaf(d, w0.x, o[q][0]); o[q][1] = fmaf(d, w0.y, o[q][1]);
            o[q][2] = fmaf(d, w0.z, o[q][2]); o[q][3] = fmaf(d, w0.w, o[q][3]);
            o[q][4] = fmaf(d, w1.x, o[q][4]); o[q][5] = fmaf(d, w1.y, o[q][5]);
            o[q][6] = fmaf(d, w1.z, o[q][6]); o[q][7] = fmaf(d, w1.w, o[q][7]);
          }
        }
#pragma unroll
        for (int q = 0; q < kHfPix; ++q)
          *reinterpret_cast<bf16x8*>(t_a + (q * 32 + lane) * pitch + cv * 8) = pack8(o[q]);
      }
      __syncwarp();
      // 6. coalesced tile store
#pragma unroll 8
      for (int i0 = 0; i0 < kHfTile * CV; i0 += 32) {
        const int i = i0 + lane;
        const int px = i / CV, cv = i - px * CV;
        if (p0 + px < P)
          st8(da + (p0 + px) * ld_da + cv * 8, *reinterpret_cast<const bf16x8*>(t_a + px * pitch + cv * 8));
      }
    }
    __syncwarp();
  }
  // ---- block reduction -> per-CTA partial -----------------------------------------------------
  loss_local = warp_sum(loss_local);
  correct_local = warp_sum(correct_local);
#pragma unroll
  for (int k = 0; k < KT; ++k) {
    accb[k] = warp_sum(accb[k]);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
      for (int o = CV; o < 32; o <<= 1) accw[k][j] += __shfl_xor_sync(0xffffffffu, accw[k][j], o);
    }
  }
  for (int wv = 0; wv < kHfWarps; ++wv) {               // fixed warp order: deterministic sums
    if (warp == wv) {
      if (bwd) {
#pragma unroll
        for (int k = 0; k < KT; ++k)
          if (k < K) {
            if (lane < CV) {
#pragma unroll
              for (int j = 0; j < 8; ++j) s_red[k * CT + lane * 8 + j] += accw[k][j];
            }
            if (lane == 0) s_red[K * CT + k] += accb[k];
          }
      }
      if (lane == 0) {
        s_red[K * CT + K] += loss_local;
        s_red[K * CT + K + 1] += correct_local;
      }
    }
    __syncthreads();
  }
  float* mine = workspace + (size_t)blockIdx.x * nred;
  for (int i = threadIdx.x; i < nred; i += blockDim.x) mine[i] = s_red[i];
  __threadfence();
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < nred; i += blockDim.x) {
    float t = 0.f;
    for (unsigned int b = 0; b < gridDim.x; ++b) t += workspace[(size_t)b * nred + i];
    if (i < K * CT) { if (bwd) dw[i] += t; }
    else if (i < K * CT + K) { if (bwd) db[i - K * CT] += t; }
    else loss_acc[i - K * CT - K] = t;
  }
  if (threadIdx.x == 0) *counter = 0u;
}

template <int KT, int CT>
static int launch_head_fast(const void* a, long long ld_a, const float* w, const float* bias,
                            const unsigned char* labels, long long P, int K, float grad_scale, void* da,
                            long long ld_da, float* dw, float* db, unsigned char* pred, float* loss_acc, float* ws,
                            unsigned int* counter, cudaStream_t st) {
  constexpr int kHfTile = HfCfg<CT>::kTile;
  const long long chunks = (P + kHfTile - 1) / kHfTile;
  int grid = (int)((chunks + kHfWarps - 1) / kHfWarps);
  if (grid > kNumSMs) grid = kNumSMs;
  if (grid < 1) grid = 1;
  const int nred_max = 8 * CT + 8 + 2;
  const size_t smem = sizeof(float) * ((size_t)KT * CT + 8 + ((nred_max + 3) & ~3) + (size_t)kHfWarps * kHfTile * 8) +
                      (size_t)kHfWarps * kHfTile * (CT + 8) * 2;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(head_ce_fast_kernel<KT, CT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  return (int)launch_k(head_ce_fast_kernel<KT, CT>, grid, kHfWarps * 32, smem, st, (const __nv_bfloat16*)a, ld_a, w,
                       bias, (const uint8_t*)labels, P, K, grad_scale, (__nv_bfloat16*)da, ld_da, dw, db,
                       (uint8_t*)pred, loss_acc, ws, counter);
}

static int head_fast_env() {           // B200SEG_HEAD_FAST=0 keeps the generic kernel (A/B runs)
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_HEAD_FAST"); v = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return v;
}

}  // namespace b200

using namespace b200;

// The general kernels' dynamic shared memory limit is raised ONCE per device to the most the device allows
// (not per call to the size of that call: two host threads could interleave so that one lowers the limit
// just before the other launches with a larger request).
#include <mutex>
template <int KMAX>
static cudaError_t head_ce_raise_smem_limit() {
  static std::mutex mu;
  static bool done[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  if (dev >= 0 && dev < 64 && done[dev]) return cudaSuccess;
  int optin = 0;
  e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  cudaFuncAttributes fa{};
  e = cudaFuncGetAttributes(&fa, b200::head_ce_kernel<KMAX>);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(b200::head_ce_kernel<KMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           optin - (int)fa.sharedSizeBytes);   // the opt-in limit covers static + dynamic
  if (e == cudaSuccess && dev >= 0 && dev < 64) done[dev] = true;
  return e;
}

extern "C" long long b200seg_head_ce_workspace_bytes(int C, int K) {
  return (long long)(2 * kNumSMs) * (K * C + K + 2) * (long long)sizeof(float) + 256;
}

// workspace: b200seg_head_ce_workspace_bytes(C, K) bytes; its LAST 256 bytes hold the CTA counter
// and must be zero before the first launch (the kernel leaves them zero afterwards).
extern "C" int b200seg_head_ce(const void* a, long long ld_a, const float* w, const float* bias,
                               const unsigned char* labels, long long P, int C, int K, float grad_scale,
                               void* da, long long ld_da, float* dw, float* db, unsigned char* pred,
                               float* loss_acc, void* workspace, cudaStream_t st) {
  if (C % 8 || C > 256 || K < 1 || K > 16) return (int)cudaErrorInvalidValue;
  const int KMAX = K <= 8 ? 8 : 16;
  const int warps = C <= 128 ? 8 : 4;
  const int threads = warps * 32;
  long long chunks = (P + 31) / 32;
  int grid = (int)((chunks + warps - 1) / warps);
  if (grid > 2 * kNumSMs) grid = 2 * kNumSMs;
  if (grid < 1) grid = 1;
  const int nred = K * C + K + 2;
  size_t smem = sizeof(float) * ((size_t)K * C + KMAX + ((nred + 3) & ~3) + (size_t)warps * 32 * KMAX) +
                (size_t)warps * 2 * 32 * (C + 8) * 2;
  float* ws = (float*)workspace;
  unsigned int* counter = (unsigned int*)((char*)workspace + b200seg_head_ce_workspace_bytes(C, K) - 256);
  if (K <= 8 && (C == 32 || C == 64 || C == 128) && head_fast_env()) {
#define B200_HEAD_FAST(KT, CT)                                                                                  \
  return launch_head_fast<KT, CT>(a, ld_a, w, bias, labels, P, K, grad_scale, da, ld_da, dw, db, pred, loss_acc, \
                                  ws, counter, st)
    if (K <= 6) {
      if (C == 32) B200_HEAD_FAST(6, 32);
      if (C == 64) B200_HEAD_FAST(6, 64);
      B200_HEAD_FAST(6, 128);
    }
    if (C == 32) B200_HEAD_FAST(8, 32);
    if (C == 64) B200_HEAD_FAST(8, 64);
    B200_HEAD_FAST(8, 128);
#undef B200_HEAD_FAST
  }
  cudaError_t e;
  if (KMAX == 8) {
    e = head_ce_raise_smem_limit<8>();
    if (e != cudaSuccess) return (int)e;
    head_ce_kernel<8><<<grid, threads, smem, st>>>((const __nv_bfloat16*)a, ld_a, w, bias, labels, P, C, K,
                                                   grad_scale, (__nv_bfloat16*)da, ld_da, dw, db, pred, loss_acc,
                                                   ws, counter, warps);
  } else {
    e = head_ce_raise_smem_limit<16>();
    if (e != cudaSuccess) return (int)e;
    head_ce_kernel<16><<<grid, threads, smem, st>>>((const __nv_bfloat16*)a, ld_a, w, bias, labels, P, C, K,
                                                    grad_scale, (__nv_bfloat16*)da, ld_da, dw, db, pred, loss_acc,
                                                    ws, counter, warps);
  }
  return (int)cudaGetLastError();
}

// =================================================================================================
// Allocation-free variant of the head kernel.
//
// Who this is for: callers that are not allowed to (or cannot afford to) obtain device memory at
// call time.  Typical cases: a host application that hands the whole GPU to its own memory pool at
// start-up and forbids later driver allocations by the code it loads; a latency-critical loop,
// where cudaMalloc / cudaFree would serialise the entire device and stall every other stream; a
// plain C program that has no framework allocator to borrow a workspace from.  b200seg_head_ce()
// needs a caller-provided workspace for its per-CTA partial sums; this entry point needs none,
// because the partials live in a fixed array that is part of the library image itself.
//
// The price: that array (kHeadStaticBytes, about 0.6 MB) is resident on every GPU that loads the
// library, for the life of the process, whether or not this function is ever called; and the
// problem size is capped -- C <= 128 channels and K <= 8 classes (any number of pixels).  Larger
// sizes are refused with cudaErrorInvalidValue and a message on stderr; nothing is launched.
//
// All calls on one device share the one array and it keeps whatever the previous call left in it.
// The only part that has to start at zero, the CTA counter, is cleared on the stream at the start
// of every call.  Two calls that can run on the GPU at the same time (different streams, or
// different host threads on different streams) would trample each other's partials: keep calls on
// one device ordered.  The function never calls cudaMalloc, cudaFree, or any library routine that
// allocates.
//
// First call on each device: the device address of the array is looked up once (with that device
// current) and cached.  Make that first call during set-up, before any stream capture is begun.
namespace b200 {

constexpr int kHeadStaticMaxC = 128;
constexpr int kHeadStaticMaxK = 8;
constexpr int kHeadStaticBlocks = kNumSMs;                      // one CTA per SM
constexpr size_t kHeadStaticPartials =
    (size_t)kHeadStaticBlocks * (kHeadStaticMaxK * kHeadStaticMaxC + kHeadStaticMaxK + 2) * sizeof(float);
constexpr size_t kHeadStaticBytes = kHeadStaticPartials + 256;  // + the CTA counter

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device pointer.
static __device__ __align__(256) unsigned char g_head_static_scratch[kHeadStaticBytes];

constexpr int kMaxDevices = 64;
static void* g_head_static_ptr[kMaxDevices] = {nullptr};
static std::mutex g_head_static_mutex;

// Device address of the scratch array on the CURRENT device (resolved once per device).
static cudaError_t head_static_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_head_static_mutex);
  if (g_head_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_head_static_scratch);
    if (e != cudaSuccess) return e;
    g_head_static_ptr[dev] = ptr;
  }
  *out = g_head_static_ptr[dev];
  return cudaSuccess;
}

}  // namespace b200

extern "C" long long b200seg_head_ce_static_scratch_bytes() { return (long long)kHeadStaticBytes; }

extern "C" int b200seg_head_ce_static(const void* a, long long ld_a, const float* w, const float* bias,
                                      const unsigned char* labels, long long P, int C, int K,
                                      float grad_scale, void* da, long long ld_da, float* dw, float* db,
                                      unsigned char* pred, float* loss_acc, cudaStream_t st) {
  if (C % 8 || C < 8 || K < 1 || P < 1) return (int)cudaErrorInvalidValue;
  if (C > kHeadStaticMaxC || K > kHeadStaticMaxK) {
    fprintf(stderr,
            "b200seg_head_ce_static: C=%d, K=%d do not fit the built-in scratch (limits: C <= %d, K <= %d); "
            "use b200seg_head_ce with a caller-provided workspace\n",
            C, K, kHeadStaticMaxC, kHeadStaticMaxK);
    return (int)cudaErrorInvalidValue;
  }
  void* scratch = nullptr;
  cudaError_t e = head_static_scratch(&scratch);
  if (e != cudaSuccess) return (int)e;
  float* partials = reinterpret_cast<float*>(scratch);
  unsigned int* counter = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(scratch) + kHeadStaticPartials);
  e = cudaMemsetAsync(counter, 0, 256, st);             // the one thing that must start at zero
  if (e != cudaSuccess) return (int)e;
  const int warps = 8, threads = warps * 32;
  long long chunks = (P + 31) / 32;
  int grid = (int)((chunks + warps - 1) / warps);
  if (grid > kHeadStaticBlocks) grid = kHeadStaticBlocks;
  const int nred = K * C + K + 2;
  const size_t smem = sizeof(float) * ((size_t)K * C + 8 + ((nred + 3) & ~3) + (size_t)warps * 32 * 8) +
                      (size_t)warps * 2 * 32 * (C + 8) * 2;
  e = head_ce_raise_smem_limit<8>();
  if (e != cudaSuccess) return (int)e;
  head_ce_kernel<8><<<grid, threads, smem, st>>>((const __nv_bfloat16*)a, ld_a, w, bias, labels, P, C, K, grad_scale,
                                                 (__nv_bfloat16*)da, ld_da, dw, db, pred, loss_acc, partials,
                                                 counter, warps);
  return (int)cudaGetLastError();
}
