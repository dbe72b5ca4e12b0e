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
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

constexpr int kHeadThreads = 256;
constexpr int kHeadWarps = kHeadThreads / 32;

template <int KMAX>
__global__ void __launch_bounds__(kHeadThreads)
head_ce_kernel(const __nv_bfloat16* __restrict__ a, long long ld_a, const float* __restrict__ w,
               const float* __restrict__ bias, const uint8_t* __restrict__ labels, long long P, int C,
               int K, float grad_scale, __nv_bfloat16* __restrict__ da, long long ld_da,
               float* __restrict__ dw, float* __restrict__ db, uint8_t* __restrict__ pred,
               float* __restrict__ loss_acc, float* __restrict__ workspace,
               unsigned int* __restrict__ counter) {
  extern __shared__ float sm[];
  float* s_w = sm;                                  // [K][C]
  float* s_b = s_w + K * C;                         // [K]
  float* s_dl = s_b + KMAX;                         // [warps][32][KMAX]
  float* s_red = s_dl + kHeadWarps * 32 * KMAX;     // [K*C + K + 2] block reduction
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nred = K * C + K + 2;
  for (int i = threadIdx.x; i < K * C; i += kHeadThreads) s_w[i] = w[i];
  for (int i = threadIdx.x; i < K; i += kHeadThreads) s_b[i] = bias[i];
  for (int i = threadIdx.x; i < nred; i += kHeadThreads) s_red[i] = 0.f;
  __syncthreads();

  const bool bwd = (da != nullptr);
  const int CV = C >> 3;
  const int cpl = (C + 31) / 32;                    // channels per lane in the dW phase (<= 8)
  float accw[KMAX][8];
  float accb[KMAX];
#pragma unroll
  for (int k = 0; k < KMAX; ++k) {
    accb[k] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) accw[k][j] = 0.f;
  }
  float loss_local = 0.f, correct_local = 0.f;
  float* my_dl = s_dl + warp * 32 * KMAX;

  const long long chunks = (P + 31) / 32;
  for (long long ch = (long long)blockIdx.x * kHeadWarps + warp; ch < chunks;
       ch += (long long)gridDim.x * kHeadWarps) {
    const long long p = ch * 32 + lane;
    const bool valid = p < P;
    float z[KMAX];
#pragma unroll
    for (int k = 0; k < KMAX; ++k) z[k] = (k < K) ? s_b[k] : -INFINITY;
    if (valid) {
      for (int cv = 0; cv < CV; ++cv) {
        float f[8];
        unpack8(ld8(a + p * ld_a + cv * 8), f);
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < K) {
#pragma unroll
            for (int j = 0; j < 8; ++j) z[k] = fmaf(f[j], s_w[k * C + cv * 8 + j], z[k]);
          }
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
      if (valid) {
        for (int cv = 0; cv < CV; ++cv) {
          float o[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = 0.f;
#pragma unroll
          for (int k = 0; k < KMAX; ++k)
            if (k < K) {
#pragma unroll
              for (int j = 0; j < 8; ++j) o[j] = fmaf(dl[k], s_w[k * C + cv * 8 + j], o[j]);
            }
          st8(da + p * ld_da + cv * 8, pack8(o));
        }
      }
      __syncwarp();
      // dW phase: lane owns channels lane, lane+32, ...; walks the 32 pixels of this chunk.
      const long long p0 = ch * 32;
      const int npx = (int)((P - p0) < 32 ? (P - p0) : 32);
      for (int i = 0; i < npx; ++i) {
        float av[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int c = lane + 32 * j;
          av[j] = (j < cpl && c < C) ? __bfloat162float(a[(p0 + i) * ld_a + c]) : 0.f;
        }
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
          const float d = my_dl[i * KMAX + k];
#pragma unroll
          for (int j = 0; j < 8; ++j) accw[k][j] = fmaf(d, av[j], accw[k][j]);
        }
      }
      __syncwarp();
    }
  }
  // ---- block reduction -> per-CTA partial ----------------------------------------------------
  loss_local = warp_sum(loss_local);
  correct_local = warp_sum(correct_local);
#pragma unroll
  for (int k = 0; k < KMAX; ++k) accb[k] = warp_sum(accb[k]);
  for (int wv = 0; wv < kHeadWarps; ++wv) {          // fixed warp order: deterministic sums
    if (warp == wv) {
      if (bwd) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < K) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int c = lane + 32 * j;
              if (j < cpl && c < C) s_red[k * C + c] += accw[k][j];
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
  __syncthreads();
  float* mine = workspace + (size_t)blockIdx.x * nred;
  for (int i = threadIdx.x; i < nred; i += kHeadThreads) mine[i] = s_red[i];
  __threadfence();
  __shared__ unsigned int s_last;
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < nred; i += kHeadThreads) {
    float t = 0.f;
    for (unsigned int b = 0; b < gridDim.x; ++b) t += workspace[(size_t)b * nred + i];
    if (i < K * C) { if (bwd) dw[i] += t; }
    else if (i < K * C + K) { if (bwd) db[i - K * C] += t; }
    else loss_acc[i - K * C - K] = t;
  }
  if (threadIdx.x == 0) *counter = 0u;               // ready for the next launch
}

}  // namespace b200

using namespace b200;

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
  long long chunks = (P + 31) / 32;
  int grid = (int)((chunks + kHeadWarps - 1) / kHeadWarps);
  if (grid > 2 * kNumSMs) grid = 2 * kNumSMs;
  if (grid < 1) grid = 1;
  const int nred = K * C + K + 2;
  size_t smem = sizeof(float) * ((size_t)K * C + KMAX + (size_t)kHeadWarps * 32 * KMAX + nred);
  float* ws = (float*)workspace;
  unsigned int* counter = (unsigned int*)((char*)workspace + b200seg_head_ce_workspace_bytes(C, K) - 256);
  if (KMAX == 8) {
    head_ce_kernel<8><<<grid, kHeadThreads, smem, st>>>((const __nv_bfloat16*)a, ld_a, w, bias, labels, P, C, K,
                                                        grad_scale, (__nv_bfloat16*)da, ld_da, dw, db, pred,
                                                        loss_acc, ws, counter);
  } else {
    cudaFuncSetAttribute(head_ce_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    head_ce_kernel<16><<<grid, kHeadThreads, smem, st>>>((const __nv_bfloat16*)a, ld_a, w, bias, labels, P, C, K,
                                                         grad_scale, (__nv_bfloat16*)da, ld_da, dw, db, pred,
                                                         loss_acc, ws, counter);
  }
  return (int)cudaGetLastError();
}
