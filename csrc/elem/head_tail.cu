// The tail of the network in ONE pass over the last conv's raw output:
//
//   y (raw output of up_conv1's second conv, bf16 NHWC)
//     -> a = relu(y * scale + shift)                 BatchNorm(train) + ReLU  (R:583-584), never written to HBM
//     -> logits = a W^T + b                          1x1 head                 (R:641, R:655)
//     -> softmax cross-entropy, arg-max, accuracy    (R:703, R:755, R:775)
//     -> dl = (softmax - onehot) * grad_scale
//     -> dW += dl^T a, db += sum dl, g = dl W        head backward; g = d(loss)/d(a) is written (bf16)
//     -> sum g*[a>0], sum g*[a>0]*y per channel      the two reductions of that BatchNorm's backward
//
// It replaces, for the last conv stage, bn_relu_pool_fwd (1.07 GB of traffic at the headline shape), the
// head kernel's fp32 CUDA-core GEMMs (0.54 ms, issue-bound at 0.33 of the HBM rate) and bn_bwd_reduce
// (another 1.07 GB): what remains of that stage's BN backward is finalize + apply.
//
// The three GEMMs are tiny in N (classes <= 8) and run on the tensor cores through warp-level
// mma.sync (m16n8k16 / m16n8k8, bf16 in, fp32 accumulate): a tcgen05 instruction needs N >= 16 per
// 128 rows and a TMEM round trip, and this kernel is bound by HBM, not by math -- what matters is that the
// ~1150 FMAs per pixel of the CUDA-core version shrink to 20 warp instructions per 16 pixels.
// fp32 operands are split into bf16 hi + lo parts where they enter a GEMM that the step's numerics
// depend on (head weights for the logits, dl for dW), so those keep ~16 mantissa bits.
#define B200SEG_PDL_CLASS 2
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

constexpr int kHtWarps = 8;
constexpr int kHtTile = 32;                     // pixels per warp tile (two 16-row MMA tiles)
// Two CTAs of 8 warps per SM (one for 128 channels); every warp streams its own 32-pixel tiles.  A variant with one
// 14-warp CTA per SM that prefetched the next tile with cp.async while the current one was in the MMAs measured
// SLOWER (0.365 vs 0.349 ms at 16 x 512 x 512 x 64): the kernel is bound by instruction issue (~1700 per tile, most
// of them the fp32 BatchNorm arithmetic around the MMAs), not by load latency, so resident warps matter more.

__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr)
               : "memory");
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr)
               : "memory");
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t (&r)[2], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];"
               : "=r"(r[0]), "=r"(r[1])
               : "r"(addr)
               : "memory");
}
// D(16x8, fp32) += A(16x16, row) * B(16x8, col)
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
      "{%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// D(16x8, fp32) += A(16x8, row) * B(8x8, col)
__device__ __forceinline__ void mma_1688(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%0, %1, %2, %3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(b0));
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
// v = hi + lo with hi, lo bf16: the pair carries ~16 mantissa bits through a bf16 MMA
__device__ __forceinline__ void split2(float v0, float v1, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(v0, v1);
  const float2 hf = __bfloat1622float2(h);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = pack2(v0 - hf.x, v1 - hf.y);
}

// Fragment coordinates used throughout (PTX mma.m16n8k16 layouts): g = lane / 4, t = lane % 4.
//   A (16 x 16, row):  a0 = (g, 2t..2t+1)  a1 = (g+8, 2t..)  a2 = (g, 2t+8..)  a3 = (g+8, 2t+8..)
//   B (16 x 8, col):   b0 = (k = 2t..2t+1, n = g)  b1 = (k = 2t+8.., n = g)
//   C (16 x 8):        c0, c1 = (g, 2t), (g, 2t+1)   c2, c3 = (g+8, 2t), (g+8, 2t+1)
template <int CT>
__global__ void __launch_bounds__(kHtWarps * 32, CT > 64 ? 1 : 2)
head_tail_kernel(const __nv_bfloat16* __restrict__ y, long long ld_y, const float* __restrict__ ss,
                 const float* __restrict__ w, const float* __restrict__ bias, const uint8_t* __restrict__ labels,
                 long long P, int K, float grad_scale, __nv_bfloat16* __restrict__ g_out, long long ld_g,
                 float* __restrict__ dw, float* __restrict__ db, uint8_t* __restrict__ pred,
                 float* __restrict__ loss_acc, float* __restrict__ bn_partial, float* __restrict__ workspace,
                 unsigned int* __restrict__ counter) {
  constexpr int CV = CT / 8;                    // 16-byte pieces per pixel
  constexpr int pitch = CT + 8;                 // bf16 elements per tile row (ldmatrix rows fall on distinct banks)
  constexpr int KS = CT / 16;                   // k steps of the logits GEMM = m tiles of the dW GEMM
  constexpr int NT = CT / 8;                    // n tiles of the g GEMM
  constexpr int nredp = 8 * CT + 16;
  constexpr int warp_elems = 2 * kHtTile * pitch + 2 * kHtTile * 8;
  extern __shared__ __align__(16) uint8_t smraw[];
  pdl_wait();
  pdl_trigger();
  float* s_red = reinterpret_cast<float*>(smraw);                    // [K*CT dW | K db | loss, correct]
  float* s_bn = s_red + nredp;                                        // [2][CT]: sum g, sum g*y
  uint4* s_wB = reinterpret_cast<uint4*>(s_bn + 2 * CT);              // [KS][32]: logits B fragments (hi0, hi1, lo0, lo1)
  uint32_t* s_wT = reinterpret_cast<uint32_t*>(s_wB + KS * 32);       // [NT][32]: g-GEMM B fragments
  __nv_bfloat16* s_tiles = reinterpret_cast<__nv_bfloat16*>(s_wT + NT * 32);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int nred = K * CT + K + 2;
  const bool bwd = (g_out != nullptr);

  for (int i = threadIdx.x; i < nredp + 2 * CT; i += blockDim.x) s_red[i] = 0.f;      // s_red and s_bn
  for (int i = threadIdx.x; i < KS * 32; i += blockDim.x) {
    const int ks = i >> 5, l = i & 31, n = l >> 2, k0 = ks * 16 + 2 * (l & 3);
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (n < K) { v[0] = w[n * CT + k0]; v[1] = w[n * CT + k0 + 1]; v[2] = w[n * CT + k0 + 8]; v[3] = w[n * CT + k0 + 9]; }
    uint4 f;
    split2(v[0], v[1], f.x, f.z);
    split2(v[2], v[3], f.y, f.w);
    s_wB[i] = f;
  }
  for (int i = threadIdx.x; i < NT * 32; i += blockDim.x) {
    const int nt = i >> 5, l = i & 31, c = nt * 8 + (l >> 2), k0 = 2 * (l & 3);
    s_wT[i] = pack2(k0 < K ? w[k0 * CT + c] : 0.f, k0 + 1 < K ? w[(k0 + 1) * CT + c] : 0.f);
  }
  __syncthreads();

  __nv_bfloat16* s_a = s_tiles + (size_t)warp * warp_elems;          // [32][pitch] activation, later g
  __nv_bfloat16* s_y = s_a + kHtTile * pitch;                        // [32][pitch] raw conv output
  __nv_bfloat16* s_dl = s_y + kHtTile * pitch;                       // [2][32][8]  dl hi / lo
  const uint32_t a_u32 = smem_addr(s_a), dl_u32 = smem_addr(s_dl);
  const int cv = lane % CV;                                          // this lane's channel piece in the row-wise phases
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { sc[j] = __ldg(ss + cv * 8 + j); sh[j] = __ldg(ss + CT + cv * 8 + j); }
  const int c0 = 2 * t, c1 = 2 * t + 1;                              // this lane's classes in the C fragments
  const float bias0 = c0 < K ? __ldg(bias + c0) : 0.f, bias1 = c1 < K ? __ldg(bias + c1) : 0.f;

  float dwT[KS][4];                   // dW^T fragments: (channel 16 mc + g (+8), class 2t (+1))
#pragma unroll
  for (int mc = 0; mc < KS; ++mc) dwT[mc][0] = dwT[mc][1] = dwT[mc][2] = dwT[mc][3] = 0.f;
  float accb0 = 0.f, accb1 = 0.f, loss_local = 0.f, correct_local = 0.f;
  float bs[8], bq[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bs[j] = bq[j] = 0.f;

  const long long chunks = (P + kHtTile - 1) / kHtTile;
  // A lane always handles the same (pixel, 16-byte piece) slots of a tile -- when it loads y, when it turns y into
  // a, and when it sums and stores g.
  for (long long ch = (long long)blockIdx.x * kHtWarps + warp; ch < chunks; ch += (long long)gridDim.x * kHtWarps) {
    const long long p0 = ch * kHtTile;
    const int npx = (P - p0) < (long long)kHtTile ? (int)(P - p0) : kHtTile;      // pixels of this tile that exist
    // ---- 1. coalesced load of the y tile; a = relu(y * scale + shift), rounded as the forward BN kernel rounds it
    constexpr int LB = CV < 8 ? CV : 8;                              // loads in flight per lane
#pragma unroll
    for (int ib = 0; ib < CV; ib += LB) {
      bf16x8 yv[LB];
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        const int px = ((ib + u) * 32 + lane) / CV;
        if (px < npx) yv[u] = ld8(y + (p0 + px) * ld_y + cv * 8);
        else yv[u].v[0] = yv[u].v[1] = yv[u].v[2] = yv[u].v[3] = __floats2bfloat162_rn(0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < LB; ++u) {
        const int px = ((ib + u) * 32 + lane) / CV;
        float f[8];
        unpack8(yv[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(fmaf(f[j], sc[j], sh[j]), 0.f);
        *reinterpret_cast<bf16x8*>(s_y + px * pitch + cv * 8) = yv[u];
        *reinterpret_cast<bf16x8*>(s_a + px * pitch + cv * 8) = pack8(f);
      }
    }
    const int lab_mine = (lane < npx) ? (int)labels[p0 + lane] : 0;
    __syncwarp();

    // ---- 2. logits: z[mt] = C fragment of pixels 16 mt .. 16 mt + 15
    float z[2][4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) { z[mt][0] = z[mt][2] = bias0; z[mt][1] = z[mt][3] = bias1; }
    {
      const uint32_t a_lane = a_u32 + (uint32_t)((((lane & 7) + ((lane >> 3) & 1) * 8) * pitch + (lane >> 4) * 8) * 2);
#pragma unroll
      for (int ks = 0; ks < KS; ++ks) {
        const uint4 wf = s_wB[ks * 32 + lane];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          uint32_t af[4];
          ldsm_x4(af, a_lane + (uint32_t)((mt * 16 * pitch + ks * 16) * 2));
          mma_16816(z[mt], af, wf.x, wf.y);
          mma_16816(z[mt], af, wf.z, wf.w);
        }
      }
    }

    // ---- 3. softmax / loss / arg-max per pixel (a pixel's classes live in one lane quad)
    float dl[2][4];
    int argv[4];
#pragma unroll
    for (int mt = 0; mt < 2; ++mt) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int pi = mt * 16 + h * 8 + g;
        const float z0 = c0 < K ? z[mt][2 * h] : -INFINITY, z1 = c1 < K ? z[mt][2 * h + 1] : -INFINITY;
        float mx = fmaxf(z0, z1);
        int arg = (z1 > z0) ? c1 : c0;                               // first maximum wins, as ATen
#pragma unroll
        for (int o = 1; o <= 2; o <<= 1) {
          const float om = __shfl_xor_sync(0xffffffffu, mx, o);
          const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
          if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
        }
        const float e0 = c0 < K ? __expf(z0 - mx) : 0.f, e1 = c1 < K ? __expf(z1 - mx) : 0.f;
        float sum = e0 + e1;
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        sum += __shfl_xor_sync(0xffffffffu, sum, 2);
        const int lab = __shfl_sync(0xffffffffu, lab_mine, pi);
        const bool valid = pi < npx;
        if (valid) {
          if (t == 0) {
            loss_local += __logf(sum) + mx;
            correct_local += (arg == lab) ? 1.f : 0.f;
          }
          if (lab == c0) loss_local -= z0;
          if (lab == c1) loss_local -= z1;
        }
        const float inv = grad_scale / sum;
        const float d0 = (valid && c0 < K) ? (e0 * inv - (lab == c0 ? grad_scale : 0.f)) : 0.f;
        const float d1 = (valid && c1 < K) ? (e1 * inv - (lab == c1 ? grad_scale : 0.f)) : 0.f;
        accb0 += d0;
        accb1 += d1;
        dl[mt][2 * h] = d0;
        dl[mt][2 * h + 1] = d1;
        argv[mt * 2 + h] = arg;
      }
    }
    if (pred != nullptr) {                                           // lane i stores the prediction of pixel i
      const int src = (lane & 7) * 4;
      const int a0 = __shfl_sync(0xffffffffu, argv[0], src), a1 = __shfl_sync(0xffffffffu, argv[1], src);
      const int a2 = __shfl_sync(0xffffffffu, argv[2], src), a3 = __shfl_sync(0xffffffffu, argv[3], src);
      const int sel = lane >> 3;
      const int mine = sel == 0 ? a0 : (sel == 1 ? a1 : (sel == 2 ? a2 : a3));
      if (lane < npx) pred[p0 + lane] = (uint8_t)mine;
    }

    if (bwd) {
      // ---- 4. dW^T[channel][class] += a^T dl  (A = a^T through ldmatrix.trans, B = dl hi + lo)
#pragma unroll
      for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int pi = mt * 16 + h * 8 + g;
          uint32_t hi, lo;
          split2(dl[mt][2 * h], dl[mt][2 * h + 1], hi, lo);
          *reinterpret_cast<uint32_t*>(s_dl + pi * 8 + 2 * t) = hi;
          *reinterpret_cast<uint32_t*>(s_dl + kHtTile * 8 + pi * 8 + 2 * t) = lo;
        }
      __syncwarp();
#pragma unroll
      for (int kp = 0; kp < 2; ++kp) {
        uint32_t bh[2], bl[2];
        ldsm_x2_t(bh, dl_u32 + (uint32_t)((kp * 16 + (lane & 15)) * 16));
        ldsm_x2_t(bl, dl_u32 + (uint32_t)(kHtTile * 16 + (kp * 16 + (lane & 15)) * 16));
        const uint32_t at_lane =
            a_u32 + (uint32_t)(((kp * 16 + (lane & 7) + (lane >> 4) * 8) * pitch + ((lane >> 3) & 1) * 8) * 2);
#pragma unroll
        for (int mc = 0; mc < KS; ++mc) {
          uint32_t af[4];
          ldsm_x4_t(af, at_lane + (uint32_t)(mc * 32));
          mma_16816(dwT[mc], af, bh[0], bh[1]);
          mma_16816(dwT[mc], af, bl[0], bl[1]);
        }
      }
      __syncwarp();                                                  // every lane is done reading the a tile
      // ---- 5. g = dl W, written over the a tile (bf16, as it goes to memory)
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const uint32_t da0 = pack2(dl[mt][0], dl[mt][1]), da1 = pack2(dl[mt][2], dl[mt][3]);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float o[4] = {0.f, 0.f, 0.f, 0.f};
          mma_1688(o, da0, da1, s_wT[nt * 32 + lane]);
          *reinterpret_cast<uint32_t*>(s_a + (mt * 16 + g) * pitch + nt * 8 + 2 * t) = pack2(o[0], o[1]);
          *reinterpret_cast<uint32_t*>(s_a + (mt * 16 + g + 8) * pitch + nt * 8 + 2 * t) = pack2(o[2], o[3]);
        }
      }
      __syncwarp();
      // ---- 6. coalesced store of g + the two BatchNorm-backward sums of this stage
#pragma unroll
      for (int it = 0; it < CV; ++it) {
        const int px = (it * 32 + lane) / CV;
        const bf16x8 gv = *reinterpret_cast<const bf16x8*>(s_a + px * pitch + cv * 8);
        float gf[8], yf[8];
        unpack8(gv, gf);
        unpack8(*reinterpret_cast<const bf16x8*>(s_y + px * pitch + cv * 8), yf);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float gm = (fmaf(yf[j], sc[j], sh[j]) > 0.f) ? gf[j] : 0.f;
          bs[j] += gm;
          bq[j] = fmaf(gm, yf[j], bq[j]);
        }
        if (px < npx) st8(g_out + (p0 + px) * ld_g + cv * 8, gv);
      }
    }
    __syncwarp();
  }

  // ---- CTA reduction (fixed warp order: the sums do not depend on scheduling) ---------------------------
  loss_local = warp_sum(loss_local);
  correct_local = warp_sum(correct_local);
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) {
    accb0 += __shfl_xor_sync(0xffffffffu, accb0, o);
    accb1 += __shfl_xor_sync(0xffffffffu, accb1, o);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int o = CV; o < 32; o <<= 1) {
      bs[j] += __shfl_xor_sync(0xffffffffu, bs[j], o);
      bq[j] += __shfl_xor_sync(0xffffffffu, bq[j], o);
    }
  }
  for (int wv = 0; wv < kHtWarps; ++wv) {
    if (warp == wv) {
      if (bwd) {
#pragma unroll
        for (int mc = 0; mc < KS; ++mc) {
          const int ch0 = mc * 16 + g;
          if (c0 < K) { s_red[c0 * CT + ch0] += dwT[mc][0]; s_red[c0 * CT + ch0 + 8] += dwT[mc][2]; }
          if (c1 < K) { s_red[c1 * CT + ch0] += dwT[mc][1]; s_red[c1 * CT + ch0 + 8] += dwT[mc][3]; }
        }
        if (g == 0) {
          if (c0 < K) s_red[K * CT + c0] += accb0;
          if (c1 < K) s_red[K * CT + c1] += accb1;
        }
        if (lane < CV) {
#pragma unroll
          for (int j = 0; j < 8; ++j) { s_bn[lane * 8 + j] += bs[j]; s_bn[CT + lane * 8 + j] += bq[j]; }
        }
      }
      if (lane == 0) {
        s_red[K * CT + K] += loss_local;
        s_red[K * CT + K + 1] += correct_local;
      }
    }
    __syncthreads();
  }
  if (bwd && bn_partial != nullptr)
    for (int i = threadIdx.x; i < 2 * CT; i += blockDim.x) bn_partial[(size_t)blockIdx.x * 2 * CT + i] = s_bn[i];
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
    float tsum = 0.f;
    for (unsigned int b = 0; b < gridDim.x; ++b) tsum += workspace[(size_t)b * nred + i];
    if (i < K * CT) { if (bwd) dw[i] += tsum; }
    else if (i < K * CT + K) { if (bwd) db[i - K * CT] += tsum; }
    else loss_acc[i - K * CT - K] = tsum;
  }
  if (threadIdx.x == 0) *counter = 0u;               // ready for the next launch
}

template <int CT>
static size_t head_tail_smem() {
  return sizeof(float) * ((size_t)(8 * CT + 16) + 2 * CT) + sizeof(uint4) * (CT / 16) * 32 +
         sizeof(uint32_t) * (CT / 8) * 32 +
         (size_t)kHtWarps * (2 * kHtTile * (CT + 8) + 2 * kHtTile * 8) * sizeof(__nv_bfloat16);
}

template <int CT>
static int launch_head_tail(const void* y, long long ld_y, const float* ss, const float* w, const float* bias,
                            const unsigned char* labels, long long P, int K, float grad_scale, void* g_out,
                            long long ld_g, float* dw, float* db, unsigned char* pred, float* loss_acc,
                            float* bn_partial, float* ws, unsigned int* counter, int grid, cudaStream_t st) {
  const size_t smem = head_tail_smem<CT>();
  static bool attr_done[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64 || !attr_done[dev]) {
    e = cudaFuncSetAttribute(head_tail_kernel<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    if (dev >= 0 && dev < 64) attr_done[dev] = true;
  }
  return (int)launch_k(head_tail_kernel<CT>, grid, kHtWarps * 32, smem, st, (const __nv_bfloat16*)y, ld_y, ss, w, bias,
                       (const uint8_t*)labels, P, K, grad_scale, (__nv_bfloat16*)g_out, ld_g, dw, db, (uint8_t*)pred,
                       loss_acc, bn_partial, ws, counter);
}

}  // namespace b200

using namespace b200;

// Rows of `bn_partial` (= CTAs) that b200seg_head_tail writes for P pixels of C channels.
extern "C" int b200seg_head_tail_rows(long long P, int C) {
  const long long chunks = (P + kHtTile - 1) / kHtTile;
  long long grid = (chunks + kHtWarps - 1) / kHtWarps;
  const long long cap = (long long)kNumSMs * (C > 64 ? 1 : 2);
  if (grid > cap) grid = cap;
  return grid < 1 ? 1 : (int)grid;
}

extern "C" int b200seg_head_tail_supported(int C, int K) {
  return (C == 32 || C == 64 || C == 128) && K >= 1 && K <= 8;
}

// ss: float[4][C] of the stage's BatchNorm (scale, shift, mean, invstd; bn_finalize's output).
// bn_partial: float[b200seg_head_tail_rows(P, C)][2][C], written when g_out != nullptr.
// workspace: b200seg_head_ce_workspace_bytes(C, K) bytes, same contract as b200seg_head_ce.
extern "C" int b200seg_head_tail(const void* y, long long ld_y, const float* ss, const float* w, const float* bias,
                                 const unsigned char* labels, long long P, int C, int K, float grad_scale,
                                 void* g_out, long long ld_g, float* dw, float* db, unsigned char* pred,
                                 float* loss_acc, float* bn_partial, void* workspace, cudaStream_t st) {
  if (!b200seg_head_tail_supported(C, K) || P < 1) return (int)cudaErrorInvalidValue;
  if (g_out != nullptr && (dw == nullptr || db == nullptr || bn_partial == nullptr)) return (int)cudaErrorInvalidValue;
  const int grid = b200seg_head_tail_rows(P, C);
  float* ws = (float*)workspace;
  unsigned int* counter = (unsigned int*)((char*)workspace + b200seg_head_ce_workspace_bytes(C, K) - 256);
  if (C == 32)
    return launch_head_tail<32>(y, ld_y, ss, w, bias, labels, P, K, grad_scale, g_out, ld_g, dw, db, pred, loss_acc,
                                bn_partial, ws, counter, grid, st);
  if (C == 64)
    return launch_head_tail<64>(y, ld_y, ss, w, bias, labels, P, K, grad_scale, g_out, ld_g, dw, db, pred, loss_acc,
                                bn_partial, ws, counter, grid, st);
  return launch_head_tail<128>(y, ld_y, ss, w, bias, labels, P, K, grad_scale, g_out, ld_g, dw, db, pred, loss_acc,
                               bn_partial, ws, counter, grid, st);
}
