// Input preparation and the bilinear x2 up-sampler.
//
// prep_input: the reference turns uint8 tiles into float32 / 255, NHWC -> NCHW on the host and
// copies 4-byte floats over PCIe every micro-batch (R:737-741, R:754).  Here uint8
// tiles cross PCIe and one kernel normalises, casts to bf16 and zero-pads 3 -> 16 channels (one
// UMMA K step) in NHWC.
// upsample_bilinear2x: nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True) (R:609;
// SURVEY K11) written straight into the first channels of the decoder's concat buffer, plus its
// adjoint for the backward pass.
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

__global__ void __launch_bounds__(256) prep_input_kernel(const uint8_t* __restrict__ img, long long P,
                                                         int Cin, __nv_bfloat16* __restrict__ out,
                                                         int Cpad) {
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < P;
       p += (long long)gridDim.x * blockDim.x) {
    for (int c0 = 0; c0 < Cpad; c0 += 8) {
      float f[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = (c0 + j < Cin) ? (float)img[p * Cin + c0 + j] / 255.0f : 0.f;
      st8(out + p * Cpad + c0, pack8(f));
    }
  }
}

__device__ __forceinline__ void src_index(int dst, int n_src, int n_dst, int& i0, int& i1, float& w1) {
  const float r = (n_dst > 1) ? (float)(n_src - 1) / (float)(n_dst - 1) : 0.f;
  const float s = r * (float)dst;
  i0 = min((int)s, n_src - 1);
  i1 = min(i0 + 1, n_src - 1);
  w1 = s - (float)i0;
}

__global__ void __launch_bounds__(256)
upsample2x_fwd_kernel(const __nv_bfloat16* __restrict__ x, long long ld_x, __nv_bfloat16* __restrict__ out,
                      long long ld_o, int N, int h, int w, int C) {
  const int CV = C >> 3, H = 2 * h, W = 2 * w;
  const long long items = (long long)N * H * W * CV;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < items;
       idx += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(idx % CV);
    long long t = idx / CV;
    const int X = (int)(t % W); t /= W;
    const int Y = (int)(t % H);
    const long long n = t / H;
    int y0, y1, x0, x1; float wy, wx;
    src_index(Y, h, H, y0, y1, wy);
    src_index(X, w, W, x0, x1, wx);
    float a[8], b[8], c[8], d[8], o[8];
    unpack8(ld8(x + ((n * h + y0) * w + x0) * ld_x + cv * 8), a);
    unpack8(ld8(x + ((n * h + y0) * w + x1) * ld_x + cv * 8), b);
    unpack8(ld8(x + ((n * h + y1) * w + x0) * ld_x + cv * 8), c);
    unpack8(ld8(x + ((n * h + y1) * w + x1) * ld_x + cv * 8), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float top = a[j] + (b[j] - a[j]) * wx, bot = c[j] + (d[j] - c[j]) * wx;
      o[j] = top + (bot - top) * wy;
    }
    st8(out + ((n * H + Y) * W + X) * ld_o + cv * 8, pack8(o));
  }
}

// Adjoint: every low-res pixel gathers from the up-sampled pixels it contributed to.
__global__ void __launch_bounds__(256)
upsample2x_bwd_kernel(const __nv_bfloat16* __restrict__ g, long long ld_g, __nv_bfloat16* __restrict__ dx,
                      long long ld_x, int N, int h, int w, int C) {
  const int CV = C >> 3, H = 2 * h, W = 2 * w;
  const long long items = (long long)N * h * w * CV;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < items;
       idx += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(idx % CV);
    long long t = idx / CV;
    const int xs = (int)(t % w); t /= w;
    const int ys = (int)(t % h);
    const long long n = t / h;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const int Ylo = max(0, 2 * ys - 3), Yhi = min(H - 1, 2 * ys + 3);
    const int Xlo = max(0, 2 * xs - 3), Xhi = min(W - 1, 2 * xs + 3);
    for (int Y = Ylo; Y <= Yhi; ++Y) {
      int y0, y1; float wy;
      src_index(Y, h, H, y0, y1, wy);
      float cy = (y0 == ys ? 1.f - wy : 0.f) + (y1 == ys ? wy : 0.f);
      if (cy == 0.f) continue;
      for (int X = Xlo; X <= Xhi; ++X) {
        int x0, x1; float wx;
        src_index(X, w, W, x0, x1, wx);
        float cx = (x0 == xs ? 1.f - wx : 0.f) + (x1 == xs ? wx : 0.f);
        if (cx == 0.f) continue;
        float f[8];
        unpack8(ld8(g + ((n * H + Y) * W + X) * ld_g + cv * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(cy * cx, f[j], acc[j]);
      }
    }
    st8(dx + ((n * h + ys) * w + xs) * ld_x + cv * 8, pack8(acc));
  }
}

// ---- small glue kernels that keep eager ATen ops out of the captured training step -----------------------
// rows x cin bf16 -> rows x cpad bf16, zero padded (the stem's 3-channel weights as one UMMA K step)
__global__ void pad_rows_kernel(const __nv_bfloat16* __restrict__ src, int rows, int cin, __nv_bfloat16* __restrict__ dst,
                                int cpad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cpad) return;
  const int r = i / cpad, c = i - r * cpad;
  dst[i] = c < cin ? src[r * cin + c] : __float2bfloat16(0.f);
}
// dst[r][c] += src[r][c] for c < cin, then src := 0 (the padded stem weight gradient folded back and re-armed)
__global__ void unpad_add_zero_kernel(float* __restrict__ src, int rows, int cpad, float* __restrict__ dst, int cin) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * cpad) return;
  const int r = i / cpad, c = i - r * cpad;
  if (c < cin) dst[r * cin + c] += src[i];
  src[i] = 0.f;
}
// out[c] += sum over rows of part[r * row_stride + c]  (per-CTA column sums -> the transposed conv's bias gradient)
__global__ void rows_sum_add_kernel(const float* __restrict__ part, int rows, long long row_stride, int C,
                                    float* __restrict__ out) {
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (c >= C) return;
  float s = 0.f;
  for (int r = lane; r < rows; r += 32) s += part[(size_t)r * row_stride + c];
  s = warp_sum(s);
  if (lane == 0) out[c] += s;
}

}  // namespace b200

using namespace b200;

extern "C" int b200seg_pad_rows(const void* src, int rows, int cin, void* dst, int cpad, cudaStream_t st) {
  if (cin > cpad || rows < 1) return (int)cudaErrorInvalidValue;
  pad_rows_kernel<<<(rows * cpad + 255) / 256, 256, 0, st>>>((const __nv_bfloat16*)src, rows, cin, (__nv_bfloat16*)dst, cpad);
  return (int)cudaGetLastError();
}
extern "C" int b200seg_unpad_add_zero(float* src, int rows, int cpad, float* dst, int cin, cudaStream_t st) {
  if (cin > cpad || rows < 1) return (int)cudaErrorInvalidValue;
  unpad_add_zero_kernel<<<(rows * cpad + 255) / 256, 256, 0, st>>>(src, rows, cpad, dst, cin);
  return (int)cudaGetLastError();
}
extern "C" int b200seg_rows_sum_add(const float* part, int rows, long long row_stride, int C, float* out, cudaStream_t st) {
  if (rows < 1 || C < 1) return (int)cudaErrorInvalidValue;
  rows_sum_add_kernel<<<(C * 32 + 127) / 128, 128, 0, st>>>(part, rows, row_stride, C, out);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_prep_input(const unsigned char* img, long long P, int Cin, void* out, int Cpad,
                                  cudaStream_t st) {
  if (Cpad % 8) return (int)cudaErrorInvalidValue;
  prep_input_kernel<<<grid_for(P, 256, 8), 256, 0, st>>>(img, P, Cin, (__nv_bfloat16*)out, Cpad);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_upsample2x_fwd(const void* x, long long ld_x, void* out, long long ld_o, int N, int h,
                                      int w, int C, cudaStream_t st) {
  if (C % 8) return (int)cudaErrorInvalidValue;
  upsample2x_fwd_kernel<<<grid_for((long long)N * 4 * h * w * (C / 8), 256, 8), 256, 0, st>>>(
      (const __nv_bfloat16*)x, ld_x, (__nv_bfloat16*)out, ld_o, N, h, w, C);
  return (int)cudaGetLastError();
}

extern "C" int b200seg_upsample2x_bwd(const void* g, long long ld_g, void* dx, long long ld_x, int N, int h,
                                      int w, int C, cudaStream_t st) {
  if (C % 8) return (int)cudaErrorInvalidValue;
  upsample2x_bwd_kernel<<<grid_for((long long)N * h * w * (C / 8), 256, 8), 256, 0, st>>>(
      (const __nv_bfloat16*)g, ld_g, (__nv_bfloat16*)dx, ld_x, N, h, w, C);
  return (int)cudaGetLastError();
}
