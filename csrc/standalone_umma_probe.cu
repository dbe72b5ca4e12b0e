// Stand-alone probe for the tcgen05 shared-memory descriptor conventions used by csrc/conv.
// One CTA computes D[128 x N] = A[128 x K] * B[N x K]^T (K = 128) with every operand loaded by
// TMA with 128-byte swizzle, for several operand layouts, and compares against the host:
//   exp 0  A K-major,  B K-major                      (plain GEMM tiles)
//   exp 1  A K-major,  B MN-major ([K][N] in memory)  (dgrad: weights used un-transposed)
//   exp 2  A MN-major, B MN-major                     (wgrad: both operands pixel-major)
//   exp 3  A K-major read through a window shifted by `shift` rows of a taller smem tile
//          (conv taps served from one halo strip), descriptor base_offset = 0
//   exp 4  same with base_offset = shift & 7
//   exp 5  A MN-major with the window shifted by `shift` K-rows, base_offset = 0
//   exp 6  same with base_offset = shift & 7
// Usage: umma_probe <exp> <N> [shift] [lbo_override] [sbo_override]
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o umma_probe standalone_umma_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cmath>
#include "conv/umma.cuh"

using namespace b200;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                CUtensorMapFloatOOBfill);

static EncodeTiled get_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  return reinterpret_cast<EncodeTiled>(fn);
}

struct Params {
  int N;            // columns of D
  int a_mn, b_mn;   // operand majors
  int shift;        // window shift in smem rows (of 128 B)
  int base_off;     // descriptor base_offset for the shifted operand
  int a_rows;       // rows of the A smem block (128 + extra when shifted, K-major)
  uint32_t lbo_a, sbo_a, lbo_b, sbo_b;
  uint32_t a_kb_stride, b_kb_stride;   // bytes between consecutive 64-wide K blocks in smem
  uint32_t a_k16_step, b_k16_step;     // bytes to advance the start address per UMMA (K = 16)
};

constexpr int KTOT = 128;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
             float* __restrict__ d_out, Params p, int* __restrict__ status) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar_full, bar_mma;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5;
  uint8_t* sA = smem;                       // 2 K-blocks
  uint8_t* sB = smem + 2 * p.a_kb_stride;

  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_mma, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(&tmem_base_s);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_base_s;

  if (threadIdx.x == 0) {
    uint32_t bytes = (KTOT / 64) * ((p.a_mn ? 2 : 1) * p.a_rows * 128 + p.N * 128);
    mbar_arrive_expect_tx(&bar_full, bytes);
    for (int kb = 0; kb < KTOT / 64; ++kb) {
      if (!p.a_mn) {   // A[rows][K]: box = (64 k, a_rows)
        tma_load_2d(sA + kb * p.a_kb_stride, &map_a, &bar_full, kb * 64, 0);
      } else {         // A^T[K][128]: two boxes (64 m, 64(+shift rows) k)
        for (int mb = 0; mb < 2; ++mb) {
          tma_load_2d(sA + kb * p.a_kb_stride + mb * p.lbo_a, &map_a, &bar_full, mb * 64, kb * 64);
          }
      }
      if (!p.b_mn) {   // B[N][K]: box = (64 k, N)
        tma_load_2d(sB + kb * p.b_kb_stride, &map_b, &bar_full, kb * 64, 0);
      } else {         // B^T[K][N]: N/64 boxes (64 n, 64 k)
        for (int nb = 0; nb < p.N / 64; ++nb) {
          tma_load_2d(sB + kb * p.b_kb_stride + nb * p.lbo_b, &map_b, &bar_full, nb * 64, kb * 64);
        }
      }
    }
    if (!mbar_wait(&bar_full, 0, 20000000u)) { status[0] = 1; asm volatile("trap;"); }
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, p.N, p.a_mn, p.b_mn);
    uint32_t acc = 0;
    for (int kb = 0; kb < KTOT / 64; ++kb) {
      for (int k4 = 0; k4 < 4; ++k4) {
        uint32_t a_addr = smem_u32(sA + kb * p.a_kb_stride) + k4 * p.a_k16_step + p.shift * 128;
        uint32_t b_addr = smem_u32(sB + kb * p.b_kb_stride) + k4 * p.b_k16_step;
        uint64_t da = make_smem_desc(a_addr, p.lbo_a, p.sbo_a, LAYOUT_SW128, p.base_off);
        uint64_t db = make_smem_desc(b_addr, p.lbo_b, p.sbo_b, LAYOUT_SW128, 0);
        umma_bf16(tmem_d, da, db, idesc, acc);
        acc = 1;
      }
    }
    umma_commit(&bar_mma);
  }
  __syncwarp();
  if (!mbar_wait(&bar_mma, 0, 20000000u)) { status[0] = 2; asm volatile("trap;"); }
  tc_fence_after();
  for (int c0 = 0; c0 < p.N; c0 += 32) {
    uint32_t r[32];
    tmem_ld_32x32b_x32(tmem_d + ((uint32_t)(warp * 32) << 16) + c0, r);
    tmem_ld_wait();
    int row = warp * 32 + (threadIdx.x & 31);
    for (int j = 0; j < 32; ++j) d_out[row * p.N + c0 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_d);
}

static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

int main(int argc, char** argv) {
  int exp = argc > 1 ? atoi(argv[1]) : 0;
  int N = argc > 2 ? atoi(argv[2]) : 128;
  int shift = argc > 3 ? atoi(argv[3]) : 0;
  long lbo_o = argc > 4 ? atol(argv[4]) : -1, sbo_o = argc > 5 ? atol(argv[5]) : -1;
  Params p{};
  p.N = N;
  p.a_mn = (exp == 2 || exp == 5 || exp == 6);
  p.b_mn = (exp == 1 || exp == 2);
  p.shift = (exp >= 3) ? shift : 0;
  p.base_off = (exp == 4 || exp == 6) ? (shift & 7) : 0;
  const int extra = (exp >= 3) ? 8 : 0;            // extra smem rows loaded below the window
  const int M = 128, K = KTOT;
  // host data: A_full has (M + extra) rows (K-major) or (K + extra rows per block) handled below
  std::vector<float> A((size_t)(M + 64) * (K + 64)), B((size_t)N * K);
  srand(1234);
  auto rnd = []() { return bf((float)(rand() % 2001 - 1000) / 1000.0f); };
  for (auto& v : A) v = rnd();
  for (auto& v : B) v = rnd();

  EncodeTiled enc = get_encode();
  CUtensorMap ma, mb;
  __nv_bfloat16 *dA, *dB;
  float* dD;
  int* dS;
  CK(cudaMalloc(&dD, sizeof(float) * M * N));
  CK(cudaMalloc(&dS, sizeof(int)));
  CK(cudaMemset(dS, 0, sizeof(int)));
  std::vector<float> ref((size_t)M * N, 0.f);
  cuuint32_t estr[2] = {1, 1};

  if (!p.a_mn) {
    // A stored [M+extra][K]; window rows [shift, shift+128)
    int R = M + extra;
    std::vector<__nv_bfloat16> h((size_t)R * K);
    for (int r = 0; r < R; ++r) for (int k = 0; k < K; ++k) h[(size_t)r * K + k] = __float2bfloat16(A[(size_t)r * K + k]);
    CK(cudaMalloc(&dA, h.size() * 2));
    CK(cudaMemcpy(dA, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)R};
    CUresult r = enc(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dA, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode A failed %d\n", (int)r); return 2; }
    p.a_rows = R;
    p.a_kb_stride = ((R * 128 + 1023) / 1024) * 1024;
    p.a_k16_step = 32;
    p.lbo_a = 16;       // unused for swizzled K-major
    p.sbo_a = 1024;
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
      float s = 0; for (int k = 0; k < K; ++k) s += A[(size_t)(m + p.shift) * K + k] * B[(size_t)n * K + k];
      ref[(size_t)m * N + n] = s;
    }
  } else {
    // A^T stored [K + extra][M] (M contiguous).  Each 64-wide K block is loaded as box (64 m, 64+extra k)
    // so the window can start `shift` K-rows into the block: rows [kb*64+shift, kb*64+shift+64).
    int KR = K + extra;
    std::vector<__nv_bfloat16> h((size_t)KR * M);
    for (int k = 0; k < KR; ++k) for (int m = 0; m < M; ++m) h[(size_t)k * M + m] = __float2bfloat16(A[(size_t)k * M + m]);
    CK(cudaMalloc(&dA, h.size() * 2));
    CK(cudaMemcpy(dA, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)M, (cuuint64_t)KR};
    cuuint64_t strides[1] = {(cuuint64_t)M * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(64 + extra)};
    CUresult r = enc(&ma, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dA, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode A^T failed %d\n", (int)r); return 2; }
    p.a_rows = 64 + extra;
    uint32_t blk = ((p.a_rows * 128 + 1023) / 1024) * 1024;   // one (64 m) block
    p.lbo_a = blk;                 // between the two 64-wide M blocks
    p.sbo_a = 1024;                // between 8-row K groups
    p.a_kb_stride = 2 * blk;
    p.a_k16_step = 16 * 128;       // 16 K-rows
    // with a shift, block kb covers K rows kb*64+shift .. +64 of the stored matrix
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
      float s = 0;
      for (int k = 0; k < K; ++k) {
        int kb = k / 64, kk = k % 64;
        int krow = kb * 64 + kk + p.shift;       // row of the stored A^T that the MMA reads
        s += A[(size_t)krow * M + m] * B[(size_t)n * K + k];
      }
      ref[(size_t)m * N + n] = s;
    }
  }
  if (!p.b_mn) {
    std::vector<__nv_bfloat16> h((size_t)N * K);
    for (size_t i = 0; i < h.size(); ++i) h[i] = __float2bfloat16(B[i]);
    CK(cudaMalloc(&dB, h.size() * 2));
    CK(cudaMemcpy(dB, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)N};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)N};
    CUresult r = enc(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dB, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode B failed %d\n", (int)r); return 2; }
    p.b_kb_stride = N * 128;
    p.b_k16_step = 32;
    p.lbo_b = 16;
    p.sbo_b = 1024;
  } else {
    // B^T stored [K][N]
    std::vector<__nv_bfloat16> h((size_t)K * N);
    for (int k = 0; k < K; ++k) for (int n = 0; n < N; ++n) h[(size_t)k * N + n] = __float2bfloat16(B[(size_t)n * K + k]);
    CK(cudaMalloc(&dB, h.size() * 2));
    CK(cudaMemcpy(dB, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
    cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)K};
    cuuint64_t strides[1] = {(cuuint64_t)N * 2};
    cuuint32_t box[2] = {64, 64};
    CUresult r = enc(&mb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dB, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode B^T failed %d\n", (int)r); return 2; }
    p.lbo_b = 64 * 128;            // between 64-wide N blocks
    p.sbo_b = 1024;
    p.b_kb_stride = (N / 64) * 64 * 128;
    p.b_k16_step = 16 * 128;
  }
  if (lbo_o >= 0) { if (p.a_mn) p.lbo_a = (uint32_t)lbo_o; if (p.b_mn) p.lbo_b = (uint32_t)lbo_o; }
  if (sbo_o >= 0) { if (p.a_mn) p.sbo_a = (uint32_t)sbo_o; if (p.b_mn) p.sbo_b = (uint32_t)sbo_o; }

  size_t smem = 2 * (size_t)p.a_kb_stride + 2 * (size_t)p.b_kb_stride + 2048;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  probe_kernel<<<1, 128, smem>>>(ma, mb, dD, p, dS);
  cudaError_t e = cudaDeviceSynchronize();
  int st = -1;
  if (e != cudaSuccess) {
    printf("exp %d N %d shift %d: KERNEL ERROR %s\n", exp, N, shift, cudaGetErrorString(e));
    return 1;
  }
  CK(cudaMemcpy(&st, dS, sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<float> out((size_t)M * N);
  CK(cudaMemcpy(out.data(), dD, out.size() * 4, cudaMemcpyDeviceToHost));
  double maxerr = 0, maxref = 0;
  int bad = 0;
  for (size_t i = 0; i < out.size(); ++i) {
    double d = fabs((double)out[i] - ref[i]);
    if (d > maxerr) maxerr = d;
    if (fabs(ref[i]) > maxref) maxref = fabs(ref[i]);
    if (d > 0.05) ++bad;
  }
  printf("exp %d N %d shift %d base_off %d lbo_a %u sbo_a %u lbo_b %u sbo_b %u: max_err %.5f (max_ref %.3f) bad %d/%zu -> %s\n",
         exp, N, p.shift, p.base_off, p.lbo_a, p.sbo_a, p.lbo_b, p.sbo_b, maxerr, maxref, bad, out.size(),
         bad == 0 ? "PASS" : "FAIL");
  return bad == 0 ? 0 : 3;
}
