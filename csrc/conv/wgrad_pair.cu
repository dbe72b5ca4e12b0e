// Weight gradient of the 3x3 conv for 64-channel output tiles, "row-pair" form (W >= 64):
//     dW[co][(r,s)][ci] += sum_p dy[p][co] * x[p + (r-1, s-1)][ci]                 (autograd of R:579)
//
// Why another kernel: measured on B200 across three kernel forms (profiles/umma_cost_model.txt), one
// tcgen05.mma with M = 128, K = 16 occupies the tensor pipe for about N/2 + 32 cycles: N/2 is the math, the 32
// are a fixed cost per instruction.  wgrad_strip_kernel<64> (wgrad_strip.cu) issues M = two taps x 64 Cin,
// N = 64 Cout: 20 instructions x 64 cycles per 64-pixel step for 576 cycles of math.  Fewer, wider
// instructions are what pays, so here the roles are swapped and N is made as wide as the data allows:
//
//     A (M = 128) = the dy tiles of TWO consecutive image rows stacked ([64 px][64 co] each; the second block
//                   is found through the descriptor's leading-offset field = one ring stage),
//     B (N = 192) = ONE input strip (66 px x 64 ci) seen through three windows shifted by one pixel
//                   (leading offset = one 128-byte pixel row): the three horizontal taps.
//
//   For input row s:   G1:  [dy(s-1) ; dy(s)]  x  strip(s)   ->  D1[(j, co)][(dw, ci)]   j=0: tap dh=+1, j=1: tap dh=0
//                      G2:  [dy(s+1) ;  --  ]  x  strip(s)   ->  D2[co][(dw, ci)]        tap dh=-1 (upper 64 lanes unused)
//   8 instructions x (96 + 32) cycles = 1024 cycles per step instead of 1280 (measured: 1080).  A variant with
//   both operands as row pairs (12 instructions of 128 x 128, three of four quadrants useful) measured 1145, as
//   the cost model predicts (12 x 96); it is not kept.
//
// A CTA walks "spans" of rows exactly like wgrad_strip_kernel (same stream-K split, same ring of 8 stages;
// stage j of a span carries input row h0-1+j and, from j = 2, the dy tile of row h0+j-2).  At the ends of
// a span one of the two stacked dy rows does not belong to the span: there the missing block is a tile of
// zeros (one below and one above the ring, so that the leading offset stays positive).  The pair
// (stage 7, stage 0) wraps around the ring; the dy tile of stage 0 is therefore also loaded into a
// mirror stage behind stage 7.
// Epilogue: accumulator row = (tap row, co), 64 consecutive columns = 64 consecutive ci of one tap, which
// is 256 contiguous bytes of dW: 16-byte vector reductions (red.global.add.v4.f32) with no transposition.
#include <cstdlib>

#include "umma.cuh"
#include "../common.cuh"
#include "b200seg.h"
#include "tma_host.h"

namespace b200 {

constexpr int kWpThreads = 192;
constexpr int kWpStages = 8;
constexpr uint32_t kWpStrip = 9216;                    // 66 rows x 128 B, padded to 9 KiB
constexpr uint32_t kWpDy = 8192;                       // 64 px x 64 co
constexpr uint32_t kWpStage = kWpStrip + kWpDy;
// [zero tile][stage 0 .. 7][mirror stage (dy part only)][zero tile]
constexpr uint32_t kWpRing0 = kWpDy;
constexpr uint32_t kWpZeroHi = kWpRing0 + (kWpStages + 1) * kWpStage;
constexpr size_t kWpSmem = (size_t)kWpZeroHi + kWpDy + 1024;

struct WgradPairParams {
  int N, H, W, Cin, Cout;
  int chunks, n_tiles, groups;         // Cin/64 (rounded up), Cout/64 (rounded up), chunks * n_tiles
  int wsegs, rows_per_piece, pieces_per_col, pieces;
  float* dw;
  long long mc_delta;                  // != 0: add through the multicast alias (fused all-reduce)
  const GradArm* arm;                  // != NULL: this launch completes a gradient bucket (all-reduce trigger)
};

struct WpSpan { int group, n, w0, h0, h1, next; };
__device__ __forceinline__ WpSpan wp_span(const WgradPairParams& p, int gid, int g1) {
  WpSpan s;
  s.group = gid / p.pieces;
  const int piece = gid - s.group * p.pieces;
  const int col = piece / p.pieces_per_col, seg = piece - col * p.pieces_per_col;
  int nseg = p.pieces_per_col - seg;
  if (gid + nseg > g1) nseg = g1 - gid;
  s.n = col / p.wsegs;
  s.w0 = (col - s.n * p.wsegs) * 64;
  s.h0 = seg * p.rows_per_piece;
  s.h1 = min(p.H, (seg + nseg) * p.rows_per_piece);
  s.next = gid + nseg;
  return s;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__global__ void __launch_bounds__(kWpThreads, 1)
wgrad_pair_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                  const WgradPairParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar_full[kWpStages], bar_empty[kWpStages];
  __shared__ __align__(8) uint64_t bar_done, bar_drained;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const long long total = (long long)p.groups * p.pieces;
  const int g0 = (int)(total * blockIdx.x / gridDim.x);
  const int g1 = (int)(total * (blockIdx.x + 1) / gridDim.x);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_dy);
    for (int s = 0; s < kWpStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_done, 1);
    mbar_init(&bar_drained, 128);
    fence_barrier_init();
  }
  // the two zero tiles (generic-proxy writes, made visible to the tensor core's async proxy below)
  for (uint32_t i = threadIdx.x * 16; i < kWpDy; i += kWpThreads * 16) {
    *reinterpret_cast<uint4*>(smem + i) = make_uint4(0u, 0u, 0u, 0u);
    *reinterpret_cast<uint4*>(smem + kWpZeroHi + i) = make_uint4(0u, 0u, 0u, 0u);
  }
  fence_proxy_async();
  if (warp == 1) tmem_alloc<512>(&tmem_base_s);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  pdl_trigger();
  const uint32_t tmem_base = tmem_base_s;
  const bool has_work = g1 > g0;
  uint8_t* const ring = smem + kWpRing0;

  if (warp == 0) {
    // ------------------------------------------------ producer ------------------------------------
    if (elect_one() && has_work) {
      uint32_t q = 0;                                      // global stage counter
      for (int gid = g0; gid < g1;) {
        const WpSpan sp = wp_span(p, gid, g1);
        gid = sp.next;
        const int chunk = sp.group % p.chunks, n_tile = sp.group / p.chunks;
        const int nst = (sp.h1 - sp.h0) + 2;
        for (int j = 0; j < nst; ++j, ++q) {
          const int st = q % kWpStages;
          MBAR_WAIT_OR_TRAP(&bar_empty[st], ((q / kWpStages) & 1) ^ 1);
          uint8_t* sx = ring + (size_t)st * kWpStage;
          const bool with_dy = j >= 2;
          const bool mirror = with_dy && st == 0;          // the dy tile of stage 0 also lives behind stage 7
          mbar_arrive_expect_tx(&bar_full[st], 66 * 128 + (with_dy ? (mirror ? 2 : 1) * kWpDy : 0));
          tma_load_4d(sx, &map_x, &bar_full[st], chunk * 64, sp.w0 - 1, sp.h0 - 1 + j, sp.n);
          if (with_dy) {
            tma_load_4d(sx + kWpStrip, &map_dy, &bar_full[st], n_tile * 64, sp.w0, sp.h0 + j - 2, sp.n);
            if (mirror)
              tma_load_4d(ring + (size_t)kWpStages * kWpStage + kWpStrip, &map_dy, &bar_full[st], n_tile * 64, sp.w0,
                          sp.h0 + j - 2, sp.n);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer ----------------------------------
    // One elected lane; shared-window addresses computed once (see tap_gemm_kernel in conv_tc.cu).
    if (has_work && elect_one()) {
      const uint32_t idesc = make_idesc_bf16(128, 192, 1, 1);
      const uint32_t zlo = hold(smem_u32(smem)), zhi = zlo + kWpZeroHi;
      const uint32_t rbase = zlo + kWpRing0;
      const uint32_t a_full = hold(smem_u32(bar_full)), a_empty = hold(smem_u32(bar_empty));
      const uint32_t a_done = hold(smem_u32(&bar_done)), a_drained = hold(smem_u32(&bar_drained));
      const uint32_t d1 = hold(tmem_base), d2 = d1 + 256;
      constexpr uint32_t HI = desc_hi(1024, LAYOUT_SW128);
      uint32_t q = 0;                                       // ring position of stage 0 of the current span
      uint32_t first1 = 0, first2 = 0;                      // 0: the next G1 / G2 overwrites its accumulator
      int cur_group = g0 / p.pieces;
      uint32_t seg_no = 0;
      for (int gid = g0; gid < g1;) {
        const WpSpan sp = wp_span(p, gid, g1);
        gid = sp.next;
        if (sp.group != cur_group) {                        // group boundary: hand the accumulators over
          umma_commit_a(a_done);
          mbar_wait_a(a_drained, seg_no & 1);               // ... and wait until they have been read out
          tc_fence_after();
          ++seg_no;
          cur_group = sp.group;
          first1 = first2 = 0;
        }
        const int rows = sp.h1 - sp.h0;
        const int nst = rows + 2;
        int loaded = 0;
        for (int i = 0; i <= rows + 1; ++i) {
          const int need = min(i + 2, nst - 1);
          while (loaded <= need) {
            const uint32_t qq = q + loaded;
            mbar_wait_a(a_full + 8 * (qq % kWpStages), (qq / kWpStages) & 1);
            ++loaded;
          }
          tc_fence_after();
          const uint32_t s_i = (q + i) % kWpStages;
          const uint32_t st_i = rbase + s_i * kWpStage;                   // strip(s) and dy(s-1)
          const uint32_t b_lo = desc_lo(st_i, 128);
          if (i >= 1) {                                     // G1: [dy(s-1) ; dy(s)] -> taps dh = +1, 0
            uint32_t a0, lbo;
            if (i == 1) {                                   // dy(s-1) is not part of the span: zeros below the ring
              a0 = zlo;
              lbo = rbase + ((q + 2) % kWpStages) * kWpStage + kWpStrip - zlo;
            } else if (i == rows + 1) {                     // dy(s) is not part of the span: zeros above the ring
              a0 = st_i + kWpStrip;
              lbo = zhi - a0;
            } else {                                        // next stage, or its mirror behind stage 7
              a0 = st_i + kWpStrip;
              lbo = kWpStage;
            }
            const uint32_t a_lo = desc_lo(a0, lbo);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16_lh(d1, a_lo + k4 * 128, HI, b_lo + k4 * 128, HI, idesc, k4 == 0 ? first1 : 1u);
            first1 = 1u;
          }
          if (i <= rows - 1) {                              // G2: [dy(s+1) ; (ignored)] -> tap dh = -1
            const uint32_t a0 = rbase + ((q + i + 2) % kWpStages) * kWpStage + kWpStrip;
            const uint32_t a_lo = desc_lo(a0, 1024);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16_lh(d2, a_lo + k4 * 128, HI, b_lo + k4 * 128, HI, idesc, k4 == 0 ? first2 : 1u);
            first2 = 1u;
          }
          umma_commit_a(a_empty + 8 * s_i);                 // stage i (its strip, and dy(s-1)) is no longer needed
        }
        q += nst;
      }
      umma_commit_a(a_done);
    }
  } else if (has_work) {
    // ------------------------------------------------ epilogue (once per group touched) -----------
    const int rb = warp & 3;
    const int row = rb * 32 + lane;                         // accumulator row: (j, co)
    const int jrow = row >> 6, co_l = row & 63;
    const size_t co_stride = (size_t)9 * p.Cin;
    const int grp_first = g0 / p.pieces, grp_last = (g1 - 1) / p.pieces;
    uint32_t seg_no = 0;
    for (int group = grp_first; group <= grp_last; ++group, ++seg_no) {
      const int chunk = group % p.chunks, n_tile = group / p.chunks;
      const int co = n_tile * 64 + co_l;
      const int ci0 = chunk * 64;
      MBAR_WAIT_OR_TRAP(&bar_done, seg_no & 1);
      tc_fence_after();
#pragma unroll 1
      for (int a = 0; a < 2; ++a) {                         // a = 0: D1 (filter rows 2, 1); a = 1: D2 (filter row 0)
        if (a == 1 && jrow) continue;                       // upper 64 lanes of D2 hold nothing
        const int r = a == 0 ? (jrow == 0 ? 2 : 1) : 0;
        const uint32_t t_addr = tmem_base + a * 256 + (static_cast<uint32_t>(rb * 32) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < 192; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_addr + c0, v);
          tmem_ld_wait();
          const int s = c0 >> 6;                            // horizontal tap
          const int ci = ci0 + (c0 & 63);
          if (co >= p.Cout) continue;
          float* dst = p.dw + (size_t)co * co_stride + (size_t)(r * 3 + s) * p.Cin + ci;
          if (ci + 32 <= p.Cin) {
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (p.mc_delta != 0)
                grad_red_add4_mc(dst + 4 * g, __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
                                 __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]), p.mc_delta);
              else
                red_add_v4(dst + 4 * g, __uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1]),
                           __uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3]));
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (ci + j < p.Cin) grad_red_add(dst + j, __uint_as_float(v[j]), p.mc_delta);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&bar_drained);                            // the MMA thread may overwrite the accumulators
    }
  }
  grad_arm_publish(p.arm);                 // all-reduce trigger (csrc/common.cuh)
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

}  // namespace b200

using namespace b200;

extern "C" int b200conv_wgrad_pair_supported(int N, int H, int W, int Cin, int Cout) {
  (void)N; (void)H;
  return (W >= 64 && Cin % 8 == 0 && Cout % 8 == 0) ? 1 : 0;
}

extern "C" int b200conv_wgrad_pair(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw,
                                   int N, int H, int W, int Cin, int Cout, cudaStream_t st) {
  if (!b200conv_wgrad_pair_supported(N, H, W, Cin, Cout)) return (int)cudaErrorInvalidValue;
  WgradPairParams p{};
  p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
  p.chunks = (Cin + 63) / 64;
  p.n_tiles = (Cout + 63) / 64;
  p.groups = p.chunks * p.n_tiles;
  p.wsegs = (W + 63) / 64;
  p.rows_per_piece = H < 8 ? H : 8;                      // granularity of the split between CTAs only
  p.pieces_per_col = (H + p.rows_per_piece - 1) / p.rows_per_piece;
  p.pieces = N * p.wsegs * p.pieces_per_col;
  const long long total = (long long)p.groups * p.pieces;
  if (total > 0x7fffffffll) return (int)cudaErrorInvalidValue;
  const int grid = total < kNumSMs ? (int)total : kNumSMs;
  p.dw = dw;
  p.mc_delta = b200seg_grad_multicast_delta(dw);
  p.arm = static_cast<const GradArm*>(b200seg_take_grad_arm());
  CUtensorMap mx, mdy;
  {
    long long dims[4] = {Cin, W, H, N}, str[4] = {1, ld_x, ld_x * W, ld_x * W * H};
    int box[4] = {64, 66, 1, 1};
    if (!make_map(&mx, x, 4, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  }
  {
    long long dims[4] = {Cout, W, H, N}, str[4] = {1, ld_dy, ld_dy * W, ld_dy * W * H};
    int box[4] = {64, 64, 1, 1};
    if (!make_map(&mdy, dy, 4, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  }
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWpSmem);
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  return (int)launch_k(wgrad_pair_kernel, grid, kWpThreads, kWpSmem, st, mx, mdy, p);
}

// Points this translation unit's error record (umma.cuh mbar_timeout_trap) at mapped host memory.
extern "C" int b200conv_set_error_record_pair(unsigned int* host_mapped) {
  return (int)cudaMemcpyToSymbol(b200::g_error_record, &host_mapped, sizeof(host_mapped));
}
