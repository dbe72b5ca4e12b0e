// Weight gradient of the 3x3 conv, halo-strip version (W >= 64):
//     dW[co][(r,s)][ci] += sum_p dy[p][co] * x[p + (r-1, s-1)][ci]                 
//
// The plain split-K kernel in conv_tc.cu re-reads the input once per filter tap (9x) from L2 and is
// bandwidth-bound for the wide, shallow layers that dominate the U-Net.  Here one CTA owns ONE
// 64-channel chunk of Cin and ONE 64-channel tile of Cout and marches down image columns:
//
//   * a K step is 64 consecutive pixels of one image row.  The producer loads ONE new input "strip"
//     per step -- 66 pixels (one halo pixel each side) x 64 channels, 128-byte swizzled -- and the
//     matching 64 x 64 dy tile.  The three strips of rows h-1, h, h+1 live in a ring, so every input
//     byte is fetched once per (Cin chunk, Cout tile) instead of nine times.
//   * all nine taps are UMMA operands that point INTO those strips: tap (r,s) is the window of 64
//     pixel-rows starting s rows into strip r.  tcgen05 applies the 128-byte swizzle as a function
//     of the absolute shared-memory address, so a window that starts on any 128-byte row is a valid
//     MN-major operand (probed on hardware, csrc/standalone_umma_probe.cu exp 5).  Two taps form one
//     M = 128 tile; the descriptor's leading byte offset is the distance between them:
//         tiles 0..2 : {(r,0),(r,1)}   same strip, one row apart (LBO = 128 B)
//         tile  3    : {(0,2),(1,2)}   consecutive strips (LBO = one ring stage)
//         tile  4    : {(2,2), unused} (64-column variant only)
//     Tile 3 needs strip r+1 exactly one stage after strip r, which fails where the ring wraps; so
//     whatever is loaded into slot 0 is ALSO loaded into a mirror slot after the last one (1/8 more
//     strip traffic) and slot 7 + one stage is always the right partner.
//   * BN = 64 output channels per CTA: five accumulators of 64 columns, all nine taps (9/10 of the
//     issued MMA rows are useful; with N = 64 the MMA is shared-memory-bound at 2/3 of peak).
//     BN = 128 (Cout % 128 == 0): four accumulators of 128 columns = all of TMEM, taps 0..7 at the
//     full tensor-core rate with no padding; tap (2,2) is left to the split-K kernel (1/9 of the work).
//   * the accumulators persist across all the pixels the CTA visits; the epilogue runs once and adds
//     the 9 x 64 x 64 block into the fp32 gradient buffer with red.global.add (the buffer accumulates
//     over micro-batches anyway, R:756).
// Out-of-image strips (row -1, row H, the halo columns) are zero-filled by TMA = the conv padding.
#include <cstdlib>

#include "umma.cuh"
#include "../common.cuh"
#include "b200seg.h"
#include "tma_host.h"

namespace b200 {

constexpr int kWsThreads = 192;
constexpr int kWsStages = 8;
constexpr uint32_t kStripBytes = 9216;       // 66 rows x 128 B = 8448, padded to 9 KiB (72 rows)
template <int BN> struct WsCfg {
  static constexpr int kTiles = BN == 64 ? 5 : 4;
  static constexpr uint32_t kDyBytes = BN * 128;               // 64 pixels x BN channels
  static constexpr uint32_t kStageBytes = kStripBytes + kDyBytes;
  static constexpr size_t kSmem = (size_t)kWsStages * kStageBytes + kStripBytes + 1024;   // + mirror slot
};

struct WgradStripParams {
  int N, H, W, Cin, Cout;
  int chunks, n_tiles, groups;         // Cin/64 (rounded up), Cout/BN (rounded up), chunks * n_tiles
  int wsegs, rows_per_piece, pieces_per_col, pieces;   // pieces per group
  float* dw;
  long long mc_delta;                  // != 0: add through the multicast alias (fused all-reduce)
  const GradArm* arm;                  // != NULL: this launch completes a gradient bucket (all-reduce trigger)
};

// Work distribution ("stream-K" over pixels): the (group, piece) pairs -- group = (Cin chunk, Cout tile),
// piece = rows_per_piece rows of one 64-pixel column -- are numbered group-major and the grid's CTAs
// take equal contiguous ranges.  A CTA therefore usually works for one group, sometimes two or three;
// at a group boundary it flushes its accumulators (epilogue) and starts over.  Vertically adjacent
// pieces of a range are walked as ONE span, so the two halo rows are paid once per span, not per piece.
struct WsSpan { int group, n, w0, h0, h1, next; };
__device__ __forceinline__ WsSpan ws_span(const WgradStripParams& p, int gid, int g1) {
  WsSpan s;
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

template <int BN>
__global__ void __launch_bounds__(kWsThreads, 1)
wgrad_strip_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
                   const WgradStripParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar_full[kWsStages], bar_empty[kWsStages];
  __shared__ __align__(8) uint64_t bar_done, bar_drained;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr uint32_t kDyBytes = WsCfg<BN>::kDyBytes, kWsStageBytes = WsCfg<BN>::kStageBytes;
  constexpr int kTiles = WsCfg<BN>::kTiles;

  const long long total = (long long)p.groups * p.pieces;
  const int g0 = (int)(total * blockIdx.x / gridDim.x);
  const int g1 = (int)(total * (blockIdx.x + 1) / gridDim.x);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_dy);
    for (int s = 0; s < kWsStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_done, 1);
    mbar_init(&bar_drained, 128);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(&tmem_base_s);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  pdl_trigger();
  const uint32_t tmem_base = tmem_base_s;
  const bool has_work = g1 > g0;

  if (warp == 0) {
    // ------------------------------------------------ producer ------------------------------------
    if (elect_one() && has_work) {
      uint32_t q = 0;                                      // global stage counter
      for (int gid = g0; gid < g1;) {
        const WsSpan sp = ws_span(p, gid, g1);
        gid = sp.next;
        const int chunk = sp.group % p.chunks, n_tile = sp.group / p.chunks;
        // stage j of the span carries input row (h0 - 1 + j); from j = 2 on it also carries the dy
        // tile of output row (h0 + j - 2).
        const int nst = (sp.h1 - sp.h0) + 2;
        for (int j = 0; j < nst; ++j, ++q) {
          const int st = q % kWsStages;
          MBAR_WAIT_OR_TRAP(&bar_empty[st], ((q / kWsStages) & 1) ^ 1);
          uint8_t* sx = smem + (size_t)st * kWsStageBytes;
          const bool with_dy = j >= 2;
          const bool mirror = st == 0;                     // slot 0 is duplicated behind the last slot
          mbar_arrive_expect_tx(&bar_full[st], (mirror ? 2 : 1) * 66 * 128 + (with_dy ? kDyBytes : 0));
          tma_load_4d(sx, &map_x, &bar_full[st], chunk * 64, sp.w0 - 1, sp.h0 - 1 + j, sp.n);
          if (mirror)
            tma_load_4d(smem + (size_t)kWsStages * kWsStageBytes, &map_x, &bar_full[st], chunk * 64, sp.w0 - 1,
                        sp.h0 - 1 + j, sp.n);
          if (with_dy) {
#pragma unroll
            for (int nb = 0; nb < BN / 64; ++nb)
              tma_load_4d(sx + kStripBytes + nb * 8192, &map_dy, &bar_full[st], n_tile * BN + nb * 64, sp.w0,
                          sp.h0 + j - 2, sp.n);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------ MMA issuer ----------------------------------
    // One elected lane; shared-window addresses computed once (see tap_gemm_kernel).
    if (has_work && elect_one()) {
      constexpr bool issuer = true;
      const uint32_t idesc = make_idesc_bf16(128, BN, 1, 1);
      const uint32_t sbase = hold(smem_u32(smem));
      const uint32_t a_full = hold(smem_u32(bar_full)), a_empty = hold(smem_u32(bar_empty));
      const uint32_t a_done = hold(smem_u32(&bar_done)), a_drained = hold(smem_u32(&bar_drained));
      const uint32_t tmem_b = hold(tmem_base);
      constexpr uint32_t HI = desc_hi(1024, LAYOUT_SW128);
      uint32_t q = 0;
      // ring positions of the strips holding filter rows 0, 1, 2 of the current step
      uint32_t st0 = kWsStages - 2, st1 = kWsStages - 1, st2 = 0, ph2 = 0;
      uint32_t first_step = 0;                              // 0: the next step overwrites the accumulators
      int cur_group = g0 / p.pieces;
      uint32_t seg_no = 0;
      for (int gid = g0; gid < g1;) {
        const WsSpan sp = ws_span(p, gid, g1);
        gid = sp.next;
        if (sp.group != cur_group) {                        // group boundary: hand the accumulators over
          if (issuer) umma_commit_a(a_done);
          mbar_wait_a(a_drained, seg_no & 1);               // ... and wait until they have been read out
          tc_fence_after();
          ++seg_no;
          cur_group = sp.group;
          first_step = 0;
        }
        const int nst = (sp.h1 - sp.h0) + 2;
        for (int j = 0; j < nst; ++j, ++q) {
          mbar_wait_a(a_full + 8 * st2, ph2);
          if (j >= 2) {                                     // j < 2: halo rows of the span, loaded only
            tc_fence_after();
            const uint32_t strip0 = sbase + st0 * kWsStageBytes;   // filter row 0
            const uint32_t strip1 = sbase + st1 * kWsStageBytes;
            const uint32_t strip2 = sbase + st2 * kWsStageBytes;
            const uint32_t b_lo = desc_lo(strip2 + kStripBytes, 8192);
            const uint32_t a_t[5] = {desc_lo(strip0, 128), desc_lo(strip1, 128), desc_lo(strip2, 128),
                                     desc_lo(strip0 + 256, kWsStageBytes),      // + mirror slot across the wrap
                                     desc_lo(strip2 + 256, 128)};
            if (issuer) {
              // k-step outer, accumulator inner: consecutive UMMAs never hit the same accumulator (a chain
              // of dependent N = 64 instructions is latency-bound, not throughput-bound)
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4) {
#pragma unroll
                for (int t = 0; t < kTiles; ++t)
                  umma_bf16_lh(tmem_b + t * BN, a_t[t] + k4 * 128, HI, b_lo + k4 * 128, HI, idesc,
                               k4 == 0 ? first_step : 1u);
              }
              umma_commit_a(a_empty + 8 * st0);             // oldest strip is no longer needed
              if (j == nst - 1) {                           // end of the span: free the other two
                umma_commit_a(a_empty + 8 * st1);
                umma_commit_a(a_empty + 8 * st2);
              }
            }
              first_step = 1u;
          }
          st0 = st1; st1 = st2;
          if (++st2 == kWsStages) { st2 = 0; ph2 ^= 1; }
        }
      }
      if (issuer) umma_commit_a(a_done);
    }
  } else if (has_work) {
    // ------------------------------------------------ epilogue (once per group touched) -----------
    const int rb = warp & 3;
    const int row = rb * 32 + lane;                         // accumulator row: tap half + channel
    const size_t co_stride = (size_t)9 * p.Cin;
    const int grp_first = g0 / p.pieces, grp_last = (g1 - 1) / p.pieces;
    uint32_t seg_no = 0;
    for (int group = grp_first; group <= grp_last; ++group, ++seg_no) {
      const int chunk = group % p.chunks, n_tile = group / p.chunks;
      const int ci = chunk * 64 + (row & 63);
      MBAR_WAIT_OR_TRAP(&bar_done, seg_no & 1);
      tc_fence_after();
#pragma unroll 1
      for (int t = 0; t < kTiles; ++t) {
        const int hi = row >> 6;                            // which tap of the tile (warp-uniform)
        const int r = t < 3 ? t : (t == 3 ? hi : 2);
        const int s = t < 3 ? hi : 2;
        if (t == 4 && hi) continue;                         // padding half of the last tile
        const bool row_ok = ci < p.Cin;
        float* base = p.dw + (size_t)(r * 3 + s) * p.Cin + ci;
        const uint32_t t_addr = tmem_base + t * BN + (static_cast<uint32_t>(rb * 32) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_addr + c0, v);
          tmem_ld_wait();
          if (p.mc_delta != 0) {                            // fused all-reduce: 16-byte multicast reductions
            quad_transpose32(v, lane);
            float* qbase = base - (lane & 3);               // first of the quad's four channels
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const int co = n_tile * BN + c0 + 4 * g + (lane & 3);
              if (row_ok && co < p.Cout)
                grad_red_add4_mc(qbase + (size_t)co * co_stride, __uint_as_float(v[4 * g]),
                                 __uint_as_float(v[4 * g + 1]), __uint_as_float(v[4 * g + 2]),
                                 __uint_as_float(v[4 * g + 3]), p.mc_delta);
            }
            continue;
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int co = n_tile * BN + c0 + j;
            if (row_ok && co < p.Cout) atomicAdd(base + (size_t)co * co_stride, __uint_as_float(v[j]));
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&bar_drained);                            // the MMA thread may overwrite the accumulators
    }
  }
  grad_arm_publish(p.arm);               // all-reduce trigger (csrc/common.cuh)
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

}  // namespace b200

using namespace b200;

// Supported whenever a row holds at least one 64-pixel strip.  `preferred`: rows narrower than 64
// pixels (and channel counts the split-K kernel tiles better) stay on the split-K kernel.
// B200SEG_WGRAD_BN=64 forces the 64-column variant (A/B runs).
extern "C" int b200conv_wgrad_strip_supported(int N, int H, int W, int Cin, int Cout) {
  (void)N; (void)H;
  return (W >= 64 && Cin % 8 == 0 && Cout % 8 == 0) ? 1 : 0;
}
// 128-column variant: taps 0-7 at the full MMA rate, but tap (2,2) goes to the split-K kernel, which is
// memory-bound for a single tap.  Measured per layer (profiles/conv_layers_*.txt) it pays when the rows
// are long and the input is at least as wide as the output (W >= 128, Cin >= Cout); otherwise the
// 64-column variant (all nine taps, five tiles) or, for Cout > 128, the split-K kernel is faster.
// B200SEG_WGRAD_BN=64|128 forces one.
static int strip_bn(int W, int Cin, int Cout) {
  static int force = -1;
  if (force < 0) {
    const char* e = getenv("B200SEG_WGRAD_BN");
    force = (e == nullptr) ? 0 : (e[0] == '6' ? 64 : 128);
  }
  if (Cout % 128 != 0 || force == 64) return 64;
  if (force == 128) return 128;
  return (W >= 128 && Cin >= Cout) ? 128 : 64;
}
// Wide outputs that do not qualify for the 128-column variant stay on the split-K kernel (256-column
// tiles beat five 64-column ones there).
extern "C" int b200conv_wgrad_strip_preferred(int N, int H, int W, int Cin, int Cout) {
  if (!b200conv_wgrad_strip_supported(N, H, W, Cin, Cout)) return 0;
  return (Cout <= 128 || strip_bn(W, Cin, Cout) == 128) ? 1 : 0;
}

extern "C" int b200conv_wgrad_taps(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                                   int H, int W, int Cin, int Cout, int tap0, int ntaps, cudaStream_t st);

template <int BN>
static int launch_strip(const CUtensorMap& mx, const CUtensorMap& mdy, const WgradStripParams& p, int grid,
                        cudaStream_t st) {
  static bool attr_done = false;
  const size_t smem = WsCfg<BN>::kSmem;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_strip_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  return (int)launch_k(wgrad_strip_kernel<BN>, grid, kWsThreads, smem, st, mx, mdy, p);
}

extern "C" int b200conv_wgrad_pair(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw,
                                   int N, int H, int W, int Cin, int Cout, cudaStream_t st);

extern "C" int b200conv_wgrad_strip(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw,
                                    int N, int H, int W, int Cin, int Cout, cudaStream_t st) {
  if (!b200conv_wgrad_strip_supported(N, H, W, Cin, Cout)) return (int)cudaErrorInvalidValue;
  const int bn = strip_bn(W, Cin, Cout);
  // 64-column tiles: the row-pair kernel (wgrad_pair.cu) issues N = 192 instructions where this one
  // issues N = 64 ones, which hold the tensor pipe just as long as N = 128.  B200SEG_WGRAD_PAIR=0 keeps
  // the old kernel (A/B runs).
  static int pair_env = -1;
  if (pair_env < 0) { const char* e = getenv("B200SEG_WGRAD_PAIR"); pair_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
  if (bn == 64 && pair_env) return b200conv_wgrad_pair(x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, st);
  WgradStripParams p{};
  p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
  p.chunks = (Cin + 63) / 64;
  p.n_tiles = (Cout + bn - 1) / bn;
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
  // the bucket trigger belongs to the LAST launch of this call: here, or the tap-(2,2) split-K launch below
  p.arm = nullptr;
  if (bn == 64) {
    p.arm = static_cast<const GradArm*>(b200seg_take_grad_arm());
    return launch_strip<64>(mx, mdy, p, grid, st);
  }
  int rc = launch_strip<128>(mx, mdy, p, grid, st);
  if (rc != 0) return rc;
  return b200conv_wgrad_taps(x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, 8, 1, st);   // tap (2,2)
}

// Points this translation unit's error record (umma.cuh mbar_timeout_trap) at mapped host memory.
extern "C" int b200conv_set_error_record_strip(unsigned int* host_mapped) {
  return (int)cudaMemcpyToSymbol(b200::g_error_record, &host_mapped, sizeof(host_mapped));
}
