// Implicit-GEMM convolutions on the 5th-generation tensor cores (tcgen05.mma, TMEM accumulators,
// TMA-fed shared memory), bf16 NHWC, fp32 accumulation.
//
// What this replaces: the reference's nn.Conv2d(k=3, pad=1) forward (R:579, R:582) and its
// autograd data-gradient, and nn.ConvTranspose2d(k=2, s=2) forward (R:607) -- cuDNN calls made by
// PyTorch eager.  One kernel template, the "tap GEMM":
//
//     D[p][n] = sum_{t < taps} sum_{k < Kc}  A[p + off(t)][k] * B_t[n][k]      (+ bias[n])
//
//   fprop   : A = x,  taps = 9, off = (r-1, s-1), B_t[n][k] = W[n][t][k]   (K-major  B operand)
//   dgrad   : A = dy, taps = 9, off = (1-r, 1-s), B_t[n][k] = W[k][t][n]   (MN-major B operand: the
//             SAME bf16 weight buffer, read un-transposed -- no transposed copy is ever made)
//   convT2x2: A = x,  taps = 1, B[n=(q,co)][k] = Wt[q][co][k]; the epilogue scatters column block q
//             to output pixel (2y+dy, 2x+dx) of the concat buffer's first channels (removes
//             torch.cat, R:616).
//
// Structure (persistent, 1 CTA / SM, 192 threads):
//   warp 0      TMA producer: per (tap, 64-channel chunk) one 4-D box of the activation (128 pixels
//               = TN images x TH rows x TW columns; out-of-image coordinates are zero-filled by TMA,
//               which IS the conv padding) + the weight tile, into a ring of smem stages.
//   warp 1      allocates TMEM, then a single elected thread issues tcgen05.mma (128 x BLOCK_N x 16)
//               over the ring; tcgen05.commit releases smem stages and publishes the accumulator.
//   warps 2-5   epilogue: tcgen05.ld accumulator -> registers, + bias, -> bf16 -> swizzled smem
//               staging tile -> TMA store (clips at the image edge, honours the channel-slice
//               stride).  For fprop the BatchNorm batch statistics (sum, sum of squares of the
//               rounded outputs) are accumulated from the staging tile, so BN costs no extra pass
//               over the conv output.
//   TMEM holds two accumulators (2 x 256 columns) so the epilogue of tile i overlaps the MMAs of
//   tile i+1.
#include "umma.cuh"
#include "../common.cuh"
#include "b200seg.h"
#include "tma_host.h"
#include <cstdlib>

namespace b200 {

constexpr int kConvThreads = 192;     // wgrad kernels: producer, MMA, 4 epilogue warps
constexpr int kTapThreads = 320;      // tap GEMM: producer, MMA, 8 epilogue warps (two per TMEM lane quarter)
constexpr int kEpiThreads = 256;
constexpr int kMaxStages = 8;
constexpr int kMaxGroups = 8;
constexpr int kClusterSize = 1;       // 1 | 2 (experiment, slower: profiles/variants_in_situ.txt)
constexpr int kMaxXSlots = 16;        // strip ring: TMA latency (~1.5 us) needs several tiles of strips in flight
constexpr uint32_t kSmemBudget = 222 * 1024;

struct TapGemmParams {
  int tiles_w, tiles_h, tiles_g, m_tiles;
  int tw_log2, th_log2, tn;        // TW = 1 << tw_log2, TH = 1 << th_log2, TN images per tile
  int H, W, N;
  int taps, flip;
  int kc_per_tap;                  // K chunks (of BLOCK_K channels) per tap
  int kc_total;                    // channels per tap (K-major B: column stride between taps)
  int n_tiles, block_n, n_total;
  int b_tap_stride;                // MN-major B: column offset between taps
  int b_row_tap_stride;            // MN-major B: row (K) offset between taps (convT dgrad)
  int a_mode;                      // 0: 4-D activation box shifted per tap; 1: 5-D 2x2 quadrant gather
  int ocb, groups;                 // output channel box (64/32/16) and groups per N tile
  int out_mode;                    // 0: NHWC store, 1: convT 2x2 scatter
  int convt_cout;
  int stages;
  uint32_t a_bytes, b_bytes, stage_bytes, staging_bytes;
  int staging_bufs;                // 1 or 2 output staging tiles (2: TMA store of tile i overlaps tile i+1)
  int cluster;                     // 1, or 2: CTA pairs (adjacent M tiles) share the weight tile via TMA multicast
  int rows2;                       // STRIP mode: a work unit is TWO vertically adjacent row tiles sharing every weight tile
  int rows4;                       // STRIP mode, resident bank, N = 64: FOUR-row units, vertical taps merged along N
  int reg_stats;                   // N = 64: BN statistics accumulated in registers (B200SEG_REG_STATS=0 disables)
  int epi_private;                 // STRIP, N = 64: warp-private epilogue pipelines (no CTA-wide barriers), stores through map_o2
  int b_resident;                  // STRIP mode: all 9 * kc weight tiles stay in smem for the whole kernel
  int xslots;                      // STRIP mode: number of input-strip slots
  uint32_t xslot_bytes;            // STRIP mode: bytes per strip slot (130 rows x 128 B, padded)
  const float* bias;
  float* partials;
};

template <int BLOCK_K>
struct KCfg {
  static constexpr uint32_t kRowBytes = BLOCK_K * 2;
  static constexpr uint32_t kLayout = BLOCK_K == 64 ? LAYOUT_SW128 : (BLOCK_K == 32 ? LAYOUT_SW64 : LAYOUT_SW32);
  static constexpr uint32_t kSBO = 8 * kRowBytes;
};

// 16-byte shared-memory load through the shared window (a generic pointer makes the compiler emit four
// 4-byte loads, which conflict 4-way on the swizzled staging tile)
__device__ __forceinline__ bf16x8 lds128(uint32_t saddr) {
  bf16x8 v;
  uint32_t* w = reinterpret_cast<uint32_t*>(&v);
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(saddr));
  return v;
}

__device__ __forceinline__ void named_bar_sync(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}

// STRIP = true (3x3 only, rows >= 128 pixels wide, BLOCK_K = 64): the M tile is 128 consecutive pixels
// of ONE image row and the activation is loaded as 130-pixel strips (one halo pixel each side), one
// strip per (filter row, channel chunk).  The three taps of a filter row are UMMA operand windows
// that start 0, 1 or 2 rows into the strip (the swizzle is a function of the absolute smem address,
// so any 128-byte row is a legal start) -> activations are fetched 3x instead of 9x from L2.  Strips
// and weight tiles travel in separate rings.
template <int BLOCK_K, bool B_MN, bool STRIP>
__global__ void __launch_bounds__(kTapThreads, 1)
tap_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const __grid_constant__ CUtensorMap map_o, const __grid_constant__ CUtensorMap map_o2,
                const TapGemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + (size_t)p.stages * p.stage_bytes + (STRIP ? (size_t)p.xslots * p.xslot_bytes : 0);
  uint8_t* const xring = smem + (size_t)p.stages * p.stage_bytes;     // STRIP: strip slots after the B ring
  __shared__ __align__(8) uint64_t bar_full[kMaxStages], bar_empty[kMaxStages];
  __shared__ __align__(8) uint64_t bar_xfull[kMaxXSlots], bar_xempty[kMaxXSlots];
  __shared__ __align__(8) uint64_t bar_tfull[4], bar_tempty[4];
  __shared__ uint32_t tmem_base_s;
  __shared__ __align__(16) float s_bias[256];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Scheduling unit = cluster (1 or 2 CTAs).  A cluster walks "units" of `csz` adjacent M tiles with one
  // N tile; CTA `crank` takes tile unit*csz + crank.  Every CTA of a cluster runs the same number of
  // iterations (a CTA past the end re-computes the last tile and stores nothing) so that the shared
  // weight-tile pipeline stays in lock step.
  // Compile-time switch: 2 enables the CTA-pair weight multicast experiment (see run_tap_gemm).  As a
  // run-time value the extra branches cost the MN-major (dgrad) producer 12-30 % (A/B in
  // profiles/variants_in_situ.txt), so it is a constant.
  constexpr int csz = kClusterSize;
  const int crank = (csz == 2) ? (int)cluster_ctarank() : 0;
  const uint16_t cmask = (uint16_t)((1u << csz) - 1);
  const int unit0 = blockIdx.x / csz;
  const int n_tile = unit0 % p.n_tiles;
  const int m_first = unit0 / p.n_tiles;
  const int m_step = (gridDim.x / csz) / p.n_tiles;
  const int m_units = (p.m_tiles + csz - 1) / csz;
  const int prow = m_first * csz + crank;                 // this CTA's row of the statistics partials
  const int TW = 1 << p.tw_log2, TH = 1 << p.th_log2;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
    tma_prefetch_desc(&map_o);
    tma_prefetch_desc(&map_o2);
    for (int s = 0; s < p.stages && s < kMaxStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], csz); }
    for (int s = 0; s < 4; ++s) { mbar_init(&bar_tfull[s], 1); mbar_init(&bar_tempty[s], kEpiThreads); }
    for (int s = 0; s < kMaxXSlots; ++s) { mbar_init(&bar_xfull[s], 1); mbar_init(&bar_xempty[s], 1); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(&tmem_base_s);
  pdl_wait();                 // everything above overlapped the previous kernel's tail (see common.cuh)
  pdl_trigger();
  if (p.bias != nullptr) {
    for (int c = threadIdx.x; c < p.block_n; c += kTapThreads) {
      const int col = (blockIdx.x % p.n_tiles) * p.block_n + c;
      s_bias[c] = (col < p.n_total) ? p.bias[p.out_mode == 1 ? (col % p.convt_cout) : col] : 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (csz == 2) cluster_sync_all();                       // peers' barriers exist before anything multicasts
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  const int k_iters = p.taps * p.kc_per_tap;
  // weight-tile loader: a lone CTA fetches the whole tile; in a pair each CTA fetches half and
  // multicasts it into both (the full barrier of each CTA still expects the whole tile)
  auto load_b = [&](uint8_t* sb, uint64_t* bar, int t, int kc) {
    if (!B_MN) {
      const int half = p.block_n / csz;
      if (csz == 1) tma_load_2d(sb, &map_b, bar, t * p.kc_total + kc * BLOCK_K, n_tile * p.block_n);
      else tma_load_2d_mc(sb + (size_t)crank * half * (BLOCK_K * 2), &map_b, bar, t * p.kc_total + kc * BLOCK_K,
                          n_tile * p.block_n + crank * half, cmask);
    } else {
      const int nblocks = (p.block_n + 63) >> 6;
      for (int nb = 0; nb < nblocks; ++nb) {
        const int c0 = t * p.b_tap_stride + n_tile * p.block_n + nb * 64;
        const int c1 = t * p.b_row_tap_stride + kc * BLOCK_K;
        if (csz == 1) tma_load_2d(sb + nb * (BLOCK_K * 128), &map_b, bar, c0, c1);
        else if ((nb & 1) == crank || (nblocks == 1 && crank == 0))
          tma_load_2d_mc(sb + nb * (BLOCK_K * 128), &map_b, bar, c0, c1, cmask);
      }
    }
  };
  auto release_b = [&](uint64_t* bar) {
    if (csz == 1) umma_commit(bar); else umma_commit_mc(bar, cmask);
  };

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (STRIP && elect_one()) {
      int stage = 0, xs = 0;
      uint32_t phase = 0, xphase = 0;
      if (p.b_resident) {                                   // the whole filter bank, once
        mbar_arrive_expect_tx(&bar_full[0], 9 * p.kc_per_tap * p.b_bytes);
        for (int t = 0; t < 9; ++t)
          for (int kc = 0; kc < p.kc_per_tap; ++kc) {
            // rows4: the three filter rows of one horizontal tap sit next to each other, in the order
            // in which their output rows ascend (fprop: r = 2,1,0; dgrad: r = 0,1,2), so that a stack
            // of them is ONE N = 128 / 192 operand
            const int r = t / 3, sft = t % 3;
            const int slot = p.rows4 ? (sft * 3 + (p.flip > 0 ? 2 - r : r)) : (t * p.kc_per_tap + kc);
            load_b(smem + (size_t)slot * p.stage_bytes, &bar_full[0], t, kc);
          }
      }
      if (p.rows4) {
        // work unit = output rows h0 .. h0+3 of one 128-pixel column segment: six input strips
        const int hq_n = p.H >> 2;
        for (int q = m_first; q < (p.m_tiles >> 2); q += m_step) {
          const int twi = q % p.tiles_w;
          const int hq = (q / p.tiles_w) % hq_n;
          const int n0 = q / (p.tiles_w * hq_n);
          const int w0 = twi << p.tw_log2, h0 = hq * 4;
          for (int j = 0; j < 6; ++j) {
            MBAR_WAIT_OR_TRAP(&bar_xempty[xs], xphase ^ 1);
            mbar_arrive_expect_tx(&bar_xfull[xs], 130 * KCfg<BLOCK_K>::kRowBytes);
            tma_load_4d(xring + (size_t)xs * p.xslot_bytes, &map_a, &bar_xfull[xs], 0, w0 - 1, h0 - 1 + j, n0);
            if (++xs == p.xslots) { xs = 0; xphase ^= 1; }
          }
        }
      } else if (p.rows2) {
        // work unit = output rows (h0, h0+1) of one 128-pixel column segment: 4 input strips per K chunk
        // (rows h0-1 .. h0+2) and every weight tile ONCE for both rows
        const int hp_n = p.H >> 1;
        for (int q = m_first; q < (p.m_tiles >> 1); q += m_step) {
          const int twi = q % p.tiles_w;
          const int hp = (q / p.tiles_w) % hp_n;
          const int n0 = q / (p.tiles_w * hp_n);
          const int w0 = twi << p.tw_log2, h0 = hp * 2;
          for (int kc = 0; kc < p.kc_per_tap; ++kc) {
            for (int j = 0; j < 4; ++j) {
              MBAR_WAIT_OR_TRAP(&bar_xempty[xs], xphase ^ 1);
              mbar_arrive_expect_tx(&bar_xfull[xs], 130 * KCfg<BLOCK_K>::kRowBytes);
              tma_load_4d(xring + (size_t)xs * p.xslot_bytes, &map_a, &bar_xfull[xs], kc * BLOCK_K, w0 - 1,
                          h0 - 1 + j, n0);
              if (++xs == p.xslots) { xs = 0; xphase ^= 1; }
            }
            for (int t = 0; t < 9; ++t) {
              MBAR_WAIT_OR_TRAP(&bar_empty[stage], phase ^ 1);
              mbar_arrive_expect_tx(&bar_full[stage], p.b_bytes);
              load_b(smem + (size_t)stage * p.stage_bytes, &bar_full[stage], t, kc);
              if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      } else
      for (int mu = m_first; mu < m_units; mu += m_step) {
        const int m = min(mu * csz + crank, p.m_tiles - 1);
        const int twi = m % p.tiles_w;
        const int thi = (m / p.tiles_w) % p.tiles_h;
        const int g = m / (p.tiles_w * p.tiles_h);
        const int w0 = twi << p.tw_log2, h0 = thi, n0 = g;
        for (int kc = 0; kc < p.kc_per_tap; ++kc) {
          for (int r = 0; r < 3; ++r) {
            MBAR_WAIT_OR_TRAP(&bar_xempty[xs], xphase ^ 1);
            mbar_arrive_expect_tx(&bar_xfull[xs], 130 * KCfg<BLOCK_K>::kRowBytes);
            tma_load_4d(xring + (size_t)xs * p.xslot_bytes, &map_a, &bar_xfull[xs], kc * BLOCK_K, w0 - 1,
                        h0 + p.flip * (r - 1), n0);
            if (++xs == p.xslots) { xs = 0; xphase ^= 1; }
            if (p.b_resident) continue;
            for (int sft = 0; sft < 3; ++sft) {
              const int t = r * 3 + sft;
              MBAR_WAIT_OR_TRAP(&bar_empty[stage], phase ^ 1);
              mbar_arrive_expect_tx(&bar_full[stage], p.b_bytes);
              load_b(smem + (size_t)stage * p.stage_bytes, &bar_full[stage], t, kc);
              if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    } else if (!STRIP && elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int mu = m_first; mu < m_units; mu += m_step) {
        const int m = min(mu * csz + crank, p.m_tiles - 1);
        const int twi = m % p.tiles_w;
        const int thi = (m / p.tiles_w) % p.tiles_h;
        const int g = m / (p.tiles_w * p.tiles_h);
        const int w0 = twi << p.tw_log2, h0 = thi << p.th_log2, n0 = g * p.tn;
        for (int t = 0; t < p.taps; ++t) {
          const int dh = (p.taps == 9) ? p.flip * (t / 3 - 1) : 0;
          const int dw = (p.taps == 9) ? p.flip * (t % 3 - 1) : 0;
          for (int kc = 0; kc < p.kc_per_tap; ++kc) {
            MBAR_WAIT_OR_TRAP(&bar_empty[stage], phase ^ 1);
            uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
            uint8_t* sb = sa + p.a_bytes;
            mbar_arrive_expect_tx(&bar_full[stage], p.a_bytes + p.b_bytes);
            if (p.a_mode == 0) tma_load_4d(sa, &map_a, &bar_full[stage], kc * BLOCK_K, w0 + dw, h0 + dh, n0);
            else tma_load_5d(sa, &map_a, &bar_full[stage], kc * BLOCK_K, t & 1, w0, t >> 1, n0 * p.H + h0);
            load_b(sb, &bar_full[stage], t, kc);
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer ========================================
    // ONE elected lane runs the loops (ptxas then knows the stream is single-threaded and keeps barrier
    // addresses, descriptors and counters in uniform registers; a lane-0 predicate inside a warp-wide
    // loop made it wrap every UMMA in an ELECT / R2UR.BROADCAST waterfall instead).  All shared-window
    // addresses are computed once, here.
    if (elect_one()) {
    constexpr bool issuer = true;
    const uint32_t a_full = hold(smem_u32(bar_full)), a_empty = hold(smem_u32(bar_empty));
    const uint32_t a_xfull = hold(smem_u32(bar_xfull)), a_xempty = hold(smem_u32(bar_xempty));
    const uint32_t a_tfull = hold(smem_u32(bar_tfull)), a_tempty = hold(smem_u32(bar_tempty));
    const uint32_t a_ring = hold(smem_u32(smem)), a_xring = hold(smem_u32(xring));
    const uint32_t stage_bytes = hold(p.stage_bytes), xslot_bytes = hold(p.xslot_bytes);
    const int n_stages = (int)hold((uint32_t)p.stages), n_xslots = (int)hold((uint32_t)p.xslots);
    const int kc_per_tap = (int)hold((uint32_t)p.kc_per_tap);
    const bool resident = hold((uint32_t)p.b_resident) != 0;
    const uint32_t tmem_b = hold(tmem_base);
    constexpr uint32_t kRowB = KCfg<BLOCK_K>::kRowBytes;
    constexpr uint32_t A_HI = desc_hi(KCfg<BLOCK_K>::kSBO, KCfg<BLOCK_K>::kLayout);
    constexpr uint32_t B_HI = B_MN ? desc_hi(1024, LAYOUT_SW128) : A_HI;
    constexpr uint32_t B_LBO = B_MN ? BLOCK_K * 128 : 16;
    constexpr uint32_t A_K4 = 32 >> 4, B_K4 = B_MN ? (2048 >> 4) : (32 >> 4);
    constexpr int K4N = BLOCK_K / 16;
    auto release_a = [&](uint32_t bar_addr) {
      if (csz == 1) umma_commit_a(bar_addr); else umma_commit_mc_a(bar_addr, cmask);
    };
    if (STRIP) {
      const uint32_t idesc = make_idesc_bf16(128, p.block_n, 0, B_MN ? 1 : 0);
      int stage = 0, xs = 0;
      uint32_t phase = 0, xphase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      const uint32_t win0 = (p.flip > 0) ? 0u : 2u * kRowB;              // operand window of sft = 0 ...
      const uint32_t win_step = (p.flip > 0) ? kRowB : (0u - kRowB);     // ... and the step to the next tap
      if (resident) mbar_wait_a(a_full, 0);
      if (p.rows4) {
        // Four output rows per unit, 64 columns each, side by side in one 256-column accumulator.  Input
        // strip j (image row h0-1+j) feeds output rows j-2 .. j through the three filter rows; because the
        // weights of those filter rows are adjacent in shared memory, ONE instruction with N = 128 / 192
        // does all of them: the activation window (the operand that made N = 64 shared-memory-bound) is
        // read once per 2-3 output rows.  The row that a strip touches first (row j, filter row 0 in
        // fprop) needs accumulate = 0, so the first horizontal tap issues it separately.
        const uint32_t idesc_n[4] = {0u, make_idesc_bf16(128, 64, 0, B_MN ? 1 : 0), make_idesc_bf16(128, 128, 0, B_MN ? 1 : 0),
                                     make_idesc_bf16(128, 192, 0, B_MN ? 1 : 0)};
        for (int q = m_first; q < (p.m_tiles >> 2); q += m_step) {
          mbar_wait_a(a_tempty + 8 * acc, acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d0 = tmem_b + acc * 256;
#pragma unroll
          for (int j = 0; j < 6; ++j) {
            mbar_wait_a(a_xfull + 8 * xs, xphase);
            tc_fence_after();
            const uint32_t sx = a_xring + xs * xslot_bytes;
            const int sub_min = j > 2 ? j - 2 : 0, sub_max = j < 3 ? j : 3;
            const int nb = sub_max - sub_min + 1;
            const int first_blk = j < 2 ? 2 - j : 0;
            const bool has_new = j <= 3;                   // output row j gets its first contribution here
            if (issuer) {
#pragma unroll
              for (int sft = 0; sft < 3; ++sft) {
                const uint32_t a_lo = desc_lo(sx + win0 + sft * win_step, 16);
                const uint32_t b_lo = desc_lo(a_ring + (sft * 3 + first_blk) * stage_bytes, B_LBO);
                if (sft == 0 && has_new) {
                  const int nb_old = nb - 1;
                  if (nb_old > 0) {
#pragma unroll
                    for (int k4 = 0; k4 < K4N; ++k4)
                      umma_bf16_lh(d0 + sub_min * 64, a_lo + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc_n[nb_old], 1u);
                  }
                  const uint32_t bn_lo = b_lo + ((nb_old * stage_bytes) >> 4);
#pragma unroll
                  for (int k4 = 0; k4 < K4N; ++k4)
                    umma_bf16_lh(d0 + j * 64, a_lo + k4 * A_K4, A_HI, bn_lo + k4 * B_K4, B_HI, idesc_n[1], k4 == 0 ? 0u : 1u);
                } else {
#pragma unroll
                  for (int k4 = 0; k4 < K4N; ++k4)
                    umma_bf16_lh(d0 + sub_min * 64, a_lo + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc_n[nb], 1u);
                }
              }
              umma_commit_a(a_xempty + 8 * xs);
            }
              if (++xs == n_xslots) { xs = 0; xphase ^= 1; }
          }
          if (issuer) umma_commit_a(a_tfull + 8 * acc);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      } else if (p.rows2) {
        for (int q = m_first; q < (p.m_tiles >> 1); q += m_step) {
          // accumulator slots acc, acc+1 (of 4, 128 columns each) hold output rows h0 and h0+1
          mbar_wait_a(a_tempty + 8 * acc, acc_phase ^ 1);
          mbar_wait_a(a_tempty + 8 * (acc + 1), acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d0 = tmem_b + acc * 128, d1 = d0 + 128;
          for (int kc = 0; kc < kc_per_tap; ++kc) {
            uint32_t sxs[4];
            int xs0 = xs;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              mbar_wait_a(a_xfull + 8 * xs, xphase);
              sxs[j] = a_xring + xs * xslot_bytes;
              if (++xs == n_xslots) { xs = 0; xphase ^= 1; }
            }
            tc_fence_after();
#pragma unroll
            for (int t = 0; t < 9; ++t) {
              const int r = t / 3, sft = t % 3;
              mbar_wait_a(a_full + 8 * stage, phase);
              tc_fence_after();
              // input strip of output row `sub` for filter row r: sub + r (fprop) or sub + 2 - r (dgrad)
              const uint32_t woff = win0 + sft * win_step;
              const uint32_t s0 = (p.flip > 0) ? sxs[r] : sxs[2 - r];
              const uint32_t s1 = (p.flip > 0) ? sxs[r + 1] : sxs[3 - r];
              const uint32_t a0 = desc_lo(s0 + woff, 16), a1 = desc_lo(s1 + woff, 16);
              const uint32_t b_lo = desc_lo(a_ring + stage * stage_bytes, B_LBO);
              const uint32_t accf = (kc | t) != 0 ? 1u : 0u;
              if (issuer) {
#pragma unroll
                for (int k4 = 0; k4 < K4N; ++k4) {
                  umma_bf16_lh(d0, a0 + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc, k4 == 0 ? accf : 1u);
                  umma_bf16_lh(d1, a1 + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc, k4 == 0 ? accf : 1u);
                }
                release_a(a_empty + 8 * stage);
              }
                  if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
            if (issuer) {
#pragma unroll
              for (int j = 0; j < 4; ++j) umma_commit_a(a_xempty + 8 * (xs0 + j));   // four slots never straddle the ring end
            }
            }
          if (issuer) {
            umma_commit_a(a_tfull + 8 * acc);
            umma_commit_a(a_tfull + 8 * (acc + 1));
          }
          acc += 2;
          if (acc == 4) { acc = 0; acc_phase ^= 1; }
        }
      } else {
        const int kr_n = kc_per_tap * 3;
        const uint32_t tap_pitch = kc_per_tap * stage_bytes;          // resident bank: distance between taps
        for (int mu = m_first; mu < m_units; mu += m_step) {
          mbar_wait_a(a_tempty + 8 * acc, acc_phase ^ 1);
          tc_fence_after();
          const uint32_t d_tmem = tmem_b + acc * 256;
          int kc_ = 0, r_ = 0;
          for (int kr = 0; kr < kr_n; ++kr) {
            mbar_wait_a(a_xfull + 8 * xs, xphase);
            tc_fence_after();
            const uint32_t sx = a_xring + xs * xslot_bytes;
            const uint32_t bank0 = a_ring + (r_ * 3 * kc_per_tap + kc_) * stage_bytes;
#pragma unroll
            for (int sft = 0; sft < 3; ++sft) {
              uint32_t sb = bank0 + sft * tap_pitch;
              if (!resident) {
                mbar_wait_a(a_full + 8 * stage, phase);
                tc_fence_after();
                sb = a_ring + stage * stage_bytes;
              }
              // window of 128 pixel rows starting (1 + flip*(sft-1)) rows into the strip
              const uint32_t a_lo = desc_lo(sx + win0 + sft * win_step, 16);
              const uint32_t b_lo = desc_lo(sb, B_LBO);
              const uint32_t accf = (kr | sft) != 0 ? 1u : 0u;
              if (issuer) {
#pragma unroll
                for (int k4 = 0; k4 < K4N; ++k4)
                  umma_bf16_lh(d_tmem, a_lo + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc, k4 == 0 ? accf : 1u);
                if (!resident) release_a(a_empty + 8 * stage);
              }
                  if (!resident) { if (++stage == n_stages) { stage = 0; phase ^= 1; } }
            }
            if (issuer) umma_commit_a(a_xempty + 8 * xs);     // all three taps of this strip are issued
              if (++xs == n_xslots) { xs = 0; xphase ^= 1; }
            if (++r_ == 3) { r_ = 0; ++kc_; }
          }
          if (issuer) umma_commit_a(a_tfull + 8 * acc);
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    } else {
      const uint32_t idesc = make_idesc_bf16(128, p.block_n, 0, B_MN ? 1 : 0);
      const uint32_t a_bytes = p.a_bytes;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int mu = m_first; mu < m_units; mu += m_step) {
        mbar_wait_a(a_tempty + 8 * acc, acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_b + acc * 256;
        for (int it = 0; it < k_iters; ++it) {
          mbar_wait_a(a_full + 8 * stage, phase);
          tc_fence_after();
          const uint32_t sa = a_ring + stage * stage_bytes;
          const uint32_t a_lo = desc_lo(sa, 16), b_lo = desc_lo(sa + a_bytes, B_LBO);
          const uint32_t accf = it != 0 ? 1u : 0u;
          if (issuer) {
#pragma unroll
            for (int k4 = 0; k4 < K4N; ++k4)
              umma_bf16_lh(d_tmem, a_lo + k4 * A_K4, A_HI, b_lo + k4 * B_K4, B_HI, idesc, k4 == 0 ? accf : 1u);
            release_a(a_empty + 8 * stage);                 // frees the smem stage (in both CTAs of a pair)
          }
          if (++stage == n_stages) { stage = 0; phase ^= 1; }
        }
        if (issuer) umma_commit_a(a_tfull + 8 * acc);       // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
    }   // elect_one
  } else {
    // ===================================== epilogue ==========================================
    if (STRIP && p.epi_private) {
      // Warp-private pipelines for the strip layers with 64-channel column groups.  ncu showed the shared
      // epilogue below taking ~2300 cycles per 128 x 64 group, most of it in its two CTA-wide barriers and in all
      // 256 threads waiting for one thread's TMA store (profiles/ncu_mma_issue.txt).  Here warp (q, half) owns
      // rows 32q..32q+31 x columns 32half..32half+31 of every 64-column group of every tile: it reads its TMEM
      // block, packs it into its OWN 2 KB slab (two slabs, double-buffered) and its lane 0 issues its own
      // 32-pixel x 32-channel TMA store.  Nothing is shared between warps until the statistics are folded at the
      // very end.  fprop: one group (N = 64), sum and sum of squares in registers.  dgrad: up to three groups
      // (N = 64 / 128 / 192), column sums only (the transposed conv's bias gradient; the sum-of-squares row of
      // the partials is written as zero).
      constexpr int kMaxGrp = B_MN ? 3 : 1;
      const int q = warp & 3, half = (warp - 2) >> 2;
      uint8_t* const slab0 = staging + (size_t)(warp - 2) * 4096;
      const uint32_t slab_a = smem_u32(slab0) + lane * 64;
      const uint32_t a_tfull = smem_u32(bar_tfull);
      const bool stats = p.partials != nullptr;
      const bool has_bias = p.bias != nullptr;
      const int ngrp = p.block_n >> 6;
      float rs_s[kMaxGrp][32], rs_q[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        rs_q[j] = 0.f;
#pragma unroll
        for (int g = 0; g < kMaxGrp; ++g) rs_s[g][j] = 0.f;
      }
      int acc = 0, buf = 0;
      uint32_t acc_phase = 0;
      const int acc_n = p.rows2 ? 4 : 2;
      const int acc_cols = p.rows2 ? 128 : 256;
      const int subs = p.rows4 ? 4 : (p.rows2 ? 2 : 1);
      const int sh = p.rows4 ? 2 : (p.rows2 ? 1 : 0);
      const int n_units = p.m_tiles >> sh;
      const int hp_n = p.H >> sh;
      const int ch0 = half * 32;
      for (int mu = m_first; mu < n_units; mu += m_step) {
        // tile coordinates, off the critical path (needed by lane 0 only; computed before the wait)
        const int w0 = ((mu % p.tiles_w) << p.tw_log2) + q * 32;
        const int hb = ((mu / p.tiles_w) % hp_n) << sh;
        const int n0 = mu / (p.tiles_w * hp_n);
#pragma unroll 1
        for (int sub = 0; sub < subs; ++sub) {
          if (!p.rows4 || sub == 0) {
            mbar_wait_a(a_tfull + 8 * acc, acc_phase);
            tc_fence_after();
          }
          const uint32_t t_base = tmem_base + acc * acc_cols + (p.rows4 ? sub * 64 : 0) +
                                  (static_cast<uint32_t>(q * 32) << 16) + ch0;
#pragma unroll
          for (int g = 0; g < kMaxGrp; ++g) {
            if (g < ngrp) {
              uint32_t r[32];
              tmem_ld_32x32b_x32(t_base + g * 64, r);
              tmem_ld_wait();
              if (g == ngrp - 1) {
                tc_fence_before();
                if (!p.rows4 || sub == 3) {
                  mbar_arrive(&bar_tempty[acc]);            // accumulator drained (it is in registers now)
                  if (++acc == acc_n) { acc = 0; acc_phase ^= 1; }
                }
              }
              if (has_bias) {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + s_bias[g * 64 + ch0 + j]);
              }
              if (stats) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                  const float v = __uint_as_float(r[j]);
                  rs_s[g][j] += v;
                  if (!B_MN) rs_q[j] = fmaf(v, v, rs_q[j]);
                }
              }
              const uint32_t dst = slab_a + buf * 2048;
#pragma unroll
              for (int v = 0; v < 4; ++v) {
                float f[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[v * 8 + j]);
                const bf16x8 pk = pack8(f);
                const uint32_t* pw = reinterpret_cast<const uint32_t*>(&pk);
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + v * 16), "r"(pw[0]), "r"(pw[1]),
                             "r"(pw[2]), "r"(pw[3])
                             : "memory");
              }
              fence_proxy_async();                          // slab writes -> visible to TMA
              __syncwarp();
              if (lane == 0) {
                tma_store_4d(&map_o2, slab0 + buf * 2048, n_tile * p.block_n + g * 64 + ch0, w0, hb + sub, n0);
                tma_store_commit();
                tma_store_wait_read<1>();                   // the slab written NEXT (stored one block ago) is free
              }
              __syncwarp();
              buf ^= 1;
            }
          }
        }
      }
      if (lane == 0) tma_store_wait_read<0>();
      if (stats) {
        // fold the 128 rows: butterfly inside each warp, then the four warps of a column half through smem
        named_bar_sync(1, kEpiThreads);                     // every slab is free (they are the scratch below)
        float* red = reinterpret_cast<float*>(staging);    // [group][8 warps][2][32]
#pragma unroll
        for (int g = 0; g < kMaxGrp; ++g) {
          if (g < ngrp) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
#pragma unroll
              for (int o = 16; o > 0; o >>= 1) {
                rs_s[g][j] += __shfl_xor_sync(0xffffffffu, rs_s[g][j], o);
                if (!B_MN) rs_q[j] += __shfl_xor_sync(0xffffffffu, rs_q[j], o);
              }
            }
            if (lane == 0) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                red[((g * 8 + (warp - 2)) * 2 + 0) * 32 + j] = rs_s[g][j];
                red[((g * 8 + (warp - 2)) * 2 + 1) * 32 + j] = B_MN ? 0.f : rs_q[j];
              }
            }
          }
        }
        named_bar_sync(1, kEpiThreads);
        const int et = threadIdx.x - 64;
        if (et < ngrp * 64) {                               // column et: group et / 64, half (et % 64) / 32
          const int g = et >> 6, hf = (et & 63) >> 5, j = et & 31;
          float sv = 0.f, qv = 0.f;
          for (int wq = 0; wq < 4; ++wq) {
            sv += red[((g * 8 + hf * 4 + wq) * 2 + 0) * 32 + j];
            qv += red[((g * 8 + hf * 4 + wq) * 2 + 1) * 32 + j];
          }
          const int col = n_tile * p.block_n + et;
          if (col < p.n_total) {
            p.partials[((size_t)prow * 2 + 0) * p.n_total + col] = sv;
            p.partials[((size_t)prow * 2 + 1) * p.n_total + col] = qv;
          }
        }
      }
      if (lane == 0) tma_store_wait_all<0>();
    } else {
    // The epilogue is instruction-issue bound for the narrow (N = 64) layers, so everything that does
    // not depend on the tile is hoisted: per-thread staging addresses, the statistics addresses, and
    // the tile-coordinate divisions (only the one thread that issues the TMA store needs them unless
    // the layer has ragged tiles).
    const int rb = warp & 3;                              // TMEM lane quarter this warp may read
    const int row = rb * 32 + lane;                       // tile row (pixel) owned by this thread
    const int et = threadIdx.x - 64;                      // 0..255 within the epilogue group
    const int colhalf = (warp - 2) >> 2;                  // two warps share a lane quarter: even / odd 32-column blocks
    const int ocb = p.ocb;
    const int ocb_log2 = (ocb == 64) ? 6 : (ocb == 32 ? 5 : 4);
    const uint32_t row_bytes = ocb * 2;
    const uint32_t swz_mask = (row_bytes >> 4) - 1;       // 16-byte chunks per row - 1
    const uint32_t group_bytes = 128 * row_bytes;
    // swizzle<B,4,3>: chunk index XOR (row-derived constant); the XOR term only depends on the row
    const uint32_t my_row_off = row * row_bytes;
    const uint32_t my_rx = (my_row_off >> 7) & swz_mask;
    // statistics: this thread owns ONE 16-byte chunk column (8 channels) of the N tile and walks every
    // st_rstep-th of the 128 rows; when block_n / 8 does not divide the thread count (block_n = 192)
    // the last few threads sit out.
    const int cpr = p.block_n >> 3;                       // chunks per row
    const int st_cc = et % cpr;                           // my chunk column
    const int st_rg = et / cpr;                           // my row group
    const int st_rstep = kEpiThreads / cpr;
    const bool st_active = st_rg < st_rstep;
    const int st_rows = st_active ? (128 - st_rg + st_rstep - 1) / st_rstep : 0;   // rows this thread sums
    const int st_grp = (st_cc * 8) >> ocb_log2;
    const uint32_t st_chunk = ((st_cc * 8) & (ocb - 1)) >> 3;
    float st_s[8], st_q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) st_s[j] = st_q[j] = 0.f;
    const bool has_bias = p.bias != nullptr;
    const bool ragged = ((p.W & (TW - 1)) | (p.H & (TH - 1))) != 0 || (p.N % p.tn) != 0;
    const int block_n = p.block_n;
    // N = 64 (where this epilogue, not the MMAs, bounds the kernel): a thread always meets the same 32
    // channels, so it keeps their sum / sum of squares in registers across ALL its tiles (2 instructions
    // per element, on the fp32 accumulator values) instead of re-reading the staging tile (3 + a shared
    // load); the rows are folded once, at the end of the kernel.
    const bool reg_stats = p.partials != nullptr && p.reg_stats && block_n == 64 && !ragged && csz == 1;
    const bool want_stats = p.partials != nullptr && !reg_stats;
    float rs_s[32], rs_q[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) rs_s[j] = rs_q[j] = 0.f;
    int acc = 0;
    uint32_t acc_phase = 0;
    int sbuf = 0;
    uint8_t* const staging0 = staging;
    const int acc_n = p.rows2 ? 4 : 2;                    // accumulator slots in TMEM ...
    const int acc_cols = p.rows2 ? 128 : 256;             // ... and their column pitch
    const int subs = p.rows4 ? 4 : (p.rows2 ? 2 : 1);     // output rows (tiles) per work unit
    const int n_units = p.rows4 ? (p.m_tiles >> 2) : (p.rows2 ? (p.m_tiles >> 1) : m_units);
    for (int mu = m_first; mu < n_units; mu += m_step)
    for (int sub = 0; sub < subs; ++sub) {
      const bool live = (p.rows2 | p.rows4) ? true : (mu * csz + crank < p.m_tiles);
      uint8_t* const staging = staging0 + (size_t)sbuf * p.staging_bytes;     // this tile's staging buffer
      if (p.staging_bufs == 2) sbuf ^= 1;
      // rows4: the unit's four rows share ONE accumulator slot (64 columns each)
      if (!p.rows4 || sub == 0) {
        MBAR_WAIT_OR_TRAP(&bar_tfull[acc], acc_phase);
        tc_fence_after();
      }
      const uint32_t t_addr = tmem_base + acc * acc_cols + (p.rows4 ? sub * 64 : 0) +
                              (static_cast<uint32_t>(rb * 32) << 16);
      for (int c0 = colhalf * 32; c0 < block_n; c0 += 64) {
        uint32_t r[32];
        const bool wide = block_n - c0 >= 32;
        if (wide) {
          tmem_ld_32x32b_x32(t_addr + c0, r);
        } else {
          uint32_t r16[16];
          tmem_ld_32x32b_x16(t_addr + c0, r16);
#pragma unroll
          for (int j = 0; j < 16; ++j) { r[j] = r16[j]; r[j + 16] = 0u; }
        }
        tmem_ld_wait();
        if (reg_stats) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float v = __uint_as_float(r[j]) + (has_bias ? s_bias[c0 + j] : 0.f);
            rs_s[j] += v;
            rs_q[j] = fmaf(v, v, rs_q[j]);
          }
        }
        // the 32 columns lie inside one group when ocb >= 32; for ocb == 16 they span two groups
        const uint32_t gbase = smem_u32(staging + (c0 >> ocb_log2) * group_bytes + my_row_off);
        const uint32_t ch0 = (c0 & (ocb - 1)) >> 3;
#pragma unroll
        for (int v = 0; v < 4; ++v) {                     // four 16-byte chunks of 8 channels
          if (wide || v < 2) {
            float f[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(r[v * 8 + j]);
            if (has_bias) {                               // two 16-byte shared loads per 8 channels
              const float4 b0 = *reinterpret_cast<const float4*>(s_bias + c0 + v * 8);
              const float4 b1 = *reinterpret_cast<const float4*>(s_bias + c0 + v * 8 + 4);
              f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
              f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
            }
            uint32_t chunk = ch0 + v;
            uint32_t gb = gbase;
            if (ocb == 16 && v >= 2) { chunk -= 2; gb += group_bytes; }
            // one st.shared.v4 (through a generic pointer the compiler emitted four 4-byte stores with
            // 4-way bank conflicts: 37 % of the kernel's shared-memory wavefronts were excess)
            const bf16x8 pk = pack8(f);
            const uint32_t* pw = reinterpret_cast<const uint32_t*>(&pk);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(gb + ((chunk ^ my_rx) << 4)), "r"(pw[0]),
                         "r"(pw[1]), "r"(pw[2]), "r"(pw[3])
                         : "memory");
          }
        }
      }
      tc_fence_before();
      if (!p.rows4 || sub == 3) {
        mbar_arrive(&bar_tempty[acc]);                    // accumulator drained: MMA may reuse it
        if (++acc == acc_n) { acc = 0; acc_phase ^= 1; }
      }
      fence_proxy_async();                                // staging writes -> visible to TMA
      named_bar_sync(1, kEpiThreads);
      int w0 = 0, h0 = 0, n0 = 0;
      if (et == 0 || (ragged && want_stats)) {            // tile coordinates: integer divisions, off the hot path
        if (p.rows2 | p.rows4) {
          const int sh = p.rows4 ? 2 : 1;
          const int hp_n = p.H >> sh;
          w0 = (mu % p.tiles_w) << p.tw_log2;
          h0 = (((mu / p.tiles_w) % hp_n) << sh) + sub;
          n0 = mu / (p.tiles_w * hp_n);
        } else {
          const int m = min(mu * csz + crank, p.m_tiles - 1);
          const int twi = m % p.tiles_w;
          const int thi = (m / p.tiles_w) % p.tiles_h;
          const int g_img = m / (p.tiles_w * p.tiles_h);
          w0 = twi << p.tw_log2; h0 = thi << p.th_log2; n0 = g_img * p.tn;
        }
      }
      if (want_stats && live) {
        const uint32_t gb = smem_u32(staging + st_grp * group_bytes);
        if (!ragged) {
#pragma unroll 4
          for (int i = 0; i < st_rows; ++i) {
            const uint32_t ro = (st_rg + i * st_rstep) * row_bytes;
            float f[8];
            unpack8(lds128(gb + ro + ((st_chunk ^ ((ro >> 7) & swz_mask)) << 4)), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) { st_s[j] += f[j]; st_q[j] = fmaf(f[j], f[j], st_q[j]); }
          }
        } else {
          for (int i = 0; i < st_rows; ++i) {
            const int rr = st_rg + i * st_rstep;
            const int tw_ = rr & (TW - 1), th_ = (rr >> p.tw_log2) & (TH - 1), tn_ = rr >> (p.tw_log2 + p.th_log2);
            const bool valid = (w0 + tw_ < p.W) && (h0 + th_ < p.H) && (n0 + tn_ < p.N);
            const uint32_t ro = rr * row_bytes;
            float f[8];
            unpack8(lds128(gb + ro + ((st_chunk ^ ((ro >> 7) & swz_mask)) << 4)), f);
            if (valid) {
#pragma unroll
              for (int j = 0; j < 8; ++j) { st_s[j] += f[j]; st_q[j] = fmaf(f[j], f[j], st_q[j]); }
            }
          }
        }
      }
      if (et == 0 && live) {
        for (int g = 0; g < p.groups; ++g) {
          const int ch0 = n_tile * block_n + g * ocb;
          if (p.out_mode == 0) {
            tma_store_4d(&map_o, staging + g * group_bytes, ch0, w0, h0, n0);
          } else {
            const int q = ch0 / p.convt_cout, co0 = ch0 % p.convt_cout;
            tma_store_5d(&map_o, staging + g * group_bytes, co0, q & 1, w0, q >> 1, n0 * p.H + h0);
          }
        }
      }
      if (et == 0) {
        tma_store_commit();
        // the buffer written NEXT must be free: with two buffers that is the store issued one tile ago
        if (p.staging_bufs == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>();
      }
      named_bar_sync(1, kEpiThreads);
    }
    if (reg_stats) {
      // fold the 128 rows: butterfly inside each warp, then the four warps of a column half through smem
      if (et == 0) tma_store_wait_read<0>();
      named_bar_sync(1, kEpiThreads);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          rs_s[j] += __shfl_xor_sync(0xffffffffu, rs_s[j], o);
          rs_q[j] += __shfl_xor_sync(0xffffffffu, rs_q[j], o);
        }
      }
      float* red = reinterpret_cast<float*>(staging0);   // [8 warps][2][32]
      if (lane == 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          red[((warp - 2) * 2 + 0) * 32 + j] = rs_s[j];
          red[((warp - 2) * 2 + 1) * 32 + j] = rs_q[j];
        }
      }
      named_bar_sync(1, kEpiThreads);
      if (et < 64) {                                      // channel et: column half et / 32, warps 4*half .. 4*half+3
        const int half = et >> 5, j = et & 31;
        float sv = 0.f, qv = 0.f;
        for (int wq = 0; wq < 4; ++wq) {
          sv += red[((half * 4 + wq) * 2 + 0) * 32 + j];
          qv += red[((half * 4 + wq) * 2 + 1) * 32 + j];
        }
        const int col = n_tile * 64 + et;
        if (col < p.n_total) {
          p.partials[((size_t)prow * 2 + 0) * p.n_total + col] = sv;
          p.partials[((size_t)prow * 2 + 1) * p.n_total + col] = qv;
        }
      }
    } else if (p.partials != nullptr) {
      // combine the (128 / cpr) threads that share a chunk column, then one row of partials per CTA
      if (et == 0) tma_store_wait_read<0>();
      named_bar_sync(1, kEpiThreads);
      float* red = reinterpret_cast<float*>(staging0);   // [256/cpr][2][block_n]
      const int ncol = p.block_n;
      if (st_active) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          red[(st_rg * 2 + 0) * ncol + st_cc * 8 + j] = st_s[j];
          red[(st_rg * 2 + 1) * ncol + st_cc * 8 + j] = st_q[j];
        }
      }
      named_bar_sync(1, kEpiThreads);
      for (int c = et; c < ncol; c += kEpiThreads) {
        float s = 0.f, q = 0.f;
        for (int i = 0; i < st_rstep; ++i) { s += red[(i * 2 + 0) * ncol + c]; q += red[(i * 2 + 1) * ncol + c]; }
        const int col = n_tile * p.block_n + c;
        if (col < p.n_total) {
          p.partials[((size_t)prow * 2 + 0) * p.n_total + col] = s;
          p.partials[((size_t)prow * 2 + 1) * p.n_total + col] = q;
        }
      }
    }
    if (et == 0) tma_store_wait_all<0>();
    }   // shared epilogue
  }
  tc_fence_before();
  __syncthreads();
  if (csz == 2) cluster_sync_all();                       // no CTA leaves while its peer may still write to it
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// =================================================================================================
// Weight gradient:  dW[co][t][ci] += sum_p dy[p][co] * x[p + off(t)][ci]     
// GEMM view: D[(t, ci)][co], M = 9*Cin walked in 64-channel blocks (an M tile of 128 rows pairs two
// blocks, possibly of different taps), N = Cout, K = pixels (64 per stage).  Both operands are
// "MN-major" (the reduction index, the pixel, is the slow one in NHWC memory), which tcgen05 reads
// directly -- no transposes.  K is split across CTAs; partial tiles are added into the fp32
// gradient buffer with red.global.add (the buffer accumulates across micro-batches anyway, R:756).
struct WgradParams {
  int tiles_w, tiles_h, tiles_g, p_tiles;   // pixel tiles of 64
  int tw_log2, th_log2, tn;
  int Cin, Cout;
  int chunks_per_tap;                       // ceil(Cin / 64)
  int m_blocks, m_tiles;                    // 64-row blocks (9 * chunks_per_tap) and 128-row tiles
  int n_tiles, block_n;
  int k_slices;
  int stages;
  uint32_t a_bytes, b_bytes, stage_bytes;
  float* dw;
  int mt;                                   // M tiles (of 128 rows) per CTA sharing one dy tile: 1 or 2
  int convt;                                // 1: dW[(q, co)][ci] of the 2x2 transposed conv
  int nt_per_q;                             // convT: N tiles per quadrant
  int H;                                    // convT: low-res height (Y = n*H + y)
  long long co_stride;                      // floats between consecutive output channels in dw
  int tap0;                                 // first filter tap handled (m_blocks counts taps from here)
  long long mc_delta;                       // != 0: add through the multicast alias (fused all-reduce)
  int quads;                                // convT: output quadrants per CTA (4 = all share one x tile), else 1
  const GradArm* arm;                       // != NULL: this launch completes a gradient bucket (all-reduce trigger)
  int vec_red;                              // 16-byte vector reductions (needs Cin % 4 == 0 and whole quads of rows)
};

// CSZ = 2: CTA pairs (two adjacent M tiles of the same N tile and K slice) share every dy tile: each CTA
// fetches half of its 64-channel blocks and TMA-multicasts them into both, which takes a quarter to a
// third off the L2 traffic of a kernel that is L2-bound on the deep layers (64 B/clk/SM demanded).
// CSZ = 4: a 2 x 2 cluster (two M tiles x two N tiles); the x tile is shared along N the same way, so
// every CTA fetches half of each operand (32 B/clk/SM, below what L2 delivers).  A stage may be
// refilled once the consumers of every CTA that is written into have released it: the MMA thread's
// commit goes to itself and to the two peers it shares an operand with.
template <int CSZ>
__global__ void __launch_bounds__(kConvThreads, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_dy,
             const WgradParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar_full[kMaxStages], bar_empty[kMaxStages];
  __shared__ __align__(8) uint64_t bar_tfull;
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  int id = blockIdx.x;
  int n_tile, m_tile;
  const int crank = (CSZ > 1) ? (int)cluster_ctarank() : 0;
  const int m_bit = crank & 1, n_bit = crank >> 1;
  if (CSZ == 4) {                        // cluster = (M pair) x (N pair)
    id >>= 2;
    const int mp = p.m_tiles >> 1, np = p.n_tiles >> 1;
    m_tile = 2 * (id % mp) + m_bit; id /= mp;
    n_tile = 2 * (id % np) + n_bit; id /= np;
  } else if (CSZ == 2) {                 // the two CTAs of a cluster are adjacent M tiles
    m_tile = id % p.m_tiles; id /= p.m_tiles;
    n_tile = id % p.n_tiles; id /= p.n_tiles;
  } else {
    n_tile = id % p.n_tiles; id /= p.n_tiles;
    m_tile = id % p.m_tiles; id /= p.m_tiles;
  }
  const int ks = id;
  const uint16_t mask_b = (uint16_t)(3u << (2 * n_bit));                 // the M pair of this N tile
  const uint16_t mask_a = (uint16_t)((1u << m_bit) | (1u << (m_bit + 2)));   // the N pair of this M tile
  const uint16_t mask_c = (CSZ == 4) ? (uint16_t)((1u << crank) | (1u << (crank ^ 1)) | (1u << (crank ^ 2)))
                                     : (uint16_t)3;
  const int per = (p.p_tiles + p.k_slices - 1) / p.k_slices;
  const int pt0 = ks * per;
  const int pt1 = min(p.p_tiles, pt0 + per);
  const int k_iters = pt1 - pt0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&map_x);
    tma_prefetch_desc(&map_dy);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], CSZ == 4 ? 3 : CSZ); }
    mbar_init(&bar_tfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(&tmem_base_s);
  tc_fence_before();
  __syncthreads();
  if (CSZ > 1) cluster_sync_all();       // the peers' barriers exist before anything is multicast
  tc_fence_after();
  pdl_wait();                 // everything above overlapped the previous kernel's tail (see common.cuh)
  pdl_trigger();
  const uint32_t tmem_base = tmem_base_s;
  const int nblocks = (p.block_n + 63) >> 6;

  if (k_iters > 0) {
    if (warp == 0) {
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        for (int pt = pt0; pt < pt1; ++pt) {
          const int twi = pt % p.tiles_w;
          const int thi = (pt / p.tiles_w) % p.tiles_h;
          const int g = pt / (p.tiles_w * p.tiles_h);
          const int w0 = twi << p.tw_log2, h0 = thi << p.th_log2, n0 = g * p.tn;
          MBAR_WAIT_OR_TRAP(&bar_empty[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * p.stage_bytes;
          uint8_t* sb = sa + p.a_bytes;
          mbar_arrive_expect_tx(&bar_full[stage], p.a_bytes + p.b_bytes);
          for (int hb = 0; hb < 2 * p.mt; ++hb) {
            int blk = m_tile * 2 * p.mt + hb;
            if (blk >= p.m_blocks) blk = p.m_blocks - 1;     // padding block: loaded, never stored
            const int t = p.tap0 + blk / p.chunks_per_tap, cc = blk % p.chunks_per_tap;
            const int dw_ = p.convt ? 0 : (t % 3 - 1), dh_ = p.convt ? 0 : (t / 3 - 1);
            if (CSZ < 4)
              tma_load_4d(sa + hb * 8192, &map_x, &bar_full[stage], cc * 64, w0 + dw_, h0 + dh_, n0);
            else if ((hb & 1) == n_bit)
              tma_load_4d_mc(sa + hb * 8192, &map_x, &bar_full[stage], cc * 64, w0 + dw_, h0 + dh_, n0, mask_a);
          }
          if (!p.convt) {
            for (int nb = 0; nb < nblocks; ++nb) {
              if (CSZ == 1)
                tma_load_4d(sb + nb * 8192, &map_dy, &bar_full[stage], n_tile * p.block_n + nb * 64, w0, h0, n0);
              else if ((nb & 1) == m_bit)
                tma_load_4d_mc(sb + nb * 8192, &map_dy, &bar_full[stage], n_tile * p.block_n + nb * 64, w0, h0, n0,
                               mask_b);
            }
          } else {
            for (int qi = 0; qi < p.quads; ++qi) {
              const int q = p.quads == 4 ? qi : n_tile / p.nt_per_q;
              const int nt = p.quads == 4 ? n_tile : n_tile % p.nt_per_q;
              for (int nb = 0; nb < nblocks; ++nb) {
                const int bi = qi * nblocks + nb;
                if (CSZ == 1)
                  tma_load_5d(sb + bi * 8192, &map_dy, &bar_full[stage], nt * p.block_n + nb * 64, q & 1, w0, q >> 1,
                              n0 * p.H + h0);
                else if ((bi & 1) == m_bit)
                  tma_load_5d_mc(sb + bi * 8192, &map_dy, &bar_full[stage], nt * p.block_n + nb * 64, q & 1, w0,
                                 q >> 1, n0 * p.H + h0, mask_b);
              }
            }
          }
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      // single elected lane, addresses computed once (see tap_gemm_kernel)
      if (elect_one()) {
      constexpr bool issuer = true;
      const uint32_t a_full = hold(smem_u32(bar_full)), a_empty = hold(smem_u32(bar_empty));
      const uint32_t a_tfull = hold(smem_u32(&bar_tfull));
      const uint32_t a_ring = hold(smem_u32(smem));
      const uint32_t stage_bytes = hold(p.stage_bytes), a_bytes = hold(p.a_bytes);
      const int n_stages = (int)hold((uint32_t)p.stages), mt_n = (int)hold((uint32_t)p.mt);
      const uint32_t tmem_b = hold(tmem_base);
      const bool quads4 = p.quads == 4;
      const uint32_t idesc = make_idesc_bf16(128, p.block_n, 1, 1);
      constexpr uint32_t HI = desc_hi(1024, LAYOUT_SW128);
      const uint32_t q_step = (uint32_t)(nblocks * 8192) >> 4;
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < k_iters; ++it) {
        mbar_wait_a(a_full + 8 * stage, phase);
        tc_fence_after();
        const uint32_t sa = a_ring + stage * stage_bytes;
        const uint32_t b_lo = desc_lo(sa + a_bytes, 8192);
        const uint32_t accf = it != 0 ? 1u : 0u;
        if (issuer) {
          if (quads4) {                                     // convT: the four quadrants share the x tile
            const uint32_t a_lo = desc_lo(sa, 8192);
#pragma unroll
            for (int qi = 0; qi < 4; ++qi) {
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                umma_bf16_lh(tmem_b + qi * 128, a_lo + k4 * 128, HI, b_lo + qi * q_step + k4 * 128, HI, idesc,
                             k4 == 0 ? accf : 1u);
            }
          } else {
            for (int t = 0; t < mt_n; ++t) {                // the M tiles of this CTA share the dy tile
              const uint32_t a_lo = desc_lo(sa + t * 16384, 8192);
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                umma_bf16_lh(tmem_b + t * 256, a_lo + k4 * 128, HI, b_lo + k4 * 128, HI, idesc, k4 == 0 ? accf : 1u);
            }
          }
          if (CSZ == 1) umma_commit_a(a_empty + 8 * stage);
          else umma_commit_mc_a(a_empty + 8 * stage, mask_c);      // frees the stage in every CTA written into
        }
        if (++stage == n_stages) { stage = 0; phase ^= 1; }
      }
      if (issuer) umma_commit_a(a_tfull);
      }   // elect_one
    } else {
      const int rb = warp & 3;
      const int row = rb * 32 + lane;
      MBAR_WAIT_OR_TRAP(&bar_tfull, 0);
      tc_fence_after();
      const size_t co_stride = (size_t)p.co_stride;
      const int n_acc = p.quads == 4 ? 4 : p.mt;          // accumulators: quadrants (128 cols apart) or M tiles (256)
      for (int acc = 0; acc < n_acc; ++acc) {
        const int mt = p.quads == 4 ? 0 : acc;
        const int q = !p.convt ? 0 : (p.quads == 4 ? acc : n_tile / p.nt_per_q);
        const int nt = p.quads == 4 ? n_tile : n_tile % p.nt_per_q;
        const int co_base = p.convt ? q * p.Cout + nt * p.block_n : n_tile * p.block_n;
        const int co_lim = p.convt ? (q + 1) * p.Cout : p.Cout;
        int blk = (m_tile * p.mt + mt) * 2 + (row >> 6);
        const bool blk_ok = blk < p.m_blocks;
        if (!blk_ok) blk = p.m_blocks - 1;
        const int t = p.tap0 + blk / p.chunks_per_tap, cc = blk % p.chunks_per_tap;
        const int ci = cc * 64 + (row & 63);
        const bool row_ok = blk_ok && ci < p.Cin;
        const uint32_t t_addr = tmem_base + (p.quads == 4 ? acc * 128 : mt * 256) + (static_cast<uint32_t>(rb * 32) << 16);
        float* base = p.dw + (p.convt ? (size_t)0 : (size_t)t * p.Cin) + ci;
        for (int c0 = 0; c0 < p.block_n; c0 += 32) {
          uint32_t r[32];
          if (p.block_n - c0 >= 32) {
            tmem_ld_32x32b_x32(t_addr + c0, r);
          } else {
            uint32_t r16[16];
            tmem_ld_32x32b_x16(t_addr + c0, r16);
#pragma unroll
            for (int j = 0; j < 16; ++j) { r[j] = r16[j]; r[j + 16] = 0u; }
          }
          tmem_ld_wait();
          if (p.vec_red) {
            // 16-byte reductions: after the quad transpose a lane holds four CONSECUTIVE input channels of
            // one output channel, i.e. 16 contiguous bytes of dW -> red.global.add.v4.f32 (a quarter of the
            // reduction instructions; the split-K epilogue was atomics-bound on the deep layers), or
            // multimem.red through the multicast alias in fused mode.
            quad_transpose32(r, lane);
            float* qbase = base - (lane & 3);
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              const int cj = c0 + 4 * g + (lane & 3);
              const int co = co_base + cj;
              if (row_ok && cj < p.block_n && co < co_lim) {
                float* dst = qbase + (size_t)co * co_stride;
                if (p.mc_delta != 0)
                  grad_red_add4_mc(dst, __uint_as_float(r[4 * g]), __uint_as_float(r[4 * g + 1]),
                                   __uint_as_float(r[4 * g + 2]), __uint_as_float(r[4 * g + 3]), p.mc_delta);
                else
                  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(__uint_as_float(r[4 * g])),
                               "f"(__uint_as_float(r[4 * g + 1])), "f"(__uint_as_float(r[4 * g + 2])),
                               "f"(__uint_as_float(r[4 * g + 3]))
                               : "memory");
              }
            }
          } else if (row_ok) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int co = co_base + c0 + j;
              if (c0 + j < p.block_n && co < co_lim) atomicAdd(base + (size_t)co * co_stride, __uint_as_float(r[j]));
            }
          }
        }
      }
    }
  }
  grad_arm_publish(p.arm);               // all-reduce trigger: the last CTA tells every rank the bucket is ready
  tc_fence_before();
  __syncthreads();
  if (CSZ > 1) cluster_sync_all();       // no CTA leaves while a peer may still multicast into it
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// =================================================================================================
// Host side

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another
// GPU's memory hung the GPU.  Operands of the cluster launches are activations and weights that the
// engine allocates locally; a pointer that is not plain device memory of the CURRENT device (a peer
// mapping, an IPC or cuMemMap'd range owned by another GPU, host memory) takes the non-multicast kernel.
static bool local_device_memory(const void* p) {
  cudaPointerAttributes a{};
  int dev = -1;
  if (p == nullptr || cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice && a.device == dev;
}

static int ilog2_floor(int x) { int l = 0; while ((2 << l) <= x) ++l; return l; }

struct TileGeom { int tw_log2, th_log2, tn, tiles_w, tiles_h, tiles_g; };

static TileGeom geom_for(int N, int H, int W, int pixels) {
  TileGeom g;
  const int pl = ilog2_floor(pixels);
  g.tw_log2 = ilog2_floor(W < 16 ? W : 16);
  int rem = pl - g.tw_log2;
  g.th_log2 = ilog2_floor(H);
  if (g.th_log2 > rem) g.th_log2 = rem;
  g.tn = 1 << (pl - g.tw_log2 - g.th_log2);
  g.tiles_w = (W + (1 << g.tw_log2) - 1) >> g.tw_log2;
  g.tiles_h = (H + (1 << g.th_log2) - 1) >> g.th_log2;
  g.tiles_g = (N + g.tn - 1) / g.tn;
  return g;
}

static int pick_block_k(int kc) { return kc % 64 == 0 ? 64 : (kc % 32 == 0 ? 32 : (kc % 16 == 0 ? 16 : 0)); }
static int pick_ocb(int n) { return n % 64 == 0 ? 64 : (n % 32 == 0 ? 32 : 16); }
static int pick_block_n(int n_total, int ocb) {
  if (n_total <= 256) return n_total;
  if (ocb == 64) {
    if (n_total % 256 == 0) return 256;
    if (n_total % 128 == 0) return 128;
    return 64;
  }
  return ocb * (256 / ocb >= kMaxGroups ? kMaxGroups : 256 / ocb);
}

template <int BK, bool BMN, bool STRIP = false>
static int launch_tap(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mo, const CUtensorMap& mo2,
                      const TapGemmParams& p, int grid, size_t smem, cudaStream_t st) {
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(tap_gemm_kernel<BK, BMN, STRIP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kSmemBudget + 2048));
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  if (p.cluster == 1) {
    return (int)launch_k(tap_gemm_kernel<BK, BMN, STRIP>, grid, kTapThreads, smem, st, ma, mb, mo, mo2, p);
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kTapThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return (int)cudaLaunchKernelEx(&cfg, tap_gemm_kernel<BK, BMN, STRIP>, ma, mb, mo, mo2, p);
}

static int cluster_env() {             // B200SEG_CLUSTER=1 enables CTA-pair weight multicast (off by default)
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_CLUSTER"); v = (e != nullptr && e[0] == '1') ? 1 : 0; }
  return v;
}

static int rows4_env() {               // four-row N-merged units (N = 64, resident bank): 0 off, 1 dgrad only, 2 also fprop (default)
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_ROWS4"); v = (e == nullptr) ? 2 : (e[0] == '0' ? 0 : (e[0] == '1' ? 1 : 2)); }
  return v;
}

static int rows2_env() {               // B200SEG_ROWS2=0 disables the two-row strip work units
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_ROWS2"); v = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return v;
}

static int strip_mode_env() {          // B200SEG_CONV_STRIP=0 disables the strip path (benchmarking)
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_CONV_STRIP"); v = (e != nullptr && e[0] == '0') ? 0 : 1; }
  return v;
}

// mode: 0 fprop, 1 dgrad, 2 convT forward, 3 convT data gradient
// What run_tap_gemm decided for the last call of this thread (b200conv_plan reads it after a dry run).
static thread_local B200ConvPlan g_last_plan;

static int run_tap_gemm(int mode, const void* a, long long ld_a, const void* w, const float* bias, void* out,
                        long long ld_o, int N, int H, int W, int Kc, int n_total, int conv_cin, float* partials,
                        int* stat_rows_out, bool dry, cudaStream_t st) {
  const int BK = pick_block_k(Kc);
  if (BK == 0 || n_total % 16) return (int)cudaErrorInvalidValue;
  TapGemmParams p{};
  TileGeom g = geom_for(N, H, W, 128);
  // strip mode: 3x3, full 64-channel chunks, rows made of whole 128-pixel segments
  bool strip = (mode == 0 || mode == 1) && (BK == 64 || (BK == 16 && mode == 0)) && (W % 128 == 0) &&
               strip_mode_env();
  if (strip) { g.tw_log2 = 7; g.th_log2 = 0; g.tn = 1; g.tiles_w = W / 128; g.tiles_h = H; g.tiles_g = N; }
  p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h; p.tiles_g = g.tiles_g;
  p.m_tiles = g.tiles_w * g.tiles_h * g.tiles_g;
  p.tw_log2 = g.tw_log2; p.th_log2 = g.th_log2; p.tn = g.tn;
  p.H = H; p.W = W; p.N = N;
  p.taps = (mode == 2) ? 1 : (mode == 3 ? 4 : 9);
  p.flip = (mode == 1) ? -1 : 1;
  p.kc_per_tap = Kc / BK;
  p.kc_total = Kc;
  p.n_total = n_total;
  p.ocb = pick_ocb(mode == 2 ? n_total / 4 : n_total);
  p.block_n = pick_block_n(n_total, p.ocb);
  {
    // Strip mode with streamed weights: a 256-column tile re-reads the whole 9 x Kc x 256 filter slice from L2
    // for every 128-pixel row (0.6 MB per tile at Kc = 128) and cannot use two-row units (TMEM holds two such
    // accumulators, and they are the double buffer).  Two 128-column N tiles with two-row units read each
    // weight byte a quarter as often at the price of reading the strips twice.  B200SEG_STRIP_N=256 keeps 256.
    static int sn_env = -1;
    if (sn_env < 0) { const char* e = getenv("B200SEG_STRIP_N"); sn_env = (e != nullptr && e[0] == '2') ? 256 : 128; }
    if (strip && sn_env == 128 && p.ocb == 64 && n_total % 128 == 0 && n_total > 128 && (H % 2 == 0)) p.block_n = 128;
    // dgrad of a 64-channel-output conv with a wide input (the decoder's first conv, 64 -> 192): three N tiles of
    // 64 each keep their slice of the filter bank resident (9 x 8 KB) and take the four-row N-merged units, like
    // the 64 -> 64 layers; as ONE 192-column tile the 216 KB bank is streamed from L2 for every 128 pixels.
    static int dn_env = -1;
    if (dn_env < 0) { const char* e = getenv("B200SEG_DGRAD_N64"); dn_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    // The same holds for a forward conv with 64 input channels and a wider output (the encoder's 64 -> 128): two
    // resident 64-column N tiles with register statistics instead of one streamed, epilogue-bound 128-column tile.
    if (strip && dn_env && (mode == 0 || mode == 1) && Kc == 64 && p.ocb == 64 && n_total > 64 && n_total <= 256 &&
        (H % 4 == 0))
      p.block_n = 64;
  }
  {
    // Box mode (rows shorter than 128 pixels) with 128- instead of 256-column tiles: twice the N tiles to spread over
    // the SMs, at the price of N = 128 instructions and reading each pixel tile twice as often.  Measured per layer
    // (16 x 64 x 64 ... 16 x 16 x 16, 256 - 1024 channels): 25 - 55 % SLOWER wherever 256-column tiles already fill the
    // machine, faster only for the 16 x 16 bottleneck (32 pixel tiles x 2 N tiles = 64 CTAs -> 128 CTAs: fprop
    // 0.035 -> 0.028 ms).  So: only when the 256-column tiling leaves more than half of the SMs idle.
    // B200SEG_BOX_N=128 / 256 forces one of them (A/B runs).
    static int bn_env = -1;
    if (bn_env < 0) { const char* e = getenv("B200SEG_BOX_N"); bn_env = (e == nullptr) ? 0 : (e[0] == '1' ? 128 : 256); }
    const long long items256 = (long long)p.m_tiles * (n_total / 256);
    const bool few = 2 * items256 <= kNumSMs;
    if (!strip && (bn_env == 128 || (bn_env == 0 && few)) && (mode == 0 || mode == 1) && p.ocb == 64 && p.block_n == 256 &&
        n_total % 128 == 0)
      p.block_n = 128;
  }
  if (p.block_n % 32 && p.block_n != 16) return (int)cudaErrorInvalidValue;
  p.n_tiles = (n_total + p.block_n - 1) / p.block_n;
  p.groups = p.block_n / p.ocb;
  if (p.groups > kMaxGroups || p.block_n % p.ocb) return (int)cudaErrorInvalidValue;
  p.b_tap_stride = (mode == 3) ? 0 : conv_cin;
  p.b_row_tap_stride = (mode == 3) ? Kc : 0;
  p.a_mode = (mode == 3) ? 1 : 0;
  p.out_mode = (mode == 2) ? 1 : 0;
  p.convt_cout = (mode == 2) ? n_total / 4 : 1;
  if (mode == 2 && (p.convt_cout % p.ocb)) return (int)cudaErrorInvalidValue;
  if ((mode == 2 || mode == 3) && g.tn > 1 && (1 << g.th_log2) != H) return (int)cudaErrorInvalidValue;
  p.bias = bias;
  p.partials = partials;
  {
    static int rs_env = -1;
    if (rs_env < 0) { const char* e = getenv("B200SEG_REG_STATS"); rs_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    p.reg_stats = rs_env;
  }
  if (partials != nullptr && (p.block_n % 8 != 0 || p.block_n < 16 || p.block_n > 256)) return (int)cudaErrorInvalidValue;
  p.a_bytes = 128 * BK * 2;
  const bool bmn = (mode == 1 || mode == 3);
  const uint32_t b_raw = bmn ? (uint32_t)((p.block_n + 63) / 64) * 64 * BK * 2 : (uint32_t)p.block_n * BK * 2;
  p.b_bytes = b_raw;
  p.stage_bytes = p.a_bytes + ((b_raw + 1023) / 1024) * 1024;
  p.staging_bytes = 128u * p.block_n * 2;
  if (p.staging_bytes < 16384) p.staging_bytes = 16384;
  if (strip) {
    p.xslot_bytes = ((130u * BK * 2 + 1023) / 1024) * 1024;  // 130 rows of one K chunk (16640 B -> 17 KiB at BK=64)
    p.stage_bytes = ((b_raw + 1023) / 1024) * 1024;        // the ring holds weight tiles only
    // small filter banks (64 -> 64, the stem) stay resident: no weight traffic after the first tile
    const uint32_t bank = 9u * p.kc_per_tap * p.stage_bytes;
    // (a CTA keeps one N tile for its whole life, so its slice of the bank can stay resident with several N tiles)
    p.b_resident = ((p.n_tiles == 1 || p.block_n == 64) && 4 * p.xslot_bytes + 2 * p.staging_bytes + bank <= kSmemBudget) ? 1 : 0;
    auto fits = [&](int bufs, int xs, int bst) {
      return (long long)bufs * p.staging_bytes + (long long)xs * p.xslot_bytes +
             (p.b_resident ? (long long)bank : (long long)bst * p.stage_bytes) <= (long long)kSmemBudget;
    };
    p.staging_bufs = fits(2, 6, 6) ? 2 : 1;
    // two-row work units (every weight tile serves two output rows) when the filter bank is streamed:
    // needs an even height, N <= 128 (four 128-column accumulators) and room for 8 strips
    p.rows2 = (!p.b_resident && rows2_env() && (H % 2 == 0) && p.block_n <= 128 &&
               fits(1, 8, 3)) ? 1 : 0;
    if (p.rows2) p.staging_bufs = fits(2, 8, 6) ? 2 : 1;
    // four-row units with the vertical taps merged along N (resident bank, N = 64, one K chunk): the
    // bank's tiles must be exactly one 64-row block each so that stacked filter rows are contiguous.
    // An M = 128, N = 64 UMMA occupies the tensor pipe as long as an N = 128 one (64 cycles, measured:
    // profiles/ncu_mma_issue.txt), so merging is worth up to 2x on the MMA side.  With the warp-private
    // epilogue it pays for fprop as well (64 -> 64 fprop 0.351 -> 0.241 ms, stem 0.271 -> 0.154 ms).
    p.rows4 = (p.b_resident && (rows4_env() == 2 || (rows4_env() == 1 && mode == 1)) && (H % 4 == 0) &&
               p.block_n == 64 &&
               p.kc_per_tap == 1 && p.stage_bytes == 64u * BK * 2u && fits(1, 6, 0)) ? 1 : 0;
    if (p.rows4) p.staging_bufs = fits(2, 6, 0) ? 2 : 1;
    if (!fits(p.staging_bufs, 4, 3)) {
      strip = false;
    } else {
      int bst = 6;
      while (!fits(p.staging_bufs, 4, bst)) --bst;         // weight ring: 6 deep if possible, never < 3
      // every tile needs 3 strips per K chunk and a TMA round trip is ~1.5 us: give the strip ring as
      // many slots as shared memory allows
      int xs = p.rows2 ? 8 : (p.rows4 ? 6 : 4);
      while (!fits(p.staging_bufs, xs, bst) && bst > 3) --bst;
      while (xs < kMaxXSlots && fits(p.staging_bufs, xs + 1, bst)) ++xs;
      if (p.rows2) xs -= xs % 4;                            // a K chunk's four strips never straddle the ring end
      p.xslots = xs;
    }
  }
  {
    static int ep_env = -1;
    if (ep_env < 0) { const char* e = getenv("B200SEG_EPI_PRIVATE"); ep_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
    // fprop: 64-channel layers; dgrad: any strip layer with up to three 64-column groups per tile
    const bool shape_ok = p.block_n == 64 || (bmn && p.block_n % 64 == 0 && p.block_n <= 192);
    p.epi_private = (strip && ep_env && shape_ok && p.ocb == 64 && p.staging_bufs == 2 &&
                     p.staging_bytes >= 16384 && kClusterSize == 1) ? 1 : 0;
  }
  if (!strip) {
    g = geom_for(N, H, W, 128);
    p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h; p.tiles_g = g.tiles_g;
    p.m_tiles = g.tiles_w * g.tiles_h * g.tiles_g;
    p.tw_log2 = g.tw_log2; p.th_log2 = g.th_log2; p.tn = g.tn;
    p.xslots = 0; p.xslot_bytes = 0; p.rows2 = 0; p.rows4 = 0; p.b_resident = 0;
    p.stage_bytes = p.a_bytes + ((b_raw + 1023) / 1024) * 1024;
    p.staging_bufs = (2 * p.staging_bytes + 4 * p.stage_bytes <= kSmemBudget) ? 2 : 1;
  }
  int stages = (int)((kSmemBudget - p.staging_bufs * p.staging_bytes - p.xslots * p.xslot_bytes) / p.stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return (int)cudaErrorInvalidValue;
  if (strip && p.b_resident) stages = 9 * p.kc_per_tap;      // smem layout only; no ring barriers are used
  p.stages = stages;
  // CTA pairs (clusters of 2 along M) can share each weight tile through TMA multicast, halving the
  // weight traffic from L2.  Implemented and correct (tests pass with B200SEG_CLUSTER=1) but measured
  // SLOWER on B200 for these shapes (L0.dec0 fprop 1.66 -> 2.08 ms, whole step 34.4 -> 35.0 ms): the
  // pair runs in lock step and the 8 x 8 KB weight ring is too shallow to hide the coupled latency.
  // Kept as an opt-in experiment (profiles/variants_in_situ.txt).
  p.cluster = 1;
  if (kClusterSize == 2 && cluster_env() && mode != 2 && mode != 3 && p.m_tiles >= 2 * kNumSMs && (p.block_n % 32 == 0) &&
      (bmn ? true : (p.block_n / 2) % 8 == 0) && !dry && local_device_memory(w))
    p.cluster = 2;
  const int unit = p.n_tiles * p.cluster;
  int grid = kNumSMs - (kNumSMs % unit);
  const long long total = p.rows4 ? (long long)(p.m_tiles / 4) * unit
                          : p.rows2 ? (long long)(p.m_tiles / 2) * unit
                                    : (long long)((p.m_tiles + p.cluster - 1) / p.cluster) * unit;
  if (total < grid) grid = (int)total;
  if (stat_rows_out) *stat_rows_out = grid / p.n_tiles;
  const size_t smem = (size_t)p.stages * p.stage_bytes + (size_t)p.xslots * p.xslot_bytes +
                      (size_t)p.staging_bufs * p.staging_bytes + 1024;
  {
    B200ConvPlan& q = g_last_plan;
    q.kernel = strip ? 1 : 0;
    q.block_k = BK; q.block_n = p.block_n; q.n_tiles = p.n_tiles; q.ocb = p.ocb;
    q.rows_per_unit = p.rows4 ? 4 : (p.rows2 ? 2 : 1);
    q.weights_resident = p.b_resident; q.private_epilogue = p.epi_private;
    q.staging_bufs = p.staging_bufs; q.strip_slots = p.xslots; q.stages = p.stages;
    q.grid = grid; q.cluster = p.cluster; q.m_tiles = p.m_tiles;
    q.smem_bytes = (long long)smem;
  }
  if (dry) return 0;

  CUtensorMap ma, mb, mo;
  const int TW = 1 << g.tw_log2, TH = 1 << g.th_log2;
  if (strip) {
    long long dims[4] = {Kc, W, H, N}, str[4] = {1, ld_a, ld_a * W, ld_a * W * H};
    int box[4] = {BK, 130, 1, 1};
    if (!make_map(&ma, a, 4, dims, str, box, BK * 2)) return (int)cudaErrorInvalidValue;
  } else if (mode != 3) {
    long long dims[4] = {Kc, W, H, N}, str[4] = {1, ld_a, ld_a * W, ld_a * W * H};
    int box[4] = {BK, TW, TH, g.tn};
    if (!make_map(&ma, a, 4, dims, str, box, BK * 2)) return (int)cudaErrorInvalidValue;
  } else {         // d_up (N, 2H, 2W, C) viewed as (c, dx, x, dy, Y = n*H + y)
    long long dims[5] = {Kc, 2, W, 2, (long long)N * H};
    long long str[5] = {1, ld_a, 2 * ld_a, 2 * (long long)W * ld_a, 4 * (long long)W * ld_a};
    int box[5] = {BK, 1, TW, 1, TH * g.tn};
    if (!make_map(&ma, a, 5, dims, str, box, BK * 2)) return (int)cudaErrorInvalidValue;
  }
  if (!bmn) {      // weights as [n_total rows][taps*Kc cols], K contiguous
    long long dims[2] = {(long long)p.taps * Kc, n_total}, str[2] = {1, (long long)p.taps * Kc};
    int box[2] = {BK, p.block_n / p.cluster};               // a CTA pair loads half a tile each
    if (!make_map(&mb, w, 2, dims, str, box, BK * 2)) return (int)cudaErrorInvalidValue;
  } else if (mode == 3) {   // transposed-conv weights [(q, co) rows][ci cols], N = ci contiguous
    long long dims[2] = {n_total, (long long)4 * Kc}, str[2] = {1, n_total};
    int box[2] = {64, BK};
    if (!make_map(&mb, w, 2, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  } else {         // weights as [Kc rows (conv Cout)][9*conv_cin cols], N contiguous
    long long dims[2] = {(long long)9 * conv_cin, Kc}, str[2] = {1, (long long)9 * conv_cin};
    int box[2] = {64, BK};
    if (!make_map(&mb, w, 2, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  }
  if (mode != 2) {
    long long dims[4] = {n_total, W, H, N}, str[4] = {1, ld_o, ld_o * W, ld_o * W * H};
    int box[4] = {p.ocb, TW, TH, g.tn};
    if (!make_map(&mo, out, 4, dims, str, box, p.ocb * 2)) return (int)cudaErrorInvalidValue;
  } else {         // (c, dx, x, dy, Y = n*h + y) view of the (N, 2h, 2w, C) output
    long long dims[5] = {p.convt_cout, 2, W, 2, (long long)N * H};
    long long str[5] = {1, ld_o, 2 * ld_o, 2 * (long long)W * ld_o, 4 * (long long)W * ld_o};
    int box[5] = {p.ocb, 1, TW, 1, TH * g.tn};
    if (!make_map(&mo, out, 5, dims, str, box, p.ocb * 2)) return (int)cudaErrorInvalidValue;
  }
  CUtensorMap mo2 = mo;
  if (p.epi_private) {   // one 32-pixel x 32-channel slab per warp, plain (unswizzled) 64-byte rows
    long long dims[4] = {n_total, W, H, N}, str[4] = {1, ld_o, ld_o * W, ld_o * W * H};
    int box[4] = {32, 32, 1, 1};
    if (!make_map(&mo2, out, 4, dims, str, box, 0)) return (int)cudaErrorInvalidValue;
  }
  if (strip) {
    if (BK == 16) return launch_tap<16, false, true>(ma, mb, mo, mo2, p, grid, smem, st);
    if (bmn) return launch_tap<64, true, true>(ma, mb, mo, mo2, p, grid, smem, st);
    return launch_tap<64, false, true>(ma, mb, mo, mo2, p, grid, smem, st);
  }
  if (bmn) {
    if (BK == 64) return launch_tap<64, true>(ma, mb, mo, mo2, p, grid, smem, st);
    if (BK == 32) return launch_tap<32, true>(ma, mb, mo, mo2, p, grid, smem, st);
    return launch_tap<16, true>(ma, mb, mo, mo2, p, grid, smem, st);
  }
  if (BK == 64) return launch_tap<64, false>(ma, mb, mo, mo2, p, grid, smem, st);
  if (BK == 32) return launch_tap<32, false>(ma, mb, mo, mo2, p, grid, smem, st);
  return launch_tap<16, false>(ma, mb, mo, mo2, p, grid, smem, st);
}

}  // namespace b200

using namespace b200;

extern "C" int b200conv_supported(int cin, int cout) { return (cin % 16 == 0 && cout % 16 == 0) ? 1 : 0; }

extern "C" int b200conv_fprop_stat_rows(int N, int H, int W, int Cin, int Cout) {
  int rows = 0;
  int e = run_tap_gemm(0, nullptr, Cin, nullptr, nullptr, nullptr, Cout, N, H, W, Cin, Cout, Cin, nullptr, &rows,
                       true, 0);
  return e == 0 ? rows : 0;
}

extern "C" int b200conv_fprop(const void* x, long long ld_x, const void* w, const float* bias, void* y,
                              long long ld_y, int N, int H, int W, int Cin, int Cout, float* partials,
                              cudaStream_t st) {
  return run_tap_gemm(0, x, ld_x, w, bias, y, ld_y, N, H, W, Cin, Cout, Cin, partials, nullptr, false, st);
}

extern "C" int b200conv_dgrad(const void* dy, long long ld_dy, const void* w, void* dx, long long ld_dx, int N,
                              int H, int W, int Cin, int Cout, cudaStream_t st) {
  return run_tap_gemm(1, dy, ld_dy, w, nullptr, dx, ld_dx, N, H, W, Cout, Cin, Cin, nullptr, nullptr, false, st);
}

// dgrad that also returns per-CTA column sums of dx ([rows][2][Cin]: sum, sum of squares, of the
// bf16-rounded output) -- the transposed conv's bias gradient is the column sum of the slice of dx it
// reads, so this saves a separate pass over that tensor.
extern "C" int b200conv_dgrad_stat_rows(int N, int H, int W, int Cin, int Cout) {
  int rows = 0;
  int e = run_tap_gemm(1, nullptr, Cout, nullptr, nullptr, nullptr, Cin, N, H, W, Cout, Cin, Cin, nullptr, &rows, true,
                       0);
  return e == 0 ? rows : 0;
}
extern "C" int b200conv_dgrad_stats(const void* dy, long long ld_dy, const void* w, void* dx, long long ld_dx, int N,
                                    int H, int W, int Cin, int Cout, float* partials, cudaStream_t st) {
  return run_tap_gemm(1, dy, ld_dy, w, nullptr, dx, ld_dx, N, H, W, Cout, Cin, Cin, partials, nullptr, false, st);
}

extern "C" int b200conv_convT_fwd(const void* x, long long ld_x, const void* w, const float* bias, void* out,
                                  long long ld_o, int N, int h, int w_, int Cin, int Cout, cudaStream_t st) {
  return run_tap_gemm(2, x, ld_x, w, bias, out, ld_o, N, h, w_, Cin, 4 * Cout, Cin, nullptr, nullptr, false, st);
}

static int run_wgrad(int convt, const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                     int H, int W, int Cin, int Cout, cudaStream_t st, int tap0 = 0, int ntaps = 9) {
  if (Cin % 8 || Cout % 16) return (int)cudaErrorInvalidValue;
  WgradParams p{};
  TileGeom g = geom_for(N, H, W, 64);
  if (convt && g.tn > 1 && (1 << g.th_log2) != H) return (int)cudaErrorInvalidValue;
  p.tiles_w = g.tiles_w; p.tiles_h = g.tiles_h; p.tiles_g = g.tiles_g;
  p.p_tiles = g.tiles_w * g.tiles_h * g.tiles_g;
  p.tw_log2 = g.tw_log2; p.th_log2 = g.th_log2; p.tn = g.tn;
  p.Cin = Cin; p.Cout = Cout;
  p.chunks_per_tap = (Cin + 63) / 64;
  p.m_blocks = (convt ? 1 : ntaps) * p.chunks_per_tap;
  p.tap0 = convt ? 0 : tap0;
  // Two 128-row M tiles per CTA share each dy tile (halves the dy traffic per FLOP).  In isolation it is
  // neutral, inside the training step it is worth 1.5 ms of 35 (profiles/variants_in_situ.txt), so it
  // is the default; B200SEG_WGRAD_MT=1 turns it off.
  static int mt_env = -1;
  if (mt_env < 0) { const char* e = getenv("B200SEG_WGRAD_MT"); mt_env = (e != nullptr && e[0] == '1') ? 1 : 2; }
  p.mt = (mt_env == 2 && p.m_blocks >= 8) ? 2 : 1;
  p.block_n = Cout <= 256 ? Cout : (Cout % 256 == 0 ? 256 : (Cout % 128 == 0 ? 128 : 64));
  // Transposed conv with few channels: 128-column tiles put all four output quadrants next to each other
  // in TMEM, so one CTA serves them from ONE x tile (x was re-read once per quadrant and these layers are
  // L2-bound).  For wide layers the dy re-reads of the extra M tiles would cost more than that saves.
  // B200SEG_CONVT_QUADS=0 restores one quadrant per CTA.
  static int quads_env = -1;
  if (quads_env < 0) { const char* e = getenv("B200SEG_CONVT_QUADS"); quads_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
  p.quads = 1;
  if (convt && quads_env && Cout <= 256 && Cout % 64 == 0) {
    p.quads = 4;
    p.mt = 1;
    p.block_n = Cout <= 128 ? Cout : 128;
  }
  p.m_tiles = (p.m_blocks + 2 * p.mt - 1) / (2 * p.mt);
  if (p.block_n % 32 && p.block_n != 16) return (int)cudaErrorInvalidValue;
  p.nt_per_q = (Cout + p.block_n - 1) / p.block_n;
  p.n_tiles = ((convt && p.quads == 1) ? 4 : 1) * p.nt_per_q;
  p.convt = convt;
  p.H = H;
  p.co_stride = convt ? Cin : (long long)9 * Cin;
  const int nblocks = (p.block_n + 63) / 64;
  p.a_bytes = 2 * p.mt * 8192;
  p.b_bytes = p.quads * nblocks * 8192;
  p.stage_bytes = p.a_bytes + p.b_bytes;
  int stages = (int)(kSmemBudget / p.stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  p.stages = stages;
  const int base_items = p.m_tiles * p.n_tiles;
  static int waves_env = -1, vec_env = -1;
  if (waves_env < 0) { const char* e = getenv("B200SEG_WGRAD_WAVES"); waves_env = (e != nullptr && e[0] == '1') ? 1 : 2; }
  if (vec_env < 0) { const char* e = getenv("B200SEG_WGRAD_VEC"); vec_env = (e != nullptr && e[0] == '0') ? 0 : 1; }
  // a quad of rows = four consecutive input channels of one 64-channel block: whole quads need Cin % 4 == 0
  p.vec_red = (vec_env && Cin % 4 == 0) || b200seg_grad_multicast_delta(dw) != 0 ? 1 : 0;
  // split-K factor: two waves of CTAs hide the prologue / epilogue of one behind the main loop of the other,
  // but every extra slice adds a full set of partial sums to reduce; with few pixels (<= 256 tiles of 64) one
  // wave is faster (L4.c0 0.085 -> 0.080 ms, L5 0.043 -> 0.031 ms; L3 and up prefer two).
  const int waves = (waves_env == 1 || p.p_tiles <= 256) ? 1 : 2;
  int ks = (waves * kNumSMs) / base_items;
  if (ks < 1) ks = 1;
  if (ks > p.p_tiles) ks = p.p_tiles;
  p.k_slices = ks;
  p.dw = dw;
  p.mc_delta = b200seg_grad_multicast_delta(dw);
  p.arm = static_cast<const GradArm*>(b200seg_take_grad_arm());   // this is always the last launch of a wgrad call
  CUtensorMap mx, mdy;
  const int TW = 1 << g.tw_log2, TH = 1 << g.th_log2;
  {
    long long dims[4] = {Cin, W, H, N}, str[4] = {1, ld_x, ld_x * W, ld_x * W * H};
    int box[4] = {64, TW, TH, g.tn};
    if (!make_map(&mx, x, 4, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  }
  if (!convt) {
    long long dims[4] = {Cout, W, H, N}, str[4] = {1, ld_dy, ld_dy * W, ld_dy * W * H};
    int box[4] = {64, TW, TH, g.tn};
    if (!make_map(&mdy, dy, 4, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  } else {
    long long dims[5] = {Cout, 2, W, 2, (long long)N * H};
    long long str[5] = {1, ld_dy, 2 * ld_dy, 2 * (long long)W * ld_dy, 4 * (long long)W * ld_dy};
    int box[5] = {64, 1, TW, 1, TH * g.tn};
    if (!make_map(&mdy, dy, 5, dims, str, box, 128)) return (int)cudaErrorInvalidValue;
  }
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)(kSmemBudget + 2048));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(wgrad_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kSmemBudget + 2048));
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(wgrad_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kSmemBudget + 2048));
    if (e != cudaSuccess) return (int)e;
    attr_done = true;
  }
  const size_t smem = (size_t)p.stages * p.stage_bytes + 1024;
  // CTA pairs sharing the dy tile: needs an even number of M tiles (the pair) and of dy blocks (the halves)
  // B200SEG_WGRAD_CLUSTER: 0 none, 2 pairs (default), 4 also 2 x 2 clusters where the tiling allows.  The
  // 2 x 2 form is correct but measured slower (L3.c1 0.387 vs 0.259 ms, L3.dec0 0.699 vs 0.471): four CTAs
  // in lock step over a 3-stage ring lose more than the halved x traffic gains.
  static int wcl_env = -1;
  if (wcl_env < 0) { const char* e = getenv("B200SEG_WGRAD_CLUSTER"); wcl_env = (e == nullptr) ? 2 : (e[0] == '0' ? 0 : (e[0] == '4' ? 4 : 2)); }
  const int nbt = p.quads * nblocks;
  if (wcl_env && p.m_tiles % 2 == 0 && nbt % 2 == 0 && local_device_memory(x) && local_device_memory(dy)) {
    const bool quad = wcl_env == 4 && p.quads == 1 && p.n_tiles % 2 == 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(base_items * ks);
    cfg.blockDim = dim3(kConvThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = quad ? 4 : 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (quad) return (int)cudaLaunchKernelEx(&cfg, wgrad_kernel<4>, mx, mdy, p);
    return (int)cudaLaunchKernelEx(&cfg, wgrad_kernel<2>, mx, mdy, p);
  }
  return (int)launch_k(wgrad_kernel<1>, base_items * ks, kConvThreads, smem, st, mx, mdy, p);
}

extern "C" int b200conv_wgrad_pair_supported(int N, int H, int W, int Cin, int Cout);
extern "C" int b200conv_wgrad_pair(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw,
                                   int N, int H, int W, int Cin, int Cout, cudaStream_t st);
extern "C" int b200conv_wgrad_strip_supported(int N, int H, int W, int Cin, int Cout);
extern "C" int b200conv_wgrad_strip_preferred(int N, int H, int W, int Cin, int Cout);
extern "C" int b200conv_wgrad_strip(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw,
                                    int N, int H, int W, int Cin, int Cout, cudaStream_t st);

// One filter tap (or a range) with the split-K kernel: the 128-column strip kernel leaves tap (2,2) to it.
extern "C" int b200conv_wgrad_taps(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                                   int H, int W, int Cin, int Cout, int tap0, int ntaps, cudaStream_t st) {
  if (tap0 < 0 || ntaps < 1 || tap0 + ntaps > 9) return (int)cudaErrorInvalidValue;
  return run_wgrad(0, x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, st, tap0, ntaps);
}

// Dispatch: halo-strip kernel (wgrad_strip.cu) when the rows are at least 64 pixels wide, else the
// split-K kernel above.  B200SEG_WGRAD_KERNEL=splitk|strip forces one of them (for benchmarking).
// 0: split-K kernel (this file), 1: 128-column strip kernel + tap (2,2) pass, 2: row-pair kernel
static int wgrad_kernel_for(int N, int H, int W, int Cin, int Cout) {
  static int force = -1;
  if (force < 0) {
    const char* e = getenv("B200SEG_WGRAD_KERNEL");
    force = (e == nullptr) ? 0 : (e[0] == 's' && e[1] == 'p' ? 1 : 2);
  }
  static int pair_all = -1;            // B200SEG_WGRAD_PAIR=2: row-pair kernel for every layer it supports (A/B runs)
  if (pair_all < 0) { const char* e = getenv("B200SEG_WGRAD_PAIR"); pair_all = (e != nullptr && e[0] == '2') ? 1 : 0; }
  static int pair_off = -1;
  if (pair_off < 0) { const char* e = getenv("B200SEG_WGRAD_PAIR"); pair_off = (e != nullptr && e[0] == '0') ? 1 : 0; }
  // Measured per layer (profiles/conv_layers_b64_n16_512.txt): the row-pair kernel beats the 128-column strip
  // kernel + its separate tap-(2,2) pass and the split-K kernel wherever rows are >= 64 pixels wide, except
  // for the widest inputs (Cin > 512), where re-reading x once per 64-channel output tile costs more.
  if (force == 0 && !pair_off && b200conv_wgrad_pair_supported(N, H, W, Cin, Cout) && (pair_all || Cin <= 512)) return 2;
  const bool strip_ok = (force == 2) ? b200conv_wgrad_strip_supported(N, H, W, Cin, Cout) != 0
                                     : b200conv_wgrad_strip_preferred(N, H, W, Cin, Cout) != 0;
  return (strip_ok && force != 1) ? 1 : 0;
}

extern "C" int b200conv_wgrad(const void* x, long long ld_x, const void* dy, long long ld_dy, float* dw, int N,
                              int H, int W, int Cin, int Cout, cudaStream_t st) {
  const int k = wgrad_kernel_for(N, H, W, Cin, Cout);
  if (k == 2) return b200conv_wgrad_pair(x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, st);
  if (k == 1) return b200conv_wgrad_strip(x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, st);
  return run_wgrad(0, x, ld_x, dy, ld_dy, dw, N, H, W, Cin, Cout, st);
}

// Which kernel variant and tiling a conv call of this shape takes, without launching anything (no GPU needed).
// op: 0 fprop, 1 dgrad, 2 wgrad, 3 transposed-conv forward, 4 transposed-conv dgrad.
extern "C" int b200conv_plan(int op, int N, int H, int W, int Cin, int Cout, B200ConvPlan* out) {
  if (out == nullptr) return (int)cudaErrorInvalidValue;
  *out = B200ConvPlan{};
  int e = 0;
  switch (op) {
    case 0: e = run_tap_gemm(0, nullptr, Cin, nullptr, nullptr, nullptr, Cout, N, H, W, Cin, Cout, Cin, nullptr, nullptr, true, 0); break;
    case 1: e = run_tap_gemm(1, nullptr, Cout, nullptr, nullptr, nullptr, Cin, N, H, W, Cout, Cin, Cin, nullptr, nullptr, true, 0); break;
    case 2: out->kernel = 2 + wgrad_kernel_for(N, H, W, Cin, Cout); return 0;
    case 3: e = run_tap_gemm(2, nullptr, Cin, nullptr, nullptr, nullptr, Cout, N, H, W, Cin, 4 * Cout, Cin, nullptr, nullptr, true, 0); break;
    case 4: e = run_tap_gemm(3, nullptr, Cout, nullptr, nullptr, nullptr, Cin, N, H, W, Cout, Cin, Cin, nullptr, nullptr, true, 0); break;
    default: return (int)cudaErrorInvalidValue;
  }
  if (e == 0) *out = g_last_plan;
  return e;
}

// Backward of ConvTranspose2d(k=2, s=2): d_up is (N, 2h, 2w, Cout) (a channel slice of the concat
// gradient), x (N, h, w, Cin), w / dw in the kernel layout [(dy, dx, co)][ci].
extern "C" int b200conv_convT_dgrad(const void* d_up, long long ld_g, const void* w, void* dx, long long ld_dx,
                                    int N, int h, int w_, int Cin, int Cout, cudaStream_t st) {
  return run_tap_gemm(3, d_up, ld_g, w, nullptr, dx, ld_dx, N, h, w_, Cout, Cin, Cin, nullptr, nullptr, false, st);
}

extern "C" int b200conv_convT_wgrad(const void* x, long long ld_x, const void* d_up, long long ld_g, float* dw,
                                    int N, int h, int w_, int Cin, int Cout, cudaStream_t st) {
  return run_wgrad(1, x, ld_x, d_up, ld_g, dw, N, h, w_, Cin, Cout, st);
}

// Points this translation unit's error record (umma.cuh mbar_timeout_trap) at mapped host memory.
extern "C" int b200conv_set_error_record_tc(unsigned int* host_mapped) {
  return (int)cudaMemcpyToSymbol(b200::g_error_record, &host_mapped, sizeof(host_mapped));
}

extern "C" int b200conv_set_error_record_strip(unsigned int* host_mapped);
extern "C" int b200conv_set_error_record_pair(unsigned int* host_mapped);
// `host_mapped`: >= 8 words of pinned, device-mapped host memory (or NULL to disable), set per device with that
// device current.  A bounded mbarrier wait of the tcgen05 kernels that times out stores {0xB200DEAD, barrier
// address, parity, CTA, thread} there before it traps, so the host can still say WHICH wait failed after the
// context is gone.
extern "C" int b200seg_set_error_record(unsigned int* host_mapped) {
  int e = b200conv_set_error_record_tc(host_mapped);
  if (e == 0) e = b200conv_set_error_record_strip(host_mapped);
  if (e == 0) e = b200conv_set_error_record_pair(host_mapped);
  return e;
}
