// Small device helpers shared by the memory-bound kernels.
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {

constexpr int kNumSMs = 148;

struct __align__(16) bf16x8 {
  __nv_bfloat162 v[4];
};

// 16-byte accesses, spelled as uint4: copying the struct member by member made the compiler emit four
// 4-byte LDG / STG per access in every kernel that used these (found in the SASS; 4x the memory
// instructions and partial-sector stores).
__device__ __forceinline__ bf16x8 ld8(const __nv_bfloat16* p) {
  const uint4 t = *reinterpret_cast<const uint4*>(p);
  return *reinterpret_cast<const bf16x8*>(&t);
}
__device__ __forceinline__ void st8(__nv_bfloat16* p, const bf16x8& x) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&x);
}
__device__ __forceinline__ void unpack8(const bf16x8& x, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(x.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float (&f)[8]) {
  bf16x8 x;
#pragma unroll
  for (int i = 0; i < 4; ++i) x.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return x;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Gradient accumulation that may be fused with the cross-GPU reduction: with mc_delta == 0 it is a
// plain red.global.add into this GPU's gradient buffer; otherwise `p + mc_delta` is the NVSwitch
// multicast alias of the same element and ONE multimem.red adds the value into every replica's buffer
// (the switch fans the reduction out), so the weight-gradient GEMM epilogue IS the all-reduce.
__device__ __forceinline__ void grad_red_add(float* p, float v, long long mc_delta) {
  if (mc_delta == 0) {
    atomicAdd(p, v);
  } else {
    asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(reinterpret_cast<char*>(p) + mc_delta),
                 "f"(v)
                 : "memory");
  }
}

// Four consecutive floats through the multicast alias in one 16-byte multimem.red (4x fewer fabric
// packets than the scalar form; the alias must be 16-byte aligned).
__device__ __forceinline__ void grad_red_add4_mc(float* p, float a, float b, float c, float d, long long mc_delta) {
  asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(
                   reinterpret_cast<char*>(p) + mc_delta),
               "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// Register transpose inside lane quads.  In: lane holds row (lane), v[c] = column c.  Out: v[4g + i]
// (i = 0..3) is the element of row (lane & ~3) + i, column 4g + (lane & 3) -- i.e. every lane ends up
// with four CONSECUTIVE rows of one column, which is what a 16-byte vector reduction needs when rows
// are the contiguous direction in memory.  4 shuffles per 4x4 block.
__device__ __forceinline__ void quad_transpose32(uint32_t (&v)[32], int lane) {
  const bool b0 = lane & 1, b1 = lane & 2;
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    uint32_t a0 = v[4 * g], a1 = v[4 * g + 1], a2 = v[4 * g + 2], a3 = v[4 * g + 3];
    uint32_t t0 = b0 ? a0 : a1, t1 = b0 ? a2 : a3;
    t0 = __shfl_xor_sync(0xffffffffu, t0, 1);
    t1 = __shfl_xor_sync(0xffffffffu, t1, 1);
    if (b0) { a0 = t0; a2 = t1; } else { a1 = t0; a3 = t1; }
    t0 = b1 ? a0 : a2;
    t1 = b1 ? a1 : a3;
    t0 = __shfl_xor_sync(0xffffffffu, t0, 2);
    t1 = __shfl_xor_sync(0xffffffffu, t1, 2);
    if (b1) { a0 = t0; a1 = t1; } else { a2 = t0; a3 = t1; }
    v[4 * g] = a0; v[4 * g + 1] = a1; v[4 * g + 2] = a2; v[4 * g + 3] = a3;
  }
}

// ---- all-reduce trigger carried by a weight-gradient kernel (csrc/comm/overlap.cu) ------------------
// A `GradArm` lives in device memory (built by parallel/symm.py, one per gradient bucket).  The wgrad
// launch that completes a bucket gets a pointer to it; every CTA calls grad_arm_publish() after its last
// red.global.add, the LAST one to arrive tells every rank -- this one included -- that the bucket is ready
// by storing the current exchange number into flags[bucket][this rank] of each rank's signal pad.  These
// are peer stores over NVLink issued from the same kernel that ran the UMMA tiles.
constexpr int kArmMaxRanks = 16;
struct GradArm {
  unsigned int* counter;                    // local; wraps back to 0 when the last CTA arrives
  const unsigned int* epoch;                // local; exchange number, advanced by the final bucket's reduce kernel
  unsigned int* peer_flag[kArmMaxRanks];    // flags[bucket][this rank] in every rank's pad
  int world;
  int pad_;
};
// Must be reached by ALL threads of the CTA (it contains a __syncthreads).
__device__ __forceinline__ void grad_arm_publish(const GradArm* arm) {
  if (arm == nullptr) return;
  __threadfence();                                         // this thread's reductions are performed at L2
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int total = gridDim.x * gridDim.y * gridDim.z;
    const unsigned int old = atomicInc(arm->counter, total - 1);
    if (old == total - 1) {                                // every CTA of the launch has arrived
      __threadfence();
      const unsigned int e = *reinterpret_cast<const volatile unsigned int*>(arm->epoch);
      __threadfence_system();
      for (int r = 0; r < arm->world; ++r)
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(arm->peer_flag[r]), "r"(e) : "memory");
    }
  }
}
extern "C" const void* b200seg_take_grad_arm();           // one-shot arm set by b200seg_arm_next_wgrad

// Programmatic dependent launch (B200SEG_PDL=1): a kernel may be scheduled while its predecessor in the
// stream drains; pdl_wait() blocks until the predecessor has completed and its writes are visible, so
// everything before it (barrier init, TMEM allocation, descriptor prefetch, CTA ramp-up) overlaps the
// predecessor's tail.  Without the launch attribute both instructions are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// B200SEG_PDL: 0 off (default), 1 every kernel that supports it, 2 only the small elementwise kernels
// (level 1 = tensor-core kernels too).
inline int pdl_level() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200SEG_PDL"); v = (e == nullptr) ? 0 : (e[0] == '1' ? 1 : (e[0] == '2' ? 2 : 0)); }
  return v;
}
#ifndef B200SEG_PDL_CLASS
#define B200SEG_PDL_CLASS 1        // translation units of elementwise kernels define 2 before including this
#endif
static inline bool pdl_enabled() { return pdl_level() == 1 || (pdl_level() == 2 && B200SEG_PDL_CLASS == 2); }

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

inline int grid_for(long long work_items, int threads, int max_blocks_per_sm = 8) {
  long long b = (work_items + threads - 1) / threads;
  long long cap = static_cast<long long>(kNumSMs) * max_blocks_per_sm;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace b200
