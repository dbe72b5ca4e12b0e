// Bucketed gradient reduction that runs UNDER the backward pass.
//
// The reference exchanges gradients strictly after backward (R:756-763 -> grad_serv_mean R:255-440 /
// grad_client_mean R:443-556): the GPUs idle for the whole round.  Here the flat gradient buffer is cut
// into buckets in the order backward completes them, and each bucket is reduced over NVLink while the
// remaining layers are still computing:
//
//   * trigger: the weight-gradient kernel that writes the LAST tensor of a bucket carries a `GradArm`
//     (csrc/common.cuh).  Its CTAs count themselves out; the last one publishes "bucket b of rank r is
//     complete for exchange e" straight into every peer's signal pad (st.release.sys over NVLink) --
//     the all-reduce trigger is part of the wgrad kernel's epilogue.
//   * reduce: `bucket_reduce_kernel` (<= 16 CTAs, on a second captured stream, ordered after the arming
//     launch by an event, so it never waits for work queued behind it on its own GPU) waits until all
//     ranks have published bucket b, then rank r reduces slice r of the bucket: multimem.ld_reduce pulls
//     the switch-added sum of all replicas, multimem.st broadcasts it (NVLS), or -- without multicast --
//     peer loads in fixed rank order and peer stores (two-shot).  One owner per element => every replica
//     ends with identical bits (R:402-433).  Fabric traffic is 1x the gradient.
//   * the final bucket (the first encoder block and the small BN / bias / head tensors, complete only
//     when backward ends) publishes its own flag, and its kernel ends with a "done" exchange so that no
//     rank's optimizer reads a slice a peer is still writing.
//
// All waits are bounded by %globaltimer: on a timeout the kernel raises the abort word in every rank's
// pad and returns; the host turns that into PeerLost (parallel/symm.py) instead of a hung GPU.
#include "../common.cuh"
#include "b200seg.h"

namespace b200 {

constexpr int kOvlMaxRanks = 16;
constexpr int kOvlMaxBuckets = 32;
constexpr int kOvlThreads = 512;
constexpr int kOvlMaxRanges = 4;
// pad layout (uint32 words, identical on every rank)
constexpr int kOvlFlagWords = kOvlMaxBuckets * kOvlMaxRanks;     // flags[bucket][rank]
constexpr int kOvlDoneOff = kOvlFlagWords;                       // done[rank]
constexpr int kOvlAbortOff = kOvlDoneOff + kOvlMaxRanks;         // abort word
constexpr int kOvlPadWords = kOvlAbortOff + 16;

struct OvlParams {
  float* mc;                               // multicast alias of the gradient buffer (NULL: two-shot over peer pointers)
  float* grad[kOvlMaxRanks];               // every rank's gradient buffer (two-shot)
  unsigned int* pad[kOvlMaxRanks];         // every rank's signal pad
  unsigned int* epoch;                     // local: current exchange number
  unsigned int* cta_done;                  // local: CTAs of the final kernel that have finished
  int rank, world, bucket;
  int self_publish;                        // 1: this kernel publishes the bucket's ready flag itself
  int final;                               // 1: last bucket of the exchange -> done exchange, epoch += 1
  int n_ranges;
  long long v0[kOvlMaxRanges], nv[kOvlMaxRanges];   // ranges in 16-byte vectors
  float post_scale;
  unsigned long long timeout_ns;
};

__device__ __forceinline__ void ovl_st_release(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ovl_ld_acquire(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ovl_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Wait until *flag >= want (wrap-safe) or the abort word is up or the deadline passes.  Returns false on
// abort / timeout (after raising the abort word everywhere on a timeout).
__device__ bool ovl_wait(const OvlParams& p, const unsigned int* flag, unsigned int want) {
  const unsigned int* my_abort = p.pad[p.rank] + kOvlAbortOff;
  unsigned long long t0 = 0;
  for (unsigned int spin = 0;; ++spin) {
    if ((int)(ovl_ld_acquire(flag) - want) >= 0) return true;
    if ((spin & 63u) == 63u) {
      if (ovl_ld_acquire(my_abort) != 0u) return false;
      const unsigned long long now = ovl_now();
      if (t0 == 0) t0 = now;
      else if (now - t0 > p.timeout_ns) {
        for (int r = 0; r < p.world; ++r) ovl_st_release(p.pad[r] + kOvlAbortOff, 0x80000000u | (unsigned)p.rank);
        return false;
      }
      __nanosleep(200);
    }
  }
}

__global__ void __launch_bounds__(kOvlThreads)
bucket_reduce_kernel(const OvlParams p) {
  __shared__ int s_ok;
  const unsigned int e = *reinterpret_cast<const volatile unsigned int*>(p.epoch);
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    if (p.self_publish && blockIdx.x == 0) {
      __threadfence_system();
      ovl_st_release(p.pad[peer] + p.bucket * kOvlMaxRanks + p.rank, e);
    }
    if (!ovl_wait(p, p.pad[p.rank] + p.bucket * kOvlMaxRanks + peer, e)) s_ok = 0;
  }
  __syncthreads();
  const bool ok = s_ok != 0;
  if (ok) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthr = (size_t)gridDim.x * blockDim.x;
    for (int ri = 0; ri < p.n_ranges; ++ri) {
      const size_t nv = (size_t)p.nv[ri];
      const size_t per = (nv + p.world - 1) / p.world;
      const size_t b0 = (size_t)p.v0[ri] + (size_t)p.rank * per;
      size_t b1 = b0 + per;
      const size_t lim = (size_t)p.v0[ri] + nv;
      if (b1 > lim) b1 = lim;
      if (p.mc != nullptr) {
        // four independent 16-byte switch reductions in flight per thread
        for (size_t i = b0 + tid; i < b1; i += 4 * nthr) {
          float4 v[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const size_t j = i + (size_t)u * nthr;
            if (j < b1)
              asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                           : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                           : "l"(p.mc + j * 4)
                           : "memory");
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const size_t j = i + (size_t)u * nthr;
            if (j < b1) {
              v[u].x *= p.post_scale; v[u].y *= p.post_scale; v[u].z *= p.post_scale; v[u].w *= p.post_scale;
              asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.mc + j * 4),
                           "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                           : "memory");
            }
          }
        }
      } else {
        for (size_t i = b0 + tid; i < b1; i += nthr) {
          float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
          float4 f[kOvlMaxRanks];
#pragma unroll
          for (int r = 0; r < kOvlMaxRanks; ++r)
            if (r < p.world) f[r] = reinterpret_cast<const float4*>(p.grad[r])[i];
#pragma unroll
          for (int r = 0; r < kOvlMaxRanks; ++r)          // fixed order: identical bits on every rank
            if (r < p.world) { a.x += f[r].x; a.y += f[r].y; a.z += f[r].z; a.w += f[r].w; }
          a.x *= p.post_scale; a.y *= p.post_scale; a.z *= p.post_scale; a.w *= p.post_scale;
          for (int r = 0; r < p.world; ++r)
            reinterpret_cast<float4*>(p.grad[(p.rank + r) % p.world])[i] = a;
        }
      }
    }
  }
  if (!p.final) return;
  // ---- end of the exchange: every rank's stores have landed everywhere before anyone's optimizer runs
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(p.cta_done, 1u) == gridDim.x - 1) ? 1 : 0;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x < p.world) {
    const int peer = threadIdx.x;
    __threadfence_system();
    ovl_st_release(p.pad[peer] + kOvlDoneOff + p.rank, e);
    if (ok) ovl_wait(p, p.pad[p.rank] + kOvlDoneOff + peer, e);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *p.cta_done = 0u;
    *p.epoch = e + 1u;
    __threadfence();
  }
}

// Next wgrad launcher call arms this (one-shot; see take_grad_arm in conv_tc.cu / wgrad_strip.cu).
static const GradArm* g_next_arm = nullptr;

}  // namespace b200

using namespace b200;

extern "C" void b200seg_arm_next_wgrad(const void* arm) { g_next_arm = static_cast<const GradArm*>(arm); }
extern "C" const void* b200seg_take_grad_arm() {
  const GradArm* a = g_next_arm;
  g_next_arm = nullptr;
  return a;
}

extern "C" int b200seg_overlap_pad_words() { return kOvlPadWords; }
extern "C" int b200seg_overlap_max_buckets() { return kOvlMaxBuckets; }
extern "C" int b200seg_overlap_flag_index(int bucket, int rank) { return bucket * kOvlMaxRanks + rank; }
extern "C" int b200seg_overlap_abort_index() { return kOvlAbortOff; }

// Reduce one bucket (up to 4 ranges of the flat buffer, offsets / counts in floats, multiples of 4) across
// `world` ranks, in place.  `multicast_ptr` == NULL selects peer loads / stores over `grad_ptrs`.
// `state` is a local device array: [0] = exchange epoch (start it at 1), [1] = CTA counter (0).
extern "C" int b200seg_bucket_reduce(float* multicast_ptr, float* const* grad_ptrs, unsigned int* const* pad_ptrs,
                                     unsigned int* state, int rank, int world, int bucket, int self_publish,
                                     int final, int n_ranges, const long long* offs, const long long* counts,
                                     float post_scale, double timeout_s, int blocks, cudaStream_t st) {
  if (world < 2 || world > kOvlMaxRanks || bucket < 0 || bucket >= kOvlMaxBuckets || n_ranges < 1 ||
      n_ranges > kOvlMaxRanges || blocks < 1 || blocks > 64)
    return (int)cudaErrorInvalidValue;
  OvlParams p{};
  p.mc = multicast_ptr;
  for (int r = 0; r < world; ++r) {
    p.grad[r] = grad_ptrs ? grad_ptrs[r] : nullptr;
    p.pad[r] = pad_ptrs[r];
    if (multicast_ptr == nullptr && p.grad[r] == nullptr) return (int)cudaErrorInvalidValue;
  }
  p.epoch = state;
  p.cta_done = state + 1;
  p.rank = rank; p.world = world; p.bucket = bucket;
  p.self_publish = self_publish; p.final = final; p.n_ranges = n_ranges;
  for (int i = 0; i < n_ranges; ++i) {
    if ((offs[i] % 4) || (counts[i] % 4) || counts[i] < 0) return (int)cudaErrorInvalidValue;
    p.v0[i] = offs[i] / 4;
    p.nv[i] = counts[i] / 4;
  }
  p.post_scale = post_scale;
  p.timeout_ns = (unsigned long long)((timeout_s > 0 ? timeout_s : 30.0) * 1e9);
  bucket_reduce_kernel<<<blocks, kOvlThreads, 0, st>>>(p);
  return (int)cudaGetLastError();
}
