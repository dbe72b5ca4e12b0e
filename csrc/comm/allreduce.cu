// Gradient all-reduce over NVLink peer memory (no NCCL on this path).
//
// Replaces the reference's whole exchange round: quantize -> D2H -> pickle -> gzip -> TCP to the
// server, reduce at the root, re-quantize, broadcast, dequantize (grad_serv_mean R:255-440,
// grad_client_mean R:443-556).  On an NVSwitch box every GPU reaches every peer
// at full bandwidth, so the star is the wrong shape; each rank maps all peers' buffers
// ("symmetric memory", handles exchanged once at start-up) and the kernels below read / write
// them directly:
//
//   two-shot (bandwidth-optimal, in place): rank r owns slice r.  It loads slice r of EVERY rank's
//     wire buffer (peer loads), dequantizes on load, accumulates in fp32 in fixed rank order
//     (=> bit-identical result on all ranks, the reference's R:402-433 guarantee), applies the
//     1/world (and 1/K) scale, and stores the result into slice r of EVERY rank's fp32 gradient
//     buffer (peer stores).  Traffic per GPU: (W-1)/W of the wire size in, (W-1)/W * 4 B out.
//   one-shot (latency-optimal, small buffers): every rank loads everything from everyone and
//     reduces locally into a private output.
//
// Cross-GPU ordering uses per-CTA flags in a symmetric signal pad: st.release.sys to the peer,
// ld.acquire.sys spin on the local pad (bounded by %globaltimer; on expiry an abort word is raised in
// every rank's pad and the kernels return), with a monotonically increasing epoch so pads never need
// clearing.
#include "../common.cuh"
#include "b200seg.h"
#include <cuda_fp16.h>
#include <cuda_fp8.h>

namespace b200 {

constexpr int kMaxRanks = 16;
constexpr int kArThreads = 512;

struct PeerPtrs {
  const void* wire[kMaxRanks];     // each rank's wire-format gradient buffer
  float* grad[kMaxRanks];          // each rank's fp32 gradient buffer (two-shot destination)
  uint32_t* signal[kMaxRanks];     // each rank's signal pad: [max_blocks][kMaxRanks]
  const float* mult[kMaxRanks];    // each rank's dequantization multiplier (device, symmetric); NULL = 1
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Word (in uint32 units) of the abort flag inside every rank's signal pad: behind the per-CTA flags and the
// multiplier slot.  Non-zero = some rank gave up waiting (value = 0x80000000 | rank); sticky.
constexpr int kAbortWord = 256 * kMaxRanks + 8;
static unsigned long long g_comm_timeout_ns = 600ull * 1000000000ull;

__device__ __forceinline__ unsigned long long ar_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Every CTA b of every rank meets CTA b of all other ranks.  The wait is bounded by %globaltimer: when a peer
// does not arrive within `timeout_ns` (it stalled or died -- the reference would sit in recv forever, R:98)
// the waiting thread raises the abort word in EVERY rank's pad and the kernel returns without touching the
// gradients; the host turns the word into PeerLost (parallel/symm.py check_abort).  Returns false then.
__device__ __forceinline__ bool cross_gpu_barrier(const PeerPtrs& pp, int rank, int world, uint32_t epoch,
                                                  unsigned long long timeout_ns) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < world) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(pp.signal[peer] + (size_t)blockIdx.x * kMaxRanks + rank, epoch);
    const uint32_t* mine = pp.signal[rank] + (size_t)blockIdx.x * kMaxRanks + peer;
    const uint32_t* my_abort = pp.signal[rank] + kAbortWord;
    unsigned long long t0 = 0;
    for (unsigned int spin = 0; (int)(ld_acquire_sys(mine) - epoch) < 0; ++spin) {
      if ((spin & 255u) == 255u) {
        if (ld_acquire_sys(my_abort) != 0u) { s_ok = 0; break; }
        const unsigned long long now = ar_now();
        if (t0 == 0) t0 = now;
        else if (now - t0 > timeout_ns) {
          for (int r = 0; r < world; ++r) st_release_sys(pp.signal[r] + kAbortWord, 0x80000000u | (uint32_t)rank);
          s_ok = 0;
          break;
        }
      }
    }
  }
  __syncthreads();
  return s_ok != 0;
}

template <typename T> struct Wire;
template <> struct Wire<float> {
  static constexpr int kVec = 4;   // elements per 16 bytes
  __device__ static void load(const void* p, size_t i, float (&f)[16 / sizeof(float)]) {
    float4 v = reinterpret_cast<const float4*>(p)[i];
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
};
template <> struct Wire<__nv_bfloat16> {
  static constexpr int kVec = 8;
  __device__ static void load(const void* p, size_t i, float (&f)[8]) {
    const uint4 v = reinterpret_cast<const uint4*>(p)[i];      // one 16-byte load (a bf16x8 copy became four)
    unpack8(*reinterpret_cast<const bf16x8*>(&v), f);
  }
};
template <> struct Wire<__half> {
  static constexpr int kVec = 8;
  __device__ static void load(const void* p, size_t i, float (&f)[8]) {
    uint4 v = reinterpret_cast<const uint4*>(p)[i];
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) { float2 t = __half22float2(h[j]); f[2 * j] = t.x; f[2 * j + 1] = t.y; }
  }
};
template <> struct Wire<int8_t> {
  static constexpr int kVec = 16;
  __device__ static void load(const void* p, size_t i, float (&f)[16]) {
    uint4 v = reinterpret_cast<const uint4*>(p)[i];
    const int8_t* b = reinterpret_cast<const int8_t*>(&v);
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = (float)b[j];
  }
};
template <> struct Wire<__nv_fp8_e4m3> {
  static constexpr int kVec = 16;
  __device__ static void load(const void* p, size_t i, float (&f)[16]) {
    uint4 v = reinterpret_cast<const uint4*>(p)[i];
    const __nv_fp8_e4m3* b = reinterpret_cast<const __nv_fp8_e4m3*>(&v);
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = float(b[j]);
  }
};

// n_vec = number of 16-byte wire vectors in the whole buffer (the buffer is padded to that).
template <typename T, bool ONESHOT>
__global__ void __launch_bounds__(kArThreads)
allreduce_kernel(PeerPtrs pp, int rank, int world, size_t n_vec, float post_scale, uint32_t epoch,
                 float* __restrict__ oneshot_out, unsigned long long timeout_ns) {
  constexpr int V = Wire<T>::kVec;
  if (!cross_gpu_barrier(pp, rank, world, epoch, timeout_ns)) return;   // every rank's wire buffer is complete
  float mult[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    mult[r] = (r < world && pp.mult[r] != nullptr) ? *reinterpret_cast<const volatile float*>(pp.mult[r]) : 1.f;
  size_t v0 = 0, v1 = n_vec;
  if (!ONESHOT) {
    const size_t per = (n_vec + world - 1) / world;
    v0 = (size_t)rank * per;
    v1 = v0 + per < n_vec ? v0 + per : n_vec;
  }
#pragma unroll 2
  for (size_t i = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < v1;
       i += (size_t)gridDim.x * blockDim.x) {
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    float f[kMaxRanks <= 8 ? 8 : 8][V];                 // issue up to 8 peer loads before using any
#pragma unroll
    for (int r0 = 0; r0 < kMaxRanks; r0 += 8) {
      if (r0 >= world) break;
#pragma unroll
      for (int r = 0; r < 8; ++r)
        if (r0 + r < world) Wire<T>::load(pp.wire[r0 + r], i, f[r]);
#pragma unroll
      for (int r = 0; r < 8; ++r) {                     // fixed order: identical bits on every rank
        if (r0 + r < world) {
          const float m = mult[r0 + r];
#pragma unroll
          for (int j = 0; j < V; ++j) acc[j] = fmaf(f[r][j], m, acc[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] *= post_scale;
    if (ONESHOT) {
#pragma unroll
      for (int q = 0; q < V / 4; ++q)
        reinterpret_cast<float4*>(oneshot_out)[i * (V / 4) + q] =
            make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
    } else {
      for (int r = 0; r < world; ++r) {
        const int dst = (rank + r) % world;            // stagger the peers
#pragma unroll
        for (int q = 0; q < V / 4; ++q)
          reinterpret_cast<float4*>(pp.grad[dst])[i * (V / 4) + q] =
              make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
      }
    }
  }
  // two-shot: all slices have landed everywhere.  one-shot: nobody may overwrite its wire buffer (with fp32 it
  // IS the gradient buffer, which the caller copies the result into next) while a slower peer still reads it.
  cross_gpu_barrier(pp, rank, world, epoch + 1, timeout_ns);
}

template <typename T>
static int launch_ar(const PeerPtrs& pp, int rank, int world, size_t n_elems, float post_scale,
                     uint32_t epoch, float* oneshot_out, int blocks, cudaStream_t st) {
  const size_t n_vec = n_elems / Wire<T>::kVec;
  if (oneshot_out != nullptr)
    allreduce_kernel<T, true><<<blocks, kArThreads, 0, st>>>(pp, rank, world, n_vec, post_scale, epoch, oneshot_out,
                                                             g_comm_timeout_ns);
  else
    allreduce_kernel<T, false><<<blocks, kArThreads, 0, st>>>(pp, rank, world, n_vec, post_scale, epoch, nullptr,
                                                              g_comm_timeout_ns);
  return (int)cudaGetLastError();
}

}  // namespace b200

namespace b200 {

// ---- NVLS: in-switch reduction through a multicast mapping -------------------------------------
// `mc` is the multicast address of the symmetric fp32 gradient buffer (one virtual address that the
// NVSwitch fans out to every rank's copy).  multimem.ld_reduce makes the SWITCH add the element from
// all ranks and return the sum; multimem.st writes the result to all ranks.  Rank r does this for
// slice r only, so every element is reduced once (one owner => identical bits everywhere) and each
// GPU moves ~S/W bytes in and S/W out instead of (W-1)/W * S each way.
template <int U>
__global__ void __launch_bounds__(kArThreads)
allreduce_nvls_kernel(PeerPtrs pp, float* __restrict__ mc, int rank, int world, size_t n_vec, float post_scale,
                      uint32_t epoch, unsigned long long timeout_ns) {
  if (!cross_gpu_barrier(pp, rank, world, epoch, timeout_ns)) return;
  const size_t per = (n_vec + world - 1) / world;
  const size_t v0 = (size_t)rank * per;
  const size_t v1 = v0 + per < n_vec ? v0 + per : n_vec;
  const size_t nthr = (size_t)gridDim.x * blockDim.x;
  // U independent 16-byte switch reductions in flight per thread before the first store
  for (size_t i = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < v1; i += U * nthr) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = i + (size_t)u * nthr;
      if (j < v1)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                     : "l"(mc + j * 4)
                     : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = i + (size_t)u * nthr;
      if (j < v1) {
        v[u].x *= post_scale; v[u].y *= post_scale; v[u].z *= post_scale; v[u].w *= post_scale;
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + j * 4), "f"(v[u].x),
                     "f"(v[u].y), "f"(v[u].z), "f"(v[u].w)
                     : "memory");
      }
    }
  }
  cross_gpu_barrier(pp, rank, world, epoch + 1, timeout_ns);
}

__global__ void barrier_kernel(PeerPtrs pp, int rank, int world, uint32_t epoch, unsigned long long timeout_ns) {
  cross_gpu_barrier(pp, rank, world, epoch, timeout_ns);
}

// ---- quantize-on-send two-shot ---------------------------------------------------------------------
// The lossy wire codecs used to cost a separate encode pass over the whole gradient (read 4 B, write 1-2 B
// per element) before an all-reduce that then PULLED the encoded bytes from the peers.  Here the sender
// quantizes in registers and PUSHES: rank s reads slice p of its own fp32 gradient, encodes it and stores it
// straight into rank p's inbox over NVLink (slot s); after one barrier rank p dequantizes the W slots of its
// slice, adds them in fixed rank order in fp32 (identical bits everywhere, R:402-433) and stores the result
// into every rank's fp32 gradient buffer.  No encode kernel, no write_mult kernel: the only extra launch a
// lossy codec costs is the global max-abs (R:451-463 does that with one host sync per tensor).
// Bytes over NVLink per GPU: (W-1)/W * S/2 (16-bit) or S/4 (8-bit) in, (W-1)/W * S out.
template <int CODEC> struct PushWire;      // codec ids of csrc/elem/codec.cu
template <> struct PushWire<1> { static constexpr int V = 8; };    // bf16
template <> struct PushWire<2> { static constexpr int V = 8; };    // fp16
template <> struct PushWire<3> { static constexpr int V = 8; };    // q201: round(g/max*100) as fp16
template <> struct PushWire<4> { static constexpr int V = 16; };   // q21:  round(g/max*10) as int8
template <> struct PushWire<5> { static constexpr int V = 16; };   // int8s
template <> struct PushWire<6> { static constexpr int V = 16; };   // fp8 e4m3

template <int CODEC>
__device__ __forceinline__ uint4 push_encode(const float* f, float s, float levels) {
  constexpr int V = PushWire<CODEC>::V;
  uint4 out;
  if (CODEC == 1) {
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&out);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = __floats2bfloat162_rn(f[2 * j], f[2 * j + 1]);
  } else if (CODEC == 2 || CODEC == 3) {
    __half2* o = reinterpret_cast<__half2*>(&out);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float a = f[2 * j], b = f[2 * j + 1];
      if (CODEC == 3) {     // (g / max) * S, rounded, as R:487; IEEE division and product for bit parity with torch
        a = (s > 0.f) ? rintf(__fmul_rn(__fdiv_rn(a, s), levels)) : 0.f;
        b = (s > 0.f) ? rintf(__fmul_rn(__fdiv_rn(b, s), levels)) : 0.f;
      }
      o[j] = __floats2half2_rn(a, b);
    }
  } else if (CODEC == 4 || CODEC == 5) {
    int8_t* o = reinterpret_cast<int8_t*>(&out);
#pragma unroll
    for (int j = 0; j < V; ++j) o[j] = (s > 0.f) ? (int8_t)rintf(__fmul_rn(__fdiv_rn(f[j], s), levels)) : (int8_t)0;
  } else {
    __nv_fp8_e4m3* o = reinterpret_cast<__nv_fp8_e4m3*>(&out);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float q = (s > 0.f) ? __fmul_rn(__fdiv_rn(f[j], s), levels) : 0.f;
      o[j] = __nv_fp8_e4m3(fminf(fmaxf(q, -448.f), 448.f));
    }
  }
  return out;
}

template <int CODEC, typename WT>
__global__ void __launch_bounds__(kArThreads)
allreduce_push_kernel(PeerPtrs pp, const float* __restrict__ my_scale, float levels, int rank, int world, size_t n_vec,
                      float post_scale, uint32_t epoch, unsigned long long timeout_ns) {
  constexpr int V = PushWire<CODEC>::V;
  const size_t per = (n_vec + world - 1) / world;                 // vectors per slice
  const float s = my_scale != nullptr ? *my_scale : 1.f;
  // what peers multiply my payload by; EVERY CTA writes it (same value): the barrier below pairs CTA b with the
  // CTAs b of the other ranks only, so CTA b must not depend on a store made by some other CTA of a peer
  if (threadIdx.x == 0 && pp.mult[rank] != nullptr)
    *const_cast<float*>(pp.mult[rank]) = (my_scale != nullptr && levels > 0.f) ? __fdiv_rn(s, levels) : 1.f;
  const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (size_t)gridDim.x * blockDim.x;
  const float4* mine = reinterpret_cast<const float4*>(pp.grad[rank]);
  // phase A: quantize my gradient slice by slice and push slice p into rank p's inbox, slot `rank`
  for (int k = 0; k < world; ++k) {
    const int dst = (rank + k) % world;                             // stagger the peers
    const size_t v0 = (size_t)dst * per;
    const size_t v1 = v0 + per < n_vec ? v0 + per : n_vec;
    uint4* inbox = reinterpret_cast<uint4*>(const_cast<void*>(pp.wire[dst])) + (size_t)rank * per;
    for (size_t i = v0 + tid; i < v1; i += nthr) {
      float f[V];
#pragma unroll
      for (int q4 = 0; q4 < V / 4; ++q4) {
        const float4 t = mine[i * (V / 4) + q4];
        f[4 * q4] = t.x; f[4 * q4 + 1] = t.y; f[4 * q4 + 2] = t.z; f[4 * q4 + 3] = t.w;
      }
      inbox[i - v0] = push_encode<CODEC>(f, s, levels);
    }
  }
  if (!cross_gpu_barrier(pp, rank, world, epoch, timeout_ns)) return;   // every slot of my inbox has landed
  float mult[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    mult[r] = (r < world && pp.mult[r] != nullptr) ? *reinterpret_cast<const volatile float*>(pp.mult[r]) : 1.f;
  // phase B: dequantize the W slots of my slice, fp32 sum in rank order, store into every rank's gradient
  const size_t v0 = (size_t)rank * per;
  const size_t v1 = v0 + per < n_vec ? v0 + per : n_vec;
  const void* my_inbox = pp.wire[rank];
  for (size_t i = v0 + tid; i < v1; i += nthr) {
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    constexpr int G = V == 16 ? 4 : 8;                     // slots loaded before any is used (register budget)
    float f[G][V];
#pragma unroll
    for (int r0 = 0; r0 < kMaxRanks; r0 += G) {
      if (r0 >= world) break;
#pragma unroll
      for (int r = 0; r < G; ++r)
        if (r0 + r < world) Wire<WT>::load(my_inbox, (size_t)(r0 + r) * per + (i - v0), f[r]);
#pragma unroll
      for (int r = 0; r < G; ++r) {
        if (r0 + r < world) {
          const float m = mult[r0 + r];
#pragma unroll
          for (int j = 0; j < V; ++j) acc[j] = fmaf(f[r][j], m, acc[j]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] *= post_scale;
    for (int r = 0; r < world; ++r) {
      float4* dstp = reinterpret_cast<float4*>(pp.grad[(rank + r) % world]);
#pragma unroll
      for (int q4 = 0; q4 < V / 4; ++q4)
        dstp[i * (V / 4) + q4] = make_float4(acc[4 * q4], acc[4 * q4 + 1], acc[4 * q4 + 2], acc[4 * q4 + 3]);
    }
  }
  cross_gpu_barrier(pp, rank, world, epoch + 1, timeout_ns);        // all slices have landed everywhere
}

// Registered multicast alias of the gradient buffer (fused wgrad -> all-reduce, see grad_red_add).
static const char* g_mc_local = nullptr;
static long long g_mc_bytes = 0;
static long long g_mc_delta = 0;

__global__ void write_mult_kernel(const float* scale, float levels, float* out) {
  *out = (levels > 0.f) ? __fdiv_rn(*scale, levels) : 1.f;
}
}  // namespace b200

using namespace b200;

// out[0] = scale / levels (levels <= 0: 1): the multiplier peers apply to this rank's wire buffer.
extern "C" int b200seg_write_mult(const float* scale, float levels, float* out, cudaStream_t st) {
  write_mult_kernel<<<1, 1, 0, st>>>(scale, levels, out);
  return (int)cudaGetLastError();
}

// One cross-GPU barrier on the signal pad (consumes ONE epoch of the same counter the all-reduces use).
extern "C" int b200seg_barrier(unsigned int* const* signal_ptrs, int rank, int world, unsigned int epoch,
                               cudaStream_t st) {
  if (world < 2 || world > kMaxRanks) return (int)cudaErrorInvalidValue;
  PeerPtrs pp{};
  for (int r = 0; r < world; ++r) pp.signal[r] = signal_ptrs[r];
  barrier_kernel<<<1, 32, 0, st>>>(pp, rank, world, epoch, g_comm_timeout_ns);
  return (int)cudaGetLastError();
}

// Fused weight-gradient + all-reduce: register [local_base, local_base + bytes) together with the
// multicast address that aliases it on every rank.  From then on every weight-gradient launcher whose
// destination lies inside the range adds through the multicast alias (one multimem.red per element,
// fanned out by the NVSwitch into all replicas).  multicast_base == NULL unregisters.  Process-wide;
// set it before capturing CUDA graphs (the alias is baked into the kernel arguments).
extern "C" int b200seg_set_grad_multicast(const void* local_base, long long bytes, const void* multicast_base) {
  if (multicast_base == nullptr || local_base == nullptr || bytes <= 0) {
    g_mc_local = nullptr; g_mc_bytes = 0; g_mc_delta = 0;
    return 0;
  }
  g_mc_local = static_cast<const char*>(local_base);
  g_mc_bytes = bytes;
  g_mc_delta = static_cast<const char*>(multicast_base) - static_cast<const char*>(local_base);
  return 0;
}
extern "C" long long b200seg_grad_multicast_delta(const void* p) {
  const char* c = static_cast<const char*>(p);
  return (g_mc_local != nullptr && c >= g_mc_local && c < g_mc_local + g_mc_bytes) ? g_mc_delta : 0;
}

// Bound (seconds) on every device-side wait for a peer; process-wide, read at launch time.
extern "C" void b200seg_set_comm_timeout(double seconds) {
  g_comm_timeout_ns = (unsigned long long)((seconds > 0 ? seconds : 600.0) * 1e9);
}
extern "C" int b200seg_allreduce_abort_word() { return kAbortWord; }

extern "C" int b200seg_allreduce_max_blocks() { return 256; }   // all co-resident: 4 x 512 threads per SM
extern "C" long long b200seg_allreduce_signal_bytes() {
  return (long long)b200seg_allreduce_max_blocks() * kMaxRanks * sizeof(uint32_t);
}

// wire_kind: 0 fp32, 1 bf16, 2 fp16 (also q201), 3 int8 (q21, int8s), 4 fp8.  n_elems must be a
// multiple of 16.  oneshot_out == NULL selects the in-place two-shot algorithm.  Two epochs are
// consumed per call: pass epoch = 2 * call_index + 1.
extern "C" int b200seg_allreduce(const void* const* wire_ptrs, float* const* grad_ptrs,
                                 unsigned int* const* signal_ptrs, const float* const* mult_ptrs, int rank, int world,
                                 long long n_elems, int wire_kind, float post_scale, unsigned int epoch,
                                 float* oneshot_out, int blocks, cudaStream_t st) {
  if (world < 1 || world > kMaxRanks || (n_elems % 16) || blocks < 1 || blocks > b200seg_allreduce_max_blocks())
    return (int)cudaErrorInvalidValue;
  PeerPtrs pp{};
  for (int r = 0; r < world; ++r) {
    pp.wire[r] = wire_ptrs[r];
    pp.grad[r] = grad_ptrs ? grad_ptrs[r] : nullptr;
    pp.signal[r] = signal_ptrs[r];
    pp.mult[r] = mult_ptrs ? mult_ptrs[r] : nullptr;
  }
  switch (wire_kind) {
    case 0: return launch_ar<float>(pp, rank, world, (size_t)n_elems, post_scale, epoch, oneshot_out, blocks, st);
    case 1: return launch_ar<__nv_bfloat16>(pp, rank, world, (size_t)n_elems, post_scale, epoch, oneshot_out, blocks, st);
    case 2: return launch_ar<__half>(pp, rank, world, (size_t)n_elems, post_scale, epoch, oneshot_out, blocks, st);
    case 3: return launch_ar<int8_t>(pp, rank, world, (size_t)n_elems, post_scale, epoch, oneshot_out, blocks, st);
    case 4: return launch_ar<__nv_fp8_e4m3>(pp, rank, world, (size_t)n_elems, post_scale, epoch, oneshot_out, blocks, st);
  }
  return (int)cudaErrorInvalidValue;
}

// Quantize-on-send two-shot.  codec: 1 bf16, 2 fp16, 3 q201, 4 q21, 5 int8s, 6 fp8 (ids of b200seg_codec_*).
// inbox_ptrs[r]: rank r's inbox, n_elems wire elements (rank p's inbox holds slot s = rank s's copy of slice p);
// grad_ptrs[r]: rank r's fp32 gradient buffer (source on this rank, destination on all); mult_ptrs[r]: rank r's
// multiplier slot (written by this kernel); scale: this rank's global max-abs (device), NULL for bf16 / fp16.
// n_elems % 64 == 0.  Consumes two epochs.
extern "C" int b200seg_allreduce_push(void* const* inbox_ptrs, float* const* grad_ptrs, unsigned int* const* signal_ptrs,
                                      float* const* mult_ptrs, const float* scale, int codec, int rank, int world,
                                      long long n_elems, float post_scale, unsigned int epoch, int blocks,
                                      cudaStream_t st) {
  if (world < 2 || world > kMaxRanks || (n_elems % 64) || blocks < 1 || blocks > b200seg_allreduce_max_blocks() ||
      codec < 1 || codec > 6)
    return (int)cudaErrorInvalidValue;
  PeerPtrs pp{};
  for (int r = 0; r < world; ++r) {
    pp.wire[r] = inbox_ptrs[r];
    pp.grad[r] = grad_ptrs[r];
    pp.signal[r] = signal_ptrs[r];
    pp.mult[r] = mult_ptrs ? mult_ptrs[r] : nullptr;
  }
  const float levels = codec == 3 ? 100.f : (codec == 4 ? 10.f : (codec == 5 ? 127.f : (codec == 6 ? 448.f : 0.f)));
  const unsigned long long to = g_comm_timeout_ns;
#define B200_PUSH(C, WT)                                                                                         \
  allreduce_push_kernel<C, WT><<<blocks, kArThreads, 0, st>>>(pp, scale, levels, rank, world,                    \
                                                             (size_t)n_elems / PushWire<C>::V, post_scale, epoch, to)
  switch (codec) {
    case 1: B200_PUSH(1, __nv_bfloat16); break;
    case 2: B200_PUSH(2, __half); break;
    case 3: B200_PUSH(3, __half); break;
    case 4: B200_PUSH(4, int8_t); break;
    case 5: B200_PUSH(5, int8_t); break;
    case 6: B200_PUSH(6, __nv_fp8_e4m3); break;
  }
#undef B200_PUSH
  return (int)cudaGetLastError();
}

// NVLS variant (fp32 wire only).  `multicast_ptr` comes from the symmetric-memory rendezvous; 0 means
// the driver / fabric exposes no multicast and the caller must use b200seg_allreduce.
extern "C" int b200seg_allreduce_nvls(unsigned int* const* signal_ptrs, float* multicast_ptr, int rank, int world,
                                      long long n_elems, float post_scale, unsigned int epoch, int blocks,
                                      cudaStream_t st) {
  if (multicast_ptr == nullptr || world < 2 || world > kMaxRanks || (n_elems % 16) || blocks < 1 ||
      blocks > b200seg_allreduce_max_blocks())
    return (int)cudaErrorInvalidValue;
  PeerPtrs pp{};
  for (int r = 0; r < world; ++r) pp.signal[r] = signal_ptrs[r];
  static int unroll = -1;             // B200SEG_NVLS_UNROLL=4|8: switch reductions in flight per thread
  if (unroll < 0) { const char* e = getenv("B200SEG_NVLS_UNROLL"); unroll = (e != nullptr && e[0] == '8') ? 8 : 4; }
  if (unroll == 8)
    allreduce_nvls_kernel<8><<<blocks, kArThreads, 0, st>>>(pp, multicast_ptr, rank, world, (size_t)n_elems / 4, post_scale,
                                                           epoch, g_comm_timeout_ns);
  else
    allreduce_nvls_kernel<4><<<blocks, kArThreads, 0, st>>>(pp, multicast_ptr, rank, world, (size_t)n_elems / 4, post_scale,
                                                           epoch, g_comm_timeout_ns);
  return (int)cudaGetLastError();
}
