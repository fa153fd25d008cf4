// b200c_device.cuh -- device-side building blocks for the B200 (sm_100a) TP hot path.
//
// Memory-system facts this code is written around (B200, NVLink 5 / NVSwitch):
//   * peer stores are posted (fire-and-forget) while peer loads pay a full fabric round
//     trip (~1.8k cycles), so every exchange here is a PUSH: a rank reads its own input
//     from local HBM and writes into the peers' arenas;
//   * peer data bypasses the local L2, and L1 is not coherent, so every wait loop polls
//     with ld.volatile (L1 bypass) on LOCAL memory that a peer writes;
//   * the data word itself is the arrival flag (a 32-bit "not arrived" pattern that the
//     payload can never contain), so the all-reduce needs no fence and no barrier.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200c {

constexpr int kMaxRanks = 8;
// A 32-bit word that marks "not arrived yet".  For fp32 it is -0.0f; for the 16-bit
// types it is the aligned pair (+0.0, -0.0).  Payload words equal to it are rewritten to
// +0.0 before they are pushed (numerically the same value), so a receiver can never
// mistake data for the marker.  PTX models a vector access as independent accesses of
// its element type, so a .b32 element never tears.
constexpr uint32_t kNotArrived = 0x80000000u;

// ---- raw memory operations -----------------------------------------------------------

__device__ __forceinline__ uint4 ld_poll_v4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

// Streaming read of data nobody writes during the kernel (inputs, residual, gamma).
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ uint4 ld_v4(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}

// Push to a (possibly remote) arena.  volatile keeps the store where it is written and
// stops the compiler from merging or eliding it; no ordering is needed (data == flag).
__device__ __forceinline__ void st_push_v4(void* p, const uint4& v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ void st_v4(void* p, const uint4& v) {
  *reinterpret_cast<uint4*>(p) = v;
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// ---- NVLS (NVSwitch multicast) ------------------------------------------------------------------
// `mc` addresses come from a multicast mapping of a buffer that exists on every rank: a store
// to one lands in every rank's copy (the switch replicates it), a ld_reduce returns the sum of
// every rank's copy, added inside the switch (SASS: LDGMC.*ADD, STG on a multicast address).
__device__ __forceinline__ void mc_st_v4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st_release_u32(void* mc, uint32_t v) {
  asm volatile("multimem.st.release.sys.global.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Narrow (4- and 8-byte) copies for rows that are not a multiple of 16 bytes, e.g. the
// [batch, 2] fp32 (value, index) pairs of the vocab-parallel argmax.
template <int BYTES>
struct Word;
template <>
struct Word<16> {
  using type = uint4;
  static __device__ __forceinline__ uint4 ld(const void* p) { return ld_stream_v4(p); }
  static __device__ __forceinline__ uint4 poll(const void* p) { return ld_poll_v4(p); }
  static __device__ __forceinline__ void push(void* p, const uint4& v) { st_push_v4(p, v); }
  static __device__ __forceinline__ void st(void* p, const uint4& v) { st_v4(p, v); }
};
template <>
struct Word<8> {
  using type = uint2;
  static __device__ __forceinline__ uint2 ld(const void* p) { return *reinterpret_cast<const uint2*>(p); }
  static __device__ __forceinline__ uint2 poll(const void* p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
  }
  static __device__ __forceinline__ void push(void* p, const uint2& v) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
  }
  static __device__ __forceinline__ void st(void* p, const uint2& v) { *reinterpret_cast<uint2*>(p) = v; }
};
template <>
struct Word<4> {
  using type = uint32_t;
  static __device__ __forceinline__ uint32_t ld(const void* p) { return *reinterpret_cast<const uint32_t*>(p); }
  static __device__ __forceinline__ uint32_t poll(const void* p) { return ld_volatile_u32(reinterpret_cast<const uint32_t*>(p)); }
  static __device__ __forceinline__ void push(void* p, const uint32_t& v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
  }
  static __device__ __forceinline__ void st(void* p, const uint32_t& v) { *reinterpret_cast<uint32_t*>(p) = v; }
};

// ---- arrival marker helpers ----------------------------------------------------------------

__device__ __forceinline__ bool not_arrived(const uint4& v) {
  return (v.x == kNotArrived) | (v.y == kNotArrived) | (v.z == kNotArrived) |
         (v.w == kNotArrived);
}

__device__ __forceinline__ void scrub(uint4& v) {
  v.x = (v.x == kNotArrived) ? 0u : v.x;
  v.y = (v.y == kNotArrived) ? 0u : v.y;
  v.z = (v.z == kNotArrived) ? 0u : v.z;
  v.w = (v.w == kNotArrived) ? 0u : v.w;
}

// ---- bounded waiting ----------------------------------------------------------------------------
// A wait on a peer never spins forever: after timeout_ns (0 = unbounded) the sticky error
// word is set.  What happens next is the caller's policy (`trap` in the kernel parameters):
//   trap == 0  every later wait in this and following kernels falls through at once; results
//              are garbage from then on and the host must poll b200c_query_error();
//   trap != 0  the kernel executes `trap`: the CUDA context is dead, the next CUDA call of the
//              process fails, and the serving loop sees a crashed worker instead of wrong tokens.
constexpr uint32_t kErrTimeout = 1u;
constexpr uint32_t kErrClusterShape = 2u;

struct SpinGuard {
  uint32_t iters = 0;
  unsigned long long t0 = 0;
  // returns true when the caller must give up; P is RankParams or CopyParams
  template <typename P>
  __device__ __forceinline__ bool expired(const P& p) {
    if ((++iters & 0x3ffu) != 0) return false;
    uint32_t* err = &p.st->error;
    if (ld_volatile_u32(err) != 0) return true;
    if (p.timeout_ns == 0) return false;
    unsigned long long now = globaltimer_ns();
    if (t0 == 0) {
      t0 = now;
      return false;
    }
    if (now - t0 > p.timeout_ns) {
      atomicExch(err, kErrTimeout);
      if (p.trap) {
        __threadfence_system();
        asm volatile("trap;");
      }
      return true;
    }
    return false;
  }
};

// ---- element types ---------------------------------------------------------------------------------

template <typename T>
struct VecOps;

template <>
struct VecOps<__nv_bfloat16> {
  static constexpr int EPV = 8;  // elements per 16-byte vector
  static __device__ __forceinline__ void to_float(const uint4& v, float (&f)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ uint4 from_float(const float (&f)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  // round an fp32 value to T and back (the x.to(weight.dtype) step of the norm)
  static __device__ __forceinline__ float round_trip(float x) {
    return __bfloat162float(__float2bfloat16_rn(x));
  }
  // in-switch sum over all ranks of one 16-byte vector, fp32 accumulation, one rounding
  static __device__ __forceinline__ uint4 mc_ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};

template <>
struct VecOps<__half> {
  static constexpr int EPV = 8;
  static __device__ __forceinline__ void to_float(const uint4& v, float (&f)[8]) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 p = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      f[2 * i] = p.x;
      f[2 * i + 1] = p.y;
    }
  }
  static __device__ __forceinline__ uint4 from_float(const float (&f)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  static __device__ __forceinline__ float round_trip(float x) {
    return __half2float(__float2half_rn(x));
  }
  static __device__ __forceinline__ uint4 mc_ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};

template <>
struct VecOps<float> {
  static constexpr int EPV = 4;
  static __device__ __forceinline__ void to_float(const uint4& v, float (&f)[4]) {
    f[0] = __uint_as_float(v.x);
    f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z);
    f[3] = __uint_as_float(v.w);
  }
  static __device__ __forceinline__ uint4 from_float(const float (&f)[4]) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
  static __device__ __forceinline__ float round_trip(float x) { return x; }
  static __device__ __forceinline__ uint4 mc_ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
  }
};

// ---- block-wide sum (warp shuffles, one shared-memory hop) -------------------------------------
// Fixed reduction tree => the same row gives the same bits on every rank.  `smem` holds two
// banks of 32 floats used alternately (parity flips per call), so one call costs ONE barrier:
// a bank is rewritten two calls later, and reaching that call means every warp has passed the
// barrier of the call in between, i.e. has finished reading the bank.
__device__ __forceinline__ float block_sum(float v, float* smem /* 64 floats */, int& parity) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  float* bank = smem + parity * 32;
  parity ^= 1;
  if (lane == 0) bank[warp] = v;
  __syncthreads();
  float t = (lane < nwarps) ? bank[lane] : 0.0f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;  // every warp reduces the same 32 partials in the same order
}

// R independent block-wide sums behind ONE barrier (R rows per trip in the HBM-bound norm).
// smem: 2 parities x R x 32 floats.
template <int R>
__device__ __forceinline__ void block_sum_multi(float (&v)[R], float* smem, int& parity) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int k = 0; k < R; ++k) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = (blockDim.x + 31) >> 5;
  float* bank = smem + parity * (R * 32);
  parity ^= 1;
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < R; ++k) bank[k * 32 + warp] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < R; ++k) {
    float t = (lane < nwarps) ? bank[k * 32 + lane] : 0.0f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    v[k] = t;
  }
}

// ---- thread-block clusters: rows split over the CTAs of a cluster -------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// write a float into the same shared-memory variable of CTA `rank` of this cluster (DSMEM)
__device__ __forceinline__ void st_dsmem_f32(float* local_smem_ptr, uint32_t rank, float v) {
  const uint32_t local = (uint32_t)__cvta_generic_to_shared(local_smem_ptr);
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}

// ---- named barriers: producers arrive and move on, one consumer warp waits ----------------------
__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- programmatic dependent launch (no-ops when the launch did not ask for it) -----------------
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

}  // namespace b200c
