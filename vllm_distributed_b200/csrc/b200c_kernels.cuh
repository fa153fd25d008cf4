// b200c_kernels.cuh -- kernels of the TP hot path (all-reduce [+ residual + RMSNorm],
// lm_head all-gather, pipeline send/recv).  See DESIGN.md for the protocol proofs.
#pragma once

#include "b200c_device.cuh"

// Resident CTAs per SM the two-shot kernels are compiled for (register cap = 65536 / 512 / this).
#ifndef B200C_TWOSHOT_CTAS_PER_SM
#define B200C_TWOSHOT_CTAS_PER_SM 2  // measured: the kernels are bound by bytes in flight per SM
#endif

namespace b200c {

// ---- arena layout (bytes from the arena base; identical on every rank) -----------------------
constexpr size_t kCtrlBytes = 64 * 1024;
constexpr size_t kFlagStride = 128;  // one flag per 128-byte line
constexpr size_t kAgFlagOff = 0;          // [2 parity][8 src]
constexpr size_t kP2pReadyOff = 4096;     // [8 src][2 slot]   written by the sender
constexpr size_t kP2pCreditOff = 8192;    // [8 dst][2 slot]   written by the receiver

struct ArenaGeom {
  size_t lam_stride;  // bytes of one rotating receive region (lo half + hi half)
  size_t hi_off;      // offset of the hi (gather) half inside a region
  size_t ag_off;      // offset of all-gather staging [2]
  size_t ag_stride;
  size_t p2p_off;     // offset of mailboxes [world src][2 slot]
  size_t p2p_stride;
};

// Private (not peer-visible) per-rank device state.  Lives in device memory so that a
// captured CUDA graph replays correctly: the kernels advance it themselves.
struct LocalState {
  unsigned long long call;  // (epoch << 32) | tickets taken in the running call; region = epoch % 3
  uint32_t error;        // sticky: 1 = a wait on a peer timed out
  uint32_t ag_seq;       // all-gather call counter
  uint32_t ag_arrive_push;
  uint32_t ag_arrive_done;
  uint32_t p2p_arrive_send;  // send and recv kernels of one ctx may run on different streams
  uint32_t p2p_arrive_recv;
  unsigned long long dirty_lo[3];  // bytes peers wrote into each region's lo / hi half
  unsigned long long dirty_hi[3];
  uint32_t send_seq[kMaxRanks];
  uint32_t recv_seq[kMaxRanks];
  // NVLS variant: CTA arrival counters of the copy-in groups and of the reduce slices
  uint32_t nv_arrive[16];
  uint32_t nv_arrive_red[16];
};

struct RankParams {
  char* arena[kMaxRanks];  // every rank's arena as mapped in this process
  LocalState* st;
  ArenaGeom g;
  const void* in;
  void* out;  // plain all-reduce result
  const void* res_in;
  void* res_out;
  void* norm_out;
  const void* gamma;
  float eps;
  int rank, world;
  int tokens;
  int hv;                  // 16-byte vectors per row
  int hidden;              // elements per row
  unsigned long long nvec;  // total 16-byte vectors
  unsigned long long timeout_ns;
  int trap;         // policy when a wait times out (see SpinGuard)
  int link_blocks;  // two-shot: CTAs [0, link_blocks) do the link work, the rest the rows
  // NVLS variant (see nvls_body): this rank's symmetric buffer through its own (unicast)
  // mapping and through the multicast mapping, bytes per region, pipeline groups
  char* nv_uc;
  char* nv_mc;
  size_t nv_half;
  int nv_groups;
  int nv_copy_blocks;  // unused (every non-reduce CTA copies in, then consumes)
  int nv_teams;        // reduce CTAs work in this many teams, each on its own group
  int nv_debug;        // measurement only (B200C_TUNE_NVLS_DEBUG): skip phases, results are garbage
};

struct EmuParams {
  RankParams r[kMaxRanks];
};

__device__ __forceinline__ char* lam_region(const RankParams& p, int rank, uint32_t b) {
  return p.arena[rank] + kCtrlBytes + (size_t)b * p.g.lam_stride;
}

// ---- waiting on pushed data -------------------------------------------------------------------------

// Poll one local vector until the peer's push has landed.
__device__ __forceinline__ uint4 wait_one(const char* addr, const RankParams& p, SpinGuard& sg) {
  uint4 v = ld_poll_v4(addr);
  while (not_arrived(v)) {
    if (sg.expired(p)) break;
    v = ld_poll_v4(addr);
  }
  return v;
}

// Gather-and-sum of vector `idx`: slot r of my region holds rank r's copy; my own contribution
// comes straight from `in`.  Polls are issued BATCH peers at a time so their L2 round trips
// overlap (BATCH = 8: all at once, for the latency-bound one-shot; 4 keeps registers down in
// the bandwidth-bound two-shot); accumulation is fp32 in rank order 0..N-1 with one rounding
// (the reference's packed_reduce order, vllm_custom_all_reduce.cuh:213-221).
template <typename T, int BATCH>
__device__ __forceinline__ uint4 gather_sum(const char* my_lo, size_t slot_bytes, size_t idx,
                                            const uint4& own, const RankParams& p,
                                            SpinGuard& sg) {
  constexpr int EPV = VecOps<T>::EPV;
  float acc[EPV], f[EPV];
  const char* base = my_lo + idx * 16;
#pragma unroll
  for (int r0 = 0; r0 < kMaxRanks; r0 += BATCH) {
    if (r0 < p.world) {
      uint4 v[BATCH];
      bool pending;
      do {
        pending = false;
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
          const int r = r0 + k;
          if (r < p.world && r != p.rank) v[k] = ld_poll_v4(base + (size_t)r * slot_bytes);
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
          const int r = r0 + k;
          if (r < p.world && r != p.rank) pending |= not_arrived(v[k]);
        }
        if (pending && sg.expired(p)) break;
      } while (pending);
#pragma unroll
      for (int k = 0; k < BATCH; ++k) {
        const int r = r0 + k;
        if (r < p.world) {
          const uint4 x = (r == p.rank) ? own : v[k];
          if (r == 0) {
            VecOps<T>::to_float(x, acc);
          } else {
            VecOps<T>::to_float(x, f);
#pragma unroll
            for (int e = 0; e < EPV; ++e) acc[e] += f[e];
          }
        }
      }
    }
  }
  return VecOps<T>::from_float(acc);
}

// Re-arm the region the PREVIOUS call used (nobody can push into it again before this
// kernel has finished on this rank; see DESIGN.md "rotation").
__device__ __forceinline__ void rearm(char* base, unsigned long long bytes, size_t gtid,
                                      size_t gsize) {
  const uint4 m = make_uint4(kNotArrived, kNotArrived, kNotArrived, kNotArrived);
  const size_t n = (size_t)(bytes >> 4);
  for (size_t i = gtid; i < n; i += gsize) st_v4(base + i * 16, m);
}


// ---- NVLS buffer layout (bytes from its base; identical on every rank) ---------------------------
// [0, 16K)       flagA [16 groups][8 src], one per 128-byte line: flagA[g][r] = epoch+1 of the last
//                call in which rank r has finished copying group g of its input into xin
// [16K, 32K)     flagB [16 groups][8 owner]: flagB[g][o] = epoch+1 of the last call in which owner o
//                has broadcast the sums of slice g of its chunk
//                (both are written through the multicast mapping: one store updates every rank)
// 1 x nv_half    xin: this rank's input, copied in so that the switch can read it; the sums are
//                written back IN PLACE (by multicast) and consumed from here
constexpr size_t kNvCtrlBytes = 32 * 1024;
constexpr size_t kNvFlagBOff = 16 * 1024;
constexpr int kNvMaxGroups = 14;  // one named barrier id (1 + g) per group, ids 1..14

__device__ __forceinline__ char* nv_flag(char* base, int group, int src) {
  return base + ((size_t)group * kMaxRanks + src) * kFlagStride;
}

// Re-arm the arena region the previous call dirtied (nothing after an NVLS call, which keeps
// its data in the symmetric buffer and signals with flags).
__device__ __forceinline__ void rearm_previous(const RankParams& p, uint32_t pb, size_t tid,
                                               size_t nthreads) {
  char* prev = lam_region(p, p.rank, pb);
  rearm(prev, p.st->dirty_lo[pb], tid, nthreads);
  rearm(prev + p.g.hi_off, p.st->dirty_hi[pb], tid, nthreads);
}

// Call bookkeeping without an end-of-kernel rendezvous and without a fence: the epoch (high
// half) and the tickets of the running call (low half) share one 64-bit word.  A CTA reads the
// epoch, and only when all its threads have it does thread 0 take a ticket.  The CTA holding the
// last ticket therefore knows that every CTA of this rank has already read the epoch and may
// publish (epoch+1, 0 tickets) and the extents this call dirties whenever it likes -- the next
// kernel is the only reader.
struct CallState {
  uint32_t epoch;
  uint32_t ticket;  // meaningful in thread 0 only
};

__device__ __forceinline__ CallState begin_call(const RankParams& p) {
  // Every thread reads the epoch itself (one L2 hit, no shared-memory broadcast) and consumes
  // the value before the barrier, so when thread 0 takes the CTA's ticket afterwards every
  // thread of the CTA has finished reading it.  The ticket's own latency is off the critical
  // path: its value is needed only in end_call.
  CallState cs;
  cs.ticket = 0;
  const uint32_t* hi = reinterpret_cast<const uint32_t*>(&p.st->call) + 1;  // little endian
  cs.epoch = ld_volatile_u32(hi);
  // The barrier takes a predicate computed from the loaded value, so no thread can arrive
  // before its epoch load has returned (the result of the OR itself is irrelevant).
  (void)__syncthreads_or(cs.epoch == 0xffffffffu);
  if (threadIdx.x == 0) cs.ticket = (uint32_t)atomicAdd(&p.st->call, 1ull);
  return cs;
}

__device__ __forceinline__ void end_call(const RankParams& p, const CallState& cs,
                                         unsigned long long dlo, unsigned long long dhi) {
  if (threadIdx.x == 0 && cs.ticket == gridDim.x - 1) {
    const uint32_t b = cs.epoch % 3u;
    p.st->dirty_lo[b] = dlo;
    p.st->dirty_hi[b] = dhi;
    // 0xffffffff is a multiple of 3: wrapping from 0xfffffffe straight to 0 keeps the region
    // sequence ... 1, 2, 0 ... intact after 2^32 calls (about a week of 70B decode)
    const uint32_t next = (cs.epoch + 1u == 0xffffffffu) ? 0u : cs.epoch + 1u;
    p.st->call = (unsigned long long)next << 32;
  }
}

// ---- pushes ----------------------------------------------------------------------------------------
// Copy vectors [lo, hi) of `in` to dst + (i - lo)*16, `U` independent loads in flight per
// thread (the loop is latency-bound otherwise: one 16-byte load per ~1 us per thread).
template <int U>
__device__ __forceinline__ void push_range(const char* in, char* dst, size_t lo, size_t hi,
                                           size_t tid, size_t nthreads) {
  for (size_t base = lo + tid; base < hi; base += nthreads * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * nthreads;
      if (i < hi) v[u] = ld_stream_v4(in + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * nthreads;
      if (i < hi) {
        scrub(v[u]);
        st_push_v4(dst + (i - lo) * 16, v[u]);
      }
    }
  }
}

// Two-shot scatter: for each offset inside a chunk, the vectors bound for the NP peers are loaded
// together and pushed together, so every peer link is fed at the same time and a small message
// costs one load round trip instead of one per peer.  UO offsets are handled per trip so that a
// thread keeps UO*NP independent 16-byte loads in flight whatever the group size (2 ranks: 8x1,
// <=4 ranks: 3x3, else 1x7).  Slot k is the k-th peer to the right of this rank.
template <int UO, int NP>
__device__ __forceinline__ void scatter_chunks_t(const RankParams& p, uint32_t b, size_t chunk,
                                                 size_t nvec, size_t tid, size_t nthreads) {
  const size_t slot = (size_t)p.rank * chunk * 16;
  for (size_t base = tid; base < chunk; base += nthreads * UO) {
    uint4 v[UO][NP];
#pragma unroll
    for (int u = 0; u < UO; ++u) {
      const size_t off = base + (size_t)u * nthreads;
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        if (k + 1 < p.world && off < chunk) {
          int owner = p.rank + k + 1;
          owner = owner >= p.world ? owner - p.world : owner;
          const size_t idx = (size_t)owner * chunk + off;
          if (idx < nvec) v[u][k] = ld_stream_v4((const char*)p.in + idx * 16);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UO; ++u) {
      const size_t off = base + (size_t)u * nthreads;
#pragma unroll
      for (int k = 0; k < NP; ++k) {
        if (k + 1 < p.world && off < chunk) {
          int owner = p.rank + k + 1;
          owner = owner >= p.world ? owner - p.world : owner;
          const size_t idx = (size_t)owner * chunk + off;
          if (idx < nvec) {
            scrub(v[u][k]);
            st_push_v4(lam_region(p, owner, b) + slot + off * 16, v[u][k]);
          }
        }
      }
    }
  }
}

__device__ __forceinline__ void scatter_chunks(const RankParams& p, uint32_t b, size_t chunk,
                                               size_t nvec, size_t tid, size_t nthreads) {
  if (nvec > (size_t)(8u << 20) / 16 && p.world > 2) {
    // above 8 MiB (measured on 8 x B200) streaming one contiguous range per peer, peer after
    // peer starting at the right-hand neighbour, beats seven interleaved streams
    for (int k = 1; k < p.world; ++k) {
      int owner = p.rank + k;
      owner = owner >= p.world ? owner - p.world : owner;
      const size_t lo = (size_t)owner * chunk;
      const size_t hi = (lo + chunk < nvec) ? lo + chunk : nvec;
      if (lo < hi)
        push_range<4>((const char*)p.in, lam_region(p, owner, b) + (size_t)p.rank * chunk * 16, lo,
                      hi, tid, nthreads);
    }
    return;
  }
  if (p.world == 2)
    scatter_chunks_t<8, 1>(p, b, chunk, nvec, tid, nthreads);
  else if (p.world <= 4)
    scatter_chunks_t<3, 3>(p, b, chunk, nvec, tid, nthreads);
  else
    scatter_chunks_t<1, 7>(p, b, chunk, nvec, tid, nthreads);
}

// ---- two-shot phase B: reduce my chunk, push the rounded sums into every rank's hi half -----------
// UB vectors x NP peers are polled together so the local round trips overlap; the (UB, NP)
// shape is picked from the group size at run time (2 ranks: 4x1, <=4 ranks: 2x3, else 1x7).
template <typename T, int UB, int NP>
__device__ __forceinline__ void reduce_gather_range(const RankParams& p, uint32_t b,
                                                    const char* my_lo, size_t chunk, size_t lo,
                                                    size_t hi, size_t tid, size_t nthreads,
                                                    SpinGuard& sg) {
  constexpr int EPV = VecOps<T>::EPV;
  const size_t slot_bytes = chunk * 16;
  for (size_t base = lo + tid; base < hi; base += nthreads * UB) {
    uint4 own[UB], v[UB][NP];
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const size_t i = base + (size_t)u * nthreads;
      if (i < hi) own[u] = ld_stream_v4((const char*)p.in + i * 16);
    }
    bool pending;
    do {
      pending = false;
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const size_t i = base + (size_t)u * nthreads;
        if (i < hi) {
#pragma unroll
          for (int k = 0; k < NP; ++k) {
            // peer k of this rank, in rank order with the own rank skipped
            const int r = k + (k >= p.rank ? 1 : 0);
            if (r < p.world) v[u][k] = ld_poll_v4(my_lo + (size_t)r * slot_bytes + (i - lo) * 16);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UB; ++u) {
        const size_t i = base + (size_t)u * nthreads;
        if (i < hi) {
#pragma unroll
          for (int k = 0; k < NP; ++k) {
            const int r = k + (k >= p.rank ? 1 : 0);
            if (r < p.world) pending |= not_arrived(v[u][k]);
          }
        }
      }
      if (pending && sg.expired(p)) break;
    } while (pending);
#pragma unroll
    for (int u = 0; u < UB; ++u) {
      const size_t i = base + (size_t)u * nthreads;
      if (i < hi) {
        // fp32 accumulation in rank order 0..N-1: peer slots are in rank order with my own
        // rank skipped, so my contribution goes in just before peer slot k == rank.  (All
        // array indices are compile-time constants: nothing is spilled to local memory.)
        float acc[EPV], f[EPV];
        bool first = true;
#pragma unroll
        for (int k = 0; k <= NP; ++k) {
          if (k == p.rank) {
            if (first) {
              VecOps<T>::to_float(own[u], acc);
              first = false;
            } else {
              VecOps<T>::to_float(own[u], f);
#pragma unroll
              for (int e = 0; e < EPV; ++e) acc[e] += f[e];
            }
          }
          if (k < NP) {
            const int r = k + (k >= p.rank ? 1 : 0);
            if (r < p.world) {
              if (first) {
                VecOps<T>::to_float(v[u][k], acc);
                first = false;
              } else {
                VecOps<T>::to_float(v[u][k], f);
#pragma unroll
                for (int e = 0; e < EPV; ++e) acc[e] += f[e];
              }
            }
          }
        }
        uint4 s = VecOps<T>::from_float(acc);
        scrub(s);
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < p.world) st_push_v4(lam_region(p, r, b) + p.g.hi_off + i * 16, s);
      }
    }
  }
}

template <typename T>
__device__ __forceinline__ void reduce_gather(const RankParams& p, uint32_t b, const char* my_lo,
                                              size_t chunk, size_t nvec, size_t tid,
                                              size_t nthreads, SpinGuard& sg) {
  const size_t lo = (size_t)p.rank * chunk;
  const size_t hi = (lo + chunk < nvec) ? lo + chunk : nvec;
  if (lo >= hi) return;
  if (p.world == 2)
    reduce_gather_range<T, 4, 1>(p, b, my_lo, chunk, lo, hi, tid, nthreads, sg);
  else if (p.world <= 4)
    reduce_gather_range<T, 2, 3>(p, b, my_lo, chunk, lo, hi, tid, nthreads, sg);
  else
    reduce_gather_range<T, 1, 7>(p, b, my_lo, chunk, lo, hi, tid, nthreads, sg);
}


// ---- consumers walk the result in the order it ARRIVES ------------------------------------------
// Every rank reduces its own chunk front to back at the same time, so at any moment the first
// part of EVERY chunk has arrived.  A consumer that walks rows (or vectors) 0, 1, 2, ... waits
// on the tail of chunk 0 while the heads of chunks 1..N-1 sit ready, and is left with (N-1)/N of
// its work when the last sum lands.  ArrivalOrder interleaves the chunks slice by slice
// (G slices per chunk): position q -> unit = owner * per_chunk + slice * per_slice + j.
// Only the ORDER depends on it; which units exist does not (out-of-range positions return -1).
struct ArrivalOrder {
  uint32_t per_chunk, per_slice, per_round, total, units;
  __device__ __forceinline__ ArrivalOrder(uint32_t units_, uint32_t world, uint32_t slices) {
    units = units_;
    per_chunk = (units_ + world - 1) / world;
    per_slice = (per_chunk + slices - 1) / slices;
    per_round = per_slice * world;
    total = per_round * slices;
  }
  __device__ __forceinline__ long long at(uint32_t q) const {
    const uint32_t g = q / per_round, rem = q - g * per_round;
    const uint32_t o = rem / per_slice, j = rem - o * per_slice;
    const uint32_t within = g * per_slice + j;
    if (within >= per_chunk) return -1;
    const uint32_t u = o * per_chunk + within;
    return u < units ? (long long)u : -1;
  }
};
constexpr uint32_t kArrivalSlices = 8;

// ---- row epilogue: residual add + RMSNorm ------------------------------------------------------
// s[j] holds the T-rounded all-reduce sum for vector (tid + j*blockDim) of the row, res[j] the
// residual vector (prefetched before the wait), gam[j] the weight vector (loaded once per CTA).
// VLLM/ir/ops/layernorm.py:39-58.
template <typename T, int VPT>
__device__ __forceinline__ void row_epilogue(const uint4 (&s)[VPT], const uint4 (&res)[VPT],
                                             const uint4 (&gam)[VPT], size_t row_vec0,
                                             const RankParams& p, float* smem, int& parity) {
  constexpr int EPV = VecOps<T>::EPV;
  float x[VPT][EPV];
  float ss = 0.0f;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < p.hv) {
      VecOps<T>::to_float(s[j], x[j]);
      if (p.res_in != nullptr) {
        float r[EPV];
        VecOps<T>::to_float(res[j], r);
#pragma unroll
        for (int e = 0; e < EPV; ++e) x[j][e] = x[j][e] + r[e];
        st_v4((char*)p.res_out + (row_vec0 + c) * 16, VecOps<T>::from_float(x[j]));
      }
#pragma unroll
      for (int e = 0; e < EPV; ++e) ss += x[j][e] * x[j][e];
    }
  }
  const float total = block_sum(ss, smem, parity);
  const float rs = 1.0f / sqrtf(total / (float)p.hidden + p.eps);
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < p.hv) {
      float g[EPV], y[EPV];
      VecOps<T>::to_float(gam[j], g);
#pragma unroll
      for (int e = 0; e < EPV; ++e) y[e] = VecOps<T>::round_trip(x[j][e] * rs) * g[e];
      st_v4((char*)p.norm_out + (row_vec0 + c) * 16, VecOps<T>::from_float(y));
    }
  }
}

template <typename T, int VPT>
__device__ __forceinline__ void load_gamma(uint4 (&gam)[VPT], const RankParams& p) {
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < p.hv) gam[j] = ld_v4((const char*)p.gamma + (size_t)c * 16);
  }
}

template <int VPT>
__device__ __forceinline__ void load_residual(uint4 (&res)[VPT], size_t row_vec0,
                                              const RankParams& p) {
  if (p.res_in == nullptr) return;
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const int c = threadIdx.x + j * blockDim.x;
    if (c < p.hv) res[j] = ld_stream_v4((const char*)p.res_in + (row_vec0 + c) * 16);
  }
}

// ---- fused all-reduce + residual + RMSNorm ---------------------------------------------------
//
// ONE-SHOT (latency regime): every CTA pushes its rows into slot[rank] of every peer's region,
// then polls slot[r] for all r, sums, and runs the epilogue.  One fabric traversal, no
// barrier.  Link traffic (N-1)*S out per rank.
//
// TWO-SHOT (bandwidth regime): CTAs are specialised.  LINK CTAs [0, link_blocks) push vector
// chunk j of the input to rank j (scatter), then reduce this rank's chunk and push the
// T-rounded sums into the hi half of EVERY rank's region (gather).  ROW CTAs re-arm the
// previous region, then poll the gathered sums row by row and run the epilogue on all rows
// (local HBM work that overlaps the link work instead of following it).  Link traffic
// 2*(N-1)/N*S out per rank.  CTAs wait only on data whose producers never wait on them, and
// the grid is sized to be co-resident, so the kernel cannot deadlock.
template <typename T, int VPT, bool TWOSHOT>
__device__ __forceinline__ void ar_norm_body(const RankParams& p, float* smem) {
  pdl_launch_dependents();
  uint4 gam[VPT];
  load_gamma<T, VPT>(gam, p);  // weights: safe to read before the previous kernel has finished
  pdl_wait();

  const CallState cs = begin_call(p);
  const uint32_t epoch = cs.epoch;
  const uint32_t b = epoch % 3u, pb = (b + 2u) % 3u;  // from b: epoch + 2 may overflow
  const size_t nvec = (size_t)p.nvec;
  const size_t chunk = (nvec + p.world - 1) / p.world;  // two-shot: vectors owned per rank
  char* my_lo = lam_region(p, p.rank, b);
  char* my_hi = my_lo + p.g.hi_off;
  SpinGuard sg;
  int parity = 0;

  const int link_blocks = TWOSHOT ? p.link_blocks : 0;
  const bool specialised = TWOSHOT && link_blocks > 0 && link_blocks < (int)gridDim.x;
  const bool is_link = !specialised || (int)blockIdx.x < link_blocks;
  const bool is_row = !specialised || (int)blockIdx.x >= link_blocks;
  const int row0 = specialised ? (int)blockIdx.x - link_blocks : (int)blockIdx.x;
  const int row_stride = specialised ? (int)gridDim.x - link_blocks : (int)gridDim.x;
  const size_t ltid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // valid for link CTAs
  const size_t lsize = (size_t)(specialised ? link_blocks : gridDim.x) * blockDim.x;
  const size_t rtid = (size_t)row0 * blockDim.x + threadIdx.x;
  const size_t rsize = (size_t)row_stride * blockDim.x;

  // ---- push
  if (!TWOSHOT) {
    const size_t slot_off = (size_t)p.rank * nvec * 16;
    for (int row = blockIdx.x; row < p.tokens; row += gridDim.x) {
#pragma unroll
      for (int j = 0; j < VPT; ++j) {
        const int c = threadIdx.x + j * blockDim.x;
        if (c < p.hv) {
          const size_t idx = (size_t)row * p.hv + c;
          uint4 v = ld_stream_v4((const char*)p.in + idx * 16);
          scrub(v);
#pragma unroll
          for (int r = 0; r < kMaxRanks; ++r)
            if (r < p.world && r != p.rank) st_push_v4(lam_region(p, r, b) + slot_off + idx * 16, v);
        }
      }
    }
  } else if (is_link) {
    scatter_chunks(p, b, chunk, nvec, ltid, lsize);
  }

  // ---- re-arm the previous call's region while the pushes are in flight
  if (is_row) rearm_previous(p, pb, rtid, rsize);

  // ---- two-shot: reduce my chunk, push the sums to every rank's hi half
  if (TWOSHOT && is_link) reduce_gather<T>(p, b, my_lo, chunk, nvec, ltid, lsize, sg);

  // ---- rows: wait for the sums, epilogue (two-shot: in the order the sums arrive)
  if (is_row) {
    const ArrivalOrder order((uint32_t)p.tokens, TWOSHOT ? (uint32_t)p.world : 1u,
                             TWOSHOT ? kArrivalSlices : 1u);
    for (uint32_t q = (uint32_t)row0; q < order.total; q += (uint32_t)row_stride) {
      const long long rq = order.at(q);
      if (rq < 0) continue;
      const int row = (int)rq;
      const size_t row_vec0 = (size_t)row * p.hv;
      uint4 s[VPT], res[VPT];
      load_residual<VPT>(res, row_vec0, p);  // in flight while we wait on the peers
#pragma unroll
      for (int j = 0; j < VPT; ++j) {
        const int c = threadIdx.x + j * blockDim.x;
        if (c < p.hv) {
          const size_t idx = row_vec0 + c;
          if (TWOSHOT) {
            s[j] = wait_one(my_hi + idx * 16, p, sg);
          } else {
            const uint4 own = ld_v4((const char*)p.in + idx * 16);
            s[j] = gather_sum<T, 8>(my_lo, nvec * 16, idx, own, p, sg);
          }
        }
      }
      row_epilogue<T, VPT>(s, res, gam, row_vec0, p, smem, parity);
    }
  }

  end_call(p, cs, TWOSHOT ? (unsigned long long)p.world * chunk * 16
                            : (unsigned long long)p.world * nvec * 16,
           TWOSHOT ? (unsigned long long)nvec * 16 : 0ull);
}

// ---- one-shot fused kernel, rows split over a thread-block cluster ----------------------------
// The latency regime (decode batch 1..64): a row of hv vectors is handled by the P CTAs of a
// cluster, one vector per thread, so the (N-1) pushes and polls of a row are spread over P SMs
// and all in flight at once.  The row's sum of squares is combined through distributed shared
// memory in fixed CTA order (same bits on every rank).  P == 1 degenerates to one CTA per row.
//
// P is a COMPILE-TIME property of the kernel (__cluster_dims__), not a launch attribute: a
// launch path that drops launch attributes (ncu re-issuing a cooperative launch did exactly
// that in round 1: launch__cluster_dim_x = 0, every CTA its own cluster, 3/4 of each row never
// written) cannot change the shape the indexing below relies on.  The shape is checked on the
// device as well; a mismatch sets the sticky error word to kErrClusterShape and does nothing.
template <typename T, int SPLIT>
__device__ __forceinline__ void ar_norm_cluster_body(const RankParams& p, float* smem,
                                                     float (*part)[kMaxRanks] /* [2][8] */) {
  constexpr int EPV = VecOps<T>::EPV;
  constexpr uint32_t P = SPLIT;
  pdl_launch_dependents();
  if (cluster_nctarank() != P) {  // uniform over the grid: every CTA leaves, nobody waits
    pdl_wait();
    if (threadIdx.x == 0) atomicExch(&p.st->error, kErrClusterShape);
    return;
  }
  const uint32_t cta = (P > 1) ? cluster_ctarank() : 0u;
  const int c = (int)(cta * blockDim.x + threadIdx.x);  // vector index inside the row
  const bool valid = c < p.hv;
  uint4 gam = make_uint4(0, 0, 0, 0);
  if (valid) gam = ld_v4((const char*)p.gamma + (size_t)c * 16);
  pdl_wait();

  const int group = (int)(blockIdx.x / P), groups = (int)(gridDim.x / P);
  // first row's input is requested before the epoch round trip returns
  uint4 v0 = make_uint4(0, 0, 0, 0);
  if (valid && group < p.tokens)
    v0 = ld_stream_v4((const char*)p.in + ((size_t)group * p.hv + c) * 16);
  const CallState cs = begin_call(p);
  const uint32_t epoch = cs.epoch;
  const uint32_t b = epoch % 3u, pb = (b + 2u) % 3u;  // from b: epoch + 2 may overflow
  const size_t nvec = (size_t)p.nvec;
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t gsize = (size_t)gridDim.x * blockDim.x;
  char* my_lo = lam_region(p, p.rank, b);
  SpinGuard sg;
  int parity = 0;

  // push my rows
  const size_t slot_off = (size_t)p.rank * nvec * 16;
  for (int row = group; row < p.tokens; row += groups) {
    if (valid) {
      const size_t idx = (size_t)row * p.hv + c;
      uint4 v = (row == group) ? v0 : ld_stream_v4((const char*)p.in + idx * 16);
      scrub(v);
      if (row == group) v0 = v;
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < p.world && r != p.rank) st_push_v4(lam_region(p, r, b) + slot_off + idx * 16, v);
    }
  }
  rearm_previous(p, pb, gtid, gsize);
  // every CTA of the cluster must be running before its shared memory is written remotely;
  // this wait overlaps the fabric flight time of the pushes
  if (P > 1) cluster_sync_all();

  int pp = 0;
  for (int row = group; row < p.tokens; row += groups) {
    const size_t idx = (size_t)row * p.hv + c;
    float x[EPV];
    float ss = 0.0f;
    if (valid) {
      uint4 res = make_uint4(0, 0, 0, 0);
      if (p.res_in != nullptr) res = ld_stream_v4((const char*)p.res_in + idx * 16);
      // scrubbing only rewrites -0.0 words to +0.0, so the pushed copy is as good as the input
      const uint4 own = (row == group) ? v0 : ld_v4((const char*)p.in + idx * 16);
      const uint4 s = gather_sum<T, 8>(my_lo, nvec * 16, idx, own, p, sg);
      VecOps<T>::to_float(s, x);
      if (p.res_in != nullptr) {
        float r[EPV];
        VecOps<T>::to_float(res, r);
#pragma unroll
        for (int e = 0; e < EPV; ++e) x[e] = x[e] + r[e];
        st_v4((char*)p.res_out + idx * 16, VecOps<T>::from_float(x));
      }
#pragma unroll
      for (int e = 0; e < EPV; ++e) ss += x[e] * x[e];
    }
    float total = block_sum(ss, smem, parity);
    if (P > 1) {
      // each CTA deposits its partial in slot [cta] of every sibling (DSMEM stores), then one
      // cluster barrier; afterwards everybody sums its OWN copy in CTA order.  No CTA reads
      // remote memory, so nobody has to wait for siblings before exiting.
      if (threadIdx.x < P) st_dsmem_f32(&part[pp][cta], threadIdx.x, total);
      cluster_sync_all();
      total = 0.0f;
#pragma unroll
      for (uint32_t q = 0; q < P; ++q) total += part[pp][q];
      pp ^= 1;
    }
    if (valid) {
      const float rs = 1.0f / sqrtf(total / (float)p.hidden + p.eps);
      float g[EPV], y[EPV];
      VecOps<T>::to_float(gam, g);
#pragma unroll
      for (int e = 0; e < EPV; ++e) y[e] = VecOps<T>::round_trip(x[e] * rs) * g[e];
      st_v4((char*)p.norm_out + idx * 16, VecOps<T>::from_float(y));
    }
  }
  end_call(p, cs, (unsigned long long)p.world * nvec * 16, 0ull);
}

template <typename T, int SPLIT>
__global__ void __cluster_dims__(SPLIT, 1, 1) __launch_bounds__(256)
    ar_norm_cluster_kernel(const __grid_constant__ RankParams p) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  __shared__ float part[2][kMaxRanks];
  ar_norm_cluster_body<T, SPLIT>(p, smem, part);
}

template <typename T, int SPLIT>
__global__ void __cluster_dims__(SPLIT, 1, 1) __launch_bounds__(256)
    ar_norm_cluster_kernel_emu(const __grid_constant__ EmuParams ep) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  __shared__ float part[2][kMaxRanks];
  ar_norm_cluster_body<T, SPLIT>(ep.r[blockIdx.y], smem, part);
}

// ---- plain all-reduce (no epilogue): flat vector indexing -----------------------------------------
template <typename T, bool TWOSHOT>
__device__ __forceinline__ void ar_body(const RankParams& p) {
  pdl_launch_dependents();
  pdl_wait();
  const CallState cs = begin_call(p);
  const uint32_t epoch = cs.epoch;
  const uint32_t b = epoch % 3u, pb = (b + 2u) % 3u;  // from b: epoch + 2 may overflow
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t gsize = (size_t)gridDim.x * blockDim.x;
  const size_t nvec = (size_t)p.nvec;
  const size_t chunk = (nvec + p.world - 1) / p.world;
  char* my_lo = lam_region(p, p.rank, b);
  char* my_hi = my_lo + p.g.hi_off;
  SpinGuard sg;

  const int link_blocks = TWOSHOT ? p.link_blocks : 0;
  const bool specialised = TWOSHOT && link_blocks > 0 && link_blocks < (int)gridDim.x;
  const bool is_link = !specialised || (int)blockIdx.x < link_blocks;
  const bool is_out = !specialised || (int)blockIdx.x >= link_blocks;
  const size_t ltid = gtid;
  const size_t lsize = (size_t)(specialised ? link_blocks : gridDim.x) * blockDim.x;
  const size_t otid = specialised ? gtid - (size_t)link_blocks * blockDim.x : gtid;
  const size_t osize = specialised ? gsize - (size_t)link_blocks * blockDim.x : gsize;

  if (!TWOSHOT) {
    for (size_t idx = gtid; idx < nvec; idx += gsize) {
      uint4 v = ld_stream_v4((const char*)p.in + idx * 16);
      scrub(v);
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < p.world && r != p.rank)
          st_push_v4(lam_region(p, r, b) + ((size_t)p.rank * nvec + idx) * 16, v);
    }
  } else if (is_link) {
    scatter_chunks(p, b, chunk, nvec, ltid, lsize);
  }
  if (is_out) rearm_previous(p, pb, otid, osize);
  if (TWOSHOT) {
    if (is_link) reduce_gather<T>(p, b, my_lo, chunk, nvec, ltid, lsize, sg);
    if (is_out) {
      const ArrivalOrder order((uint32_t)nvec, (uint32_t)p.world, kArrivalSlices);
      for (size_t q = otid; q < order.total; q += osize) {
        const long long idx = order.at((uint32_t)q);
        if (idx >= 0) st_v4((char*)p.out + idx * 16, wait_one(my_hi + idx * 16, p, sg));
      }
    }
  } else {
    for (size_t idx = gtid; idx < nvec; idx += gsize) {
      const uint4 own = ld_v4((const char*)p.in + idx * 16);
      st_v4((char*)p.out + idx * 16, gather_sum<T, 8>(my_lo, nvec * 16, idx, own, p, sg));
    }
  }
  end_call(p, cs, TWOSHOT ? (unsigned long long)p.world * chunk * 16
                            : (unsigned long long)p.world * nvec * 16,
           TWOSHOT ? (unsigned long long)nvec * 16 : 0ull);
}

// ---- NVLS variant: the reduction happens inside the NVSwitch ----------------------------------
//
// Bandwidth regime, large groups.  Per rank and direction the push two-shot moves 2(N-1)/N * S
// over the link; here the switch PULLS each rank's copy (multimem.ld_reduce: one request, N reads,
// one reduced response) and REPLICATES the sums (multimem.st: one store, N writes), which costs
// (1 + 1/N) * S per direction -- the own copy also travels to the switch and back.  Measured
// with tools/nvls_bench.cu: ld_reduce is bound by the outbound link, multimem.st by the inbound
// one, so the two overlap; at N = 2 NVLS moves MORE than the push kernels (1.5 S vs S) and is
// never chosen, at N = 8 it moves 1.125 S vs 1.75 S.
//
// Everything here is sized for bandwidth, so arrival is signalled with FLAGS, not with the
// data itself: no region has to be re-armed (S bytes of stores per call), nothing is scrubbed,
// and consumers stream whole slices with several loads in flight instead of polling vector by
// vector (one L2 round trip per 16 bytes per thread).
//
// Two kinds of CTA; waits only ever point from later work to earlier work (copy -> reduce ->
// consume), the copy never waits, and the grid is co-resident, so the kernel cannot deadlock:
//   COPY/ROW CTAs  first stream the input into xin (the switch can only read the symmetric
//                buffer), one pipeline group after the other; after each group the last CTA to
//                finish stores flagA[g][rank] = epoch+1 on every rank with ONE multicast store.
//                Then they turn into ROW CTAs (below).
//   REDUCE CTAs  per group: N threads wait for the N local flagA, then x = ld_reduce(mc xin) for
//                my chunk's vectors of that group, written back IN PLACE on every rank with ONE
//                multimem.st per vector; the last CTA to finish the slice stores
//                flagB[g][rank] = epoch+1 everywhere
//   ROW CTAs     (the copy CTAs, once done) walk the rows in arrival order (ArrivalOrder), wait for the flagB of the slice a
//                vector lies in (remembered, so once per slice), read the sums from xin and run
//                the residual + RMSNorm epilogue (or copy them out for the plain all-reduce)
//
// Hazards.  xin is overwritten by the next call's copy-in: a rank starts call k+1 only after it
// has seen every flagB of call k, i.e. after every owner's stores were performed everywhere
// (release), which in turn follow that owner's ld_reduce of the same addresses.  Writing the sums
// over xin is safe because only the owner ever reads its chunk through the switch.  Flags carry
// epoch+1 and are compared modulo 2^32 (flag - tag >= 0), so no flag is ever reset.  The switch
// rounds ties differently from round-to-nearest-even and its summation order is unspecified, so
// this variant is tolerance-parity with the oracle (north_star's 1e-2), never bit-parity; all
// ranks still get identical bits because each sum is computed once and broadcast.
// Consumers wait per CTA, not per thread: ONE thread polls the flags of the slices a vector
// range lies in (with a short sleep between polls) and the CTA barrier hands the result to the
// others, which keeps 130 k threads off 64 flag lines.  A range usually lies in one slice (a row
// is far smaller than a slice wherever this variant is worth using); for tiny messages it can
// cross many (one row over every owner's chunk), and then EVERY slice it touches is waited for.
struct SliceGate {
  uint32_t seen = 0xffffffffu;  // the one slice this CTA last found complete
  template <typename P>
  static __device__ __forceinline__ void poll(const P& p, uint32_t id, uint32_t tag, SpinGuard& sg) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(
        nv_flag(p.nv_uc + kNvFlagBOff, (int)(id / kMaxRanks), (int)(id % kMaxRanks)));
    while ((int32_t)(ld_acquire_sys(f) - tag) < 0) {
      __nanosleep(64);
      if (sg.expired(p)) break;
    }
  }
  static __device__ __forceinline__ uint32_t slice_of(uint32_t idx, uint32_t chunk, uint32_t sub) {
    const uint32_t o = idx / chunk, g = (idx - o * chunk) / sub;
    return g * kMaxRanks + o;
  }
  // all threads of the CTA call this with the same arguments (vector range [first, last])
  template <typename P>
  __device__ __forceinline__ void wait(const P& p, uint32_t first, uint32_t last, uint32_t chunk,
                                       uint32_t sub, uint32_t tag, SpinGuard& sg) {
    const uint32_t a = slice_of(first, chunk, sub), b = slice_of(last, chunk, sub);
    if (a == seen && b == seen) return;
    if (threadIdx.x == 0) {
      // walk the range slice by slice: idx always is the first vector of the range inside a slice
      uint32_t idx = first;
      for (;;) {
        poll(p, slice_of(idx, chunk, sub), tag, sg);
        const uint32_t o = idx / chunk, g = (idx - o * chunk) / sub;
        const uint32_t next = o * chunk + min((g + 1u) * sub, chunk);  // first vector of the next slice
        if (next > last) break;
        idx = next;
      }
    }
    __syncthreads();
    seen = (a == b) ? a : 0xffffffffu;
  }
};

template <typename T, int VPT, bool FUSED>
__device__ __forceinline__ void nvls_body(const RankParams& p, float* smem) {
  pdl_launch_dependents();
  uint4 gam[VPT];
  if (FUSED) load_gamma<T, VPT>(gam, p);
  pdl_wait();

  const CallState cs = begin_call(p);
  const uint32_t epoch = cs.epoch;
  const uint32_t pb = (epoch % 3u + 2u) % 3u;
  const uint32_t tag = epoch + 1u;  // never 0: a zeroed flag can never satisfy a wait
  const uint32_t nvec = (uint32_t)p.nvec;  // <= 2^28 vectors (4 GiB) by construction
  const uint32_t world = (uint32_t)p.world;
  const uint32_t chunk = (nvec + world - 1) / world;
  const uint32_t G = (uint32_t)p.nv_groups;
  const uint32_t sub = (chunk + G - 1) / G;
  const int red_blocks = p.link_blocks;            // CTAs [0, red_blocks) reduce
  const int copy_blocks = (int)gridDim.x - red_blocks;  // everybody else copies in, then consumes
  char* xin = p.nv_uc + kNvCtrlBytes;
  char* mc_xin = p.nv_mc + kNvCtrlBytes;
  SpinGuard sg;
  constexpr int U = 4;

  if ((int)blockIdx.x >= red_blocks) {
    // ---- COPY: sub-slice g of EVERY rank's chunk, group after group.
    // The last warp of the CTA does not copy: it is the SIGNAL warp.  The copy warps only
    // ARRIVE on a named barrier when their part of a group is stored and go straight on to the
    // next group; the signal warp waits on that barrier, fences (cumulative over what the
    // barrier ordered before it), counts the CTA in, and the last CTA's signal warp raises the
    // flag on every rank.  The fence + atomic + flag chain (several microseconds per group,
    // tools/nvls_probe.py) is thereby off the copy path.  One barrier id per group (1 + g <= 8).
    const uint32_t cthreads = blockDim.x - 32;
    const bool is_signal = threadIdx.x >= cthreads;
    const uint32_t ctid = (blockIdx.x - red_blocks) * cthreads + threadIdx.x;
    const uint32_t csize = (uint32_t)copy_blocks * cthreads;
    for (uint32_t g = 0; g < G; ++g) {
      if (!is_signal) {
        const uint32_t v0 = min(g * sub, chunk), v1 = min(v0 + sub, chunk), n = v1 - v0;
        const uint32_t total = (p.nv_debug & 1) ? 0u : n * world;
        for (uint32_t base = ctid; base < total; base += csize * U) {
          uint4 v[U];
          uint32_t idx[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const uint32_t i = base + (uint32_t)u * csize;
            idx[u] = nvec;
            if (i < total) {
              const uint32_t o = i / n;
              idx[u] = o * chunk + v0 + (i - o * n);
            }
            if (idx[u] < nvec) v[u] = ld_stream_v4((const char*)p.in + (size_t)idx[u] * 16);
          }
#pragma unroll
          for (int u = 0; u < U; ++u)
            if (idx[u] < nvec) st_v4(xin + (size_t)idx[u] * 16, v[u]);
        }
        named_bar_arrive(1 + g, blockDim.x);
      } else {
        named_bar_sync(1 + g, blockDim.x);
        if (threadIdx.x == cthreads) {
          __threadfence();  // the copies are local stores: device scope orders them before the count
          const uint32_t prev = atomicAdd(&p.st->nv_arrive[g], 1u);
          if (prev == (uint32_t)copy_blocks - 1u) {
            p.st->nv_arrive[g] = 0;
            __threadfence();
            mc_st_release_u32(nv_flag(p.nv_mc, (int)g, p.rank), tag);  // release at system scope
          }
        }
      }
    }
  }
  if ((int)blockIdx.x < red_blocks) {
    // ---- REDUCE: in-switch sum of my chunk, broadcast of the rounded sums, in place.
    // The CTAs form `teams` teams; team t takes groups t, t + teams, ...  Finishing a slice
    // costs a drain of the team's multicast stores (system fence) plus an atomic and a flag
    // store, several microseconds in which that team moves nothing: with the teams on different
    // groups one team's drain hides behind the others' traffic.  (A signal warp, as on the copy
    // side, was tried here and lost: a system fence issued while the CTA's other warps keep
    // storing through the switch stalls them too -- 287 us instead of 169 us for 64 MiB.)
    // With few teams and many groups the slices still complete roughly in order, which is what
    // the consumers, walking the slices in that order, need to stay close behind.
    const uint32_t teams = (uint32_t)p.nv_teams;
    const uint32_t rb = blockIdx.x;
    const uint32_t team = rb % teams, team_size = (uint32_t)red_blocks / teams;
    const uint32_t ltid = (rb / teams) * blockDim.x + threadIdx.x;
    const uint32_t lsize = team_size * blockDim.x;
    const uint32_t my0 = (uint32_t)p.rank * chunk;
    for (uint32_t g = team; g < G; g += teams) {
      if (threadIdx.x < world) {
        const uint32_t* f = reinterpret_cast<const uint32_t*>(nv_flag(p.nv_uc, (int)g, (int)threadIdx.x));
        while ((int32_t)(ld_acquire_sys(f) - tag) < 0)
          if (sg.expired(p)) break;
      }
      __syncthreads();
      const uint32_t v0 = min(g * sub, chunk), v1 = min(v0 + sub, chunk);
      for (uint32_t base = v0 + ltid; base < v1; base += lsize * U) {
        uint4 x[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t v = base + (uint32_t)u * lsize;
          x[u] = make_uint4(v, 0, 0, 0);
          if (v < v1 && my0 + v < nvec && !(p.nv_debug & 2))
            x[u] = VecOps<T>::mc_ld_reduce(mc_xin + (size_t)(my0 + v) * 16);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t v = base + (uint32_t)u * lsize;
          if (v < v1 && my0 + v < nvec && !(p.nv_debug & 4)) mc_st_v4(mc_xin + (size_t)(my0 + v) * 16, x[u]);
        }
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();  // this CTA's multicast stores are performed everywhere
        const uint32_t prev = atomicAdd(&p.st->nv_arrive_red[g], 1u);
        if (prev == team_size - 1u) {
          p.st->nv_arrive_red[g] = 0;
          __threadfence();
          mc_st_release_u32(nv_flag(p.nv_mc + kNvFlagBOff, (int)g, p.rank), tag);
        }
      }
    }
  } else {
    // ---- ROW: re-arm what a preceding push call left, then consume the sums slice by slice
    const int row0 = (int)blockIdx.x - red_blocks;
    const int row_stride = copy_blocks;
    const size_t rtid = (size_t)row0 * blockDim.x + threadIdx.x;
    const size_t rsize = (size_t)row_stride * blockDim.x;
    rearm_previous(p, pb, rtid, rsize);
    SliceGate gate;
    if (p.nv_debug & 8) {
      // measurement only: no consumer
    } else if (FUSED) {
      int parity = 0;
      const ArrivalOrder order((uint32_t)p.tokens, world, G);
      for (uint32_t q = (uint32_t)row0; q < order.total; q += (uint32_t)row_stride) {
        const long long rq = order.at(q);
        if (rq < 0) continue;
        const size_t row_vec0 = (size_t)rq * p.hv;
        uint4 s[VPT], res[VPT];
        load_residual<VPT>(res, row_vec0, p);
        gate.wait(p, (uint32_t)row_vec0, (uint32_t)(row_vec0 + p.hv - 1), chunk, sub, tag, sg);
#pragma unroll
        for (int j = 0; j < VPT; ++j) {
          const int c = threadIdx.x + j * blockDim.x;
          if (c < p.hv) s[j] = ld_poll_v4(xin + (row_vec0 + c) * 16);  // L1 may hold last call's line
        }
        row_epilogue<T, VPT>(s, res, gam, row_vec0, p, smem, parity);
      }
    } else {
      // tiles of blockDim * U consecutive positions of the arrival order; a tile lies in one
      // slice or straddles two neighbouring ones (tiny messages: more -- then every run it
      // touches is waited for, one by one)
      const ArrivalOrder order(nvec, world, G);
      const uint32_t tile = blockDim.x * U;
      for (uint32_t t0 = (uint32_t)row0 * tile; t0 < order.total; t0 += (uint32_t)row_stride * tile) {
        const uint32_t t1 = min(t0 + tile, order.total) - 1;
        for (uint32_t run = t0 / order.per_slice; run <= t1 / order.per_slice; ++run) {
          const uint32_t g = run / world, o = run - g * world;
          const uint32_t within = g * order.per_slice;
          if (within >= order.per_chunk || o * order.per_chunk + within >= nvec) continue;
          const uint32_t idx = o * order.per_chunk + within;
          gate.wait(p, idx, idx, chunk, sub, tag, sg);
        }
        long long idx[U];
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint32_t q = t0 + (uint32_t)u * blockDim.x + threadIdx.x;
          idx[u] = q < order.total ? order.at(q) : -1;
          if (idx[u] >= 0) v[u] = ld_poll_v4(xin + idx[u] * 16);
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (idx[u] >= 0) st_v4((char*)p.out + idx[u] * 16, v[u]);
      }
    }
  }
  end_call(p, cs, 0ull, 0ull);
}

template <typename T, int VPT, bool FUSED>
__global__ void __launch_bounds__(512, B200C_TWOSHOT_CTAS_PER_SM)
    nvls_kernel(const __grid_constant__ RankParams p) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  nvls_body<T, VPT, FUSED>(p, smem);
}

// ---- kernels: a real launch carries one rank, an emulated launch carries all of them ----------

template <typename T, int VPT, bool TWOSHOT>
__global__ void __launch_bounds__(512, TWOSHOT ? B200C_TWOSHOT_CTAS_PER_SM : 1) ar_norm_kernel(const __grid_constant__ RankParams p) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  ar_norm_body<T, VPT, TWOSHOT>(p, smem);
}

template <typename T, int VPT, bool TWOSHOT>
__global__ void __launch_bounds__(512, TWOSHOT ? B200C_TWOSHOT_CTAS_PER_SM : 1) ar_norm_kernel_emu(const __grid_constant__ EmuParams ep) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  ar_norm_body<T, VPT, TWOSHOT>(ep.r[blockIdx.y], smem);
}

template <typename T, bool TWOSHOT>
__global__ void __launch_bounds__(512, TWOSHOT ? B200C_TWOSHOT_CTAS_PER_SM : 1) ar_kernel(const __grid_constant__ RankParams p) {
  ar_body<T, TWOSHOT>(p);
}

template <typename T, bool TWOSHOT>
__global__ void __launch_bounds__(512, TWOSHOT ? B200C_TWOSHOT_CTAS_PER_SM : 1) ar_kernel_emu(const __grid_constant__ EmuParams ep) {
  ar_body<T, TWOSHOT>(ep.r[blockIdx.y]);
}

// ---- local norm kernels (TP=1 and the unfused path) ---------------------------------------------
// residual add + RMSNorm on rows of `in` (res_in == nullptr: plain rms_norm).
template <typename T, int VPT>
__global__ void __launch_bounds__(512) norm_kernel(const __grid_constant__ RankParams p) {
  __shared__ float smem[64];  // 2 x 32 reduction banks
  pdl_launch_dependents();
  uint4 gam[VPT];
  load_gamma<T, VPT>(gam, p);
  pdl_wait();
  int parity = 0;
  for (int row = blockIdx.x; row < p.tokens; row += gridDim.x) {
    const size_t row_vec0 = (size_t)row * p.hv;
    uint4 s[VPT], res[VPT];
    load_residual<VPT>(res, row_vec0, p);
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int c = threadIdx.x + j * blockDim.x;
      if (c < p.hv) s[j] = ld_stream_v4((const char*)p.in + (row_vec0 + c) * 16);
    }
    row_epilogue<T, VPT>(s, res, gam, row_vec0, p, smem, parity);
  }
}

// Narrow rows (one vector per thread) in the bandwidth regime: R rows per trip.  Every thread
// keeps 2R independent 16-byte loads in flight, only the raw vectors stay in registers across
// the barrier (x = s + r is recomputed afterwards: a few adds instead of 8 live floats per row),
// and the R block reductions share one barrier.  Same per-row arithmetic and reduction tree as
// norm_kernel<T,1>, hence the same bits.
template <typename T, int R>
__global__ void __launch_bounds__(512, 2) norm_kernel_rows(const __grid_constant__ RankParams p) {
  constexpr int EPV = VecOps<T>::EPV;
  __shared__ float smem[2 * R * 32];
  pdl_launch_dependents();
  const int c = threadIdx.x;
  const bool valid = c < p.hv;
  uint4 gam = make_uint4(0, 0, 0, 0);
  if (valid) gam = ld_v4((const char*)p.gamma + (size_t)c * 16);
  pdl_wait();
  int parity = 0;
  const bool has_res = p.res_in != nullptr;
  for (int row = R * blockIdx.x; row < p.tokens; row += R * gridDim.x) {
    uint4 s[R], r[R];
    float ss[R];
#pragma unroll
    for (int k = 0; k < R; ++k) {
      ss[k] = 0.0f;
      if (valid && row + k < p.tokens) {
        const size_t idx = (size_t)(row + k) * p.hv + c;
        s[k] = ld_stream_v4((const char*)p.in + idx * 16);
        if (has_res) r[k] = ld_stream_v4((const char*)p.res_in + idx * 16);
      }
    }
#pragma unroll
    for (int k = 0; k < R; ++k) {
      if (valid && row + k < p.tokens) {
        float x[EPV];
        VecOps<T>::to_float(s[k], x);
        if (has_res) {
          float rr[EPV];
          VecOps<T>::to_float(r[k], rr);
#pragma unroll
          for (int e = 0; e < EPV; ++e) x[e] = x[e] + rr[e];
          st_v4((char*)p.res_out + ((size_t)(row + k) * p.hv + c) * 16, VecOps<T>::from_float(x));
        }
#pragma unroll
        for (int e = 0; e < EPV; ++e) ss[k] += x[e] * x[e];
      }
    }
    block_sum_multi<R>(ss, smem, parity);
#pragma unroll
    for (int k = 0; k < R; ++k) {
      if (valid && row + k < p.tokens) {
        float x[EPV], g[EPV], y[EPV];
        VecOps<T>::to_float(s[k], x);
        if (has_res) {
          float rr[EPV];
          VecOps<T>::to_float(r[k], rr);
#pragma unroll
          for (int e = 0; e < EPV; ++e) x[e] = x[e] + rr[e];
        }
        const float rs = 1.0f / sqrtf(ss[k] / (float)p.hidden + p.eps);
        VecOps<T>::to_float(gam, g);
#pragma unroll
        for (int e = 0; e < EPV; ++e) y[e] = VecOps<T>::round_trip(x[e] * rs) * g[e];
        st_v4((char*)p.norm_out + ((size_t)(row + k) * p.hv + c) * 16, VecOps<T>::from_float(y));
      }
    }
  }
}

// Bandwidth regime, many small CTAs: a row on blockDim = hv / VPT <= 256 threads, MINB resident
// CTAs per SM.  Nothing is pipelined inside a CTA; instead the SM holds many independent CTAs
// whose barriers (over 4-8 warps only) fall at different times, so the memory system always sees
// loads from some of them.  Registers are kept low for that residency: only the raw vectors
// live across the barrier, x = s + r is recomputed and gamma re-read (an L1 hit) afterwards.
// (Tried and dropped, profiles/r2_norm_sweep_variants.json: a software-pipelined 512-thread kernel
// that issues the next trip's loads before the barrier, 3-6 % slower; a 29-register kernel that
// reads each row twice so that eight CTAs per SM are resident, 5-30 % slower.)
// Same per-row arithmetic and reduction tree as norm_kernel, hence the same bits.
template <typename T, int VPT, int MINB>
__global__ void __launch_bounds__(256, MINB) norm_kernel_small(const __grid_constant__ RankParams p) {
  constexpr int EPV = VecOps<T>::EPV;
  __shared__ float smem[64];
  pdl_launch_dependents();
  pdl_wait();
  int parity = 0;
  const bool has_res = p.res_in != nullptr;
  for (int row = blockIdx.x; row < p.tokens; row += gridDim.x) {
    const size_t row_vec0 = (size_t)row * p.hv;
    uint4 s[VPT], r[VPT];
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int c = threadIdx.x + j * blockDim.x;
      if (c < p.hv) {
        s[j] = ld_stream_v4((const char*)p.in + (row_vec0 + c) * 16);
        if (has_res) r[j] = ld_stream_v4((const char*)p.res_in + (row_vec0 + c) * 16);
      }
    }
    float ss = 0.0f;
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int c = threadIdx.x + j * blockDim.x;
      if (c < p.hv) {
        float x[EPV];
        VecOps<T>::to_float(s[j], x);
        if (has_res) {
          float rr[EPV];
          VecOps<T>::to_float(r[j], rr);
#pragma unroll
          for (int e = 0; e < EPV; ++e) x[e] = x[e] + rr[e];
          st_v4((char*)p.res_out + (row_vec0 + c) * 16, VecOps<T>::from_float(x));
        }
#pragma unroll
        for (int e = 0; e < EPV; ++e) ss += x[e] * x[e];
      }
    }
    const float total = block_sum(ss, smem, parity);
    const float rs = 1.0f / sqrtf(total / (float)p.hidden + p.eps);
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const int c = threadIdx.x + j * blockDim.x;
      if (c < p.hv) {
        float x[EPV], g[EPV], y[EPV];
        VecOps<T>::to_float(s[j], x);
        if (has_res) {
          float rr[EPV];
          VecOps<T>::to_float(r[j], rr);
#pragma unroll
          for (int e = 0; e < EPV; ++e) x[e] = x[e] + rr[e];
        }
        VecOps<T>::to_float(ld_v4((const char*)p.gamma + (size_t)c * 16), g);
#pragma unroll
        for (int e = 0; e < EPV; ++e) y[e] = VecOps<T>::round_trip(x[e] * rs) * g[e];
        st_v4((char*)p.norm_out + (row_vec0 + c) * 16, VecOps<T>::from_float(y));
      }
    }
  }
}

// ---- all-gather along the last dim (byte-exact; flag protocol) -----------------------------------
// Staging layout in every rank's arena = the final [rows][world][row_bytes] layout, double
// buffered on the call parity.  Each rank pushes its slab into all peers' staging, fences,
// and its last CTA raises flag[parity][rank] = seq+1 on every peer; receivers copy a
// source's slab out as soon as that source's flag is up.
struct CopyParams {
  char* arena[kMaxRanks];
  LocalState* st;
  ArenaGeom g;
  const void* in;
  void* out;
  int rank, world;
  int rows;
  int peer;  // send: dst, recv: src
  unsigned long long row_vecs;  // 16-byte vectors per slab row
  unsigned long long nvec;      // p2p: message vectors
  unsigned long long timeout_ns;
  int trap;  // policy when a wait times out (see SpinGuard)
  int role;  // emu sendrecv: 0 = sender, 1 = receiver
};

struct EmuCopyParams {
  CopyParams r[kMaxRanks];
};

__device__ __forceinline__ uint32_t* ctrl_flag(char* arena, size_t off, int i) {
  return reinterpret_cast<uint32_t*>(arena + off + (size_t)i * kFlagStride);
}

__device__ __forceinline__ void wait_flag_ge(const uint32_t* f, uint32_t want, const CopyParams& p,
                                             SpinGuard& sg) {
  while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
    if (sg.expired(p)) break;
  }
}

template <int VEC>
__device__ __forceinline__ void all_gather_body(const CopyParams& p) {
  using W = Word<VEC>;
  const uint32_t seq = p.st->ag_seq;
  const uint32_t par = seq & 1u;
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t gsize = (size_t)gridDim.x * blockDim.x;
  const size_t rv = (size_t)p.row_vecs;  // VEC-byte words per slab row
  const size_t slab_vecs = rv * (size_t)p.rows;
  const size_t stage_off = p.g.ag_off + (size_t)par * p.g.ag_stride;
  SpinGuard sg;

  // push my slab: word (row, c) -> staging[(row*world + rank)*rv + c] on every peer;
  // my own slab goes straight to `out`.
  for (size_t i = gtid; i < slab_vecs; i += gsize) {
    const size_t row = i / rv, c = i - row * rv;
    const typename W::type v = W::ld((const char*)p.in + i * VEC);
    const size_t dst = ((row * p.world + p.rank) * rv + c) * VEC;
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < p.world && r != p.rank) W::push(p.arena[r] + stage_off + dst, v);
    W::st((char*)p.out + dst, v);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(&p.st->ag_arrive_push, 1u);
    if (prev == gridDim.x - 1) {
      p.st->ag_arrive_push = 0;
      __threadfence_system();
      for (int r = 0; r < p.world; ++r)
        if (r != p.rank)
          st_release_sys(ctrl_flag(p.arena[r], kAgFlagOff, (int)par * kMaxRanks + p.rank), seq + 1);
    }
  }
  // copy the other ranks' slabs out of my staging as they complete
  const char* stage = p.arena[p.rank] + stage_off;
  for (int k = 1; k < p.world; ++k) {
    const int src = (p.rank + k) % p.world;
    wait_flag_ge(ctrl_flag(p.arena[p.rank], kAgFlagOff, (int)par * kMaxRanks + src), seq + 1, p, sg);
    for (size_t i = gtid; i < slab_vecs; i += gsize) {
      const size_t row = i / rv, c = i - row * rv;
      const size_t off = ((row * p.world + src) * rv + c) * VEC;
      W::st((char*)p.out + off, W::poll(stage + off));
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(&p.st->ag_arrive_done, 1u);
    if (prev == gridDim.x - 1) {
      p.st->ag_arrive_done = 0;
      __threadfence();
      p.st->ag_seq = seq + 1;
    }
  }
}

template <int VEC>
__global__ void __launch_bounds__(512) all_gather_kernel(const __grid_constant__ CopyParams p) {
  all_gather_body<VEC>(p);
}
template <int VEC>
__global__ void __launch_bounds__(512)
    all_gather_kernel_emu(const __grid_constant__ EmuCopyParams ep) {
  all_gather_body<VEC>(ep.r[blockIdx.y]);
}

// ---- pipeline-stage send / recv (mailbox + ready flag + credit) ---------------------------------
// Message m between a pair uses mailbox slot m&1 in the RECEIVER's arena.  The sender waits
// for the credit of message m-2 (slot free), pushes, fences, and its last CTA raises
// ready[src][slot] = m+1.  The receiver waits for ready, copies out, and its last CTA
// returns credit[dst][slot] = m+1 into the SENDER's arena.
__device__ __forceinline__ void send_body(const CopyParams& p) {
  const uint32_t m = p.st->send_seq[p.peer];
  const uint32_t slot = m & 1u;
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t gsize = (size_t)gridDim.x * blockDim.x;
  SpinGuard sg;
  // credit[dst][slot] = index + 1 of the last message consumed from that slot: message m needs
  // m-2 done, i.e. credit >= m-1.  The comparison is modular (signed difference), so it also
  // holds for the first two messages (credit 0 vs m-1 = -1, 0) and across a 2^32 wrap.
  wait_flag_ge(ctrl_flag(p.arena[p.rank], kP2pCreditOff, p.peer * 2 + (int)slot), m - 1, p, sg);
  char* box = p.arena[p.peer] + p.g.p2p_off + ((size_t)p.rank * 2 + slot) * p.g.p2p_stride;
  for (size_t i = gtid; i < (size_t)p.nvec; i += gsize)
    st_push_v4(box + i * 16, ld_stream_v4((const char*)p.in + i * 16));
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t prev = atomicAdd(&p.st->p2p_arrive_send, 1u);
    if (prev == gridDim.x - 1) {
      p.st->p2p_arrive_send = 0;
      __threadfence_system();
      st_release_sys(ctrl_flag(p.arena[p.peer], kP2pReadyOff, p.rank * 2 + (int)slot), m + 1);
      p.st->send_seq[p.peer] = m + 1;
    }
  }
}

__device__ __forceinline__ void recv_body(const CopyParams& p) {
  const uint32_t m = p.st->recv_seq[p.peer];
  const uint32_t slot = m & 1u;
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t gsize = (size_t)gridDim.x * blockDim.x;
  SpinGuard sg;
  wait_flag_ge(ctrl_flag(p.arena[p.rank], kP2pReadyOff, p.peer * 2 + (int)slot), m + 1, p, sg);
  const char* box = p.arena[p.rank] + p.g.p2p_off + ((size_t)p.peer * 2 + slot) * p.g.p2p_stride;
  for (size_t i = gtid; i < (size_t)p.nvec; i += gsize)
    st_v4((char*)p.out + i * 16, ld_poll_v4(box + i * 16));
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t prev = atomicAdd(&p.st->p2p_arrive_recv, 1u);
    if (prev == gridDim.x - 1) {
      p.st->p2p_arrive_recv = 0;
      __threadfence_system();
      st_release_sys(ctrl_flag(p.arena[p.peer], kP2pCreditOff, p.rank * 2 + (int)slot), m + 1);
      p.st->recv_seq[p.peer] = m + 1;
    }
  }
}

__global__ void __launch_bounds__(512) send_kernel(const __grid_constant__ CopyParams p) {
  send_body(p);
}
__global__ void __launch_bounds__(512) recv_kernel(const __grid_constant__ CopyParams p) {
  recv_body(p);
}
__global__ void __launch_bounds__(512)
    sendrecv_kernel_emu(const __grid_constant__ EmuCopyParams ep) {
  const CopyParams& p = ep.r[blockIdx.y];
  if (p.role == 0)
    send_body(p);
  else
    recv_body(p);
}

// ---- arena initialisation ---------------------------------------------------------------------------
__global__ void fill_not_arrived_kernel(uint4* p, size_t nvec) {
  const uint4 m = make_uint4(kNotArrived, kNotArrived, kNotArrived, kNotArrived);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
       i += (size_t)gridDim.x * blockDim.x)
    p[i] = m;
}

}  // namespace b200c
