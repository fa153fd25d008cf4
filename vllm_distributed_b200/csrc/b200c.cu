// b200c.cu -- host side of libb200c.so: context, arena geometry, launch logic and the
// extern "C" boundary declared in include/b200c.h.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "../../include/b200c.h"
#include "b200c_kernels.cuh"

using namespace b200c;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess)                                                              \
      return fail(B200C_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                  __FILE__, __LINE__);                                                  \
  } while (0)

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bytes of one receive half / NVLS region for messages of up to max_bytes (see make_geom)
size_t region_half(size_t max_bytes) {
  return align_up(std::max<size_t>(max_bytes, 16) + 16 * (kMaxRanks - 1), 256);
}

ArenaGeom make_geom(int world, size_t max_bytes, size_t ag_bytes, size_t p2p_bytes,
                    size_t* total) {
  ArenaGeom g;
  // A two-shot call on `world` ranks rounds the message up to world equal chunks of whole
  // 16-byte vectors, i.e. by at most (world-1)*16 bytes: the slack makes every message of up to
  // max_bytes fit whatever the group size (3, 5, 6 and 7 ranks rarely divide it).
  const size_t half = region_half(max_bytes);
  g.hi_off = half;
  g.lam_stride = 2 * half;
  g.ag_off = kCtrlBytes + 3 * g.lam_stride;
  g.ag_stride = align_up(std::max<size_t>(ag_bytes, 16), 256);
  g.p2p_off = g.ag_off + 2 * g.ag_stride;
  g.p2p_stride = align_up(std::max<size_t>(p2p_bytes, 16), 256);
  if (total) *total = g.p2p_off + (size_t)world * 2 * g.p2p_stride;
  return g;
}

}  // namespace

struct b200c_ctx {
  int rank = 0, world = 1, device = 0;
  char* arena[kMaxRanks] = {};
  size_t max_bytes = 0, ag_bytes = 0, p2p_bytes = 0;
  ArenaGeom geom{};
  LocalState* st = nullptr;
  int sm_count = 148;
  // tuning
  long long oneshot_max_bytes = 256 * 1024;
  int oneshot_blocks = 148;
  int twoshot_blocks = 148;
  int copy_blocks = 64;
  long long spin_timeout_ms = 10000;
  int trap_on_timeout = 0;  // 0: sticky error word, 1: trap (kill the context) -- see SpinGuard
  int link_blocks = 0;  // two-shot CTAs doing link work; 0 = half of the grid
  int pdl = 1;          // programmatic dependent launch on the real (non-emulated) kernels
  int oneshot_split = 0;  // CTAs per row in the one-shot fused kernel: 0 auto, 1 = no cluster
  // NVLS variant (b200c_attach_nvls): the symmetric buffer through its unicast and multicast mappings
  char* nv_uc = nullptr;
  char* nv_mc = nullptr;
  size_t nv_half = 0;
  long long nvls_min_bytes = -1;  // AUTO picks NVLS at or above this size; < 0 = never
  int nvls_link_blocks = 0;       // reduce CTAs; 0 = a sixth of the grid
  int nvls_copy_blocks = 0;       // copy-in CTAs; 0 = a sixth of the grid
  int nvls_groups = 0;            // copy-in pipeline groups; 0 = by size
  int nvls_teams = 0;             // reduce teams; 0 = one per group
  int nvls_debug = 0;             // measurement only: phase-skipping bits
};

namespace {

int dtype_size(int dtype) {
  switch (dtype) {
    case B200C_BF16:
    case B200C_F16: return 2;
    case B200C_F32: return 4;
    default: return 0;
  }
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

void fill_rank_params(const b200c_ctx* c, RankParams& p) {
  memset(&p, 0, sizeof(p));
  for (int r = 0; r < kMaxRanks; ++r) p.arena[r] = c->arena[r];
  p.st = c->st;
  p.g = c->geom;
  p.rank = c->rank;
  p.world = c->world;
  p.timeout_ns = (unsigned long long)c->spin_timeout_ms * 1000000ull;
  p.trap = c->trap_on_timeout;
  p.nv_uc = c->nv_uc;
  p.nv_mc = c->nv_mc;
  p.nv_half = c->nv_half;
  p.nv_groups = 1;
}

void fill_copy_params(const b200c_ctx* c, CopyParams& p) {
  memset(&p, 0, sizeof(p));
  for (int r = 0; r < kMaxRanks; ++r) p.arena[r] = c->arena[r];
  p.st = c->st;
  p.g = c->geom;
  p.rank = c->rank;
  p.world = c->world;
  p.timeout_ns = (unsigned long long)c->spin_timeout_ms * 1000000ull;
  p.trap = c->trap_on_timeout;
}

// Every CTA of a collective kernel must be resident at once (CTAs wait on peers' CTAs).
// The occupancy answer is cached per (kernel, block size): it is asked on every launch.
template <typename K>
int coresident_blocks(K kernel, int threads, int sm_count, int* out) {
  static thread_local struct { const void* fn; int threads; int per_sm; } cache[8] = {};
  const void* key = reinterpret_cast<const void*>(kernel);
  for (auto& e : cache)
    if (e.fn == key && e.threads == threads) {
      *out = e.per_sm * sm_count;
      return B200C_OK;
    }
  int per_sm = 0;
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0));
  if (per_sm < 1) return fail(B200C_ECUDA, "kernel cannot be resident with %d threads", threads);
  static thread_local int next = 0;
  cache[next] = {key, threads, per_sm};
  next = (next + 1) % 8;
  *out = per_sm * sm_count;
  return B200C_OK;
}

template <typename K, typename P>
int launch(K kernel, dim3 grid, int threads, const P& params, cudaStream_t s, bool cooperative,
           bool pdl = false, int cluster = 1) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = dim3(threads, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[3];
  int na = 0;
  if (cooperative) {
    attr[na].id = cudaLaunchAttributeCooperative;
    attr[na].val.cooperative = 1;
    ++na;
  } else if (pdl) {
    // the kernel may start while its predecessor drains; it blocks in griddepcontrol.wait
    // before touching anything the predecessor produces
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = cluster;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = na ? attr : nullptr;
  cfg.numAttrs = na;
  CUDA_TRY(cudaLaunchKernelEx(&cfg, kernel, params));
  return B200C_OK;
}

// How many clusters of `cluster` CTAs x `threads` threads can be resident at once (cached).
template <typename K>
int coresident_clusters(K kernel, int threads, int cluster, int sm_count, int* out) {
  if (cluster <= 1) return coresident_blocks(kernel, threads, sm_count, out);
  static thread_local struct { const void* fn; int threads; int cluster; int n; } cache[8] = {};
  const void* key = reinterpret_cast<const void*>(kernel);
  for (auto& e : cache)
    if (e.fn == key && e.threads == threads && e.cluster == cluster) {
      *out = e.n;
      return B200C_OK;
    }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(cluster * sm_count, 1, 1);
  cfg.blockDim = dim3(threads, 1, 1);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int n = 0;
  CUDA_TRY(cudaOccupancyMaxActiveClusters(&n, kernel, &cfg));
  if (n < 1) return fail(B200C_ECUDA, "no cluster of %d x %d threads can be resident", cluster, threads);
  static thread_local int next = 0;
  cache[next] = {key, threads, cluster, n};
  next = (next + 1) % 8;
  *out = n;
  return B200C_OK;
}

// two-shot role split: CTAs [0, link) push over the link, the rest run the rows
int link_split(const b200c_ctx* c, int grid) {
  if (grid < 2) return 0;
  // measured (tools/sweep.py): ~32 link CTAs feed the fabric for a 148-CTA grid, 64 for 296;
  // every further CTA is worth more on the row (HBM) side
  int l = c->link_blocks > 0 ? c->link_blocks : std::min(64, std::max(1, grid / 4));
  return std::max(1, std::min(l, grid - 1));
}

// two-shot grid: one CTA per SM below 2 MiB (launch + scheduling cost dominates), the full
// two-per-SM residency above
int twoshot_grid(const b200c_ctx* c, size_t bytes) {
  return bytes >= (size_t)(2u << 20) ? c->twoshot_blocks : std::min(c->twoshot_blocks, c->sm_count);
}

// rows are split over `threads` lanes holding up to VPT vectors each
void norm_shape(int hv, int* vpt, int* threads) {
  int v = 1;
  while (v < 4 && (hv + v - 1) / v > 512) v *= 2;
  int t = (hv + v - 1) / v;
  t = std::max(32, (t + 31) / 32 * 32);
  *vpt = v;
  *threads = t;
}

int pick_variant(const b200c_ctx* c, size_t bytes, int variant) {
  if (variant == B200C_ONESHOT || variant == B200C_TWOSHOT || variant == B200C_NVLS) return variant;
  if (c->nv_mc && c->nvls_min_bytes >= 0 && (long long)bytes >= c->nvls_min_bytes) return B200C_NVLS;
  // AUTO: one-shot below the measured crossover, provided `world` copies fit one region
  const bool fits = bytes * (size_t)c->world <= c->geom.lam_stride;
  return ((long long)bytes <= c->oneshot_max_bytes && fits) ? B200C_ONESHOT : B200C_TWOSHOT;
}

int check_ar_fit(const b200c_ctx* c, size_t bytes, int variant) {
  if (bytes > c->max_bytes)
    return fail(B200C_ETOOBIG, "message of %zu bytes exceeds the arena's max_bytes %zu", bytes,
                c->max_bytes);
  if (variant == B200C_ONESHOT && bytes * (size_t)c->world > c->geom.lam_stride)
    return fail(B200C_ETOOBIG, "one-shot needs world*bytes = %zu <= %zu", bytes * c->world,
                c->geom.lam_stride);
  if (variant == B200C_TWOSHOT) {
    const size_t nvec = bytes / 16, chunk = (nvec + c->world - 1) / c->world;
    if (chunk * 16 * c->world > c->geom.hi_off)
      return fail(B200C_ETOOBIG, "two-shot scatter area too small for %zu bytes", bytes);
  }
  if (variant == B200C_NVLS) {
    if (!c->nv_mc)
      return fail(B200C_EINVAL, "the NVLS variant needs a multicast buffer (b200c_attach_nvls)");
    if (bytes > c->nv_half) return fail(B200C_ETOOBIG, "message of %zu bytes exceeds the NVLS regions", bytes);
  }
  return B200C_OK;
}

// ---- fused AR + norm dispatch ------------------------------------------------------------------------

template <typename T, int VPT, bool TWOSHOT>
int launch_ar_norm_t(b200c_ctx* const* ctxs, int nemu, const RankParams* rp, int threads,
                     int want_blocks, cudaStream_t s) {
  if (nemu == 0) {
    int cap = 0;
    int rc = coresident_blocks(ar_norm_kernel<T, VPT, TWOSHOT>, threads, ctxs[0]->sm_count, &cap);
    if (rc) return rc;
    const int g = std::max(1, std::min(want_blocks, cap));
    RankParams p = rp[0];
    p.link_blocks = TWOSHOT ? link_split(ctxs[0], g) : 0;
    return launch(ar_norm_kernel<T, VPT, TWOSHOT>, dim3(g, 1, 1), threads, p, s, false,
                  ctxs[0]->pdl != 0);
  }
  int cap = 0;
  int rc = coresident_blocks(ar_norm_kernel_emu<T, VPT, TWOSHOT>, threads, ctxs[0]->sm_count, &cap);
  if (rc) return rc;
  const int g = std::max(1, std::min(want_blocks, cap / nemu));
  EmuParams ep;
  memset(&ep, 0, sizeof(ep));
  for (int r = 0; r < nemu; ++r) {
    ep.r[r] = rp[r];
    ep.r[r].link_blocks = TWOSHOT ? link_split(ctxs[0], g) : 0;
  }
  return launch(ar_norm_kernel_emu<T, VPT, TWOSHOT>, dim3(g, nemu, 1), threads, ep, s, true);
}

template <typename T>
int launch_ar_norm_v(b200c_ctx* const* ctxs, int nemu, const RankParams* rp, int vpt, int threads,
                     bool twoshot, int want_blocks, cudaStream_t s) {
#define B200C_CASE(V)                                                                      \
  case V:                                                                                  \
    return twoshot ? launch_ar_norm_t<T, V, true>(ctxs, nemu, rp, threads, want_blocks, s) \
                   : launch_ar_norm_t<T, V, false>(ctxs, nemu, rp, threads, want_blocks, s);
  switch (vpt) {
    B200C_CASE(1)
    B200C_CASE(2)
    B200C_CASE(4)
  }
#undef B200C_CASE
  return fail(B200C_EINVAL, "unsupported vectors-per-thread %d", vpt);
}

// one-shot fused, rows split over clusters of SPLIT CTAs.  The cluster shape is compiled into
// the kernel (__cluster_dims__), so no launch path can silently drop it; the launch attribute
// is therefore NOT passed (the two must not disagree).
template <typename T, int SPLIT>
int launch_ar_norm_cluster_s(b200c_ctx* const* ctxs, int nemu, const RankParams* rp, int threads,
                             int want_groups, cudaStream_t s) {
  if (nemu == 0) {
    int cap = 0;
    int rc = coresident_clusters(ar_norm_cluster_kernel<T, SPLIT>, threads, SPLIT, ctxs[0]->sm_count, &cap);
    if (rc) return rc;
    const int g = std::max(1, std::min(want_groups, cap));
    return launch(ar_norm_cluster_kernel<T, SPLIT>, dim3(g * SPLIT, 1, 1), threads, rp[0], s, false,
                  ctxs[0]->pdl != 0);
  }
  int cap = 0;
  int rc = coresident_clusters(ar_norm_cluster_kernel_emu<T, SPLIT>, threads, SPLIT, ctxs[0]->sm_count, &cap);
  if (rc) return rc;
  const int g = std::max(1, std::min(want_groups, cap / nemu));
  EmuParams ep;
  memset(&ep, 0, sizeof(ep));
  for (int r = 0; r < nemu; ++r) ep.r[r] = rp[r];
  // all emulated ranks wait on each other inside ONE launch: cooperative = guaranteed co-residency
  return launch(ar_norm_cluster_kernel_emu<T, SPLIT>, dim3(g * SPLIT, nemu, 1), threads, ep, s, true);
}

template <typename T>
int launch_ar_norm_cluster(b200c_ctx* const* ctxs, int nemu, const RankParams* rp, int split,
                           int threads, int want_groups, cudaStream_t s) {
  switch (split) {
    case 1: return launch_ar_norm_cluster_s<T, 1>(ctxs, nemu, rp, threads, want_groups, s);
    case 2: return launch_ar_norm_cluster_s<T, 2>(ctxs, nemu, rp, threads, want_groups, s);
    case 4: return launch_ar_norm_cluster_s<T, 4>(ctxs, nemu, rp, threads, want_groups, s);
    case 8: return launch_ar_norm_cluster_s<T, 8>(ctxs, nemu, rp, threads, want_groups, s);
  }
  return fail(B200C_EINVAL, "cluster split %d is not one of 1, 2, 4, 8", split);
}

// ---- NVLS variant (real ranks only: a multicast group has one member per device) -------------
int nvls_groups_for(const b200c_ctx* c, size_t bytes) {
  if (c->nvls_groups > 0) return std::min(c->nvls_groups, kNvMaxGroups);
  // One pipeline group per 8 MiB, at most 8 (tools/nvls_probe.py on 8 x B200, profiles/
  // r2_nvls_probe_tp8_f.json): finer groups shorten the consumers' tail but every group costs the
  // reduce side a drain of its multicast stores; 8 groups x 8 teams was the best point at 64 MiB.
  return (int)std::max<size_t>(1, std::min<size_t>(8, bytes >> 23));
}

template <typename K>
int launch_nvls_k(K kernel, b200c_ctx* c, RankParams p, size_t bytes, cudaStream_t s) {
  const int threads = 512;
  int cap = 0;
  int rc = coresident_blocks(kernel, threads, c->sm_count, &cap);
  if (rc) return rc;
  const int g = std::max(2, std::min(twoshot_grid(c, bytes), cap));
  // roles (tools/nvls_bench.cu, tools/nvls_probe.py): about a sixth of the grid saturates the
  // link on the reduce side (more CTAs only add contention); every other CTA copies in and
  // then consumes
  p.nv_groups = nvls_groups_for(c, bytes);
  int red = c->nvls_link_blocks > 0 ? c->nvls_link_blocks : std::max(1, g / 6);
  red = std::max(1, std::min(red, g - 1));
  int teams = c->nvls_teams > 0 ? c->nvls_teams : p.nv_groups;  // every group has its own team
  teams = std::max(1, std::min(teams, std::min(p.nv_groups, red)));
  red = red / teams * teams;  // equal teams
  p.link_blocks = red;
  p.nv_copy_blocks = g - red;
  p.nv_teams = teams;
  p.nv_debug = c->nvls_debug;
  return launch(kernel, dim3(g, 1, 1), threads, p, s, false, c->pdl != 0);
}

template <typename T>
int launch_nvls_norm(b200c_ctx* c, const RankParams& p, int vpt, size_t bytes, cudaStream_t s) {
  switch (vpt) {
    case 1: return launch_nvls_k(nvls_kernel<T, 1, true>, c, p, bytes, s);
    case 2: return launch_nvls_k(nvls_kernel<T, 2, true>, c, p, bytes, s);
    case 4: return launch_nvls_k(nvls_kernel<T, 4, true>, c, p, bytes, s);
  }
  return fail(B200C_EINVAL, "unsupported vectors-per-thread %d", vpt);
}

template <typename T, bool TWOSHOT>
int launch_ar_t(b200c_ctx* const* ctxs, int nemu, const RankParams* rp, int want_blocks,
                cudaStream_t s) {
  const int threads = 512;
  if (nemu == 0) {
    int cap = 0;
    int rc = coresident_blocks(ar_kernel<T, TWOSHOT>, threads, ctxs[0]->sm_count, &cap);
    if (rc) return rc;
    const int g = std::max(1, std::min(want_blocks, cap));
    RankParams p = rp[0];
    p.link_blocks = TWOSHOT ? link_split(ctxs[0], g) : 0;
    return launch(ar_kernel<T, TWOSHOT>, dim3(g, 1, 1), threads, p, s, false, ctxs[0]->pdl != 0);
  }
  int cap = 0;
  int rc = coresident_blocks(ar_kernel_emu<T, TWOSHOT>, threads, ctxs[0]->sm_count, &cap);
  if (rc) return rc;
  const int g = std::max(1, std::min(want_blocks, cap / nemu));
  EmuParams ep;
  memset(&ep, 0, sizeof(ep));
  for (int r = 0; r < nemu; ++r) {
    ep.r[r] = rp[r];
    ep.r[r].link_blocks = TWOSHOT ? link_split(ctxs[0], g) : 0;
  }
  return launch(ar_kernel_emu<T, TWOSHOT>, dim3(g, nemu, 1), threads, ep, s, true);
}

// Tuning aid (tools/norm_sweep.py): B200C_NORM_VARIANT = rows | small2 | small4 | small8
// forces one bandwidth-regime kernel; unset = the measured choice below.
int norm_variant_override() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200C_NORM_VARIANT");
    v = 0;
    if (e) {
      if (!strcmp(e, "rows")) v = 1;
      else if (!strcmp(e, "small2")) v = 3;
      else if (!strcmp(e, "small4")) v = 4;
      else if (!strcmp(e, "small8")) v = 5;
    }
  }
  return v;
}

template <typename T>
int launch_norm_only(const RankParams& rp, int vpt, int threads, int sm_count, cudaStream_t s,
                     bool pdl) {
  const int force = norm_variant_override();
  const bool bw = rp.tokens >= 8 * sm_count;  // every SM has several rows of work
  // Many small CTAs, one row each (tools/norm_sweep.py, profiles/r2_norm_sweep_tp1.json): rows of
  // up to 512 vectors from two rows per SM upwards, wider rows once every SM has eight.  The grid
  // is one CTA per row: capping it and looping cost 3-4 % at 16384 rows (wave tails).
  const bool small_ok = rp.hv % 64 == 0 && rp.hv <= 2048 &&
                        (rp.hv <= 512 ? rp.tokens >= 2 * sm_count : bw);
  if (small_ok && (force == 0 || (force >= 3 && force <= 5))) {
    int v = force == 3 ? 2 : force == 4 ? 4 : force == 5 ? 8 : (rp.hv <= 512 ? 2 : rp.hv <= 1024 ? 4 : 8);
    while (rp.hv / v > 256) v *= 2;
    const int t = rp.hv / v;
    if (v <= 8 && t % 32 == 0) {
      const int g = std::max(1, rp.tokens);
      switch (v) {
        case 2: return launch(norm_kernel_small<T, 2, 6>, dim3(g, 1, 1), t, rp, s, false, pdl);
        case 4: return launch(norm_kernel_small<T, 4, 5>, dim3(g, 1, 1), t, rp, s, false, pdl);
        case 8: return launch(norm_kernel_small<T, 8, 3>, dim3(g, 1, 1), t, rp, s, false, pdl);
      }
    }
  }
  if (vpt == 1 && bw) {
    // narrow rows: four rows per trip (see norm_kernel_rows)
    const int g = std::max(1, std::min((rp.tokens + 3) / 4, sm_count * 8));
    return launch(norm_kernel_rows<T, 4>, dim3(g, 1, 1), threads, rp, s, false, pdl);
  }
  if (vpt == 1 && rp.tokens >= 4 * sm_count) {
    const int g = std::max(1, std::min((rp.tokens + 1) / 2, sm_count * 16));
    return launch(norm_kernel_rows<T, 2>, dim3(g, 1, 1), threads, rp, s, false, pdl);
  }
  const int g = std::max(1, std::min(rp.tokens, sm_count * 16));
  switch (vpt) {
    case 1: return launch(norm_kernel<T, 1>, dim3(g, 1, 1), threads, rp, s, false, pdl);
    case 2: return launch(norm_kernel<T, 2>, dim3(g, 1, 1), threads, rp, s, false, pdl);
    case 4: return launch(norm_kernel<T, 4>, dim3(g, 1, 1), threads, rp, s, false, pdl);
  }
  return fail(B200C_EINVAL, "unsupported vectors-per-thread %d", vpt);
}

int norm_only(RankParams& rp, int dtype, int sm_count, cudaStream_t s, bool pdl = true) {
  int vpt, threads;
  norm_shape(rp.hv, &vpt, &threads);
  switch (dtype) {
    case B200C_BF16: return launch_norm_only<__nv_bfloat16>(rp, vpt, threads, sm_count, s, pdl);
    case B200C_F16: return launch_norm_only<__half>(rp, vpt, threads, sm_count, s, pdl);
    case B200C_F32: return launch_norm_only<float>(rp, vpt, threads, sm_count, s, pdl);
  }
  return fail(B200C_EINVAL, "bad dtype %d", dtype);
}

int check_rows(int tokens, int hidden, int dtype, int* hv) {
  const int es = dtype_size(dtype);
  if (!es) return fail(B200C_EINVAL, "bad dtype %d", dtype);
  if (tokens < 0 || hidden <= 0) return fail(B200C_EINVAL, "bad shape [%d, %d]", tokens, hidden);
  if (((size_t)hidden * es) % 16) return fail(B200C_EINVAL, "hidden*sizeof(T) must be a multiple of 16");
  *hv = (int)((size_t)hidden * es / 16);
  if (*hv > 4 * 512) return fail(B200C_EINVAL, "hidden %d too large for the fused row kernel", hidden);
  return B200C_OK;
}

int sm_count_of_current_device(int* dev, int* sms) {
  static int cached[64] = {};  // asked on every ctx-less norm call
  CUDA_TRY(cudaGetDevice(dev));
  if (*dev >= 0 && *dev < 64 && cached[*dev] > 0) {
    *sms = cached[*dev];
    return B200C_OK;
  }
  CUDA_TRY(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, *dev));
  if (*dev >= 0 && *dev < 64) cached[*dev] = *sms;
  return B200C_OK;
}

// shared implementation of the real (nemu == 0) and emulated fused entry points
int ar_norm_impl(b200c_ctx* const* ctxs, int nemu, const void* const* ins,
                 const void* const* res_ins, void* const* res_outs, void* const* norm_outs,
                 const void* gamma, float eps, int tokens, int hidden, int dtype, int variant,
                 cudaStream_t s) {
  b200c_ctx* c0 = ctxs[0];
  int hv = 0;
  int rc = check_rows(tokens, hidden, dtype, &hv);
  if (rc) return rc;
  if (tokens == 0) return B200C_OK;
  const int n = nemu ? nemu : 1;
  RankParams rp[kMaxRanks];
  for (int r = 0; r < n; ++r) {
    if (!ins[r] || !norm_outs[r] || !gamma || !aligned16(ins[r]) || !aligned16(norm_outs[r]) ||
        !aligned16(gamma) || !aligned16(res_ins[r]) || !aligned16(res_outs[r]))
      return fail(B200C_EINVAL, "null or misaligned pointer");
    if ((res_ins[r] == nullptr) != (res_outs[r] == nullptr))
      return fail(B200C_EINVAL, "residual_in and residual_out must both be given or both be NULL");
    fill_rank_params(ctxs[r], rp[r]);
    rp[r].in = ins[r];
    rp[r].res_in = res_ins[r];
    rp[r].res_out = res_outs[r];
    rp[r].norm_out = norm_outs[r];
    rp[r].gamma = gamma;
    rp[r].eps = eps;
    rp[r].tokens = tokens;
    rp[r].hv = hv;
    rp[r].hidden = hidden;
    rp[r].nvec = (unsigned long long)tokens * hv;
  }
  if (c0->world == 1) return norm_only(rp[0], dtype, c0->sm_count, s, c0->pdl != 0);

  const size_t bytes = (size_t)tokens * hv * 16;
  variant = pick_variant(c0, bytes, variant);
  rc = check_ar_fit(c0, bytes, variant);
  if (rc) return rc;
  int vpt, threads;
  norm_shape(hv, &vpt, &threads);
  if (variant == B200C_NVLS) {
    if (nemu) return fail(B200C_EINVAL, "the NVLS variant cannot be emulated on one GPU");
    // the row CTAs run 512 threads here (one launch shape for link and row work)
    int v = 1;
    while (v < 4 && (hv + v - 1) / v > 512) v *= 2;
    switch (dtype) {
      case B200C_BF16: return launch_nvls_norm<__nv_bfloat16>(c0, rp[0], v, bytes, s);
      case B200C_F16: return launch_nvls_norm<__half>(c0, rp[0], v, bytes, s);
      case B200C_F32: return launch_nvls_norm<float>(c0, rp[0], v, bytes, s);
    }
  }
  const bool two = variant == B200C_TWOSHOT;
  if (!two && c0->oneshot_split != 1) {
    // latency regime: one vector per thread, a row spread over up to 8 CTAs of a cluster
    // (compiled shapes: 1, 2, 4, 8 CTAs per row)
    int split = 1;
    const int want_split = c0->oneshot_split > 1 ? c0->oneshot_split : (hv + 127) / 128;
    while (split < 8 && split < want_split) split <<= 1;
    if (c0->oneshot_split == 0) {
      // splitting pays while all tokens x split CTAs are resident at once (about 2 per SM);
      // beyond that one wider CTA per row is faster (measured, tools/sweep.py)
      while (split > 1 && (long long)tokens * split > 2LL * c0->sm_count) split >>= 1;
    }
    while (split < 8 && (hv + split - 1) / split > 256) split <<= 1;
    if ((hv + split - 1) / split <= 256) {
      const int cthreads = std::max(32, ((hv + split - 1) / split + 31) / 32 * 32);
      const int want_groups = std::min(tokens, c0->oneshot_blocks);
      switch (dtype) {
        case B200C_BF16: return launch_ar_norm_cluster<__nv_bfloat16>(ctxs, nemu, rp, split, cthreads, want_groups, s);
        case B200C_F16: return launch_ar_norm_cluster<__half>(ctxs, nemu, rp, split, cthreads, want_groups, s);
        case B200C_F32: return launch_ar_norm_cluster<float>(ctxs, nemu, rp, split, cthreads, want_groups, s);
      }
    }
  }
  const int want = two ? twoshot_grid(c0, bytes) : std::min(tokens, c0->oneshot_blocks);
  switch (dtype) {
    case B200C_BF16: return launch_ar_norm_v<__nv_bfloat16>(ctxs, nemu, rp, vpt, threads, two, want, s);
    case B200C_F16: return launch_ar_norm_v<__half>(ctxs, nemu, rp, vpt, threads, two, want, s);
    case B200C_F32: return launch_ar_norm_v<float>(ctxs, nemu, rp, vpt, threads, two, want, s);
  }
  return fail(B200C_EINVAL, "bad dtype %d", dtype);
}

int ar_impl(b200c_ctx* const* ctxs, int nemu, const void* const* ins, void* const* outs,
            size_t numel, int dtype, int variant, cudaStream_t s) {
  b200c_ctx* c0 = ctxs[0];
  const int es = dtype_size(dtype);
  if (!es) return fail(B200C_EINVAL, "bad dtype %d", dtype);
  const size_t bytes = numel * es;
  if (bytes % 16) return fail(B200C_EINVAL, "numel*sizeof(T) = %zu is not a multiple of 16", bytes);
  if (bytes == 0) return B200C_OK;
  const int n = nemu ? nemu : 1;
  RankParams rp[kMaxRanks];
  for (int r = 0; r < n; ++r) {
    if (!ins[r] || !outs[r] || !aligned16(ins[r]) || !aligned16(outs[r]))
      return fail(B200C_EINVAL, "null or misaligned pointer");
    fill_rank_params(ctxs[r], rp[r]);
    rp[r].in = ins[r];
    rp[r].out = outs[r];
    rp[r].nvec = bytes / 16;
  }
  if (c0->world == 1) {
    if (ins[0] != outs[0])
      CUDA_TRY(cudaMemcpyAsync(outs[0], ins[0], bytes, cudaMemcpyDeviceToDevice, s));
    return B200C_OK;
  }
  variant = pick_variant(c0, bytes, variant);
  int rc = check_ar_fit(c0, bytes, variant);
  if (rc) return rc;
  if (variant == B200C_NVLS) {
    if (nemu) return fail(B200C_EINVAL, "the NVLS variant cannot be emulated on one GPU");
    switch (dtype) {
      case B200C_BF16: return launch_nvls_k(nvls_kernel<__nv_bfloat16, 1, false>, c0, rp[0], bytes, s);
      case B200C_F16: return launch_nvls_k(nvls_kernel<__half, 1, false>, c0, rp[0], bytes, s);
      case B200C_F32: return launch_nvls_k(nvls_kernel<float, 1, false>, c0, rp[0], bytes, s);
    }
  }
  const bool two = variant == B200C_TWOSHOT;
  const size_t nvec = bytes / 16;
  int want = two ? twoshot_grid(c0, bytes) : c0->oneshot_blocks;
  want = (int)std::min<size_t>((size_t)want, (nvec + 511) / 512);
  if (two) want = std::max(want, 2);
#define B200C_AR(T) \
  (two ? launch_ar_t<T, true>(ctxs, nemu, rp, want, s) : launch_ar_t<T, false>(ctxs, nemu, rp, want, s))
  switch (dtype) {
    case B200C_BF16: return B200C_AR(__nv_bfloat16);
    case B200C_F16: return B200C_AR(__half);
    case B200C_F32: return B200C_AR(float);
  }
#undef B200C_AR
  return fail(B200C_EINVAL, "bad dtype %d", dtype);
}

int check_emu(b200c_ctx* const* ctxs, int world) {
  if (!ctxs || world < 2 || world > kMaxRanks) return fail(B200C_EINVAL, "bad emulated world %d", world);
  for (int r = 0; r < world; ++r)
    if (!ctxs[r] || ctxs[r]->world != world || ctxs[r]->rank != r)
      return fail(B200C_EINVAL, "ctxs[%d] is not rank %d of a %d-rank group", r, r, world);
  return B200C_OK;
}

template <typename K>
int copy_grid(const b200c_ctx* c, size_t nvec, int nemu, K kernel_fn, int* g) {
  int cap = 0;
  int rc = coresident_blocks(kernel_fn, 512, c->sm_count, &cap);
  if (rc) return rc;
  cap = std::max(1, cap / std::max(1, nemu));
  size_t want = std::min<size_t>((size_t)c->copy_blocks, (nvec + 511) / 512);
  *g = (int)std::max<size_t>(1, std::min<size_t>(want, (size_t)cap));
  return B200C_OK;
}

}  // namespace

// ================================ extern "C" =========================================================

extern "C" {

int b200c_version(void) { return 1; }

const char* b200c_last_error(void) { return g_err; }

int b200c_arena_bytes(int world, size_t max_bytes, size_t ag_bytes, size_t p2p_bytes,
                      size_t* out_bytes) {
  if (world < 1 || world > kMaxRanks || !out_bytes) return fail(B200C_EINVAL, "bad world %d", world);
  make_geom(world, max_bytes, ag_bytes, p2p_bytes, out_bytes);
  return B200C_OK;
}

int b200c_alloc_shared(size_t bytes, void** dptr, uint8_t handle[B200C_IPC_HANDLE_BYTES]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == B200C_IPC_HANDLE_BYTES, "IPC handle size");
  if (!dptr || !bytes) return fail(B200C_EINVAL, "bad arguments");
  CUDA_TRY(cudaMalloc(dptr, bytes));
  if (handle) {
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, *dptr);
    if (e != cudaSuccess) {
      cudaFree(*dptr);
      *dptr = nullptr;
      return fail(B200C_ECUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    memcpy(handle, &h, sizeof(h));
  }
  return B200C_OK;
}

int b200c_open_handle(const uint8_t handle[B200C_IPC_HANDLE_BYTES], void** dptr) {
  if (!handle || !dptr) return fail(B200C_EINVAL, "bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  CUDA_TRY(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200C_OK;
}

int b200c_close_handle(void* dptr) {
  if (!dptr) return B200C_OK;
  CUDA_TRY(cudaIpcCloseMemHandle(dptr));
  return B200C_OK;
}

int b200c_free_shared(void* dptr) {
  if (!dptr) return B200C_OK;
  CUDA_TRY(cudaFree(dptr));
  return B200C_OK;
}

int b200c_init(int rank, int world, void* const* arena_ptrs, size_t max_bytes, size_t ag_bytes,
               size_t p2p_bytes, b200c_ctx** out) {
  if (!out) return fail(B200C_EINVAL, "ctx out pointer is NULL");
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world)
    return fail(B200C_EINVAL, "bad rank %d / world %d (1..%d ranks)", rank, world, kMaxRanks);
  if (world > 1) {
    if (!arena_ptrs) return fail(B200C_EINVAL, "arena_ptrs is NULL");
    for (int r = 0; r < world; ++r)
      if (!arena_ptrs[r] || !aligned16(arena_ptrs[r]))
        return fail(B200C_EINVAL, "arena_ptrs[%d] is NULL or misaligned", r);
    if (max_bytes % 16 || ag_bytes % 16 || p2p_bytes % 16)
      return fail(B200C_EINVAL, "arena sizes must be multiples of 16");
  }
  b200c_ctx* c = new (std::nothrow) b200c_ctx();
  if (!c) return fail(B200C_ENOMEM, "out of host memory");
  c->rank = rank;
  c->world = world;
  int rc = sm_count_of_current_device(&c->device, &c->sm_count);
  if (rc) {
    delete c;
    return rc;
  }
  c->oneshot_blocks = c->sm_count;
  c->twoshot_blocks = 2 * c->sm_count;  // two resident CTAs per SM (clamped to what fits)
  // one-shot / two-shot crossover measured on 8 x B200 (tools/sweep.py, profiles/): the more
  // ranks, the sooner (N-1)*S of one-shot link traffic costs more than a second fabric hop
  c->oneshot_max_bytes = world <= 2 ? (2 << 20) : world <= 4 ? (512 << 10) : (384 << 10);
  if (world > 1) {
    c->max_bytes = max_bytes;
    c->ag_bytes = ag_bytes;
    c->p2p_bytes = p2p_bytes;
    size_t total = 0;
    c->geom = make_geom(world, max_bytes, ag_bytes, p2p_bytes, &total);
    for (int r = 0; r < world; ++r) c->arena[r] = static_cast<char*>(arena_ptrs[r]);
    cudaError_t e = cudaMalloc(&c->st, sizeof(LocalState));
    if (e == cudaSuccess) e = cudaMemset(c->st, 0, sizeof(LocalState));
    // own arena: flags to zero, receive regions to "not arrived"
    if (e == cudaSuccess) e = cudaMemset(c->arena[rank], 0, kCtrlBytes);
    if (e == cudaSuccess) {
      const size_t nvec = 3 * c->geom.lam_stride / 16;
      fill_not_arrived_kernel<<<c->sm_count * 4, 256>>>(
          reinterpret_cast<uint4*>(c->arena[rank] + kCtrlBytes), nvec);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      if (c->st) cudaFree(c->st);
      delete c;
      return fail(B200C_ECUDA, "b200c_init: %s", cudaGetErrorString(e));
    }
  }
  *out = c;
  return B200C_OK;
}

int b200c_destroy(b200c_ctx* c) {
  if (!c) return B200C_OK;
  if (c->st) {
    cudaDeviceSynchronize();
    cudaFree(c->st);
  }
  delete c;
  return B200C_OK;
}

int b200c_nvls_bytes(size_t max_bytes, size_t* out_bytes) {
  if (!out_bytes) return fail(B200C_EINVAL, "out_bytes is NULL");
  *out_bytes = kNvCtrlBytes + region_half(max_bytes);
  return B200C_OK;
}

int b200c_attach_nvls(b200c_ctx* c, void* uc_ptr, void* mc_ptr, size_t bytes) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  if (c->world < 2 || !c->st) return fail(B200C_EINVAL, "NVLS needs a group of at least 2 ranks");
  if (!uc_ptr || !mc_ptr || !aligned16(uc_ptr) || !aligned16(mc_ptr))
    return fail(B200C_EINVAL, "NULL or misaligned NVLS buffer");
  const size_t half = region_half(c->max_bytes);
  if (bytes < kNvCtrlBytes + half)
    return fail(B200C_ETOOBIG, "NVLS buffer of %zu bytes is smaller than the %zu this ctx needs", bytes,
                kNvCtrlBytes + half);
  char* uc = static_cast<char*>(uc_ptr);
  // own copy of the flags to zero (through the unicast mapping only); the data area needs no arming
  CUDA_TRY(cudaMemset(uc, 0, kNvCtrlBytes));
  CUDA_TRY(cudaDeviceSynchronize());
  c->nv_uc = uc;
  c->nv_mc = static_cast<char*>(mc_ptr);
  c->nv_half = half;
  // AUTO: measured on 8 x B200 (profiles/r2_sweep_tp8_h8192_final.json) the in-switch variant wins
  // from 64 MiB upwards at 8 ranks (194 us vs 215 us) and loses below; at 2 ranks it moves MORE
  // bytes than the push kernels and at 4 it was not measured to win, so it stays opt-in there
  // (B200C_TUNE_NVLS_MIN_BYTES).  It is tolerance-parity, not bit-parity (see nvls_body).
  if (c->nvls_min_bytes < 0 && c->world >= 8) c->nvls_min_bytes = 48ll << 20;
  return B200C_OK;
}

int b200c_set_tuning(b200c_ctx* c, int key, long long v) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  switch (key) {
    case B200C_TUNE_ONESHOT_MAX_BYTES: c->oneshot_max_bytes = v; break;
    case B200C_TUNE_ONESHOT_BLOCKS: c->oneshot_blocks = (int)std::max<long long>(1, v); break;
    case B200C_TUNE_TWOSHOT_BLOCKS: c->twoshot_blocks = (int)std::max<long long>(1, v); break;
    case B200C_TUNE_SPIN_TIMEOUT_MS: c->spin_timeout_ms = std::max<long long>(0, v); break;
    case B200C_TUNE_COPY_BLOCKS: c->copy_blocks = (int)std::max<long long>(1, v); break;
    case B200C_TUNE_LINK_BLOCKS: c->link_blocks = (int)std::max<long long>(0, v); break;
    case B200C_TUNE_PDL: c->pdl = v != 0; break;
    case B200C_TUNE_ONESHOT_SPLIT: c->oneshot_split = (int)std::max<long long>(0, std::min<long long>(8, v)); break;
    case B200C_TUNE_TRAP_ON_TIMEOUT: c->trap_on_timeout = v != 0; break;
    case B200C_TUNE_NVLS_MIN_BYTES: c->nvls_min_bytes = v; break;
    case B200C_TUNE_NVLS_LINK_BLOCKS: c->nvls_link_blocks = (int)std::max<long long>(0, v); break;
    case B200C_TUNE_NVLS_GROUPS: c->nvls_groups = (int)std::max<long long>(0, std::min<long long>(kNvMaxGroups, v)); break;
    case B200C_TUNE_NVLS_COPY_BLOCKS: c->nvls_copy_blocks = (int)std::max<long long>(0, v); break;
    case B200C_TUNE_NVLS_TEAMS: c->nvls_teams = (int)std::max<long long>(0, std::min<long long>(kNvMaxGroups, v)); break;
    case B200C_TUNE_NVLS_DEBUG: c->nvls_debug = (int)v; break;
    case B200C_TUNE_DEBUG_EPOCH: {
      // test hook (host-synchronous): jump the all-reduce call counter, e.g. to just below 2^32
      if (!c->st) return fail(B200C_EINVAL, "no device state (world == 1)");
      const unsigned long long w = (unsigned long long)(uint32_t)v << 32;
      CUDA_TRY(cudaDeviceSynchronize());
      CUDA_TRY(cudaMemcpy(&c->st->call, &w, sizeof(w), cudaMemcpyHostToDevice));
      break;
    }
    default: return fail(B200C_EINVAL, "unknown tuning key %d", key);
  }
  return B200C_OK;
}

int b200c_get_tuning(b200c_ctx* c, int key, long long* v) {
  if (!c || !v) return fail(B200C_EINVAL, "bad arguments");
  switch (key) {
    case B200C_TUNE_ONESHOT_MAX_BYTES: *v = c->oneshot_max_bytes; break;
    case B200C_TUNE_ONESHOT_BLOCKS: *v = c->oneshot_blocks; break;
    case B200C_TUNE_TWOSHOT_BLOCKS: *v = c->twoshot_blocks; break;
    case B200C_TUNE_SPIN_TIMEOUT_MS: *v = c->spin_timeout_ms; break;
    case B200C_TUNE_COPY_BLOCKS: *v = c->copy_blocks; break;
    case B200C_TUNE_LINK_BLOCKS: *v = c->link_blocks; break;
    case B200C_TUNE_PDL: *v = c->pdl; break;
    case B200C_TUNE_ONESHOT_SPLIT: *v = c->oneshot_split; break;
    case B200C_TUNE_TRAP_ON_TIMEOUT: *v = c->trap_on_timeout; break;
    case B200C_TUNE_NVLS_MIN_BYTES: *v = c->nvls_min_bytes; break;
    case B200C_TUNE_NVLS_LINK_BLOCKS: *v = c->nvls_link_blocks; break;
    case B200C_TUNE_NVLS_GROUPS: *v = c->nvls_groups; break;
    case B200C_TUNE_NVLS_COPY_BLOCKS: *v = c->nvls_copy_blocks; break;
    case B200C_TUNE_NVLS_TEAMS: *v = c->nvls_teams; break;
    case B200C_TUNE_NVLS_DEBUG: *v = c->nvls_debug; break;
    case B200C_TUNE_DEBUG_EPOCH: {
      if (!c->st) return fail(B200C_EINVAL, "no device state (world == 1)");
      unsigned long long w = 0;
      CUDA_TRY(cudaDeviceSynchronize());
      CUDA_TRY(cudaMemcpy(&w, &c->st->call, sizeof(w), cudaMemcpyDeviceToHost));
      *v = (long long)(w >> 32);
      break;
    }
    default: return fail(B200C_EINVAL, "unknown tuning key %d", key);
  }
  return B200C_OK;
}

int b200c_query_error(b200c_ctx* c) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  if (!c->st) return B200C_OK;
  uint32_t e = 0;
  CUDA_TRY(cudaMemcpy(&e, &c->st->error, sizeof(e), cudaMemcpyDeviceToHost));
  if (e == kErrClusterShape)
    return fail(B200C_ECUDA, "the one-shot cluster kernel ran with a cluster shape other than the "
                "one it was compiled for (rank %d); its output was not written", c->rank);
  if (e) return fail(B200C_ETIMEOUT, "a device-side wait on a peer timed out (rank %d)", c->rank);
  return B200C_OK;
}

int b200c_all_reduce(b200c_ctx* c, const void* in, void* out, size_t numel, int dtype, int variant,
                     void* stream) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  b200c_ctx* ctxs[1] = {c};
  const void* ins[1] = {in};
  void* outs[1] = {out};
  return ar_impl(ctxs, 0, ins, outs, numel, dtype, variant, static_cast<cudaStream_t>(stream));
}

int b200c_all_reduce_add_rmsnorm(b200c_ctx* c, const void* in, const void* residual_in,
                                 void* residual_out, void* norm_out, const void* gamma, float eps,
                                 int tokens, int hidden, int dtype, int variant, void* stream) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  b200c_ctx* ctxs[1] = {c};
  const void* ins[1] = {in};
  const void* ri[1] = {residual_in};
  void* ro[1] = {residual_out};
  void* no[1] = {norm_out};
  return ar_norm_impl(ctxs, 0, ins, ri, ro, no, gamma, eps, tokens, hidden, dtype, variant,
                      static_cast<cudaStream_t>(stream));
}

static int local_norm(void* out, const void* x, const void* res_in, void* res_out,
                      const void* gamma, float eps, int tokens, int hidden, int dtype,
                      void* stream) {
  int hv = 0;
  int rc = check_rows(tokens, hidden, dtype, &hv);
  if (rc) return rc;
  if (tokens == 0) return B200C_OK;
  if (!out || !x || !gamma || !aligned16(out) || !aligned16(x) || !aligned16(gamma) ||
      !aligned16(res_in) || !aligned16(res_out))
    return fail(B200C_EINVAL, "null or misaligned pointer");
  int dev = 0, sms = 0;
  rc = sm_count_of_current_device(&dev, &sms);
  if (rc) return rc;
  RankParams rp;
  memset(&rp, 0, sizeof(rp));
  rp.world = 1;
  rp.in = x;
  rp.res_in = res_in;
  rp.res_out = res_out;
  rp.norm_out = out;
  rp.gamma = gamma;
  rp.eps = eps;
  rp.tokens = tokens;
  rp.hv = hv;
  rp.hidden = hidden;
  rp.nvec = (unsigned long long)tokens * hv;
  return norm_only(rp, dtype, sms, static_cast<cudaStream_t>(stream));
}

int b200c_fused_add_rms_norm(void* x_inout, void* residual_inout, const void* gamma, float eps,
                             int tokens, int hidden, int dtype, void* stream) {
  if (!residual_inout) return fail(B200C_EINVAL, "residual is NULL");
  return local_norm(x_inout, x_inout, residual_inout, residual_inout, gamma, eps, tokens, hidden,
                    dtype, stream);
}

int b200c_rms_norm(void* out, const void* x, const void* gamma, float eps, int tokens, int hidden,
                   int dtype, void* stream) {
  return local_norm(out, x, nullptr, nullptr, gamma, eps, tokens, hidden, dtype, stream);
}

}  // extern "C" (templates below need C++ linkage)

template <int VEC>
static int all_gather_launch(b200c_ctx* const* ctxs, int nemu, CopyParams* cp, size_t slab_words,
                             cudaStream_t s) {
  b200c_ctx* c0 = ctxs[0];
  int g = 1;
  if (nemu == 0) {
    int rc = copy_grid(c0, slab_words, 0, all_gather_kernel<VEC>, &g);
    if (rc) return rc;
    return launch(all_gather_kernel<VEC>, dim3(g, 1, 1), 512, cp[0], s, false);
  }
  int rc = copy_grid(c0, slab_words, nemu, all_gather_kernel_emu<VEC>, &g);
  if (rc) return rc;
  EmuCopyParams ep;
  memset(&ep, 0, sizeof(ep));
  for (int r = 0; r < nemu; ++r) ep.r[r] = cp[r];
  return launch(all_gather_kernel_emu<VEC>, dim3(g, nemu, 1), 512, ep, s, true);
}

static int all_gather_impl(b200c_ctx* const* ctxs, int nemu, const void* const* ins,
                           void* const* outs, int rows, size_t row_bytes, cudaStream_t s) {
  b200c_ctx* c0 = ctxs[0];
  if (rows < 0 || row_bytes % 4) return fail(B200C_EINVAL, "row_bytes must be a multiple of 4");
  if (rows == 0 || row_bytes == 0) return B200C_OK;
  // widest word that divides a row (16-byte vectors for logits, 8 for [batch, 2] fp32 pairs)
  const int vec = (row_bytes % 16 == 0) ? 16 : (row_bytes % 8 == 0) ? 8 : 4;
  const size_t out_bytes = (size_t)rows * row_bytes * c0->world;
  const int n = nemu ? nemu : 1;
  for (int r = 0; r < n; ++r)
    if (!ins[r] || !outs[r] || (reinterpret_cast<uintptr_t>(ins[r]) % vec) ||
        (reinterpret_cast<uintptr_t>(outs[r]) % vec))
      return fail(B200C_EINVAL, "null or misaligned pointer");
  if (c0->world == 1) {
    if (ins[0] != outs[0])
      CUDA_TRY(cudaMemcpyAsync(outs[0], ins[0], out_bytes, cudaMemcpyDeviceToDevice, s));
    return B200C_OK;
  }
  if (out_bytes > c0->ag_bytes)
    return fail(B200C_ETOOBIG, "all-gather output of %zu bytes exceeds ag_bytes %zu", out_bytes,
                c0->ag_bytes);
  const size_t slab_words = (size_t)rows * row_bytes / vec;
  CopyParams cp[kMaxRanks];
  for (int r = 0; r < n; ++r) {
    fill_copy_params(ctxs[r], cp[r]);
    cp[r].in = ins[r];
    cp[r].out = outs[r];
    cp[r].rows = rows;
    cp[r].row_vecs = row_bytes / vec;
  }
  switch (vec) {
    case 16: return all_gather_launch<16>(ctxs, nemu, cp, slab_words, s);
    case 8: return all_gather_launch<8>(ctxs, nemu, cp, slab_words, s);
    default: return all_gather_launch<4>(ctxs, nemu, cp, slab_words, s);
  }
}

extern "C" {

int b200c_all_gather_lastdim(b200c_ctx* c, const void* in, void* out, int rows, size_t row_bytes,
                             void* stream) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  b200c_ctx* ctxs[1] = {c};
  const void* ins[1] = {in};
  void* outs[1] = {out};
  return all_gather_impl(ctxs, 0, ins, outs, rows, row_bytes, static_cast<cudaStream_t>(stream));
}

static int p2p_check(b200c_ctx* c, int peer, const void* buf, size_t bytes) {
  if (!c) return fail(B200C_EINVAL, "ctx is NULL");
  if (c->world < 2) return fail(B200C_EINVAL, "send/recv need a group of at least 2 ranks");
  if (peer < 0 || peer >= c->world || peer == c->rank) return fail(B200C_EINVAL, "bad peer %d", peer);
  if (bytes % 16 || !buf || !aligned16(buf)) return fail(B200C_EINVAL, "buffer must be 16-byte aligned and sized");
  if (bytes > c->p2p_bytes)
    return fail(B200C_ETOOBIG, "message of %zu bytes exceeds p2p_bytes %zu", bytes, c->p2p_bytes);
  return B200C_OK;
}

int b200c_send(b200c_ctx* c, int dst, const void* buf, size_t bytes, void* stream) {
  int rc = p2p_check(c, dst, buf, bytes);
  if (rc) return rc;
  if (bytes == 0) return B200C_OK;
  CopyParams cp;
  fill_copy_params(c, cp);
  cp.in = buf;
  cp.peer = dst;
  cp.nvec = bytes / 16;
  int g = 1;
  rc = copy_grid(c, cp.nvec, 0, send_kernel, &g);
  if (rc) return rc;
  return launch(send_kernel, dim3(g, 1, 1), 512, cp, static_cast<cudaStream_t>(stream), false);
}

int b200c_recv(b200c_ctx* c, int src, void* buf, size_t bytes, void* stream) {
  int rc = p2p_check(c, src, buf, bytes);
  if (rc) return rc;
  if (bytes == 0) return B200C_OK;
  CopyParams cp;
  fill_copy_params(c, cp);
  cp.out = buf;
  cp.peer = src;
  cp.nvec = bytes / 16;
  int g = 1;
  rc = copy_grid(c, cp.nvec, 0, recv_kernel, &g);
  if (rc) return rc;
  return launch(recv_kernel, dim3(g, 1, 1), 512, cp, static_cast<cudaStream_t>(stream), false);
}

// ---- single-GPU rank emulation ----------------------------------------------------------------------

int b200c_emu_all_reduce(b200c_ctx* const* ctxs, int world, const void* const* ins,
                         void* const* outs, size_t numel, int dtype, int variant, void* stream) {
  int rc = check_emu(ctxs, world);
  if (rc) return rc;
  if (!ins || !outs) return fail(B200C_EINVAL, "null pointer array");
  return ar_impl(ctxs, world, ins, outs, numel, dtype, variant, static_cast<cudaStream_t>(stream));
}

int b200c_emu_all_reduce_add_rmsnorm(b200c_ctx* const* ctxs, int world, const void* const* ins,
                                     const void* const* residual_ins, void* const* residual_outs,
                                     void* const* norm_outs, const void* gamma, float eps,
                                     int tokens, int hidden, int dtype, int variant, void* stream) {
  int rc = check_emu(ctxs, world);
  if (rc) return rc;
  if (!ins || !residual_ins || !residual_outs || !norm_outs)
    return fail(B200C_EINVAL, "null pointer array");
  return ar_norm_impl(ctxs, world, ins, residual_ins, residual_outs, norm_outs, gamma, eps, tokens,
                      hidden, dtype, variant, static_cast<cudaStream_t>(stream));
}

int b200c_emu_all_gather_lastdim(b200c_ctx* const* ctxs, int world, const void* const* ins,
                                 void* const* outs, int rows, size_t row_bytes, void* stream) {
  int rc = check_emu(ctxs, world);
  if (rc) return rc;
  if (!ins || !outs) return fail(B200C_EINVAL, "null pointer array");
  return all_gather_impl(ctxs, world, ins, outs, rows, row_bytes, static_cast<cudaStream_t>(stream));
}

int b200c_emu_sendrecv(b200c_ctx* const* ctxs, int world, int src, int dst, const void* sendbuf,
                       void* recvbuf, size_t bytes, void* stream) {
  int rc = check_emu(ctxs, world);
  if (rc) return rc;
  if (src < 0 || src >= world || dst < 0 || dst >= world || src == dst)
    return fail(B200C_EINVAL, "bad src %d / dst %d", src, dst);
  rc = p2p_check(ctxs[src], dst, sendbuf, bytes);
  if (rc) return rc;
  rc = p2p_check(ctxs[dst], src, recvbuf, bytes);
  if (rc) return rc;
  if (bytes == 0) return B200C_OK;
  EmuCopyParams ep;
  memset(&ep, 0, sizeof(ep));
  fill_copy_params(ctxs[src], ep.r[0]);
  ep.r[0].in = sendbuf;
  ep.r[0].peer = dst;
  ep.r[0].nvec = bytes / 16;
  ep.r[0].role = 0;
  fill_copy_params(ctxs[dst], ep.r[1]);
  ep.r[1].out = recvbuf;
  ep.r[1].peer = src;
  ep.r[1].nvec = bytes / 16;
  ep.r[1].role = 1;
  int g = 1;
  rc = copy_grid(ctxs[src], bytes / 16, 2, sendrecv_kernel_emu, &g);
  if (rc) return rc;
  return launch(sendrecv_kernel_emu, dim3(g, 2, 1), 512, ep, static_cast<cudaStream_t>(stream), true);
}

}  // extern "C"

// ======================= entry point that allocates nothing ==========================================
//
// b200c_emu_all_reduce_noalloc() is b200c_emu_all_reduce() for callers that are not allowed
// to allocate device memory.  Inside an inference server that situation is the normal one
// once start-up is over: vLLM measures free HBM and hands all of it to the KV-cache pool, so a
// later cudaMalloc may simply fail; with its sleep-mode (cumem) allocator every allocation has
// to go through the tracked pool, which a raw cudaMalloc bypasses; and while any thread of the
// process is capturing a CUDA graph in the default capture mode, cudaMalloc is illegal.  A
// health check or self-test of the emulated all-reduce that runs in those windows cannot build
// the usual arenas + per-rank state, so here they live in one file-scope device array.
//
// What the caller must know:
//  * the array is kNoAllocScratchBytes (about 3.6 MiB) of HBM on every GPU that loads this
//    library, held until the process exits;
//  * limit: world <= 8 and at most B200C_NOALLOC_MAX_BYTES (64 KiB) per rank and call (the
//    one-shot variant additionally needs world * bytes <= 128 KiB); larger requests are
//    rejected with B200C_ETOOBIG and nothing is launched;
//  * the first call on a device looks the array's device address up (cudaGetSymbolAddress --
//    in host code the array's name is only a stand-in, never a usable device pointer).  Make
//    that first call before starting any stream capture;
//  * all calls share the one array and it keeps whatever the previous call left, so every
//    call first re-arms, on `stream`, the parts that must start clean (per-rank state to
//    zero, receive regions to "not arrived").  Two calls whose GPU work can overlap -- on
//    different streams, or from different threads or processes' worth of streams on the same
//    device -- race on the array; serialise them;
//  * no cudaMalloc / cudaFree / allocating library routine is called, on any call.
namespace {

constexpr size_t kNoAllocMaxBytes = 64 * 1024;
constexpr size_t kNoAllocHalf = kNoAllocMaxBytes + 256;  // make_geom's half for this max_bytes (chunk-rounding slack)
constexpr size_t kNoAllocArenaBytes = kCtrlBytes + 3 * 2 * kNoAllocHalf + 2 * 256 + kMaxRanks * 2 * 256;
constexpr size_t kNoAllocStateStride = 512;
static_assert(sizeof(LocalState) <= kNoAllocStateStride, "LocalState must fit its slot");
constexpr size_t kNoAllocStateOff = kMaxRanks * kNoAllocArenaBytes;
constexpr size_t kNoAllocScratchBytes = kNoAllocStateOff + kMaxRanks * kNoAllocStateStride;

constexpr int kNoAllocMaxDevices = 64;

}  // namespace

// File scope, device memory.  Host code may only take its address through
// cudaGetSymbolAddress; device code in this file could name it directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

struct NoAllocDevice {
  char* base = nullptr;  // device address of g_noalloc_scratch on that device
  int sm_count = 0;
};
NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
std::mutex g_noalloc_mutex;  // first calls from several threads must not race on the table

// Once per device, with that device current.
int noalloc_setup(int* dev_out, NoAllocDevice** out) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(B200C_EINVAL, "device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  NoAllocDevice& d = g_noalloc_dev[dev];
  if (d.base == nullptr) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255u) != 0)
      return fail(B200C_ECUDA, "static scratch has no usable device address");
    int sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    d.sm_count = sms;
    d.base = static_cast<char*>(p);
  }
  *dev_out = dev;
  *out = &d;
  return B200C_OK;
}

}  // namespace

extern "C" int b200c_emu_all_reduce_noalloc(int world, const void* const* ins, void* const* outs,
                                            size_t numel, int dtype, int variant, void* stream) {
  if (world < 2 || world > kMaxRanks) return fail(B200C_EINVAL, "bad emulated world %d", world);
  if (!ins || !outs) return fail(B200C_EINVAL, "null pointer array");
  const int es = dtype_size(dtype);
  if (!es) return fail(B200C_EINVAL, "bad dtype %d", dtype);
  const size_t bytes = numel * es;
  if (bytes > kNoAllocMaxBytes)
    return fail(B200C_ETOOBIG,
                "message of %zu bytes exceeds the no-allocation limit of %zu bytes per rank", bytes,
                kNoAllocMaxBytes);
  int dev = 0;
  NoAllocDevice* d = nullptr;
  int rc = noalloc_setup(&dev, &d);
  if (rc) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  // Host-side descriptors only (stack objects; nothing here owns device memory).
  b200c_ctx ctx[kMaxRanks];
  b200c_ctx* ctxs[kMaxRanks];
  size_t total = 0;
  const ArenaGeom geom = make_geom(world, kNoAllocMaxBytes, 16, 16, &total);
  if (total > kNoAllocArenaBytes) return fail(B200C_EINVAL, "internal: static arena too small");
  for (int r = 0; r < world; ++r) {
    ctx[r].rank = r;
    ctx[r].world = world;
    ctx[r].device = dev;
    ctx[r].sm_count = d->sm_count;
    ctx[r].max_bytes = kNoAllocMaxBytes;
    ctx[r].geom = geom;
    for (int q = 0; q < world; ++q) ctx[r].arena[q] = d->base + (size_t)q * kNoAllocArenaBytes;
    ctx[r].st = reinterpret_cast<LocalState*>(d->base + kNoAllocStateOff + (size_t)r * kNoAllocStateStride);
    ctx[r].oneshot_blocks = d->sm_count;
    ctx[r].twoshot_blocks = d->sm_count;
    // AUTO must only pick one-shot when world copies of the message fit one receive region
    ctx[r].oneshot_max_bytes =
        std::min<long long>(ctx[r].oneshot_max_bytes, (long long)(geom.lam_stride / world));
    ctxs[r] = &ctx[r];
  }
  // Validate before touching the array (same rules as the allocating entry point).
  if (bytes % 16) return fail(B200C_EINVAL, "numel*sizeof(T) = %zu is not a multiple of 16", bytes);
  if (bytes == 0) return B200C_OK;
  rc = check_ar_fit(ctxs[0], bytes, pick_variant(ctxs[0], bytes, variant));
  if (rc) return rc;

  // The array still holds the previous call's traffic: start every rank from call 0 again.
  CUDA_TRY(cudaMemsetAsync(d->base + kNoAllocStateOff, 0, kMaxRanks * kNoAllocStateStride, s));
  const size_t arm_vecs = (size_t)world * kNoAllocArenaBytes / 16;
  fill_not_arrived_kernel<<<std::max(1, d->sm_count), 256, 0, s>>>(
      reinterpret_cast<uint4*>(d->base), arm_vecs);
  CUDA_TRY(cudaGetLastError());
  return ar_impl(ctxs, world, ins, outs, numel, dtype, variant, s);
}
