// nvls_bench.cu -- NVSwitch multicast (NVLS) ceilings on one box, single process, all visible
// GPUs: multimem.ld_reduce latency and bandwidth against requests in flight, multimem.st
// bandwidth, the combined reduce+broadcast phase, and the same with CTAs polling the result
// region as the product kernel's row CTAs do.  Feeds the design of nvls_body (b200c_kernels.cuh).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/nvls_bench tools/nvls_bench.cu -lcuda
//   tools/nvls_bench [MiB per message, default 1 8 64]
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1);} } while (0)
#define CU(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char* s_; cuGetErrorString(r_, &s_); printf("%s: %s\n", #x, s_); exit(1);} } while (0)

constexpr unsigned kMarker = 0x80808080u;  // cudaMemset-able "not arrived" word for this bench

__device__ __forceinline__ uint4 ldred(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mcst(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ldpoll(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ldstream(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// MODE 0: ld_reduce only; 1: multimem.st only; 2: ld_reduce -> multimem.st (the reduce phase)
template <int MODE, int U>
__global__ void __launch_bounds__(512) phase_kernel(const char* mc_in, char* mc_out, size_t v0, size_t v1, unsigned* sink) {
  const size_t nt = (size_t)gridDim.x * blockDim.x;
  unsigned acc = 0;
  for (size_t base = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < v1; base += nt * U) {
    uint4 x[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * nt;
      if (MODE == 1) x[u] = make_uint4((unsigned)i, 1, 2, 3);
      else if (i < v1) x[u] = ldred(mc_in + i * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t i = base + (size_t)u * nt;
      if (i < v1) {
        if (MODE == 0) acc ^= x[u].x ^ x[u].y ^ x[u].z ^ x[u].w;
        else mcst(mc_out + i * 16, x[u]);
      }
    }
  }
  if (MODE == 0 && acc == 0x12345678u) *sink = acc;
}

// dependent chain of ld_reduce from one thread: round-trip latency through the switch
__global__ void latency_kernel(const char* mc, int iters, unsigned long long* ns) {
  unsigned long long t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  size_t off = 0;
  for (int i = 0; i < iters; ++i) {
    uint4 v = ldred(mc + off);
    off = (size_t)(v.x & 0x10u) + 32 * (size_t)(i & 63);  // depends on the loaded value
  }
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  *ns = (t1 - t0) + (off == 0xffffffffu);
}

// the product kernel's shape: LINK CTAs reduce+broadcast their chunk, ROW CTAs poll the whole
// local result region (data is its own flag) and copy it out
template <int U>
__global__ void __launch_bounds__(512, 2) fused_like_kernel(const char* mc_in, char* mc_sum, const char* uc_sum, char* out,
                                                             size_t v0, size_t v1, size_t nvec, int link_blocks, int backoff) {
  if ((int)blockIdx.x < link_blocks) {
    const size_t nt = (size_t)link_blocks * blockDim.x;
    for (size_t base = v0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < v1; base += nt * U) {
      uint4 x[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { const size_t i = base + (size_t)u * nt; if (i < v1) x[u] = ldred(mc_in + i * 16); }
#pragma unroll
      for (int u = 0; u < U; ++u) { const size_t i = base + (size_t)u * nt; if (i < v1) mcst(mc_sum + i * 16, x[u]); }
    }
  } else {
    const size_t nt = (size_t)(gridDim.x - link_blocks) * blockDim.x;
    for (size_t i = (size_t)(blockIdx.x - link_blocks) * blockDim.x + threadIdx.x; i < nvec; i += nt) {
      uint4 v = ldpoll(uc_sum + i * 16);
      while (v.x == kMarker || v.y == kMarker || v.z == kMarker || v.w == kMarker) {
        if (backoff) __nanosleep(backoff);
        v = ldpoll(uc_sum + i * 16);
      }
      *reinterpret_cast<uint4*>(out + i * 16) = v;
    }
  }
}

template <int U>
__global__ void __launch_bounds__(512) copy_kernel(const char* src, char* dst, size_t n) {
  const size_t nt = (size_t)gridDim.x * blockDim.x;
  for (size_t base = (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < n; base += nt * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { size_t i = base + u * nt; if (i < n) v[u] = ldstream(src + i * 16); }
#pragma unroll
    for (int u = 0; u < U; ++u) { size_t i = base + u * nt; if (i < n) *reinterpret_cast<uint4*>(dst + i * 16) = v[u]; }
  }
}

struct Box {
  int n = 0;
  size_t size = 0;
  CUmemGenericAllocationHandle mc_handle{};
  std::vector<CUmemGenericAllocationHandle> mem;
  std::vector<CUdeviceptr> uc;
  CUdeviceptr mc = 0;
  std::vector<cudaStream_t> stream;
  std::vector<cudaEvent_t> e0, e1;
  std::vector<char*> priv;  // ordinary cudaMalloc per device (input / output of copies)
};

static void setup(Box& b, size_t bytes) {
  CK(cudaGetDeviceCount(&b.n));
  b.n = std::min(b.n, 8);
  if (b.n < 2) { printf("needs >= 2 GPUs\n"); exit(0); }
  for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaFree(0)); }
  CUmulticastObjectProp prop = {};
  prop.numDevices = b.n;
  prop.size = bytes;
  prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CU(cuMulticastGetGranularity(&gran, &prop, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  b.size = (bytes + gran - 1) / gran * gran;
  prop.size = b.size;
  CU(cuMulticastCreate(&b.mc_handle, &prop));
  for (int d = 0; d < b.n; ++d) { CUdevice dev; CU(cuDeviceGet(&dev, d)); CU(cuMulticastAddDevice(b.mc_handle, dev)); }
  b.mem.resize(b.n); b.uc.resize(b.n); b.stream.resize(b.n); b.e0.resize(b.n); b.e1.resize(b.n); b.priv.resize(b.n);
  std::vector<CUmemAccessDesc> acc(b.n);
  for (int d = 0; d < b.n; ++d) { acc[d].location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc[d].location.id = d; acc[d].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; }
  for (int d = 0; d < b.n; ++d) {
    CK(cudaSetDevice(d));
    CUmemAllocationProp ap = {};
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d;
    ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    CU(cuMemCreate(&b.mem[d], b.size, &ap, 0));
    CU(cuMulticastBindMem(b.mc_handle, 0, b.mem[d], 0, b.size, 0));
    CU(cuMemAddressReserve(&b.uc[d], b.size, gran, 0, 0));
    CU(cuMemMap(b.uc[d], b.size, 0, b.mem[d], 0));
    CU(cuMemSetAccess(b.uc[d], b.size, &acc[d], 1));
    CK(cudaStreamCreate(&b.stream[d]));
    CK(cudaEventCreate(&b.e0[d])); CK(cudaEventCreate(&b.e1[d]));
    CK(cudaMalloc(&b.priv[d], bytes));
    CK(cudaMemset(b.priv[d], 1, bytes));
    CK(cudaMemset((void*)b.uc[d], 0, b.size));
  }
  CU(cuMemAddressReserve(&b.mc, b.size, gran, 0, 0));
  CU(cuMemMap(b.mc, b.size, 0, b.mc_handle, 0));
  CU(cuMemSetAccess(b.mc, b.size, acc.data(), b.n));
  for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
}

// run `launch(d)` on every device `reps` times back to back; returns the slowest device's us per launch
template <typename F>
static double timed(Box& b, int reps, F launch) {
  for (int w = 0; w < 2; ++w) {
    for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); launch(d); }
    for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(b.stream[d])); }
  }
  double worst = 0;
  for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(b.e0[d], b.stream[d])); }
  for (int r = 0; r < reps; ++r) for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); launch(d); }
  for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(b.e1[d], b.stream[d])); }
  for (int d = 0; d < b.n; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(b.stream[d])); float ms; CK(cudaEventElapsedTime(&ms, b.e0[d], b.e1[d])); worst = std::max(worst, (double)ms * 1e3 / reps); }
  return worst;
}

int main(int argc, char** argv) {
  std::vector<size_t> sizes;
  for (int i = 1; i < argc; ++i) sizes.push_back((size_t)atoi(argv[i]) << 20);
  if (sizes.empty()) sizes = {1u << 20, 8u << 20, 64u << 20};
  const size_t maxS = *std::max_element(sizes.begin(), sizes.end());
  Box b;
  setup(b, 2 * maxS);  // [0, maxS) = xin, [maxS, 2 maxS) = sums
  const int N = b.n;
  printf("NVLS bench: %d GPUs, multicast buffer %zu MiB\n", N, b.size >> 20);
  unsigned* sink; CK(cudaSetDevice(0)); CK(cudaMalloc(&sink, 4));

  {  // latency
    unsigned long long* ns; CK(cudaSetDevice(0)); CK(cudaMalloc(&ns, 8));
    latency_kernel<<<1, 1, 0, b.stream[0]>>>((const char*)b.mc, 2000, ns);
    CK(cudaStreamSynchronize(b.stream[0]));
    latency_kernel<<<1, 1, 0, b.stream[0]>>>((const char*)b.mc, 2000, ns);
    CK(cudaStreamSynchronize(b.stream[0]));
    unsigned long long h; CK(cudaMemcpy(&h, ns, 8, cudaMemcpyDeviceToHost));
    printf("multimem.ld_reduce dependent-chain latency: %.0f ns\n", (double)h / 2000);
  }

  const char* mc_in = (const char*)b.mc;
  char* mc_sum = (char*)b.mc + maxS;
  for (size_t S : sizes) {
    const size_t nvec = S / 16, chunk = (nvec + N - 1) / N;
    const double bus = 2.0 * (N - 1) / N * S;
    printf("\n== S = %zu MiB per rank, chunk per rank %zu KiB ==\n", S >> 20, chunk * 16 >> 10);
    auto show = [&](const char* name, int blocks, int U, double us, double bytes_per_rank) {
      printf("%-28s blocks %3d U %d : %8.1f us  %7.1f GB/s per rank   (all-reduce bus-equivalent %7.1f GB/s)\n", name, blocks, U, us,
             bytes_per_rank / us * 1e-3, bus / us * 1e-3);
    };
    const int reps = S >= (32u << 20) ? 10 : 50;
    for (int blocks : {16, 37, 74, 148, 296}) {
#define RUN(MODE, UU, NAME, BYTES)                                                                        \
  {                                                                                                       \
    auto f = [&](int d) {                                                                                 \
      const size_t v0 = std::min((size_t)d * chunk, nvec), v1 = std::min(v0 + chunk, nvec);              \
      phase_kernel<MODE, UU><<<blocks, 512, 0, b.stream[d]>>>(mc_in, mc_sum, v0, v1, sink);               \
    };                                                                                                    \
    show(NAME, blocks, UU, timed(b, reps, f), BYTES);                                                     \
  }
      RUN(0, 4, "ld_reduce own chunk", chunk * 16.0)
      RUN(0, 8, "ld_reduce own chunk", chunk * 16.0)
      RUN(1, 4, "multimem.st own chunk", chunk * 16.0)
      RUN(2, 4, "ld_reduce + multimem.st", chunk * 16.0)
      RUN(2, 8, "ld_reduce + multimem.st", chunk * 16.0)
    }
    for (int blocks : {74, 148}) {
      auto f = [&](int d) { copy_kernel<4><<<blocks, 512, 0, b.stream[d]>>>(b.priv[d], (char*)b.uc[d], nvec); };
      show("copy-in (local)", blocks, 4, timed(b, reps, f), 2.0 * S);
    }
    // product shape: link CTAs + polling row CTAs, grid 296 (2 per SM), one launch per re-armed buffer
    for (int backoff : {0, 200}) {
      for (int link : {37, 74, 148, 222}) {
        auto prep = [&](int d) { CK(cudaMemsetAsync((char*)b.uc[d] + maxS, 0x80, S, b.stream[d])); };
        auto f = [&](int d) {
          const size_t v0 = std::min((size_t)d * chunk, nvec), v1 = std::min(v0 + chunk, nvec);
          fused_like_kernel<4><<<296, 512, 0, b.stream[d]>>>(mc_in, mc_sum, (const char*)b.uc[d] + maxS, b.priv[d], v0, v1, nvec, link, backoff);
        };
        // one launch per measurement: the result region is re-armed in between
        double worst = 0;
        std::vector<double> sum(N, 0.0);
        const int r2 = 10;
        for (int r = -2; r < r2; ++r) {
          for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); prep(d); }
          for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(b.stream[d])); }
          for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(b.e0[d], b.stream[d])); f(d); CK(cudaEventRecord(b.e1[d], b.stream[d])); }
          for (int d = 0; d < N; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(b.stream[d])); float ms; CK(cudaEventElapsedTime(&ms, b.e0[d], b.e1[d])); if (r >= 0) sum[d] += ms; }
        }
        for (int d = 0; d < N; ++d) worst = std::max(worst, sum[d] * 1e3 / r2);
        char name[64];
        snprintf(name, sizeof(name), "reduce+bcast+poll bk%d", backoff);
        show(name, link, 4, worst, chunk * 16.0);
      }
    }
  }
  return 0;
}
