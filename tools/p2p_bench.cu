// p2p_bench.cu -- NVLink push/pull ceilings between two GPUs of one box (single process,
// cudaDeviceEnablePeerAccess).  Feeds the roofline denominators and the kernel design:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/p2p_bench tools/p2p_bench.cu
//   tools/p2p_bench            (needs >= 2 GPUs)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1);} } while (0)

enum { ST_WEAK = 0, ST_VOLATILE = 1, ST_CG = 2, ST_CS = 3 };

template <int KIND>
__device__ __forceinline__ void st16(void* p, uint4 v) {
  if (KIND == ST_WEAK) asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  if (KIND == ST_VOLATILE) asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  if (KIND == ST_CG) asm volatile("st.global.cg.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  if (KIND == ST_CS) asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld16(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// copy n vectors src -> dst, U loads in flight per thread
template <int KIND, int U>
__global__ void __launch_bounds__(1024) copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
  const size_t nt = (size_t)gridDim.x * blockDim.x;
  for (size_t base = (size_t)blockIdx.x * blockDim.x + threadIdx.x; base < n; base += nt * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { size_t i = base + u * nt; if (i < n) v[u] = ld16(src + i); }
#pragma unroll
    for (int u = 0; u < U; ++u) { size_t i = base + u * nt; if (i < n) st16<KIND>(dst + i, v[u]); }
  }
}

// one-way latency: GPU a writes a counter into GPU b's memory, GPU b echoes it back
__global__ void ping_kernel(volatile unsigned* mine, volatile unsigned* theirs, int iters, int first, long long* cycles) {
  long long t0 = clock64();
  for (int i = 1; i <= iters; ++i) {
    if (first) { *theirs = i; while (*mine != (unsigned)i) {} }
    else { while (*mine != (unsigned)i) {} *theirs = i; }
  }
  if (cycles) *cycles = clock64() - t0;
}

template <int KIND, int U>
float run_copy(int dev, const uint4* src, uint4* dst, size_t n, int blocks, int threads, cudaStream_t s, int reps,
               int dev2 = -1, const uint4* src2 = nullptr, uint4* dst2 = nullptr, cudaStream_t s2 = 0) {
  float best = 1e30f;
  for (int r = 0; r < reps + 1; ++r) {
    cudaEvent_t a, b, a2, b2;
    CK(cudaSetDevice(dev)); CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    if (dev2 >= 0) { CK(cudaSetDevice(dev2)); CK(cudaEventCreate(&a2)); CK(cudaEventCreate(&b2)); }
    CK(cudaSetDevice(dev)); CK(cudaDeviceSynchronize());
    if (dev2 >= 0) { CK(cudaSetDevice(dev2)); CK(cudaDeviceSynchronize()); }
    CK(cudaSetDevice(dev)); CK(cudaEventRecord(a, s));
    copy_kernel<KIND, U><<<blocks, threads, 0, s>>>(src, dst, n);
    CK(cudaEventRecord(b, s));
    if (dev2 >= 0) {
      CK(cudaSetDevice(dev2)); CK(cudaEventRecord(a2, s2));
      copy_kernel<KIND, U><<<blocks, threads, 0, s2>>>(src2, dst2, n);
      CK(cudaEventRecord(b2, s2));
    }
    CK(cudaSetDevice(dev)); CK(cudaDeviceSynchronize());
    float ms, ms2 = 0; CK(cudaEventElapsedTime(&ms, a, b));
    if (dev2 >= 0) { CK(cudaSetDevice(dev2)); CK(cudaDeviceSynchronize()); CK(cudaEventElapsedTime(&ms2, a2, b2)); }
    if (r > 0) best = std::min(best, std::max(ms, ms2));
    cudaEventDestroy(a); cudaEventDestroy(b);
    if (dev2 >= 0) { cudaEventDestroy(a2); cudaEventDestroy(b2); }
  }
  return best;
}

int main(int argc, char** argv) {
  int ndev = 0; CK(cudaGetDeviceCount(&ndev));
  if (ndev < 2) { printf("need 2 GPUs\n"); return 0; }
  const size_t bytes = (argc > 1 ? atol(argv[1]) : 64) << 20;
  const size_t n = bytes / 16;
  uint4 *src0, *dst0, *src1, *dst1, *loc0;
  CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&src0, bytes)); CK(cudaMalloc(&dst0, bytes)); CK(cudaMalloc(&loc0, bytes));
  CK(cudaMemset(src0, 1, bytes));
  CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&src1, bytes)); CK(cudaMalloc(&dst1, bytes));
  CK(cudaMemset(src1, 2, bytes));
  cudaStream_t s0, s1;
  CK(cudaSetDevice(0)); CK(cudaStreamCreate(&s0));
  CK(cudaSetDevice(1)); CK(cudaStreamCreate(&s1));
  const double gb = bytes / 1e9;
  printf("message %zu MiB\n", bytes >> 20);

  // DMA engine reference
  {
    CK(cudaSetDevice(0));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
      CK(cudaEventRecord(a, s0)); CK(cudaMemcpyPeerAsync(dst1, 1, src0, 0, bytes, s0)); CK(cudaEventRecord(b, s0));
      CK(cudaStreamSynchronize(s0)); float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (r) best = std::min(best, ms);
    }
    printf("cudaMemcpyPeerAsync 0->1           : %7.1f GB/s\n", gb / (best * 1e-3));
  }
  const int blocks_list[] = {16, 32, 64, 148, 296};
#define ROW(KIND, U, NAME)                                                                           \
  for (int t : {512, 1024})                                                                          \
    for (int b : blocks_list) {                                                                      \
      float push = run_copy<KIND, U>(0, src0, dst1, n, b, t, s0, 3);                                 \
      printf("push %-9s U=%d threads=%4d blocks=%3d : %7.1f GB/s\n", NAME, U, t, b, gb / (push * 1e-3)); \
    }
  ROW(ST_WEAK, 1, "weak") ROW(ST_WEAK, 4, "weak") ROW(ST_WEAK, 8, "weak")
  ROW(ST_VOLATILE, 1, "volatile") ROW(ST_VOLATILE, 4, "volatile") ROW(ST_VOLATILE, 8, "volatile")
  ROW(ST_CG, 4, "cg") ROW(ST_CS, 4, "cs")
  // pull: GPU0 reads GPU1's memory, writes locally
  for (int b : blocks_list) {
    float pull = run_copy<ST_WEAK, 8>(0, src1, loc0, n, b, 512, s0, 3);
    printf("pull U=8 threads= 512 blocks=%3d           : %7.1f GB/s\n", b, gb / (pull * 1e-3));
  }
  // local copy for scale
  { float loc = run_copy<ST_WEAK, 4>(0, src0, loc0, n, 296, 512, s0, 3); printf("local copy (r+w counted once)         : %7.1f GB/s\n", gb / (loc * 1e-3)); }
  // bidirectional push
  for (int b : {64, 148, 296}) {
    float bi = run_copy<ST_WEAK, 4>(0, src0, dst1, n, b, 512, s0, 3, 1, src1, dst0, s1);
    printf("bidirectional push weak U=4 blocks=%3d  : %7.1f GB/s per direction\n", b, gb / (bi * 1e-3));
    float bv = run_copy<ST_VOLATILE, 4>(0, src0, dst1, n, b, 512, s0, 3, 1, src1, dst0, s1);
    printf("bidirectional push volatile U=4 b=%3d   : %7.1f GB/s per direction\n", b, gb / (bv * 1e-3));
  }
  // ping-pong latency
  {
    unsigned *f0, *f1; long long* cyc;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&f0, 256)); CK(cudaMemset(f0, 0, 256)); CK(cudaMalloc(&cyc, 8));
    CK(cudaSetDevice(1)); CK(cudaMalloc(&f1, 256)); CK(cudaMemset(f1, 0, 256));
    const int iters = 2000;
    CK(cudaSetDevice(1)); ping_kernel<<<1, 1, 0, s1>>>(f1, f0, iters, 0, nullptr);
    CK(cudaSetDevice(0));
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaEventRecord(a, s0)); ping_kernel<<<1, 1, 0, s0>>>(f0, f1, iters, 1, cyc); CK(cudaEventRecord(b, s0));
    CK(cudaStreamSynchronize(s0)); CK(cudaSetDevice(1)); CK(cudaStreamSynchronize(s1));
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    printf("ping-pong round trip: %.2f us  (one-way push + remote poll ~ %.2f us)\n", ms * 1e3 / iters, ms * 1e3 / iters / 2);
  }
  return 0;
}
