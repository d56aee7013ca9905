// Micro-benchmark 2: the histogram kernel's inner loop without global loads.  Each thread keeps 16 packed
// bins (a uint4, re-randomised with cheap ALU each iteration), computes addresses like hist_accumulate
// (PRMT + IMAD) and issues 32 RED.shared (g and h planes, conflict-free: bank = lane).
#include <cstdio>
#include <cuda_runtime.h>
#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ void red_shared(unsigned addr, int v) { asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ void red_shared_hi(unsigned addr, int v) { asm volatile("red.shared.add.s32 [%0+32768], %1;" ::"r"(addr), "r"(v) : "memory"); }

template <int RANDOMISE>
__global__ void k(int iters, int* sink, unsigned long long* cycles) {
  extern __shared__ int sh[];
  for (int i = threadIdx.x; i < 2 * 8192 * 2; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const unsigned sbase = (unsigned)__cvta_generic_to_shared(sh) + ((threadIdx.x >> 5) & 1) * 65536u;
  unsigned cs[16];
#pragma unroll
  for (int s = 0; s < 16; ++s) { unsigned v = sbase + (unsigned)((lane & 16) + ((s + lane) & 15)) * 4u; asm volatile("mov.u32 %0, %1;" : "=r"(cs[s]) : "r"(v)); }
  unsigned x = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 12345u;
  uint4 r = make_uint4(x, x * 3u + 1u, x * 5u + 7u, x * 7u + 11u);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (RANDOMISE) { r.x = r.x * 1664525u + 1013904223u; r.y ^= r.x >> 3; r.z += r.y; r.w ^= r.z << 1; }
    const unsigned rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int s = 0; s < 16; ++s) {
      const unsigned bin = __byte_perm(rr[s >> 2], 0u, 0x4440u + (unsigned)(s & 3));
      const unsigned cell = bin * 128u + cs[s];
      red_shared(cell, (int)(x & 1023));
      red_shared_hi(cell, 1);
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) atomicMax(cycles, (unsigned long long)(t1 - t0));
  int acc = 0;
  for (int i = threadIdx.x; i < 2 * 8192 * 2; i += blockDim.x) acc += sh[i];
  if (acc == 0x7fffffff) sink[0] = acc;
}

template <int RANDOMISE>
int run(const char* name, int threads) {
  cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0));
  int* sink; unsigned long long* cyc;
  CHECK(cudaMalloc(&sink, 4)); CHECK(cudaMalloc(&cyc, 8)); CHECK(cudaMemset(cyc, 0, 8));
  CHECK(cudaFuncSetAttribute(k<RANDOMISE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072));
  const int iters = 2000;
  k<RANDOMISE><<<p.multiProcessorCount, threads, 131072>>>(10, sink, cyc);
  CHECK(cudaMemset(cyc, 0, 8));
  k<RANDOMISE><<<p.multiProcessorCount, threads, 131072>>>(iters, sink, cyc);
  CHECK(cudaDeviceSynchronize());
  unsigned long long h; CHECK(cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost));
  printf("%-34s threads=%4d  %.2f lane-atomics/clk/SM (%llu clk)\n", name, threads, (double)iters * 32 * threads / (double)h, h);
  cudaFree(sink); cudaFree(cyc);
  return 0;
}

int main() {
  for (int t : {256, 512, 768, 1024}) { run<0>("hist inner loop, fixed bins", t); run<1>("hist inner loop, changing bins", t); }
  return 0;
}
