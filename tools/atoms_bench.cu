// Micro-benchmark: shared-memory int32 atomic throughput on sm_100a for the histogram kernel's access
// patterns.  Prints lane-atomics per SM per clock (clock64 deltas) for several patterns.
#include <cstdio>
#include <cuda_runtime.h>
#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

template <int MODE>
__global__ void k(int iters, int* sink, unsigned long long* cycles) {
  extern __shared__ int sh[];   // 2 planes of 256*32 ints
  for (int i = threadIdx.x; i < 2 * 8192; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  unsigned s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 12345u;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      s = s * 1664525u + 1013904223u;
      int bin = (s >> 24) & 255;
      int addr;
      if (MODE == 0) addr = bin * 32 + lane;                    // conflict-free: bank = lane
      else if (MODE == 1) addr = bin * 32 + ((s >> 8) & 31);    // random bank
      else if (MODE == 2) addr = (bin & 7) * 32 + lane;         // conflict-free, hot rows (few bins)
      else addr = bin * 32 + lane;
      atomicAdd(&sh[addr], (int)(s & 1023));
      if (MODE == 3) atomicAdd(&sh[8192 + addr], 1);            // g and h planes
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) atomicMax(cycles, (unsigned long long)(t1 - t0));
  int acc = 0;
  for (int i = threadIdx.x; i < 2 * 8192; i += blockDim.x) acc += sh[i];
  if (acc == 0x7fffffff) sink[0] = acc;
}

// pure ATOMS issue rate: addresses precomputed, nothing but atomics in the loop
template <int PLANES>
__global__ void k_pure(int iters, int* sink, unsigned long long* cycles) {
  extern __shared__ int sh[];
  for (int i = threadIdx.x; i < 2 * 8192; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  unsigned s = threadIdx.x * 2654435761u + blockIdx.x * 40503u + 12345u;
  int a[8];
#pragma unroll
  for (int u = 0; u < 8; ++u) { s = s * 1664525u + 1013904223u; a[u] = ((s >> 24) & 255) * 32 + lane; }
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      atomicAdd(&sh[a[u]], it);
      if (PLANES == 2) atomicAdd(&sh[8192 + a[u]], 1);
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) atomicMax(cycles, (unsigned long long)(t1 - t0));
  int acc = 0;
  for (int i = threadIdx.x; i < 2 * 8192; i += blockDim.x) acc += sh[i];
  if (acc == 0x7fffffff) sink[0] = acc;
}

template <int PLANES>
int run_pure(const char* name, int threads, int ctas_per_sm) {
  cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0)); int dev_sms = p.multiProcessorCount;
  int* sink; unsigned long long* cyc;
  CHECK(cudaMalloc(&sink, 4)); CHECK(cudaMalloc(&cyc, 8)); CHECK(cudaMemset(cyc, 0, 8));
  CHECK(cudaFuncSetAttribute(k_pure<PLANES>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  const int iters = 2000;
  k_pure<PLANES><<<dev_sms * ctas_per_sm, threads, 65536>>>(10, sink, cyc);
  CHECK(cudaMemset(cyc, 0, 8));
  k_pure<PLANES><<<dev_sms * ctas_per_sm, threads, 65536>>>(iters, sink, cyc);
  CHECK(cudaDeviceSynchronize());
  unsigned long long h; CHECK(cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost));
  double lane_atoms_per_sm = (double)iters * 8 * PLANES * threads * ctas_per_sm;
  printf("%-28s threads=%4d ctas/sm=%d  %.2f lane-atomics/clk/SM (%llu clk)\n", name, threads, ctas_per_sm, lane_atoms_per_sm / (double)h, h);
  cudaFree(sink); cudaFree(cyc);
  return 0;
}

template <int MODE>
int run(const char* name, int threads, int ctas_per_sm, int atoms_per_iter) {
  int dev_sms = 148;
  cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0)); dev_sms = p.multiProcessorCount;
  int* sink; unsigned long long* cyc;
  CHECK(cudaMalloc(&sink, 4)); CHECK(cudaMalloc(&cyc, 8)); CHECK(cudaMemset(cyc, 0, 8));
  CHECK(cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  const int iters = 2000;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<dev_sms * ctas_per_sm, threads, 65536>>>(10, sink, cyc);
  CHECK(cudaMemset(cyc, 0, 8));
  cudaEventRecord(a);
  k<MODE><<<dev_sms * ctas_per_sm, threads, 65536>>>(iters, sink, cyc);
  cudaEventRecord(b);
  CHECK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b);
  unsigned long long h; CHECK(cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost));
  double lane_atoms_per_sm = (double)iters * 8 * atoms_per_iter * threads * ctas_per_sm;
  printf("%-28s threads=%4d ctas/sm=%d  %.2f lane-atomics/clk/SM  (%.1f G lane-atomics/s chip, %.3f ms, %llu clk)\n", name, threads,
         ctas_per_sm, lane_atoms_per_sm / (double)h, lane_atoms_per_sm * dev_sms / (ms * 1e6), ms, h);
  cudaFree(sink); cudaFree(cyc);
  return 0;
}

int main() {
  for (int t : {256, 1024}) for (int c : {1, 2}) { run_pure<1>("pure ATOMS 1 plane", t, c); run_pure<2>("pure ATOMS 2 planes", t, c); }
  for (int t : {256, 512, 1024}) {
    for (int c : {1, 2, 3}) {
      if (t * c > 2048) continue;
      run<0>("conflict-free bank=lane", t, c, 1);
      run<1>("random bank", t, c, 1);
      run<2>("conflict-free, 8 hot bins", t, c, 1);
      run<3>("conflict-free g+h planes", t, c, 2);
    }
  }
  return 0;
}
