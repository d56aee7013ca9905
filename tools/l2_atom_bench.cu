// Micro-benchmark: how fast can the histogram flush reach a node's global histogram?
// Pattern of k_hist's flush: every CTA adds a 32768-cell partial histogram (two planes) into one of
// `nodes` global histograms of int64 pairs; `sharers` CTAs target the same node.  Variants:
//   0  red.global.add.u64 per plane (what kFlushAtomic does)
//   1  red.global.add.u32 per plane (int32 cells)
//   2  plain 16-byte stores of int64 pairs into a private slot (no sharing: the bandwidth reference)
//   3  plain read-add-write of 16-byte cells in a private slot (kFlushAddWide)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int kCells = 32768;

template <int MODE>
__global__ void __launch_bounds__(512) k(long long* hist, int sharers, int reps) {
  const int node = blockIdx.x / sharers;
  for (int r = 0; r < reps; ++r) {
    for (int i = threadIdx.x; i < kCells; i += blockDim.x) {
      const long long g = (long long)(i ^ r) + 1, h = 1;
      if (MODE == 0) {
        long long* c = hist + ((size_t)node * kCells + i) * 2;
        atomicAdd(reinterpret_cast<unsigned long long*>(c), (unsigned long long)g);
        atomicAdd(reinterpret_cast<unsigned long long*>(c + 1), (unsigned long long)h);
      } else if (MODE == 1) {
        int* c = reinterpret_cast<int*>(hist) + ((size_t)node * kCells + i) * 2;
        atomicAdd(c, (int)g);
        atomicAdd(c + 1, (int)h);
      } else if (MODE == 2) {
        reinterpret_cast<longlong2*>(hist)[(size_t)blockIdx.x * kCells + i] = make_longlong2(g, h);
      } else {
        longlong2* c = reinterpret_cast<longlong2*>(hist) + (size_t)blockIdx.x * kCells + i;
        longlong2 v = *c; v.x += g; v.y += h; *c = v;
      }
    }
  }
}

template <int MODE>
int run(const char* name, int ctas, int sharers, int reps) {
  const int slots = (MODE >= 2) ? ctas : (ctas + sharers - 1) / sharers;
  long long* hist;
  CHECK(cudaMalloc(&hist, (size_t)slots * kCells * 16));
  CHECK(cudaMemset(hist, 0, (size_t)slots * kCells * 16));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<ctas, 512>>>(hist, sharers, 1);
  cudaEventRecord(a);
  k<MODE><<<ctas, 512>>>(hist, sharers, reps);
  cudaEventRecord(b);
  CHECK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double cells = (double)ctas * kCells * reps;
  printf("%-36s ctas=%5d sharers=%3d footprint=%7.1f MB  %.3f ms  %.1f G cells/s (%.1f G adds/s)  %.2f us per CTA flush per SM\n",
         name, ctas, sharers, slots * kCells * 16.0 / 1e6, ms, cells / (ms * 1e6), 2 * cells / (ms * 1e6),
         ms * 1e3 / ((double)ctas * reps / 148.0));
  cudaFree(hist);
  return 0;
}

int main() {
  for (int sh : {1, 4, 49}) {
    run<0>("red.global.add.u64 x2", 7840, sh, 2);
    run<1>("red.global.add.u32 x2", 7840, sh, 2);
  }
  run<0>("red.global.add.u64 x2 (few nodes)", 1480, 49, 4);
  run<2>("plain 16-byte stores", 1480, 1, 4);
  run<3>("plain 16-byte read-add-write", 1480, 1, 4);
  run<2>("plain 16-byte stores (big)", 7840, 1, 2);
  return 0;
}
