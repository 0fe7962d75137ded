// Calibration only (not part of the product, not linked into libds_shuffle.so): how fast is
// CUB's onesweep DeviceRadixSort::SortPairs (the library baseline) on this GPU for the same
// shape as our 64-bit path: 100 M (uint64 key, uint32 value) pairs?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o cub_ref cub_ref.cu && ./cub_ref
#include <cub/device/device_radix_sort.cuh>
#include <cstdio>
#include <cstdlib>
#include <vector>

__global__ void fill(unsigned long long* k, unsigned* v, unsigned n, unsigned long long seed) {
  unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned long long x = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
  k[i] = x; v[i] = i;
}

int main(int argc, char** argv) {
  const unsigned n = argc > 1 ? atoi(argv[1]) : 100000000u;
  unsigned long long *k0, *k1; unsigned *v0, *v1;
  cudaMalloc(&k0, 8ull * n); cudaMalloc(&k1, 8ull * n); cudaMalloc(&v0, 4ull * n); cudaMalloc(&v1, 4ull * n);
  size_t ws_bytes = 0; void* ws = nullptr;
  cub::DoubleBuffer<unsigned long long> dk(k0, k1); cub::DoubleBuffer<unsigned> dv(v0, v1);
  for (int end_bit : {64, 32}) {
    cub::DeviceRadixSort::SortPairs(nullptr, ws_bytes, dk, dv, n, 0, end_bit);
    cudaFree(ws); cudaMalloc(&ws, ws_bytes);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    for (int it = 0; it < 5; ++it) {
      fill<<<(n + 255) / 256, 256>>>(dk.Current(), dv.Current(), n, it);
      cudaEventRecord(a);
      cub::DeviceRadixSort::SortPairs(ws, ws_bytes, dk, dv, n, 0, end_bit);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    int passes = end_bit / 8;
    printf("cub SortPairs u64+u32 n=%u bits=[0,%d): best %.3f ms total, %.3f ms per 8-bit pass (incl. histogram), "
           "%.1f GB/s per pass at 24 B/pair\n", n, end_bit, best, best / passes, 24.0 * n / (best / passes) / 1e6);
  }
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
