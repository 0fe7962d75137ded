// Cost of launching a kernel whose CTAs all return at once (the "launch every plan's kernels and
// let the unused ones exit" form of a device-side plan decision).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o empty_launch empty_launch.cu && ./empty_launch
#include <cstdio>
#include <cuda_runtime.h>

__global__ void early_exit_kernel(const int* flag, int* out) {
  extern __shared__ int sm[];
  if (*flag == 0) return;
  sm[threadIdx.x] = threadIdx.x;  // never reached
  __syncthreads();
  out[blockIdx.x] = sm[0];
}

static float time_launch(int grid, int block, int smem, const int* flag, int* out, int reps) {
  cudaFuncSetAttribute(early_exit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) early_exit_kernel<<<grid, block, smem>>>(flag, out);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) early_exit_kernel<<<grid, block, smem>>>(flag, out);
  cudaEventRecord(b);
  cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  return ms * 1000.0f / reps;
}

int main() {
  int *flag, *out;
  cudaMalloc(&flag, 4);
  cudaMemset(flag, 0, 4);
  cudaMalloc(&out, 4 << 20);
  struct { const char* what; int grid, block, smem; } cases[] = {
      {"pass kernel shape (100 M rows / 4608 per tile, 384 threads, 60 KB)", 21702, 384, 60 * 1024},
      {"bucket kernel shape (65536 CTAs, 256 threads, 52 KB)", 65536, 256, 52 * 1024},
      {"decode kernel shape (97657 CTAs, 256 threads, no dynamic smem)", 97657, 256, 0},
      {"persistent shape (148 x 4 CTAs, 256 threads)", 592, 256, 0},
  };
  for (auto& c : cases)
    printf("%-72s %8.1f us per launch (back to back, %d launches)\n", c.what,
           time_launch(c.grid, c.block, c.smem, flag, out, 50), 50);
  cudaError_t e = cudaGetLastError();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
