// Row gather by permutation: dst[j] = src[idx[j]] for rows of row_bytes bytes.
//
// Replaces Table.take / take_table of the reference
// (python/ray/data/_internal/arrow_ops/transform_pyarrow.py:16,46-67).
//
// Rows are moved in "granules" of g = 16/8/4/2/1 bytes, the largest power of two that
// divides row_bytes and the two base addresses.  A block owns a run of consecutive
// OUTPUT rows, so its writes are one contiguous, fully coalesced span; consecutive
// threads read consecutive granules of the same source row.  Algorithmic bytes per
// row: 4 (index) + 2*row_bytes.
#include "ds_common.cuh"

namespace ds {

template <typename G>
__global__ void __launch_bounds__(256) gather_single_kernel(const G* __restrict__ src,
                                                            const uint32_t* __restrict__ idx,
                                                            G* __restrict__ dst, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) st_stream(dst + i, __ldg(src + ld_stream(idx + i)));
}

constexpr int GATHER_THREADS = 256;
constexpr int GATHER_MAX_ROWS = 1024;

template <typename G>
__global__ void __launch_bounds__(GATHER_THREADS)
    gather_rows_kernel(const G* __restrict__ src, const uint32_t* __restrict__ idx,
                       G* __restrict__ dst, uint32_t n, uint32_t gpr, uint32_t rows_per_block,
                       uint32_t magic) {
  __shared__ uint32_t s_idx[GATHER_MAX_ROWS];
  const uint32_t row0 = blockIdx.x * rows_per_block;
  const uint32_t rows = min(rows_per_block, n - row0);
  for (uint32_t r = threadIdx.x; r < rows; r += GATHER_THREADS) s_idx[r] = ld_stream(idx + row0 + r);
  __syncthreads();
  const uint32_t total = rows * gpr;
  G* out = dst + static_cast<uint64_t>(row0) * gpr;
#pragma unroll 4
  for (uint32_t e = threadIdx.x; e < total; e += GATHER_THREADS) {
    const uint32_t r = __umulhi(e, magic);  // e / gpr, exact for e*gpr < 2^32
    const uint32_t c = e - r * gpr;
    st_stream(out + e, __ldg(src + static_cast<uint64_t>(s_idx[r]) * gpr + c));
  }
}

template <typename G>
static cudaError_t launch_gather(const void* src, const uint32_t* idx, void* dst, uint32_t n,
                                 uint32_t row_bytes, cudaStream_t s) {
  const uint32_t gpr = row_bytes / sizeof(G);
  if (gpr == 1) {
    ds::note_launch(); gather_single_kernel<G><<<(n + 255) / 256, 256, 0, s>>>(static_cast<const G*>(src), idx,
                                                           static_cast<G*>(dst), n);
    return cudaGetLastError();
  }
  uint32_t rows = 8192 / gpr;
  if (rows < 1) rows = 1;
  if (rows > GATHER_MAX_ROWS) rows = GATHER_MAX_ROWS;
  const uint32_t magic = static_cast<uint32_t>((0x100000000ull + gpr - 1) / gpr);
  const uint32_t blocks = (n + rows - 1) / rows;
  ds::note_launch(); gather_rows_kernel<G><<<blocks, GATHER_THREADS, 0, s>>>(static_cast<const G*>(src), idx,
                                                         static_cast<G*>(dst), n, gpr, rows, magic);
  return cudaGetLastError();
}

ds_status gather_rows_impl(const void* src, uint32_t row_bytes, const uint32_t* idx, uint64_t n64,
                           void* dst, cudaStream_t s) {
  if (n64 == 0) return DS_OK;
  if (n64 > 0xFFFFFFFFull) {
    set_error("ds_gather_rows: n=%llu exceeds 2^32-1", (unsigned long long)n64);
    return DS_ERR_TOO_LARGE;
  }
  if (row_bytes == 0 || row_bytes > 65536) {
    set_error("ds_gather_rows: row_bytes=%u unsupported (1..65536)", row_bytes);
    return DS_ERR_UNSUPPORTED;
  }
  if (!src || !dst || !idx) {
    set_error("ds_gather_rows: NULL pointer");
    return DS_ERR_INVALID_ARG;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  const uintptr_t bits = reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst) | row_bytes;
  cudaError_t e;
  if ((bits & 15) == 0) e = launch_gather<uint4>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 7) == 0) e = launch_gather<uint2>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 3) == 0) e = launch_gather<uint32_t>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 1) == 0) e = launch_gather<uint16_t>(src, idx, dst, n, row_bytes, s);
  else e = launch_gather<uint8_t>(src, idx, dst, n, row_bytes, s);
  if (e != cudaSuccess) {
    set_error("ds_gather_rows: launch failed: %s", cudaGetErrorString(e));
    return DS_ERR_CUDA;
  }
  return DS_OK;
}

}  // namespace ds

extern "C" ds_status ds_gather_rows(const void* src, uint32_t row_bytes, const uint32_t* idx,
                                    uint64_t n, void* dst, void* stream) {
  return ds::gather_rows_impl(src, row_bytes, idx, n, dst, static_cast<cudaStream_t>(stream));
}
