// Narrow numeric key columns -> the 64-bit key types the sort kernels take.  The widened column
// orders (and ties) exactly like the original, so it stands in for it in ds_sort_indices_*,
// ds_searchsorted_*, ds_partition_hist_scatter_* and ds_group_ids_fixed; the original column is
// what gets gathered into the output.
#include <cuda_fp16.h>

#include "ds_common.cuh"

namespace ds {

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) widen_kernel(const TIn* __restrict__ in, uint64_t n, TOut* __restrict__ out) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = static_cast<TOut>(ld_stream(in + i));
}

__global__ void __launch_bounds__(256) widen_half_kernel(const __half* __restrict__ in, uint64_t n, double* __restrict__ out) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = static_cast<double>(__half2float(in[i]));
}

// uint64 -> int64 with the same order: flip the top bit.
__global__ void __launch_bounds__(256) bias_u64_kernel(const uint64_t* __restrict__ in, uint64_t n, int64_t* __restrict__ out) {
  const uint64_t stride = static_cast<uint64_t>(gridDim.x) * blockDim.x;
  for (uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = static_cast<int64_t>(ld_stream(in + i) ^ 0x8000000000000000ull);
}

template <typename TIn, typename TOut>
static ds_status widen(const void* in, uint64_t n, void* out, cudaStream_t s) {
  const uint64_t want = (n + 255) / 256;
  const uint64_t cap = static_cast<uint64_t>(sm_count()) * 16;
  note_launch();
  widen_kernel<TIn, TOut><<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, s>>>(
      static_cast<const TIn*>(in), n, static_cast<TOut*>(out));
  DS_LAUNCH_CHECK();
  return DS_OK;
}

}  // namespace ds

extern "C" ds_status ds_widen_keys(const void* col, int32_t src_type, uint64_t n, void* out64, void* stream) {
  using namespace ds;
  if (n == 0) return DS_OK;
  DS_CHECK_ARG(col != nullptr && out64 != nullptr, "NULL pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (src_type) {
    case DS_KEY_INT8: return widen<int8_t, int64_t>(col, n, out64, s);
    case DS_KEY_INT16: return widen<int16_t, int64_t>(col, n, out64, s);
    case DS_KEY_INT32: return widen<int32_t, int64_t>(col, n, out64, s);
    case DS_KEY_UINT8: return widen<uint8_t, int64_t>(col, n, out64, s);
    case DS_KEY_UINT16: return widen<uint16_t, int64_t>(col, n, out64, s);
    case DS_KEY_UINT32: return widen<uint32_t, int64_t>(col, n, out64, s);
    case DS_KEY_FLOAT32: return widen<float, double>(col, n, out64, s);
    case DS_KEY_UINT64: {
      const uint64_t want = (n + 255) / 256;
      const uint64_t cap = static_cast<uint64_t>(sm_count()) * 16;
      note_launch();
      bias_u64_kernel<<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, s>>>(
          static_cast<const uint64_t*>(col), n, static_cast<int64_t*>(out64));
      DS_LAUNCH_CHECK();
      return DS_OK;
    }
    case DS_KEY_FLOAT16: {
      const uint64_t want = (n + 255) / 256;
      const uint64_t cap = static_cast<uint64_t>(sm_count()) * 16;
      note_launch();
      widen_half_kernel<<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, s>>>(
          static_cast<const __half*>(col), n, static_cast<double*>(out64));
      DS_LAUNCH_CHECK();
      return DS_OK;
    }
    default:
      set_error("%s: unknown source type %d", __func__, src_type);
      return DS_ERR_INVALID_ARG;
  }
}
