// Variable-length binary / string columns (Arrow layout: n+1 offsets + a byte buffer) on the
// sort / shuffle / groupby path, and group boundaries for keys of any supported type.
//
// A variable-length key is sorted as a sequence of stable 64-bit radix sorts, least
// significant part first (the C ABI entry points below only PRODUCE the 64-bit keys; the
// sorts themselves are ds_sort_indices_u64):
//     key part 0 (least significant): the row length            (ds_varlen_key_lengths)
//     key parts 1..C:  bytes [8(C-c), 8(C-c)+8) of the row, big-endian, zero padded
//                                                               (ds_varlen_key_chunk)
// Zero padding makes "ab" and "ab\0" equal in every chunk; the length part orders the
// shorter one first, which is memcmp-then-length order, i.e. Arrow's order for binary and
// (UTF-8 being order preserving) for string.  Constant digits are skipped by the radix
// sort, so short ASCII keys cost far fewer than 8 passes per chunk.
//
// Offsets are int64 on the device whatever the Arrow type's offset width, and may start
// anywhere in the byte buffer (a slice keeps its parent's bytes).
#include "ds_common.cuh"

namespace ds {

constexpr int VL_THREADS = 256;
constexpr int VL_ITEMS = 8;
constexpr int VL_TILE = VL_THREADS * VL_ITEMS;

// ------------------------------------------------------------------ 64-bit scan over a functor
// out[i] = sum_{j < i} f(j) (exclusive) or sum_{j <= i} f(j) (inclusive); the exclusive form
// also writes out[n] = total.  Three launches: tile sums, one-block scan of them, tile scans.
template <typename F>
__global__ void __launch_bounds__(VL_THREADS) vl_tile_sum_kernel(F f, uint32_t n, uint64_t* __restrict__ tile_sums) {
  __shared__ uint64_t s_warp[VL_THREADS / 32];
  const uint32_t base = blockIdx.x * VL_TILE + threadIdx.x * VL_ITEMS;
  uint64_t c = 0;
#pragma unroll
  for (int k = 0; k < VL_ITEMS; ++k)
    if (base + k < n) c += f(base + k);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane_id() == 0) s_warp[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint64_t t = 0;
    for (int w = 0; w < VL_THREADS / 32; ++w) t += s_warp[w];
    tile_sums[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(1024) vl_scan_tiles_kernel(uint64_t* __restrict__ tile_sums, uint32_t tiles) {
  __shared__ uint64_t s_warp[32];
  __shared__ uint64_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t c = i < tiles ? tile_sums[i] : 0;
    uint64_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint64_t t = __shfl_up_sync(0xffffffffu, v, o);
      if ((int)lane_id() >= o) v += t;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = v;
    __syncthreads();
    uint64_t add = s_carry;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) add += s_warp[w];
    if (i < tiles) tile_sums[i] = v + add - c;  // exclusive
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = v + add;
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[tiles] = s_carry;  // total
}

template <typename F, bool INCLUSIVE>
__global__ void __launch_bounds__(VL_THREADS)
    vl_tile_scan_kernel(F f, uint32_t n, const uint64_t* __restrict__ tile_sums, int64_t* __restrict__ out) {
  __shared__ uint64_t s_warp[VL_THREADS / 32];
  const uint32_t base = blockIdx.x * VL_TILE + threadIdx.x * VL_ITEMS;
  uint64_t v[VL_ITEMS];
  uint64_t c = 0;
#pragma unroll
  for (int k = 0; k < VL_ITEMS; ++k) {
    v[k] = base + k < n ? f(base + k) : 0;
    c += v[k];
  }
  uint64_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint64_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane_id() >= o) incl += t;
  }
  if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
  __syncthreads();
  uint64_t run = tile_sums[blockIdx.x] + incl - c;
  for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) run += s_warp[w];
#pragma unroll
  for (int k = 0; k < VL_ITEMS; ++k) {
    if (base + k < n) out[base + k] = static_cast<int64_t>(INCLUSIVE ? run + v[k] : run);
    run += v[k];
  }
  if (!INCLUSIVE && blockIdx.x == 0 && threadIdx.x == 0) out[n] = static_cast<int64_t>(tile_sums[gridDim.x]);
}

template <typename F>
static ds_status scan_u64(F f, uint32_t n, bool inclusive, uint64_t* tile_sums, int64_t* out,
                          cudaStream_t s) {
  const uint32_t tiles = (n + VL_TILE - 1) / VL_TILE;
  if (n == 0) {
    if (!inclusive) DS_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t), s));
    return DS_OK;
  }
  note_launch(); vl_tile_sum_kernel<F><<<tiles, VL_THREADS, 0, s>>>(f, n, tile_sums);
  DS_LAUNCH_CHECK();
  note_launch(); vl_scan_tiles_kernel<<<1, 1024, 0, s>>>(tile_sums, tiles);
  DS_LAUNCH_CHECK();
  note_launch();
  if (inclusive)
    vl_tile_scan_kernel<F, true><<<tiles, VL_THREADS, 0, s>>>(f, n, tile_sums, out);
  else
    vl_tile_scan_kernel<F, false><<<tiles, VL_THREADS, 0, s>>>(f, n, tile_sums, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

// ------------------------------------------------------------------ functors
struct TakenLength {  // length of source row idx[i] (or row i)
  const int64_t* offsets;
  const uint32_t* idx;
  __device__ uint64_t operator()(uint32_t i) const {
    const uint32_t r = idx ? idx[i] : i;
    return static_cast<uint64_t>(offsets[r + 1] - offsets[r]);
  }
};

__device__ __forceinline__ int vl_compare(const uint8_t* a, int64_t la, const uint8_t* b, int64_t lb) {
  const int64_t m = la < lb ? la : lb;
  for (int64_t k = 0; k < m; ++k) {
    const int d = static_cast<int>(a[k]) - static_cast<int>(b[k]);
    if (d) return d;
  }
  return la < lb ? -1 : (la > lb ? 1 : 0);
}

struct VarlenDiffers {  // 1 if row i starts a new run of equal keys (i > 0), else 0
  const int64_t* offsets;
  const uint8_t* data;
  __device__ uint64_t operator()(uint32_t i) const {
    if (i == 0) return 0;
    const int64_t a0 = offsets[i - 1], a1 = offsets[i], a2 = offsets[i + 1];
    if (a1 - a0 != a2 - a1) return 1;
    return vl_compare(data + a0, a1 - a0, data + a1, a2 - a1) != 0;
  }
};

struct FixedDiffers {  // the same for fixed-width rows compared bytewise
  const uint8_t* data;
  uint32_t width;
  __device__ uint64_t operator()(uint32_t i) const {
    if (i == 0) return 0;
    const uint8_t* a = data + static_cast<uint64_t>(i - 1) * width;
    const uint8_t* b = a + width;
    for (uint32_t k = 0; k < width; ++k)
      if (a[k] != b[k]) return 1;
    return 0;
  }
};

struct Word64Differs {  // 8-byte keys: one load per row
  const uint64_t* data;
  __device__ uint64_t operator()(uint32_t i) const { return i > 0 && data[i] != data[i - 1]; }
};

struct F64Differs {  // numeric inequality: -0.0 == 0.0, NaN differs from everything
  const double* data;
  __device__ uint64_t operator()(uint32_t i) const { return i > 0 && data[i] != data[i - 1]; }
};

// ------------------------------------------------------------------ sort-key parts
__global__ void __launch_bounds__(256)
    vl_key_lengths_kernel(const int64_t* __restrict__ offsets, const uint32_t* __restrict__ idx, uint32_t n,
                          uint64_t* __restrict__ out, unsigned long long* __restrict__ max_len) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t len = 0;
  if (i < n) {
    const uint32_t r = idx ? idx[i] : i;
    len = static_cast<uint64_t>(offsets[r + 1] - offsets[r]);
    out[i] = len;
  }
  uint64_t m = len;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const uint64_t t = __shfl_xor_sync(0xffffffffu, m, o);
    m = t > m ? t : m;
  }
  if (lane_id() == 0 && m > 0) atomicMax(max_len, static_cast<unsigned long long>(m));
}

// out[i] = bytes [byte0, byte0 + 8) of row idx[i], first byte most significant, zero padded.
__global__ void __launch_bounds__(256)
    vl_key_chunk_kernel(const int64_t* __restrict__ offsets, const uint8_t* __restrict__ data,
                        const uint32_t* __restrict__ idx, uint32_t n, uint64_t byte0,
                        uint64_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = idx ? idx[i] : i;
  const int64_t beg = offsets[r], end = offsets[r + 1];
  const uint8_t* p = data + beg + byte0;
  const int64_t avail = end - beg - static_cast<int64_t>(byte0);
  uint64_t k = 0;
#pragma unroll
  for (int b = 0; b < 8; ++b) k = (k << 8) | (b < avail ? static_cast<uint64_t>(p[b]) : 0ull);
  out[i] = k;
}

// ------------------------------------------------------------------ take (gather) of rows
// One warp per output row; lanes stride over the row's bytes.
__global__ void __launch_bounds__(256)
    vl_take_bytes_kernel(const int64_t* __restrict__ src_offsets, const uint8_t* __restrict__ src,
                         const uint32_t* __restrict__ idx, uint32_t n,
                         const int64_t* __restrict__ dst_offsets, uint8_t* __restrict__ dst) {
  const uint32_t warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; i < n; i += warps) {
    const uint32_t r = idx ? idx[i] : i;
    const int64_t beg = src_offsets[r];
    const int64_t len = src_offsets[r + 1] - beg;
    const uint8_t* s = src + beg;
    uint8_t* d = dst + dst_offsets[i];
    for (int64_t k = lane_id(); k < len; k += 32) d[k] = s[k];
  }
}

// ------------------------------------------------------------------ searchsorted
// out[j] = number of rows r of the sorted column with row < bound_j (ascending) or
// row > bound_j (descending): numpy.searchsorted(side="left") on the key order.
__global__ void vl_searchsorted_kernel(const int64_t* __restrict__ offsets, const uint8_t* __restrict__ data,
                                       uint32_t n, const int64_t* __restrict__ b_offsets,
                                       const uint8_t* __restrict__ b_data, uint32_t nb, bool desc,
                                       int64_t* __restrict__ out) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb) return;
  const uint8_t* b = b_data + b_offsets[j];
  const int64_t lb = b_offsets[j + 1] - b_offsets[j];
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    const int c = vl_compare(data + offsets[mid], offsets[mid + 1] - offsets[mid], b, lb);
    if (desc ? c > 0 : c < 0) lo = mid + 1; else hi = mid;
  }
  out[j] = lo;
}

}  // namespace ds

// ====================================================================== C ABI
extern "C" {

ds_status ds_varlen_take_offsets(const int64_t* src_offsets, uint64_t src_rows, const uint32_t* idx,
                                 uint64_t n64, int64_t* dst_offsets, void* ws, size_t* ws_bytes,
                                 void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS || src_rows > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  Workspace w(ws);
  uint64_t* tile_sums = w.take<uint64_t>((n + VL_TILE - 1) / VL_TILE + 1);
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, __func__, &st)) return st;
  DS_CHECK_ARG(dst_offsets != nullptr && (n == 0 || src_offsets != nullptr), "NULL pointer");
  return scan_u64(TakenLength{src_offsets, idx}, n, false, tile_sums, dst_offsets,
                  static_cast<cudaStream_t>(stream));
}

ds_status ds_varlen_take_bytes(const int64_t* src_offsets, const uint8_t* src_data, const uint32_t* idx,
                               uint64_t n64, const int64_t* dst_offsets, uint8_t* dst_data, void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  if (n64 == 0) return DS_OK;
  DS_CHECK_ARG(src_offsets && dst_offsets, "NULL pointer");
  const uint32_t n = static_cast<uint32_t>(n64);
  const uint64_t want = (static_cast<uint64_t>(n) * 32 + 255) / 256;
  const uint64_t cap = static_cast<uint64_t>(sm_count()) * 32;
  note_launch();
  vl_take_bytes_kernel<<<static_cast<unsigned>(want < cap ? want : cap), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src_offsets, src_data, idx, n, dst_offsets, dst_data);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_varlen_key_lengths(const int64_t* offsets, const uint32_t* idx, uint64_t n64,
                                uint64_t* lengths_out, uint64_t* max_len_out, void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  DS_CHECK_ARG(max_len_out != nullptr && (n64 == 0 || (offsets && lengths_out)), "NULL pointer");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DS_CUDA(cudaMemsetAsync(max_len_out, 0, sizeof(uint64_t), s));
  if (n64 == 0) return DS_OK;
  const uint32_t n = static_cast<uint32_t>(n64);
  note_launch();
  vl_key_lengths_kernel<<<(n + 255) / 256, 256, 0, s>>>(offsets, idx, n, lengths_out,
                                                         reinterpret_cast<unsigned long long*>(max_len_out));
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_varlen_key_chunk(const int64_t* offsets, const uint8_t* data, const uint32_t* idx,
                              uint64_t n64, uint64_t byte_offset, uint64_t* keys_out, void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  if (n64 == 0) return DS_OK;
  DS_CHECK_ARG(offsets && keys_out, "NULL pointer");
  const uint32_t n = static_cast<uint32_t>(n64);
  note_launch();
  vl_key_chunk_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(offsets, data, idx, n,
                                                                                     byte_offset, keys_out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_searchsorted_varlen(const int64_t* sorted_offsets, const uint8_t* sorted_data, uint64_t n64,
                                 const int64_t* bound_offsets, const uint8_t* bound_data, uint32_t nb,
                                 int32_t descending, int64_t* out, void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  if (nb == 0) return DS_OK;
  DS_CHECK_ARG(bound_offsets && out && (n64 == 0 || sorted_offsets), "NULL pointer");
  note_launch();
  vl_searchsorted_kernel<<<(nb + 63) / 64, 64, 0, static_cast<cudaStream_t>(stream)>>>(
      sorted_offsets, sorted_data, static_cast<uint32_t>(n64), bound_offsets, bound_data, nb,
      descending != 0, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

// group_ids_out[i] = number of key changes in rows (0, i]: equal for the rows of one run of
// equal keys, increasing by one from run to run -- an int64 key column with the same runs,
// which ds_segment_heads_i64 and the segmented reductions then work on.
static ds_status group_ids_common(int kind, const void* data, const int64_t* offsets, uint32_t width,
                                  uint64_t n64, int64_t* out, void* ws, size_t* ws_bytes, void* stream,
                                  const char* fn) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: row count exceeds DS_MAX_ROWS", fn);
    return DS_ERR_TOO_LARGE;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  Workspace w(ws);
  uint64_t* tile_sums = w.take<uint64_t>((n + VL_TILE - 1) / VL_TILE + 1);
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, fn, &st)) return st;
  if (n == 0) return DS_OK;
  if (!out || !data || (kind == 0 && !offsets)) {
    set_error("%s: NULL pointer", fn);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (kind) {
    case 0: return scan_u64(VarlenDiffers{offsets, static_cast<const uint8_t*>(data)}, n, true, tile_sums, out, s);
    case 1: return scan_u64(F64Differs{static_cast<const double*>(data)}, n, true, tile_sums, out, s);
    case 2: return scan_u64(Word64Differs{static_cast<const uint64_t*>(data)}, n, true, tile_sums, out, s);
    default: return scan_u64(FixedDiffers{static_cast<const uint8_t*>(data), width}, n, true, tile_sums, out, s);
  }
}

ds_status ds_group_ids_varlen(const int64_t* sorted_offsets, const uint8_t* sorted_data, uint64_t n,
                              int64_t* group_ids_out, void* ws, size_t* ws_bytes, void* stream) {
  // an all-empty column has no bytes: any non-NULL pointer stands in for the data
  const void* d = sorted_data ? static_cast<const void*>(sorted_data) : static_cast<const void*>(sorted_offsets);
  return group_ids_common(0, d, sorted_offsets, 0, n, group_ids_out, ws, ws_bytes, stream, __func__);
}

ds_status ds_group_ids_fixed(const void* sorted_col, uint32_t width, int32_t is_float64, uint64_t n,
                             int64_t* group_ids_out, void* ws, size_t* ws_bytes, void* stream) {
  if (width == 0 || (is_float64 && width != 8)) {
    ds::set_error("%s: bad width %u", __func__, width);
    return DS_ERR_INVALID_ARG;
  }
  const int kind = is_float64 ? 1 : (width == 8 && (reinterpret_cast<uintptr_t>(sorted_col) & 7) == 0 ? 2 : 3);
  return group_ids_common(kind, sorted_col, nullptr, width, n, group_ids_out, ws, ws_bytes, stream, __func__);
}

}  // extern "C"
