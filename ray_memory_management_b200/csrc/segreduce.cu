// Run detection and segmented reduction for groupby().aggregate().
//
// Replaces, for a key-sorted block, the reference's per-row Python loops:
//   ArrowBlockAccessor.combine            python/ray/data/_internal/arrow_block.py:396-467
//   aggregate_combined_blocks             python/ray/data/_internal/arrow_block.py:488-587
// with Sum/Min/Max/Mean/Count accumulate_block + merge of python/ray/data/aggregate.py:97-211.
//
// ds_segment_heads_i64: three small kernels over tiles of 2048 keys
//   count heads per tile -> scan tile counts (one block) -> write run starts.
// ds_segmented_reduce_*: tiles of 2048 sorted positions; each tile reduces the part of
//   every run that falls inside it (warp per run, or thread per run when runs are
//   short).  Runs inside one tile are written directly; a run that crosses tile
//   borders leaves one partial per tile, and a fix-up kernel folds those left to
//   right, so the result does not depend on scheduling (no floating-point atomics).
//   The value column is read THROUGH the sort permutation (vals[idx[i]]), i.e. the
//   gather is fused.  Algorithmic bytes per row: 4 (idx) + 8 (value).
#include <type_traits>

#include <atomic>

#include "ds_common.cuh"

namespace ds {

constexpr int SEG_THREADS = 256;
constexpr int SEG_TILE = 2048;

// ------------------------------------------------------------------ heads
__global__ void __launch_bounds__(SEG_THREADS)
    head_count_kernel(const int64_t* __restrict__ keys, uint32_t n, uint32_t* __restrict__ tile_counts) {
  __shared__ uint32_t s_warp[SEG_THREADS / 32];
  const uint32_t base = blockIdx.x * SEG_TILE;
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < SEG_TILE / SEG_THREADS; ++k) {
    const uint32_t i = base + k * SEG_THREADS + threadIdx.x;
    if (i < n) c += (i == 0) || (keys[i] != keys[i - 1]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (lane_id() == 0) s_warp[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < SEG_THREADS / 32; ++w) t += s_warp[w];
    tile_counts[blockIdx.x] = t;
  }
}

// Single block: exclusive scan of tile_counts in place; total -> nseg_out.
__global__ void __launch_bounds__(1024)
    tile_scan_kernel(uint32_t* __restrict__ tile_counts, uint32_t tiles, uint32_t* __restrict__ nseg_out,
                     uint32_t* __restrict__ seg_starts, uint32_t n) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < tiles; base += 1024) {
    const uint32_t i = base + threadIdx.x;
    const uint32_t c = i < tiles ? tile_counts[i] : 0;
    uint32_t v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
      if ((int)lane_id() >= o) v += t;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t add = s_carry;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) add += s_warp[w];
    if (i < tiles) tile_counts[i] = v + add - c;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = v + add;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    *nseg_out = s_carry;
    seg_starts[s_carry] = n;  // sentinel
  }
}

__global__ void __launch_bounds__(SEG_THREADS)
    head_write_kernel(const int64_t* __restrict__ keys, uint32_t n,
                      const uint32_t* __restrict__ tile_offsets, uint32_t* __restrict__ seg_starts) {
  __shared__ uint32_t s_warp[SEG_THREADS / 32];
  __shared__ uint32_t s_run;
  if (threadIdx.x == 0) s_run = tile_offsets[blockIdx.x];
  __syncthreads();
  const uint32_t base = blockIdx.x * SEG_TILE;
  for (int k = 0; k < SEG_TILE / SEG_THREADS; ++k) {
    const uint32_t i = base + k * SEG_THREADS + threadIdx.x;
    const bool head = i < n && ((i == 0) || (keys[i] != keys[i - 1]));
    const uint32_t ball = __ballot_sync(0xffffffffu, head);
    if (lane_id() == 0) s_warp[threadIdx.x >> 5] = __popc(ball);
    __syncthreads();
    uint32_t before = s_run;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) before += s_warp[w];
    if (head) seg_starts[before + __popc(ball & lanemask_lt())] = i;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t t = 0;
      for (int w = 0; w < SEG_THREADS / 32; ++w) t += s_warp[w];
      s_run += t;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ reduce
template <typename T, int OP>
__device__ __forceinline__ T red_identity() {
  if constexpr (OP == DS_OP_SUM || OP == 100 || OP == 101 /* internal sums */ || OP == DS_OP_ABSMAX) {
    return T(0);
  } else if constexpr (std::is_floating_point<T>::value) {
    return OP == DS_OP_MIN ? __longlong_as_double(0x7FF0000000000000ll)
                           : __longlong_as_double(0xFFF0000000000000ll);
  } else {
    return OP == DS_OP_MIN ? T(0x7FFFFFFFFFFFFFFFll) : T(-0x7FFFFFFFFFFFFFFFll - 1);
  }
}
template <typename T, int OP>
__device__ __forceinline__ T red_op(T a, T b) {
  if (OP == DS_OP_SUM || OP == 100 || OP == 101 /* internal sums */) return a + b;
  if (OP == DS_OP_MIN) return b < a ? b : a;
  return b > a ? b : a;
}
template <typename T>
__device__ __forceinline__ T shfl_xor_t(T v, int o) {
  long long x = __shfl_xor_sync(0xffffffffu, *reinterpret_cast<long long*>(&v), o);
  return *reinterpret_cast<T*>(&x);
}

constexpr int OP_COUNT_VALID = 100;  // internal: sum of (valid[j] != 0)
constexpr int OP_SUMSQDIFF = 101;    // internal: sum of addend[j] + w[j] * (x[j] - mean[s])^2

struct SqDiffArgs {
  const double* means;     // one per run
  const int64_t* weights;  // optional, one per element
  const double* addend;    // optional, one per element
};

// Element j of the reduction: the value, or the op's identity where the row is null.
template <typename TIn, typename T, int OP>
__device__ __forceinline__ T seg_elem(const TIn* __restrict__ vals, const uint8_t* __restrict__ valid,
                                      uint32_t j, uint32_t s, const SqDiffArgs& sq) {
  if constexpr (OP == OP_SUMSQDIFF) {
    if (valid && valid[j] == 0) return T(0);
    const double d = static_cast<double>(vals[j]) - sq.means[s];
    const double w = sq.weights ? static_cast<double>(sq.weights[j]) : 1.0;
    return (sq.addend ? sq.addend[j] : 0.0) + w * d * d;
  } else if constexpr (OP == OP_COUNT_VALID) {
    return static_cast<T>(valid ? (valid[j] != 0) : 1);
  } else {
    if (valid && valid[j] == 0) return red_identity<T, OP>();
    const T v = vals[j];
    if constexpr (OP == DS_OP_ABSMAX) return v < T(0) ? -v : v;
    return v;
  }
}

// last position p in [0, count) with starts[p] <= x   (starts[0] == 0 <= x always)
__device__ __forceinline__ uint32_t last_le(const uint32_t* __restrict__ starts, uint32_t count,
                                            uint32_t x) {
  uint32_t lo = 0, hi = count;  // invariant: starts[lo] <= x, answer in [lo, hi)
  while (hi - lo > 1) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if (starts[mid] <= x) lo = mid;
    else hi = mid;
  }
  return lo;
}

template <typename TIn, typename T, int OP>
__global__ void __launch_bounds__(SEG_THREADS)
    segreduce_tile_kernel(const TIn* __restrict__ vals, const uint8_t* __restrict__ valid,
                          const uint32_t* __restrict__ idx,
                          const uint32_t* __restrict__ seg_starts, uint32_t nseg, uint32_t n,
                          T* __restrict__ out, T* __restrict__ p_first, T* __restrict__ p_last,
                          uint32_t* __restrict__ t_first, uint32_t* __restrict__ t_last,
                          SqDiffArgs sq) {
  __shared__ uint32_t s_f, s_l;
  const uint32_t tile = blockIdx.x;
  const uint32_t t0 = tile * SEG_TILE;
  const uint32_t t1 = min(t0 + SEG_TILE, n);
  if (threadIdx.x == 0) {
    s_f = last_le(seg_starts, nseg, t0);
    t_first[tile] = s_f;
  }
  if (threadIdx.x == 32) {
    s_l = last_le(seg_starts, nseg, t1 - 1);
    t_last[tile] = s_l;
  }
  __syncthreads();
  const uint32_t f = s_f, l = s_l;
  const uint32_t nsegs_here = l - f + 1;
  const uint32_t warp = threadIdx.x >> 5, lane = lane_id();

  if (nsegs_here >= SEG_THREADS) {
    // short runs: one thread per run, sequential left-to-right
    for (uint32_t s = f + threadIdx.x; s <= l; s += SEG_THREADS) {
      const uint32_t a0 = seg_starts[s], b0 = seg_starts[s + 1];
      const uint32_t a = max(a0, t0), b = min(b0, t1);
      T acc = red_identity<T, OP>();
      for (uint32_t i = a; i < b; ++i) acc = red_op<T, OP>(acc, seg_elem<TIn, T, OP>(vals, valid, idx ? idx[i] : i, s, sq));
      if (a0 >= t0 && b0 <= t1) out[s] = acc;
      else if (s == f) p_first[tile] = acc;
      else p_last[tile] = acc;
    }
  } else {
    for (uint32_t s = f + warp; s <= l; s += SEG_THREADS / 32) {
      const uint32_t a0 = seg_starts[s], b0 = seg_starts[s + 1];
      const uint32_t a = max(a0, t0), b = min(b0, t1);
      T acc = red_identity<T, OP>();
      for (uint32_t i = a + lane; i < b; i += 32) acc = red_op<T, OP>(acc, seg_elem<TIn, T, OP>(vals, valid, idx ? idx[i] : i, s, sq));
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc = red_op<T, OP>(acc, shfl_xor_t(acc, o));
      if (lane == 0) {
        if (a0 >= t0 && b0 <= t1) out[s] = acc;
        else if (s == f) p_first[tile] = acc;
        else p_last[tile] = acc;
      }
    }
  }
}

// One thread per tile: if a run starts in this tile and runs past its end, fold the
// per-tile partials of that run left to right.
template <typename T, int OP>
__global__ void segreduce_fixup_kernel(const uint32_t* __restrict__ seg_starts, uint32_t n,
                                       uint32_t tiles, const T* __restrict__ p_first,
                                       const T* __restrict__ p_last,
                                       const uint32_t* __restrict__ t_first,
                                       const uint32_t* __restrict__ t_last, T* __restrict__ out) {
  const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= tiles) return;
  const uint32_t t0 = tile * SEG_TILE;
  const uint32_t t1 = min(t0 + SEG_TILE, n);
  const uint32_t s = t_last[tile];
  const uint32_t a0 = seg_starts[s], b0 = seg_starts[s + 1];
  if (a0 < t0 || b0 <= t1) return;  // not the tile where a border-crossing run starts
  T acc = (t_first[tile] == s) ? p_first[tile] : p_last[tile];
  for (uint32_t t = tile + 1; t < tiles; ++t) {
    acc = red_op<T, OP>(acc, p_first[t]);
    if (b0 <= min((t + 1) * (uint32_t)SEG_TILE, n)) break;
  }
  out[s] = acc;
}

template <typename TIn, typename T>
static ds_status segreduce_impl(const TIn* vals, const uint8_t* valid, const uint32_t* idx,
                                const uint32_t* seg_starts, uint32_t nseg, uint64_t n64, int32_t op,
                                T* out, void* ws, size_t* ws_bytes, void* stream, const char* fn,
                                SqDiffArgs sq = SqDiffArgs{nullptr, nullptr, nullptr}) {
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: n exceeds DS_MAX_ROWS", fn);
    return DS_ERR_TOO_LARGE;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  const uint32_t tiles = (n + SEG_TILE - 1) / SEG_TILE;
  Workspace w(ws);
  T* p_first = w.take<T>(tiles);
  T* p_last = w.take<T>(tiles);
  uint32_t* t_first = w.take<uint32_t>(tiles);
  uint32_t* t_last = w.take<uint32_t>(tiles);
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, fn, &st)) return st;
  if (n == 0 || nseg == 0) return DS_OK;
  if ((!vals && op != OP_COUNT_VALID) || !seg_starts || !out) {
    set_error("%s: NULL pointer", fn);
    return DS_ERR_INVALID_ARG;
  }
  if (!((op >= DS_OP_SUM && op <= DS_OP_ABSMAX) || op == OP_COUNT_VALID || op == OP_SUMSQDIFF)) {
    set_error("%s: unknown op %d", fn, op);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
#define DS_SEG(OP)                                                                               \
  ds::note_launch(); segreduce_tile_kernel<TIn, T, OP><<<tiles, SEG_THREADS, 0, s>>>(vals, valid, idx, seg_starts, nseg, n, out, \
                                                             p_first, p_last, t_first, t_last, sq);  \
  ds::note_launch(); segreduce_fixup_kernel<T, OP><<<(tiles + 127) / 128, 128, 0, s>>>(                             \
      seg_starts, n, tiles, p_first, p_last, t_first, t_last, out)
  if constexpr (std::is_same<TIn, T>::value) {
    if (op == DS_OP_SUM) { DS_SEG(DS_OP_SUM); }
    else if (op == DS_OP_MIN) { DS_SEG(DS_OP_MIN); }
    else if (op == DS_OP_MAX) { DS_SEG(DS_OP_MAX); }
    else if (op == DS_OP_ABSMAX) { DS_SEG(DS_OP_ABSMAX); }
    else if (op == OP_SUMSQDIFF) { DS_SEG(OP_SUMSQDIFF); }
    else { DS_SEG(OP_COUNT_VALID); }
  } else {
    DS_SEG(OP_SUMSQDIFF);  // int64 values, double accumulator
  }
#undef DS_SEG
  DS_LAUNCH_CHECK();
  return DS_OK;
}

__global__ void segment_counts_kernel(const uint32_t* __restrict__ seg_starts, uint32_t nseg,
                                      int64_t* __restrict__ out) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < nseg) out[s] = static_cast<int64_t>(seg_starts[s + 1] - seg_starts[s]);
}

__global__ void scale_kernel(const double* __restrict__ a, const int64_t* __restrict__ w, uint32_t n,
                             double* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] * static_cast<double>(w[i]);
}

// aggregate.py:270-276: 0.0 for fewer than two values, else sqrt(M2 / (count - ddof))
__global__ void finalize_std_kernel(const double* __restrict__ m2, const int64_t* __restrict__ count,
                                    uint32_t n, int ddof, double* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = count[i] < 2 ? 0.0 : sqrt(m2[i] / static_cast<double>(count[i] - ddof));
}

// Null policy of python/ray/data/_internal/null_aggregate.py:12-26 per group:
// ignore_nulls: a value unless every row was null; otherwise: a value only if no row was.
__global__ void finalize_validity_kernel(const int64_t* __restrict__ nvalid,
                                         const int64_t* __restrict__ nrows, uint32_t n,
                                         bool ignore_nulls, uint8_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = ignore_nulls ? (nvalid[i] > 0) : (nvalid[i] == nrows[i] && nrows[i] > 0);
}

__global__ void finalize_mean_kernel(const void* __restrict__ sum, bool sum_is_i64,
                                     const int64_t* __restrict__ count, uint32_t n,
                                     double* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double s = sum_is_i64 ? static_cast<double>(static_cast<const int64_t*>(sum)[i])
                              : static_cast<const double*>(sum)[i];
  out[i] = count[i] == 0 ? 0.0 : s / static_cast<double>(count[i]);  // all-null group: value is masked out
}

}  // namespace ds

extern "C" {

ds_status ds_segment_heads_i64(const int64_t* sorted_keys, uint64_t n64, uint32_t* seg_starts_out,
                               uint32_t* nseg_out, void* ws, size_t* ws_bytes, void* stream) {
  using namespace ds;
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: n exceeds DS_MAX_ROWS", __func__);
    return DS_ERR_TOO_LARGE;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  const uint32_t tiles = (n + SEG_TILE - 1) / SEG_TILE;
  Workspace w(ws);
  uint32_t* tile_counts = w.take<uint32_t>(tiles + 1);
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, __func__, &st)) return st;
  if (!seg_starts_out || !nseg_out || (n > 0 && !sorted_keys)) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (n > 0) {
    ds::note_launch(); head_count_kernel<<<tiles, SEG_THREADS, 0, s>>>(sorted_keys, n, tile_counts);
    DS_LAUNCH_CHECK();
  }
  ds::note_launch(); tile_scan_kernel<<<1, 1024, 0, s>>>(tile_counts, tiles, nseg_out, seg_starts_out, n);
  DS_LAUNCH_CHECK();
  if (n > 0) {
    ds::note_launch(); head_write_kernel<<<tiles, SEG_THREADS, 0, s>>>(sorted_keys, n, tile_counts, seg_starts_out);
    DS_LAUNCH_CHECK();
  }
  return DS_OK;
}

ds_status ds_segmented_reduce_f64(const double* vals, const uint8_t* valid, const uint32_t* idx,
                                  const uint32_t* seg_starts, uint32_t nseg, uint64_t n, int32_t op,
                                  double* out, void* ws, size_t* ws_bytes, void* stream) {
  if (op >= ds::OP_COUNT_VALID) op = -1;
  return ds::segreduce_impl<double, double>(vals, valid, idx, seg_starts, nseg, n, op, out, ws,
                                            ws_bytes, stream, __func__);
}

ds_status ds_segmented_reduce_i64(const int64_t* vals, const uint8_t* valid, const uint32_t* idx,
                                  const uint32_t* seg_starts, uint32_t nseg, uint64_t n, int32_t op,
                                  int64_t* out, void* ws, size_t* ws_bytes, void* stream) {
  if (op >= ds::OP_COUNT_VALID) op = -1;
  return ds::segreduce_impl<int64_t, int64_t>(vals, valid, idx, seg_starts, nseg, n, op, out, ws,
                                              ws_bytes, stream, __func__);
}

ds_status ds_segmented_count_valid(const uint8_t* valid, const uint32_t* idx,
                                   const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                   int64_t* out, void* ws, size_t* ws_bytes, void* stream) {
  return ds::segreduce_impl<int64_t, int64_t>(nullptr, valid, idx, seg_starts, nseg, n,
                                              ds::OP_COUNT_VALID, out, ws, ws_bytes, stream, __func__);
}

ds_status ds_segment_counts(const uint32_t* seg_starts, uint32_t nseg, int64_t* out, void* stream) {
  if (nseg == 0) return DS_OK;
  if (!seg_starts || !out) {
    ds::set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  ds::note_launch(); ds::segment_counts_kernel<<<(nseg + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      seg_starts, nseg, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_segmented_sumsqdiff_f64(const double* vals, const uint8_t* valid, const int64_t* weights,
                                     const double* addend, const uint32_t* idx,
                                     const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                     const double* means, double* out, void* ws, size_t* ws_bytes,
                                     void* stream) {
  if (ws && !means) {
    ds::set_error("%s: means must not be NULL", __func__);
    return DS_ERR_INVALID_ARG;
  }
  return ds::segreduce_impl<double, double>(vals, valid, idx, seg_starts, nseg, n, ds::OP_SUMSQDIFF, out,
                                            ws, ws_bytes, stream, __func__,
                                            ds::SqDiffArgs{means, weights, addend});
}

ds_status ds_segmented_sumsqdiff_i64(const int64_t* vals, const uint8_t* valid, const uint32_t* idx,
                                     const uint32_t* seg_starts, uint32_t nseg, uint64_t n,
                                     const double* means, double* out, void* ws, size_t* ws_bytes,
                                     void* stream) {
  if (ws && !means) {
    ds::set_error("%s: means must not be NULL", __func__);
    return DS_ERR_INVALID_ARG;
  }
  return ds::segreduce_impl<int64_t, double>(vals, valid, idx, seg_starts, nseg, n, ds::OP_SUMSQDIFF, out,
                                             ws, ws_bytes, stream, __func__,
                                             ds::SqDiffArgs{means, nullptr, nullptr});
}

ds_status ds_scale_f64_by_i64(const double* a, const int64_t* w, uint64_t n, double* out, void* stream) {
  if (n == 0) return DS_OK;
  if (!a || !w || !out || n > 0xFFFFFFFFull) {
    ds::set_error("%s: bad argument", __func__);
    return DS_ERR_INVALID_ARG;
  }
  ds::note_launch();
  ds::scale_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      a, w, static_cast<uint32_t>(n), out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_finalize_std(const double* m2, const int64_t* count, uint64_t n, int32_t ddof, double* out,
                          void* stream) {
  if (n == 0) return DS_OK;
  if (!m2 || !count || !out || n > 0xFFFFFFFFull) {
    ds::set_error("%s: bad argument", __func__);
    return DS_ERR_INVALID_ARG;
  }
  ds::note_launch();
  ds::finalize_std_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      m2, count, static_cast<uint32_t>(n), ddof, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_finalize_validity(const int64_t* nvalid, const int64_t* nrows, uint64_t n,
                               int32_t ignore_nulls, uint8_t* out, void* stream) {
  if (n == 0) return DS_OK;
  if (!nvalid || !nrows || !out || n > 0xFFFFFFFFull) {
    ds::set_error("%s: bad argument", __func__);
    return DS_ERR_INVALID_ARG;
  }
  ds::note_launch();
  ds::finalize_validity_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      nvalid, nrows, static_cast<uint32_t>(n), ignore_nulls != 0, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

ds_status ds_finalize_mean(const void* sum, int32_t sum_is_i64, const int64_t* count, uint64_t n,
                           double* out, void* stream) {
  if (n == 0) return DS_OK;
  if (!sum || !count || !out || n > 0xFFFFFFFFull) {
    ds::set_error("%s: bad argument", __func__);
    return DS_ERR_INVALID_ARG;
  }
  ds::note_launch(); ds::finalize_mean_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      sum, sum_is_i64 != 0, count, static_cast<uint32_t>(n), out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ no-allocation variant
// ds_segment_heads_i64_static: the same three kernels as ds_segment_heads_i64, but the
// per-tile head counts live in a fixed array that is part of the loaded module instead
// of a caller-provided workspace.  Every GPU that loads this library carries the array
// (DS_SEGMENT_HEADS_STATIC_MAX_ROWS / 2048 + 1 counters, about 256 KiB) until the
// process exits.
namespace ds {

constexpr uint32_t HEADS_STATIC_TILES = DS_SEGMENT_HEADS_STATIC_MAX_ROWS / SEG_TILE;
static __device__ __align__(256) uint32_t g_heads_scratch[HEADS_STATIC_TILES + 1];

// In host code `g_heads_scratch` is only a handle for the runtime, not an address a
// kernel can use, so the device address is looked up once per device (with that
// device current) and cached.
static ds_status heads_scratch_address(uint32_t** out, const char* fn) {
  // one slot per device; relaxed atomics: concurrent first calls store the same address
  static std::atomic<uint32_t*> cached[64];
  int dev = 0;
  DS_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("%s: device ordinal %d out of range", fn, dev);
    return DS_ERR_UNSUPPORTED;
  }
  uint32_t* addr = cached[dev].load(std::memory_order_relaxed);
  if (addr == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_heads_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("%s: cudaGetSymbolAddress failed: %s", fn, cudaGetErrorString(e));
      return DS_ERR_CUDA;
    }
    addr = static_cast<uint32_t*>(p);
    cached[dev].store(addr, std::memory_order_relaxed);
  }
  *out = addr;
  return DS_OK;
}

}  // namespace ds

extern "C" ds_status ds_segment_heads_i64_static(const int64_t* sorted_keys, uint64_t n64,
                                                 uint32_t* seg_starts_out, uint32_t* nseg_out,
                                                 void* stream) {
  using namespace ds;
  if (n64 > DS_SEGMENT_HEADS_STATIC_MAX_ROWS) {
    set_error("%s: n=%llu exceeds DS_SEGMENT_HEADS_STATIC_MAX_ROWS=%u; use "
              "ds_segment_heads_i64 with a caller workspace for larger columns",
              __func__, (unsigned long long)n64, (unsigned)DS_SEGMENT_HEADS_STATIC_MAX_ROWS);
    return DS_ERR_TOO_LARGE;
  }
  if (!seg_starts_out || !nseg_out || (n64 > 0 && !sorted_keys)) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  uint32_t* tile_counts = nullptr;
  ds_status st = heads_scratch_address(&tile_counts, __func__);
  if (st != DS_OK) return st;
  const uint32_t n = static_cast<uint32_t>(n64);
  const uint32_t tiles = (n + SEG_TILE - 1) / SEG_TILE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Nothing in the array has to start at zero: head_count_kernel overwrites entries
  // [0, tiles) before tile_scan_kernel and head_write_kernel read them, and entries
  // beyond `tiles` (left over from an earlier, larger call) are never read.
  if (n > 0) {
    ds::note_launch(); head_count_kernel<<<tiles, SEG_THREADS, 0, s>>>(sorted_keys, n, tile_counts);
    DS_LAUNCH_CHECK();
  }
  ds::note_launch(); tile_scan_kernel<<<1, 1024, 0, s>>>(tile_counts, tiles, nseg_out, seg_starts_out, n);
  DS_LAUNCH_CHECK();
  if (n > 0) {
    ds::note_launch(); head_write_kernel<<<tiles, SEG_THREADS, 0, s>>>(sorted_keys, n, tile_counts, seg_starts_out);
    DS_LAUNCH_CHECK();
  }
  return DS_OK;
}
