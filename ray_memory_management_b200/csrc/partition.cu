// Range partitioning against splitters, and cut positions of sorted key columns.
//
// Replaces the partition step of ArrowBlockAccessor.sort_and_partition
// (python/ray/data/_internal/arrow_block.py:373-394 of the reference):
//   * ds_searchsorted_*            np.searchsorted on a sorted block (:380-386)
//   * ds_partition_hist_scatter_*  stable bucket -> permutation for plans that
//                                  partition BEFORE sorting (multi-GPU exchange).
//
// All comparisons happen on the radix-order encoding of ds_keys.cuh, in which both
// ascending and descending sorts are ascending, so
//   cut position       asc: lower_bound(col, b)      desc: upper_bound(col, b)
//   bucket of a key    asc: upper_bound(bounds, x)   desc: lower_bound(bounds, x)
// which is exactly the reference's rule "a key equal to a splitter goes to the
// partition holding the larger values" (SURVEY.md 8 a3).
#include "ds_common.cuh"
#include "ds_keys.cuh"

namespace ds {

constexpr int PART_THREADS = 256;
constexpr int PART_ITEMS = 16;
constexpr int PART_TILE = PART_THREADS * PART_ITEMS;
constexpr uint32_t PLB_AGGREGATE = 1u << 30;
constexpr uint32_t PLB_INCLUSIVE = 2u << 30;
constexpr uint32_t PLB_VALUE = (1u << 30) - 1;

// ------------------------------------------------------------ searchsorted
template <int KIND>
__global__ void searchsorted_kernel(const void* __restrict__ col, uint32_t width, uint64_t n,
                                    const void* __restrict__ bounds, uint32_t nb, bool desc,
                                    uint64_t* __restrict__ out) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb) return;
  uint64_t bh;
  uint32_t bl;
  load_key<KIND>(bounds, width, j, desc, bh, bl);
  uint64_t lo = 0, hi = n;
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    uint64_t kh;
    uint32_t kl;
    load_key<KIND>(col, width, mid, desc, kh, kl);
    // asc: first position with key >= b; desc: first position with key > b (encoded)
    const bool go_right = desc ? !key_less(bh, bl, kh, kl) : key_less(kh, kl, bh, bl);
    if (go_right) lo = mid + 1;
    else hi = mid;
  }
  out[j] = lo;
}

template <int KIND>
static ds_status searchsorted_impl(const void* col, uint32_t width, uint64_t n, const void* bounds,
                                   uint32_t nb, int32_t descending, uint64_t* out, void* stream,
                                   const char* fn) {
  if (nb == 0) return DS_OK;
  if (!bounds || !out || (n > 0 && !col)) {
    set_error("%s: NULL pointer", fn);
    return DS_ERR_INVALID_ARG;
  }
  if (KIND == KIND_FSB && (width < 1 || width > DS_MAX_KEY_BYTES)) {
    set_error("%s: key width %u unsupported", fn, width);
    return DS_ERR_UNSUPPORTED;
  }
  ds::note_launch(); searchsorted_kernel<KIND><<<(nb + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      col, width, n, bounds, nb, descending != 0, out);
  DS_LAUNCH_CHECK();
  return DS_OK;
}

// ------------------------------------------------------------ bucket + histogram
// Encoded splitters live in shared memory; one binary search per key.
template <int KIND>
__global__ void __launch_bounds__(PART_THREADS)
    bucket_hist_kernel(const void* __restrict__ col, uint32_t width, uint32_t n,
                       const void* __restrict__ bounds, uint32_t nb, bool desc,
                       uint8_t* __restrict__ bucket_out, unsigned long long* __restrict__ counts) {
  __shared__ uint64_t s_bh[DS_MAX_PARTITIONS];
  __shared__ uint32_t s_bl[DS_MAX_PARTITIONS];
  __shared__ uint32_t s_cnt[DS_MAX_PARTITIONS];
  for (uint32_t j = threadIdx.x; j < DS_MAX_PARTITIONS; j += blockDim.x) {
    s_cnt[j] = 0;
    if (j < nb) {
      uint64_t h;
      uint32_t l;
      load_key<KIND>(bounds, width, j, desc, h, l);
      s_bh[j] = h;
      s_bl[j] = l;
    }
  }
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t kh;
    uint32_t kl;
    load_key<KIND>(col, width, i, desc, kh, kl);
    uint32_t lo = 0, hi = nb;
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      // asc: count splitters <= key; desc: count splitters strictly before key (encoded)
      const bool go_right = desc ? key_less(s_bh[mid], s_bl[mid], kh, kl)
                                 : !key_less(kh, kl, s_bh[mid], s_bl[mid]);
      if (go_right) lo = mid + 1;
      else hi = mid;
    }
    bucket_out[i] = static_cast<uint8_t>(lo);
    const unsigned m = __activemask();
    int same;
    __match_all_sync(m, lo, &same);
    if (same) {
      if ((int)lane_id() == __ffs(m) - 1) atomicAdd(&s_cnt[lo], __popc(m));
    } else {
      atomicAdd(&s_cnt[lo], 1u);
    }
  }
  __syncthreads();
  for (uint32_t j = threadIdx.x; j <= nb; j += blockDim.x)
    if (s_cnt[j]) atomicAdd(&counts[j], static_cast<unsigned long long>(s_cnt[j]));
}

__global__ void partition_offsets_kernel(const unsigned long long* __restrict__ counts, uint32_t nparts,
                                         uint64_t* __restrict__ counts_out,
                                         uint64_t* __restrict__ offsets_out,
                                         uint32_t* __restrict__ base32) {
  if (threadIdx.x == 0) {
    uint64_t run = 0;
    for (uint32_t j = 0; j < nparts; ++j) {
      counts_out[j] = counts[j];
      offsets_out[j] = run;
      base32[j] = static_cast<uint32_t>(run);
      run += counts[j];
    }
  }
}

// Stable scatter of row indices by bucket: a single onesweep-style pass whose "digit"
// is the bucket id.  Writes go straight to perm_out: with few buckets every warp's
// stores already fall into a handful of contiguous runs.
__global__ void __launch_bounds__(PART_THREADS)
    partition_scatter_kernel(const uint8_t* __restrict__ bucket, uint32_t n,
                             const uint32_t* __restrict__ base32, uint32_t* __restrict__ perm_out,
                             volatile uint32_t* lookback, uint32_t* tile_counter) {
  constexpr int WARPS = PART_THREADS / 32;
  __shared__ uint32_t warp_hist[WARPS * 256];
  __shared__ uint32_t s_gofs[256];
  __shared__ uint32_t s_tile;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t lane = tid & 31;
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
  for (int i = tid; i < WARPS * 256; i += PART_THREADS) warp_hist[i] = 0;
  __syncthreads();
  const uint32_t tile = s_tile;
  const uint32_t tile_base = tile * PART_TILE;
  const uint32_t wofs = warp * 32 * PART_ITEMS + lane;
  uint32_t d[PART_ITEMS], rank[PART_ITEMS];
#pragma unroll
  for (int i = 0; i < PART_ITEMS; ++i) {
    const uint32_t gp = tile_base + wofs + i * 32;
    d[i] = gp < n ? bucket[gp] : 0xFFFFFFFFu;
  }
  uint32_t* wh = warp_hist + warp * 256;
  const uint32_t lt_mask = lanemask_lt();
#pragma unroll
  for (int i = 0; i < PART_ITEMS; ++i) {
    const bool valid = d[i] != 0xFFFFFFFFu;
    const uint32_t peers = __match_any_sync(0xffffffffu, d[i]);
    const uint32_t lt = __popc(peers & lt_mask);
    uint32_t prev = 0;
    if (valid) prev = wh[d[i]];
    __syncwarp();
    if (valid && lt == 0) wh[d[i]] = prev + __popc(peers);
    __syncwarp();
    rank[i] = prev + lt;
  }
  __syncthreads();
  {
    uint32_t sum = 0;  // PART_THREADS == 256: thread tid owns bucket tid
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = warp_hist[w * 256 + tid];
      warp_hist[w * 256 + tid] = sum;
      sum += c;
    }
    volatile uint32_t* my = lookback + static_cast<size_t>(tile) * 256 + tid;
    uint32_t excl = 0;
    if (tile == 0) {
      *my = PLB_INCLUSIVE | sum;
    } else {
      *my = PLB_AGGREGATE | sum;
      int t = static_cast<int>(tile) - 1;
      while (true) {
        const uint32_t s = lookback[static_cast<size_t>(t) * 256 + tid];
        if ((s & ~PLB_VALUE) == 0) continue;
        excl += s & PLB_VALUE;
        if (s & PLB_INCLUSIVE) break;
        --t;
      }
      *my = PLB_INCLUSIVE | (excl + sum);
    }
    s_gofs[tid] = base32[tid] + excl;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < PART_ITEMS; ++i) {
    if (d[i] != 0xFFFFFFFFu)
      perm_out[s_gofs[d[i]] + wh[d[i]] + rank[i]] = tile_base + wofs + i * 32;
  }
}

template <int KIND>
static ds_status partition_impl(const void* col, uint32_t width, uint64_t n64, const void* bounds,
                                uint32_t nb, int32_t descending, uint64_t* counts_out,
                                uint64_t* offsets_out, uint32_t* perm_out, void* ws, size_t* ws_bytes,
                                void* stream, const char* fn) {
  if (n64 > DS_MAX_ROWS) {
    set_error("%s: n=%llu exceeds DS_MAX_ROWS", fn, (unsigned long long)n64);
    return DS_ERR_TOO_LARGE;
  }
  if (nb + 1 > DS_MAX_PARTITIONS) {
    set_error("%s: %u partitions exceed DS_MAX_PARTITIONS", fn, nb + 1);
    return DS_ERR_UNSUPPORTED;
  }
  if (KIND == KIND_FSB && (width < 1 || width > DS_MAX_KEY_BYTES)) {
    set_error("%s: key width %u unsupported", fn, width);
    return DS_ERR_UNSUPPORTED;
  }
  const uint32_t n = static_cast<uint32_t>(n64);
  const uint32_t tiles = (n + PART_TILE - 1) / PART_TILE;
  Workspace w(ws);
  uint8_t* bucket = w.take<uint8_t>(n);
  unsigned long long* counts = w.take<unsigned long long>(256);
  uint32_t* base32 = w.take<uint32_t>(256);
  uint32_t* tile_counter = w.take<uint32_t>(64);
  uint32_t* lookback = w.take<uint32_t>(static_cast<size_t>(tiles) * 256);
  ds_status st;
  if (ws_resolve(w, ws, ws_bytes, fn, &st)) return st;
  if (!counts_out || !offsets_out || (n > 0 && (!col || !perm_out)) || (nb > 0 && !bounds)) {
    set_error("%s: NULL pointer", fn);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // counts | base32 | tile_counter are contiguous up to alignment padding: clear each
  DS_CUDA(cudaMemsetAsync(counts, 0, 256 * sizeof(unsigned long long), s));
  DS_CUDA(cudaMemsetAsync(base32, 0, 256 * sizeof(uint32_t), s));
  DS_CUDA(cudaMemsetAsync(tile_counter, 0, 64 * sizeof(uint32_t), s));
  if (n > 0) {
    DS_CUDA(cudaMemsetAsync(lookback, 0, sizeof(uint32_t) * static_cast<size_t>(tiles) * 256, s));
    int blocks = sm_count() * 8;
    const int need = (n + PART_THREADS - 1) / PART_THREADS;
    if (blocks > need) blocks = need;
    ds::note_launch(); bucket_hist_kernel<KIND><<<blocks, PART_THREADS, 0, s>>>(col, width, n, bounds, nb,
                                                             descending != 0, bucket, counts);
    DS_LAUNCH_CHECK();
  }
  ds::note_launch(); partition_offsets_kernel<<<1, 32, 0, s>>>(counts, nb + 1, counts_out, offsets_out, base32);
  DS_LAUNCH_CHECK();
  if (n > 0) {
    ds::note_launch(); partition_scatter_kernel<<<tiles, PART_THREADS, 0, s>>>(bucket, n, base32, perm_out, lookback,
                                                           tile_counter);
    DS_LAUNCH_CHECK();
  }
  return DS_OK;
}

}  // namespace ds

extern "C" {

ds_status ds_searchsorted_i64(const int64_t* sorted_col, uint64_t n, const int64_t* boundaries,
                              uint32_t nb, int32_t descending, uint64_t* bounds_out, void* stream) {
  return ds::searchsorted_impl<ds::KIND_I64>(sorted_col, 8, n, boundaries, nb, descending,
                                             bounds_out, stream, __func__);
}
ds_status ds_searchsorted_f64(const double* sorted_col, uint64_t n, const double* boundaries,
                              uint32_t nb, int32_t descending, uint64_t* bounds_out, void* stream) {
  return ds::searchsorted_impl<ds::KIND_F64>(sorted_col, 8, n, boundaries, nb, descending,
                                             bounds_out, stream, __func__);
}
ds_status ds_searchsorted_fsb(const uint8_t* sorted_col, uint32_t width, uint64_t n,
                              const uint8_t* boundaries, uint32_t nb, int32_t descending,
                              uint64_t* bounds_out, void* stream) {
  return ds::searchsorted_impl<ds::KIND_FSB>(sorted_col, width, n, boundaries, nb, descending,
                                             bounds_out, stream, __func__);
}
ds_status ds_partition_hist_scatter_i64(const int64_t* col, uint64_t n, const int64_t* boundaries,
                                        uint32_t nb, int32_t descending, uint64_t* counts_out,
                                        uint64_t* offsets_out, uint32_t* perm_out, void* ws,
                                        size_t* ws_bytes, void* stream) {
  return ds::partition_impl<ds::KIND_I64>(col, 8, n, boundaries, nb, descending, counts_out,
                                          offsets_out, perm_out, ws, ws_bytes, stream, __func__);
}
ds_status ds_partition_hist_scatter_fsb(const uint8_t* col, uint32_t width, uint64_t n,
                                        const uint8_t* boundaries, uint32_t nb, int32_t descending,
                                        uint64_t* counts_out, uint64_t* offsets_out,
                                        uint32_t* perm_out, void* ws, size_t* ws_bytes,
                                        void* stream) {
  return ds::partition_impl<ds::KIND_FSB>(col, width, n, boundaries, nb, descending, counts_out,
                                          offsets_out, perm_out, ws, ws_bytes, stream, __func__);
}

}  // extern "C"
