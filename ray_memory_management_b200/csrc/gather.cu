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
#include <atomic>

#include "ds_common.cuh"
#include "gather_bulk.cuh"

namespace ds {

static std::atomic<int> g_gather_mode{0};  // 0 auto (8 where eligible, else 2), 1 granule loads,
// 2 granule loads + L2::64B hint, 7 asynchronous bulk copies per row (gather_bulk.cuh, UBLKCP),
// 8 the same with 16-byte cp.async.  (Modes 3..6 of round 1 -- staged, direct, 32-bit words --
// were slower and are gone; their numbers are in profiles/r01_gather_staged_summary.txt.)
// Measured on 100 M x 90-byte rows (ms per sort step): 2: 10.46, 5: 11.0, 3: 12.7, 6: 14.0

// Read-only loads that ask L2 to fetch only 64 bytes around a miss: random rows gain
// nothing from the default 128-byte fetch, it only multiplies DRAM traffic.
__device__ __forceinline__ uint16_t ld_nc_64B(const uint16_t* p) {
  uint16_t v;
  asm volatile("ld.global.nc.L2::64B.u16 %0, [%1];" : "=h"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint8_t ld_nc_64B(const uint8_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L2::64B.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return static_cast<uint8_t>(v);
}
__device__ __forceinline__ uint32_t ld_nc_64B(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ uint2 ld_nc_64B(const uint2* p) {
  uint2 v;
  asm volatile("ld.global.nc.L2::64B.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ld_nc_64B(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

template <typename G, bool HINT>
__global__ void __launch_bounds__(256) gather_single_kernel(const G* __restrict__ src,
                                                            const uint32_t* __restrict__ idx,
                                                            G* __restrict__ dst, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const G* p = src + ld_stream(idx + i);
    st_stream(dst + i, HINT ? ld_nc_64B(p) : __ldg(p));
  }
}

constexpr int GATHER_THREADS = 256;
constexpr int GATHER_MAX_ROWS = 1024;

template <typename G, bool HINT>
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
  // The loop is latency bound (2..16 useful bytes per load): keep UNROLL independent loads
  // in flight per thread before the first store.
  constexpr int UNROLL = 16;
  uint32_t e = threadIdx.x;
  for (; e + (UNROLL - 1) * GATHER_THREADS < total; e += UNROLL * GATHER_THREADS) {
    G v[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const uint32_t ee = e + k * GATHER_THREADS;
      const uint32_t r = __umulhi(ee, magic);  // ee / gpr, exact for ee*gpr < 2^32
      const G* p = src + static_cast<uint64_t>(s_idx[r]) * gpr + (ee - r * gpr);
      v[k] = HINT ? ld_nc_64B(p) : __ldg(p);
    }
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) st_stream(out + e + k * GATHER_THREADS, v[k]);
  }
  for (; e < total; e += GATHER_THREADS) {
    const uint32_t r = __umulhi(e, magic);
    const G* p = src + static_cast<uint64_t>(s_idx[r]) * gpr + (e - r * gpr);
    st_stream(out + e, HINT ? ld_nc_64B(p) : __ldg(p));
  }
}

// ---- helpers of the staged peer scatter below (rows that are not a multiple of 16 bytes are
// pulled in as ALIGNED 16-byte vectors and re-read from shared memory at their misalignment)
constexpr int STAGE_THREADS = 256;

__device__ __forceinline__ uint32_t smem_word_unaligned(const uint8_t* base, uint32_t byte_off) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(base + (byte_off & ~3u));
  return __byte_perm(w[0], w[1], 0x3210u + 0x1111u * (byte_off & 3u));
}

template <typename G>
static cudaError_t launch_gather(const void* src, const uint32_t* idx, void* dst, uint32_t n,
                                 uint32_t row_bytes, cudaStream_t s) {
  const uint32_t gpr = row_bytes / sizeof(G);
  if (gpr == 1) {
    ds::note_launch();
    if (g_gather_mode == 1)
      gather_single_kernel<G, false><<<(n + 255) / 256, 256, 0, s>>>(static_cast<const G*>(src), idx,
                                                                     static_cast<G*>(dst), n);
    else
      gather_single_kernel<G, true><<<(n + 255) / 256, 256, 0, s>>>(static_cast<const G*>(src), idx,
                                                                    static_cast<G*>(dst), n);
    return cudaGetLastError();
  }
  uint32_t rows = 8192 / gpr;
  if (rows < 1) rows = 1;
  if (rows > GATHER_MAX_ROWS) rows = GATHER_MAX_ROWS;
  const uint32_t magic = static_cast<uint32_t>((0x100000000ull + gpr - 1) / gpr);
  const uint32_t blocks = (n + rows - 1) / rows;
  ds::note_launch();
  if (g_gather_mode == 1)
    gather_rows_kernel<G, false><<<blocks, GATHER_THREADS, 0, s>>>(
        static_cast<const G*>(src), idx, static_cast<G*>(dst), n, gpr, rows, magic);
  else
    gather_rows_kernel<G, true><<<blocks, GATHER_THREADS, 0, s>>>(
        static_cast<const G*>(src), idx, static_cast<G*>(dst), n, gpr, rows, magic);
  return cudaGetLastError();
}

// ---- partition scatter to peers: the map-side output of the exchange written where it
// is consumed.  Rows are visited in partitioned order (perm, as produced by
// ds_partition_hist_scatter_*); row j of partition d is stored at
// dst_bases[d] + (j - part_offsets[d]) * row_bytes, where dst_bases[d] may point into
// another GPU's memory (peer-mapped over NVLink) or into this GPU's own receive buffer.
constexpr int MAX_PEER_PARTS = 16;
struct PeerParts {
  uint32_t count;
  uint32_t offsets[MAX_PEER_PARTS + 1];
  unsigned char* bases[MAX_PEER_PARTS];
};

template <typename G>
__global__ void __launch_bounds__(GATHER_THREADS)
    scatter_rows_to_parts_kernel(const G* __restrict__ src, const uint32_t* __restrict__ perm,
                                 uint32_t n, uint32_t gpr, uint32_t rows_per_block, uint32_t magic,
                                 PeerParts parts) {
  __shared__ uint32_t s_idx[GATHER_MAX_ROWS];
  const uint32_t row0 = blockIdx.x * rows_per_block;
  const uint32_t rows = min(rows_per_block, n - row0);
  for (uint32_t r = threadIdx.x; r < rows; r += GATHER_THREADS) s_idx[r] = ld_stream(perm + row0 + r);
  __syncthreads();
  const uint32_t total = rows * gpr;
  constexpr int UNROLL = 8;
  uint32_t e = threadIdx.x;
  auto dst_of = [&](uint32_t ee, uint32_t r) -> G* {
    const uint32_t j = row0 + r;
    uint32_t d = 0;
#pragma unroll
    for (int k = 1; k < MAX_PEER_PARTS; ++k)
      if (k < (int)parts.count && j >= parts.offsets[k]) d = k;
    return reinterpret_cast<G*>(parts.bases[d]) + static_cast<uint64_t>(j - parts.offsets[d]) * gpr +
           (ee - r * gpr);
  };
  for (; e + (UNROLL - 1) * GATHER_THREADS < total; e += UNROLL * GATHER_THREADS) {
    G v[UNROLL];
    uint32_t rr[UNROLL];
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) {
      const uint32_t ee = e + k * GATHER_THREADS;
      rr[k] = gpr == 1 ? ee : __umulhi(ee, magic);
      v[k] = __ldg(src + static_cast<uint64_t>(s_idx[rr[k]]) * gpr + (ee - rr[k] * gpr));
    }
#pragma unroll
    for (int k = 0; k < UNROLL; ++k) *dst_of(e + k * GATHER_THREADS, rr[k]) = v[k];
  }
  for (; e < total; e += GATHER_THREADS) {
    const uint32_t r = gpr == 1 ? e : __umulhi(e, magic);
    *dst_of(e, r) = __ldg(src + static_cast<uint64_t>(s_idx[r]) * gpr + (e - r * gpr));
  }
}

// Staged variant for rows that are not a multiple of 16 bytes: every block serves rows of ONE
// partition, pulls them in as aligned 16-byte vectors (as gather_staged_kernel does) and
// writes its contiguous destination span with 16-byte stores -- full-size NVLink writes when
// the destination is a peer -- whatever the span's byte alignment (head and tail bytes are
// stored one by one).
struct PeerBlocks {
  uint32_t first_block[MAX_PEER_PARTS + 1];  // prefix sum of blocks per partition
};

__global__ void __launch_bounds__(STAGE_THREADS)
    scatter_staged_kernel(const uint8_t* __restrict__ src, uint32_t W, uint64_t src_bytes,
                          const uint32_t* __restrict__ perm, uint32_t rows_per_block, uint32_t chunks,
                          uint32_t magic_w, uint32_t magic_c, PeerParts parts, PeerBlocks pblk) {
  extern __shared__ __align__(16) uint8_t sm[];
  const uint32_t slot = chunks * 16 + 16;
  uint64_t* s_a0 = reinterpret_cast<uint64_t*>(sm);
  uint32_t* s_shift = reinterpret_cast<uint32_t*>(sm + rows_per_block * 8);
  uint8_t* data = sm + rows_per_block * 12;
  uint32_t d = 0;
#pragma unroll
  for (int k = 1; k < MAX_PEER_PARTS; ++k)
    if (k < (int)parts.count && blockIdx.x >= pblk.first_block[k]) d = k;
  const uint32_t local0 = (blockIdx.x - pblk.first_block[d]) * rows_per_block;  // row inside the partition
  const uint32_t part_rows = parts.offsets[d + 1] - parts.offsets[d];
  const uint32_t rows = min(rows_per_block, part_rows - local0);
  const uint32_t row0 = parts.offsets[d] + local0;  // position in partitioned order
  const uintptr_t base = reinterpret_cast<uintptr_t>(src);
  const uintptr_t safe_lo = (base + 15) & ~uintptr_t(15);
  const uintptr_t safe_hi = (base + src_bytes) & ~uintptr_t(15);

  for (uint32_t r = threadIdx.x; r < rows; r += STAGE_THREADS) {
    const uintptr_t a = base + static_cast<uint64_t>(ld_stream(perm + row0 + r)) * W;
    s_a0[r] = a & ~uintptr_t(15);
    s_shift[r] = static_cast<uint32_t>(a & 15);
  }
  __syncthreads();
  const uint32_t items = rows * chunks;
#pragma unroll 4
  for (uint32_t e = threadIdx.x; e < items; e += STAGE_THREADS) {
    const uint32_t r = __umulhi(e, magic_c);
    const uint32_t c = e - r * chunks;
    if (c * 16 >= s_shift[r] + W) continue;
    const uintptr_t a = s_a0[r] + c * 16;
    uint4 v;
    if (a >= safe_lo && a + 16 <= safe_hi) {
      v = __ldg(reinterpret_cast<const uint4*>(a));
    } else {
      uint8_t t[16];
#pragma unroll
      for (int b = 0; b < 16; ++b)
        t[b] = (a + b >= base && a + b < base + src_bytes) ? *reinterpret_cast<const uint8_t*>(a + b) : 0;
      v = *reinterpret_cast<uint4*>(t);
    }
    *reinterpret_cast<uint4*>(data + r * slot + c * 16) = v;
  }
  __syncthreads();

  const uint32_t nbytes = rows * W;
  uint8_t* out = parts.bases[d] + static_cast<uint64_t>(local0) * W;
  auto byte_at = [&](uint32_t o) -> uint8_t {
    const uint32_t r = __umulhi(o, magic_w);
    return data[r * slot + s_shift[r] + (o - r * W)];
  };
  const uint32_t lead = min(nbytes, static_cast<uint32_t>((16 - (reinterpret_cast<uintptr_t>(out) & 15)) & 15));
  for (uint32_t o = threadIdx.x; o < lead; o += STAGE_THREADS) out[o] = byte_at(o);
  const uint32_t nvec = (nbytes - lead) >> 4;
  uint4* vout = reinterpret_cast<uint4*>(out + lead);
  for (uint32_t q = threadIdx.x; q < nvec; q += STAGE_THREADS) {
    uint32_t w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t o = lead + q * 16 + k * 4;
      const uint32_t r = __umulhi(o, magic_w);
      const uint32_t off = o - r * W;
      if (off + 4 <= W) {
        w[k] = smem_word_unaligned(data + r * slot, s_shift[r] + off);
      } else {
        uint32_t acc = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) acc |= static_cast<uint32_t>(byte_at(o + b)) << (8 * b);
        w[k] = acc;
      }
    }
    vout[q] = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (uint32_t o = lead + (nvec << 4) + threadIdx.x; o < nbytes; o += STAGE_THREADS) out[o] = byte_at(o);
}

static cudaError_t launch_scatter_staged(const void* src, uint32_t W, uint64_t src_rows,
                                         const uint32_t* perm, const PeerParts& parts, cudaStream_t s) {
  const uint32_t chunks = (W + 15 + 15) / 16;
  const uint32_t slot = chunks * 16 + 16;
  uint32_t rows = 128;
  while (rows > 16 && rows * (slot + 12) > 40 * 1024) rows >>= 1;
  const uint32_t smem = rows * (slot + 12);
  PeerBlocks pblk{};
  uint32_t blocks = 0;
  for (uint32_t d = 0; d < parts.count; ++d) {
    pblk.first_block[d] = blocks;
    blocks += (parts.offsets[d + 1] - parts.offsets[d] + rows - 1) / rows;
  }
  pblk.first_block[parts.count] = blocks;
  if (blocks == 0) return cudaSuccess;
  const uint32_t magic_w = static_cast<uint32_t>((0x100000000ull + W - 1) / W);
  const uint32_t magic_c = static_cast<uint32_t>((0x100000000ull + chunks - 1) / chunks);
  ds::note_launch();
  scatter_staged_kernel<<<blocks, STAGE_THREADS, smem, s>>>(static_cast<const uint8_t*>(src), W,
                                                           src_rows * W, perm, rows, chunks, magic_w,
                                                           magic_c, parts, pblk);
  return cudaGetLastError();
}

template <typename G>
static cudaError_t launch_scatter_parts(const void* src, const uint32_t* perm, uint32_t n,
                                        uint32_t row_bytes, const PeerParts& parts, cudaStream_t s) {
  const uint32_t gpr = row_bytes / sizeof(G);
  uint32_t rows = 8192 / gpr;
  if (rows < 1) rows = 1;
  if (rows > GATHER_MAX_ROWS) rows = GATHER_MAX_ROWS;
  const uint32_t magic = gpr == 1 ? 0u : static_cast<uint32_t>((0x100000000ull + gpr - 1) / gpr);
  const uint32_t blocks = (n + rows - 1) / rows;
  ds::note_launch();
  scatter_rows_to_parts_kernel<G><<<blocks, GATHER_THREADS, 0, s>>>(static_cast<const G*>(src), perm, n,
                                                                    gpr, rows, magic, parts);
  return cudaGetLastError();
}

ds_status gather_rows_impl(const void* src, uint32_t row_bytes, uint64_t src_rows,
                           const uint32_t* idx, uint64_t n64, void* dst, cudaStream_t s) {
  const int g_gather_mode = ds::g_gather_mode.load(std::memory_order_relaxed);  // one snapshot per call
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
  LaunchTimer timer(TIMED_GATHER, s, static_cast<double>(n) * (4.0 + 2.0 * row_bytes));
  // auto: even rows that are not 16-byte multiples take the asynchronous-copy kernel with
  // 16-byte cp.async + 64-byte L2 fetch hint (4.54 ms per 100 M x 90 B against 4.96 ms for the
  // granule kernel and 5.07 ms for per-row bulk copies, whose misses fetch 128-byte lines:
  // profiles/r02_gather_final_summary.txt)
  if ((g_gather_mode == 0 || g_gather_mode == 7 || g_gather_mode == 8) &&
      gather_bulk_eligible(src, row_bytes, src_rows, dst)) {
    e = launch_gather_bulk(g_gather_mode == 7 ? FETCH_BULK : FETCH_LDGSTS, src, row_bytes, src_rows, idx, n, dst, s);
  } else if ((bits & 15) == 0) e = launch_gather<uint4>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 7) == 0) e = launch_gather<uint2>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 3) == 0) e = launch_gather<uint32_t>(src, idx, dst, n, row_bytes, s);
  else if ((bits & 1) == 0) e = launch_gather<uint16_t>(src, idx, dst, n, row_bytes, s);
  else e = launch_gather<uint8_t>(src, idx, dst, n, row_bytes, s);
  timer.stop();
  if (e != cudaSuccess) {
    set_error("ds_gather_rows: launch failed: %s", cudaGetErrorString(e));
    return DS_ERR_CUDA;
  }
  return DS_OK;
}

}  // namespace ds

extern "C" ds_status ds_gather_rows(const void* src, uint32_t row_bytes, uint64_t src_rows,
                                    const uint32_t* idx, uint64_t n, void* dst, void* stream) {
  return ds::gather_rows_impl(src, row_bytes, src_rows, idx, n, dst, static_cast<cudaStream_t>(stream));
}

extern "C" ds_status ds_gather_rows_multi(const void* const* srcs, const uint64_t* src_rows, uint32_t nsrc,
                                          uint32_t row_bytes, const uint32_t* idx, uint64_t n, void* dst,
                                          void* stream) {
  using namespace ds;
  if (n == 0) return DS_OK;
  if (n > 0xFFFFFFFFull) {
    set_error("%s: n=%llu exceeds 2^32-1", __func__, (unsigned long long)n);
    return DS_ERR_TOO_LARGE;
  }
  if (!srcs || !src_rows || !idx || !dst || nsrc == 0) {
    set_error("%s: NULL pointer or no source", __func__);
    return DS_ERR_INVALID_ARG;
  }
  if (nsrc > GB_MAX_SOURCES) {
    set_error("%s: %u sources exceed the limit of %d", __func__, nsrc, GB_MAX_SOURCES);
    return DS_ERR_UNSUPPORTED;
  }
  GatherSources g{};
  g.nsrc = nsrc;
  uint64_t total = 0;
  bool ok = true;
  for (uint32_t b = 0; b < nsrc; ++b) {
    g.start[b] = static_cast<uint32_t>(total);
    g.base[b] = static_cast<const uint8_t*>(srcs[b]);
    total += src_rows[b];
    ok = ok && (src_rows[b] == 0 || gather_bulk_eligible(srcs[b], row_bytes, src_rows[b], dst));
  }
  if (total > 0xFFFFFFFFull) {
    set_error("%s: %llu source rows exceed 2^32-1", __func__, (unsigned long long)total);
    return DS_ERR_TOO_LARGE;
  }
  g.start[nsrc] = static_cast<uint32_t>(total);
  if (!ok) {
    set_error("%s: row_bytes=%u / alignment not supported (even rows of 18..1024 bytes that are not a multiple "
              "of 16, sources 2-byte and dst 16-byte aligned); gather from one concatenated column instead",
              __func__, row_bytes);
    return DS_ERR_UNSUPPORTED;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  LaunchTimer timer(TIMED_GATHER, s, static_cast<double>(n) * (4.0 + 2.0 * row_bytes));
  cudaError_t e = launch_gather_bulk(FETCH_LDGSTS, g, row_bytes, idx, static_cast<uint32_t>(n), dst, s);
  timer.stop();
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", __func__, cudaGetErrorString(e));
    return DS_ERR_CUDA;
  }
  return DS_OK;
}

extern "C" ds_status ds_scatter_rows_to_parts(const void* src, uint32_t row_bytes, uint64_t src_rows,
                                              const uint32_t* perm, uint64_t n, uint32_t nparts,
                                              const uint64_t* part_offsets, void* const* dst_bases,
                                              void* stream) {
  using namespace ds;
  if (n == 0) return DS_OK;
  if (n > DS_MAX_ROWS || row_bytes == 0 || row_bytes > 65536) {
    set_error("%s: n or row_bytes out of range", __func__);
    return DS_ERR_TOO_LARGE;
  }
  if (nparts == 0 || nparts > MAX_PEER_PARTS) {
    set_error("%s: %u partitions unsupported (1..%d)", __func__, nparts, MAX_PEER_PARTS);
    return DS_ERR_UNSUPPORTED;
  }
  if (!src || !perm || !part_offsets || !dst_bases) {
    set_error("%s: NULL pointer", __func__);
    return DS_ERR_INVALID_ARG;
  }
  PeerParts parts{};
  parts.count = nparts;
  uintptr_t bits = reinterpret_cast<uintptr_t>(src) | row_bytes;
  for (uint32_t d = 0; d < nparts; ++d) {
    parts.offsets[d] = static_cast<uint32_t>(part_offsets[d]);
    parts.bases[d] = static_cast<unsigned char*>(dst_bases[d]);
    if (part_offsets[d + 1] > part_offsets[d]) bits |= reinterpret_cast<uintptr_t>(dst_bases[d]);
  }
  parts.offsets[nparts] = static_cast<uint32_t>(part_offsets[nparts]);
  if (part_offsets[nparts] != n) {
    set_error("%s: part_offsets[nparts] must equal n", __func__);
    return DS_ERR_INVALID_ARG;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const uint32_t n32 = static_cast<uint32_t>(n);
  cudaError_t e;
  if ((row_bytes & 15) != 0 && row_bytes >= 16 && row_bytes <= 2048 && src_rows > 0 && g_gather_mode != 1)
    e = launch_scatter_staged(src, row_bytes, src_rows, perm, parts, s);
  else if ((bits & 15) == 0) e = launch_scatter_parts<uint4>(src, perm, n32, row_bytes, parts, s);
  else if ((bits & 7) == 0) e = launch_scatter_parts<uint2>(src, perm, n32, row_bytes, parts, s);
  else if ((bits & 3) == 0) e = launch_scatter_parts<uint32_t>(src, perm, n32, row_bytes, parts, s);
  else if ((bits & 1) == 0) e = launch_scatter_parts<uint16_t>(src, perm, n32, row_bytes, parts, s);
  else e = launch_scatter_parts<uint8_t>(src, perm, n32, row_bytes, parts, s);
  if (e != cudaSuccess) {
    set_error("%s: launch failed: %s", __func__, cudaGetErrorString(e));
    return DS_ERR_CUDA;
  }
  return DS_OK;
}

// Tuning hook, see g_gather_mode.
extern "C" int32_t ds_debug_set_gather_mode(int32_t mode) {
  return ds::g_gather_mode.exchange(mode);
}

// Tuning hook for the asynchronous-copy gather: rows per tile, ring depth, CTAs per SM (0 = default).
extern "C" void ds_debug_set_gather_bulk(uint32_t rows, uint32_t stages, uint32_t ctas_per_sm) {
  ds::g_gb_rows = rows;
  ds::g_gb_stages = stages;
  ds::g_gb_ctas = ctas_per_sm;
}

// Tuning hook: cudaLimitMaxL2FetchGranularity of the current device (32 / 64 / 128).
extern "C" int32_t ds_debug_set_l2_fetch_granularity(int32_t bytes) {
  size_t old = 0;
  cudaDeviceGetLimit(&old, cudaLimitMaxL2FetchGranularity);
  if (bytes > 0) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, static_cast<size_t>(bytes));
  return static_cast<int32_t>(old);
}
